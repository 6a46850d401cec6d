"""Single-launch rollouts of small scenes (csrc/small.cu, aether_small_rollout) against the golden rollouts recorded
from the reference, the oracle, and the multi-launch path (aether_rollout_step per step).

Bars: fp32 arithmetic, so the fp32 ones -- rel <= 2e-5 of the trajectory scale against the golden vectors (the bar of
tests/test_step_gpu.py::test_rollout_golden), <= 5e-5 against the oracle over longer kNN rollouts (neighbour choice is
discrete), <= 1e-5 against the multi-launch fp32 path; kNN edge lists bit-identical to aether_knn_edges.
"""
import pytest
import torch

from conftest import golden_params, load_golden, rel_err
from oracle import aether_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ab():
    import aether_b200
    return aether_b200


def make(ab, D, H=64, seed=1, precision="fp32"):
    torch.manual_seed(seed)
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda", precision=precision).eval()
    p = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    return m, p


def dev(a):
    return torch.as_tensor(a).cuda()


@pytest.mark.parametrize("case,sizes", [("step3d_h64", [5, 5]), ("step3d_h64", [10]), ("step2d_h64_n12", [12, 12]),
                                        ("step2d_h64_n12", [24])])
def test_small_rollout_golden(ab, case, sizes):
    """Golden rollouts of the unmodified reference; one scene per graph, and both graphs in one scene (264 edges:
    three row chunks)."""
    g = load_golden(case)
    D, H = int(g["D"]), int(g["H"])
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda").eval()
    m.load_state_dict(golden_params(g))
    edges = (dev(g["send"]), dev(g["recv"]))
    T = g["rollout"].shape[0]
    r = ab.Rollout(m, g["x"].shape[0], T, edges=edges, sizes=sizes, persistent=True)
    assert r.persistent
    n0 = r.lib.aether_launch_count()
    xs = r.run(dev(g["x"]), dev(g["vel"]), dev(g["charges"])).clone()
    assert r.lib.aether_launch_count() - n0 == 2          # weight pack + the whole T-step rollout in one kernel
    ref = ab.rollout(m, dev(g["x"]), dev(g["vel"]), dev(g["charges"]), T, edges=edges, persistent=False)
    assert rel_err(xs.cpu(), g["rollout"]) <= 2e-5
    assert rel_err(xs.cpu(), ref.cpu()) <= 1e-5
    assert torch.equal(xs, r.run(dev(g["x"]), dev(g["vel"]), dev(g["charges"])))     # run-to-run deterministic


@pytest.mark.parametrize("D,N,B,T", [(2, 5, 16, 20), (3, 5, 7, 6), (2, 21, 3, 4)])
def test_small_rollout_fc_vs_oracle(ab, D, N, B, T):
    """Config-1 shape (2-D, 5 particles, batch 16, 20 steps), a 3-D batch, and the largest fully connected scene the
    kernel takes (21 particles, 420 edges)."""
    m, p = make(ab, D)
    x, v, q = O.synth_state(B, N, D, seed=3)
    send, recv = O.fc_edges(N, B)
    xs = ab.rollout(m, x.cuda(), v.cuda(), q.cuda(), T, edges=(send.cuda(), recv.cuda()), sizes=[N] * B, persistent=True)
    ref = ab.rollout(m, x.cuda(), v.cuda(), q.cuda(), T, edges=(send.cuda(), recv.cuda()), persistent=False)
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: (send, recv), T)
    assert rel_err(xs.cpu(), want) <= 5e-5
    assert rel_err(xs.cpu(), ref.cpu()) <= 1e-5


@pytest.mark.parametrize("D", [2, 3])
def test_small_rollout_knn_ragged(ab, D):
    """inD-style scenes: ragged sizes (1 and 2 agents included), edges rebuilt from the positions every step."""
    k, T = 10, 5
    sizes = [10, 40, 17, 3, 25, 1, 2, 33]
    m, p = make(ab, D)
    g = torch.Generator().manual_seed(5)
    n = sum(sizes)
    x = torch.rand(n, D, generator=g)
    v = (torch.rand(n, D, generator=g) - 0.5) * 0.1
    q = torch.randint(0, 2, (n, 1), generator=g).float() * 2 - 1
    plan = ab.KnnPlan(sizes, k, "cuda")
    r = ab.Rollout(m, n, T, knn_plan=plan, persistent=True)
    xs = r.run(x.cuda(), v.cuda(), q.cuda(), dump_edges=True).clone()
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: O.knn_edges_batched(xx, sizes, k), T)
    ref = ab.rollout(m, x.cuda(), v.cuda(), q.cuda(), T, knn_plan=plan, persistent=False)
    assert rel_err(xs.cpu(), want) <= 5e-5
    assert rel_err(xs.cpu(), ref.cpu()) <= 1e-5
    # the edge list of the last step (built inside the kernel from x_{T-1}) against the stand-alone builder: bit-identical
    s, rr = ab.knn_edges(xs[T - 2].contiguous(), plan)
    assert torch.equal(r.last_edges[0], s) and torch.equal(r.last_edges[1], rr)
    so, ro = O.knn_edges_batched(xs[T - 2].cpu(), sizes, k)
    assert torch.equal(s.cpu(), so) and torch.equal(rr.cpu(), ro)


def test_small_rollout_edge_cases(ab):
    """Isolated particles (no in-edges: the mean is 0), an empty scene in the middle, E = 0 scenes, a shuffled edge
    order inside the scenes, zero velocities and coincident particles."""
    D, T = 3, 3
    m, p = make(ab, D)
    sizes = [4, 0, 1, 6, 2]
    n = sum(sizes)
    g = torch.Generator().manual_seed(11)
    x = torch.rand(n, D, generator=g)
    v = (torch.rand(n, D, generator=g) - 0.5) * 0.2
    x[1] = x[0]                       # coincident
    v[2] = 0.0                        # zero velocity: heading frame at theta = 0
    q = torch.randint(0, 2, (n, 1), generator=g).float() * 2 - 1
    send, recv, off = [], [], 0
    for s in sizes:
        for i in range(s):
            for j in range(s):
                if i != j and not (s == 6 and j == 3):      # particle 3 of the 6-scene receives nothing
                    send.append(off + i); recv.append(off + j)
        off += s
    send, recv = torch.tensor(send), torch.tensor(recv)
    # shuffle inside each scene (edges stay grouped by scene)
    sc = torch.bucketize(send, torch.tensor(sizes).cumsum(0), right=True)
    order = torch.argsort(sc * 1000 + torch.randperm(send.numel(), generator=g) % 1000, stable=True)
    send, recv = send[order], recv[order]
    xs = ab.rollout(m, x.cuda(), v.cuda(), q.cuda(), T, edges=(send.cuda(), recv.cuda()), sizes=sizes, persistent=True)
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: (send, recv), T)
    assert torch.isfinite(xs).all()
    assert rel_err(xs.cpu(), want) <= 2e-5


@pytest.mark.parametrize("D,N,B,T", [(2, 5, 16, 20), (3, 5, 7, 6), (2, 24, 2, 4), (3, 29, 2, 3)])
def test_small_rollout_bf16_fc(ab, D, N, B, T):
    """Tensor-core mode (precision="bf16": edge MLP on mma.sync, bf16 edge state in shared memory) against the fp32 oracle
    at the bf16 bar (2e-2); 24 and 29 particles: 552 / 812 edges per scene, three / four 256-row chunks -- more than the
    fp32 mode holds."""
    m, p = make(ab, D, precision="bf16")
    x, v, q = O.synth_state(B, N, D, seed=3)
    send, recv = O.fc_edges(N, B)
    r = ab.Rollout(m, N * B, T, edges=(send.cuda(), recv.cuda()), sizes=[N] * B, persistent=True)
    xs = r.run(x.cuda(), v.cuda(), q.cuda()).clone()
    ref = ab.rollout(m, x.cuda(), v.cuda(), q.cuda(), T, edges=(send.cuda(), recv.cuda()), persistent=False)
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: (send, recv), T)
    e_or, e_ml = rel_err(xs.cpu(), want), rel_err(ref.cpu(), want)
    print("bf16 single-launch vs oracle %.2e; multi-launch bf16 vs oracle %.2e" % (e_or, e_ml))
    assert e_or <= 2e-2
    assert torch.equal(xs, r.run(x.cuda(), v.cuda(), q.cuda()))


def test_small_rollout_bf16_golden_and_knn(ab):
    """Tensor-core mode on the reference's golden rollouts (2e-2) and on ragged kNN scenes (edge lists stay bit-identical:
    they depend on positions only through the fp32 distances)."""
    for case, sizes in [("step3d_h64", [5, 5]), ("step2d_h64_n12", [24])]:
        g = load_golden(case)
        D, H = int(g["D"]), int(g["H"])
        m = ab.Aether(2 * D, H, 0.0, D, device="cuda", precision="bf16").eval()
        m.load_state_dict(golden_params(g))
        T = g["rollout"].shape[0]
        xs = ab.rollout(m, dev(g["x"]), dev(g["vel"]), dev(g["charges"]), T, edges=(dev(g["send"]), dev(g["recv"])), sizes=sizes,
                        persistent=True)
        assert rel_err(xs.cpu(), g["rollout"]) <= 2e-2
    D, k, T = 2, 10, 5
    sizes = [10, 40, 17, 3, 25, 1, 2, 33]
    m, p = make(ab, D, precision="bf16")
    gen = torch.Generator().manual_seed(5)
    n = sum(sizes)
    x = torch.rand(n, D, generator=gen)
    v = (torch.rand(n, D, generator=gen) - 0.5) * 0.1
    q = torch.randint(0, 2, (n, 1), generator=gen).float() * 2 - 1
    plan = ab.KnnPlan(sizes, k, "cuda")
    r = ab.Rollout(m, n, T, knn_plan=plan, persistent=True)
    xs = r.run(x.cuda(), v.cuda(), q.cuda(), dump_edges=True).clone()
    s, rr = ab.knn_edges(xs[T - 2].contiguous(), plan)
    assert torch.equal(r.last_edges[0], s) and torch.equal(r.last_edges[1], rr)
    # one step from the same state: no neighbour flips between the two arithmetic modes
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: O.knn_edges_batched(xx, sizes, k), 1)
    assert rel_err(xs[:1].cpu(), want) <= 2e-2


def test_small_rollout_selection(ab):
    """The single-launch path is chosen only where it applies; persistent=True says why when it does not."""
    m32, _ = make(ab, 2, H=32)
    x, v, q = O.synth_state(2, 5, 2, seed=3)
    send, recv = O.fc_edges(5, 2)
    e = (send.cuda(), recv.cuda())
    with pytest.raises(ValueError):
        ab.Rollout(m32, 10, 2, edges=e, sizes=[5, 5], persistent=True)
    assert not ab.Rollout(m32, 10, 2, edges=e, sizes=[5, 5]).persistent           # hidden 32: multi-launch path
    m, _ = make(ab, 2)
    assert ab.Rollout(m, 10, 2, edges=e, sizes=[5, 5]).persistent
    assert not ab.Rollout(m, 10, 2, edges=e).persistent                           # scene sizes unknown
    with pytest.raises(ValueError):
        ab.Rollout(m, 10, 2, edges=e, sizes=[4, 6], persistent=True)              # edges cross the scene boundary
    with pytest.raises(ValueError):
        ab.Rollout(m, 10, 2, edges=e, sizes=[5, 4], persistent=True)              # sizes do not add up
    big = O.fc_edges(40, 1)
    with pytest.raises(ValueError):                                               # 1560 edges do not fit one SM
        ab.Rollout(m, 40, 2, edges=(big[0].cuda(), big[1].cuda()), sizes=[40], persistent=True)
    lib = ab._lib.load()
    assert lib.aether_small_rollout_supported(2, 64, 40, 400, 0) == 1
