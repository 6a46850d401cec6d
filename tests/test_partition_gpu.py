"""GPU tests of the large-scene path (BASELINE config 5): the cell-list radius graph is bit-exact with the
brute-force oracle, and the receiver-partitioned rollout equals the unpartitioned one and the oracle."""
import pytest
import torch

from conftest import rel_err
from oracle import aether_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ab():
    import aether_b200
    return aether_b200


def make(ab, D, H, seed=1, precision="fp32"):
    torch.manual_seed(seed)
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda", precision=precision)
    m.eval()
    p = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    return m, p


@pytest.mark.parametrize("D,N,radius,k", [(3, 3000, 0.45, 32), (3, 1500, 0.2, 32), (2, 2000, 0.12, 32), (3, 700, 5.0, 5),
                                          (2, 40, 0.01, 8), (3, 1, 1.0, 4)])
def test_radius_graph_bit_exact(ab, D, N, radius, k):
    g = torch.Generator().manual_seed(N + k)
    pos = torch.randn(N, D, generator=g)
    s, r = ab.radius_graph(pos.cuda(), radius, k)
    so, ro = O.radius_knn_edges(pos, radius, k)
    assert s.dtype == torch.int64 and torch.equal(s.cpu(), so) and torch.equal(r.cpu(), ro)


def test_radius_graph_ties_and_duplicates(ab):
    # lattice points: many exactly equal distances, plus coincident particles
    ax = torch.arange(6, dtype=torch.float32)
    pos = torch.stack(torch.meshgrid(ax, ax, ax, indexing="ij"), -1).reshape(-1, 3)
    pos = torch.cat([pos, pos[:5]])
    s, r = ab.radius_graph(pos.cuda(), 1.5, 7)
    so, ro = O.radius_knn_edges(pos, 1.5, 7)
    assert torch.equal(s.cpu(), so) and torch.equal(r.cpu(), ro)


def _scene(D, N, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(N, D, generator=g) * (N / 40.0) ** (1.0 / D)
    v = torch.randn(N, D, generator=g)
    v = 0.05 * v / v.norm(dim=-1, keepdim=True)
    q = torch.randint(0, 2, (N, 1), generator=g).float() * 2 - 1
    return x, v, q


@pytest.mark.parametrize("D,N,H,parts", [(3, 900, 64, 3), (2, 500, 32, 4), (3, 640, 128, 8)])
def test_partitioned_rollout_fp32(ab, D, N, H, parts):
    m, p = make(ab, D, H)
    x, v, q = _scene(D, N, 7)
    send, recv = ab.radius_graph(x.cuda(), 1.0, 12)
    T = 3
    whole = ab.rollout(m, x.cuda(), v.cuda(), q.cuda(), T, edges=(send, recv))
    pr = ab.PartitionedRollout(m, send, recv, N, T, parts=parts)
    got = pr.run(x.cuda(), v.cuda(), q.cuda())
    assert rel_err(got.cpu(), whole.cpu()) <= 1e-6
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: (send.cpu(), recv.cpu()), T)
    assert rel_err(got.cpu(), want) <= 5e-5


def test_partitioned_rollout_empty_part_and_isolated_nodes(ab):
    D, N, H = 3, 64, 32
    m, p = make(ab, D, H)
    x, v, q = _scene(D, N, 9)
    # only the first 20 nodes have edges: the last partitions own no edges at all
    send, recv = O.radius_knn_edges(x[:20], 10.0, 6)
    pr = ab.PartitionedRollout(m, send.cuda(), recv.cuda(), N, 2, parts=4)
    got = pr.run(x.cuda(), v.cuda(), q.cuda())
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: (send, recv), 2)
    assert rel_err(got.cpu(), want) <= 5e-5


@pytest.mark.parametrize("H", [64, 256])
def test_partitioned_rollout_bf16(ab, H):
    D, N = 3, 1200
    m, p = make(ab, D, H, precision="bf16")
    x, v, q = _scene(D, N, 11)
    send, recv = ab.radius_graph(x.cuda(), 1.0, 16)
    whole = ab.rollout(m, x.cuda(), v.cuda(), q.cuda(), 2, edges=(send, recv))
    got = ab.PartitionedRollout(m, send, recv, N, 2, parts=3).run(x.cuda(), v.cuda(), q.cuda())
    # same kernels on row subsets: identical up to nothing at all (bf16 rounding happens per row)
    assert rel_err(got.cpu(), whole.cpu()) <= 1e-6
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: (send.cpu(), recv.cpu()), 2)
    assert rel_err(got.cpu(), want) <= 2e-2


def test_partitioned_rollout_cuda_graph(ab):
    D, N, H = 3, 800, 64
    m, _ = make(ab, D, H)
    x, v, q = _scene(D, N, 13)
    send, recv = ab.radius_graph(x.cuda(), 1.0, 12)
    eager = ab.PartitionedRollout(m, send, recv, N, 3, parts=2).run(x.cuda(), v.cuda(), q.cuda()).clone()
    gr = ab.PartitionedRollout(m, send, recv, N, 3, parts=2, cuda_graph=True)
    a = gr.run(x.cuda(), v.cuda(), q.cuda()).clone()
    b = gr.run(x.cuda(), v.cuda(), q.cuda()).clone()          # second call = pure replay
    assert torch.equal(a, eager) and torch.equal(b, eager) and gr.launches_per_replay > 0
    x2 = x + 0.01
    c = gr.run(x2.cuda(), v.cuda(), q.cuda())
    want = ab.PartitionedRollout(m, send, recv, N, 3, parts=2).run(x2.cuda(), v.cuda(), q.cuda())
    assert torch.equal(c, want)
