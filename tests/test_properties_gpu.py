"""GPU tests beyond the golden vectors: differential against the oracle on fresh seeded inputs,
edge cases, and size-independent properties at larger sizes (edge-order invariance, batch
independence, run-to-run determinism, drop-in behaviour of the module)."""
import io
import os

import numpy as np
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
    p = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    return m, p


def run_gpu(ab, m, x, v, q, send, recv, grad=False):
    xc, vc = x.cuda().requires_grad_(grad), v.cuda().requires_grad_(grad)
    h, ea = ab.caller_prep(xc.detach(), vc.detach(), q.cuda().reshape(-1), send.cuda(), recv.cuda())
    out = m(h, xc, [send.cuda(), recv.cuda()], vc, ea, q.cuda())
    return out, xc, vc


@pytest.mark.parametrize("D,N,B,H", [(2, 5, 16, 64), (3, 5, 128, 64), (3, 20, 8, 64), (2, 30, 4, 128), (2, 12, 3, 256)])
def test_forward_backward_vs_oracle(ab, D, N, B, H):
    m, p = make(ab, D, H)
    x, v, q = O.synth_state(B, N, D, seed=3)
    send, recv = O.fc_edges(N, B)
    out, xc, vc = run_gpu(ab, m, x, v, q, send, recv, grad=True)
    target = x + 0.05
    torch.nn.functional.mse_loss(out, target.cuda()).backward()

    po = {k: t.clone().requires_grad_(True) for k, t in p.items()}
    xo, vo = x.clone().requires_grad_(True), v.clone().requires_grad_(True)
    _, ea = O.caller_prep(x, v, q, send, recv)
    want = O.aether_step(po, xo, (send, recv), vo, ea, q)
    torch.nn.functional.mse_loss(want, target).backward()
    assert rel_err(out.detach().cpu(), want.detach()) <= 1e-5
    assert rel_err(xc.grad.cpu(), xo.grad) <= 1e-4
    assert rel_err(vc.grad.cpu(), vo.grad) <= 1e-4
    for k, prm in m.named_parameters():
        w = po[k].grad if po[k].grad is not None else torch.zeros_like(po[k])
        if float(w.abs().max()) == 0:
            assert float(prm.grad.abs().max()) == 0, k
        else:
            assert rel_err(prm.grad.cpu(), w) <= 1e-4, k


def test_knn_ragged_rollout_vs_oracle(ab):
    D, H, k = 2, 64, 10
    sizes = [10, 40, 17, 3, 25, 1, 2, 33]
    m, p = make(ab, D, H)
    g = torch.Generator().manual_seed(5)
    x = torch.rand(sum(sizes), D, generator=g)
    v = (torch.rand(sum(sizes), D, generator=g) - 0.5) * 0.1
    q = torch.randint(0, 2, (sum(sizes), 1), generator=g).float() * 2 - 1
    plan = ab.KnnPlan(sizes, k, "cuda")
    s, r = ab.knn_edges(x.cuda(), plan)
    so, ro = O.knn_edges_batched(x, sizes, k)
    assert torch.equal(s.cpu(), so) and torch.equal(r.cpu(), ro)
    m.eval()
    xs = ab.rollout(m, x.cuda(), v.cuda(), q.cuda(), 4, knn_plan=plan)
    with torch.no_grad():
        want = O.rollout(p, x, v, q, lambda xx: O.knn_edges_batched(xx, sizes, k), 4)
    assert rel_err(xs.cpu(), want) <= 5e-5


def test_edge_cases(ab):
    m, p = make(ab, 3, 32)
    # zero velocity, +-z headings, coincident particles, an isolated node, E = 0
    x = torch.tensor([[0., 0, 0], [0, 0, 0], [1, 1, 1], [2, 0, 1], [5, 5, 5]])
    v = torch.tensor([[0., 0, 0], [0, 0, 2], [0, 0, -2], [1, 1, 0], [0.3, 0.1, 0.2]])
    q = torch.tensor([[1.], [-1], [1], [-1], [1]])
    send = torch.tensor([0, 1, 2, 3, 0, 2])
    recv = torch.tensor([1, 0, 3, 2, 2, 0])        # node 4 isolated
    out, _, _ = run_gpu(ab, m, x, v, q, send, recv)
    _, ea = O.caller_prep(x, v, q, send, recv)
    with torch.no_grad():
        want = O.aether_step(p, x, (send, recv), v, ea, q)
    assert torch.isfinite(out).all()
    assert rel_err(out.detach().cpu(), want) <= 1e-5
    e0 = torch.zeros(0, dtype=torch.int64)
    out0, _, _ = run_gpu(ab, m, x, v, q, e0, e0, grad=True)
    with torch.no_grad():
        want0 = O.aether_step(p, x, (e0, e0), v, torch.zeros(0, 2), q)
    assert rel_err(out0.detach().cpu(), want0) <= 1e-5
    out0.sum().backward()
    assert all(torch.isfinite(prm.grad).all() for prm in m.parameters())
    # int32 indices are accepted
    out32 = m(None, x.cuda(), [send.int().cuda(), recv.int().cuda()], v.cuda(), ea.cuda(), q.cuda())
    assert torch.equal(out32, out.detach())


def test_properties_large(ab):
    """100-particle fully connected graphs (config 4 shape, smaller batch): properties that need no oracle."""
    D, N, B, H = 2, 100, 32, 64
    m, _ = make(ab, D, H)
    m.eval()
    x, v, q = O.synth_state(B, N, D, seed=11)
    send, recv = ab.fc_edges(N, B, "cuda")
    with torch.no_grad():
        out, _, _ = run_gpu(ab, m, x, v, q, send.cpu(), recv.cpu())
        out2, _, _ = run_gpu(ab, m, x, v, q, send.cpu(), recv.cpu())
        assert torch.equal(out, out2)                                  # deterministic (no atomics)
        # batch independence: graphs 3..4 computed alone give the same rows
        sl = slice(3 * N, 5 * N)
        s2, r2 = ab.fc_edges(N, 2, "cuda")
        outs, _, _ = run_gpu(ab, m, x[sl], v[sl], q[sl], s2.cpu(), r2.cpu())
        assert rel_err(outs.cpu(), out[sl].cpu()) <= 1e-6
        # edge-order invariance: a shuffled edge list gives the same step (sum order inside a segment changes)
        perm = torch.randperm(send.numel(), generator=torch.Generator().manual_seed(0))
        outp, _, _ = run_gpu(ab, m, x, v, q, send.cpu()[perm], recv.cpu()[perm])
        assert rel_err(outp.cpu(), out.cpu()) <= 1e-5
        # oracle on one graph of the batch
        p = {k: t.detach().cpu() for k, t in m.state_dict().items()}
        so, ro = O.fc_edges(N, 1)
        _, ea = O.caller_prep(x[:N], v[:N], q[:N], so, ro)
        want = O.aether_step(p, x[:N], (so, ro), v[:N], ea, q[:N])
        assert rel_err(out[:N].cpu(), want) <= 1e-5


def test_module_dropin(ab):
    m, p = make(ab, 3, 64)
    assert m.params == "132822"
    buf = io.BytesIO()
    torch.save(m, buf)                                                   # experiments/lorentz/main.py:187
    buf.seek(0)
    m2 = torch.load(buf, weights_only=False)
    x, v, q = O.synth_state(4, 5, 3, seed=2)
    s, r = O.fc_edges(5, 4)
    a, _, _ = run_gpu(ab, m, x, v, q, s, r)
    b, _, _ = run_gpu(ab, m2, x, v, q, s, r)
    assert torch.equal(a.detach(), b.detach())
    # one AdamW step as the reference's loop does (main.py:164,289-292)
    opt = torch.optim.AdamW(m.parameters(), lr=5e-4, weight_decay=1e-12)
    opt.zero_grad()
    out, _, _ = run_gpu(ab, m, x, v, q, s, r)
    loss = torch.nn.functional.mse_loss(out, x.cuda())
    loss.backward()
    opt.step()
    assert np.isfinite(loss.item())
    with pytest.raises(RuntimeError):
        m(None, x, [s, r], v, torch.zeros(s.numel(), 2), q)             # CPU tensors: no fallback
    # dropout is honoured in train mode and off in eval mode
    torch.manual_seed(0)
    md = ab.Aether(6, 64, 0.5, 3, device="cuda")
    md.train()
    o1, _, _ = run_gpu(ab, md, x, v, q, s, r)
    o2, _, _ = run_gpu(ab, md, x, v, q, s, r)
    assert not torch.equal(o1, o2)
    md.eval()
    o3, _, _ = run_gpu(ab, md, x, v, q, s, r)
    o4, _, _ = run_gpu(ab, md, x, v, q, s, r)
    assert torch.equal(o3, o4)


@pytest.mark.parametrize("fused", [False, True])
def test_graphed_train_step_matches_eager(ab, fused):
    """The captured train step (CUDA graph) takes exactly the steps the eager loop takes -- with torch's capturable AdamW
    and with the library's FusedAdamW (csrc/optim.cu) inside the graph; the eager side is always torch.optim.AdamW."""
    D, N, B, H = 3, 5, 16, 64
    x, v, q = O.synth_state(B, N, D, seed=21)
    tgt = x + 0.05
    send, recv = ab.fc_edges(N, B, "cuda")
    xs, vs, qs, ts = x.cuda(), v.cuda(), q.cuda(), tgt.cuda()
    m1, _ = make(ab, D, H, seed=4)
    m2, _ = make(ab, D, H, seed=4)
    m1.train(); m2.train()
    o1 = torch.optim.AdamW(m1.parameters(), lr=5e-4, weight_decay=1e-12, fused=True)
    o2 = (ab.FusedAdamW(m2.parameters(), lr=5e-4, weight_decay=1e-12) if fused else
          torch.optim.AdamW(m2.parameters(), lr=5e-4, weight_decay=1e-12, fused=True, capturable=True))
    g = ab.PreparedGraph(send, recv, B * N)
    losses1 = []
    for _ in range(4):
        o1.zero_grad(set_to_none=True)
        h, ea = ab.caller_prep(xs, vs, qs.reshape(-1), send, recv)
        loss = torch.nn.functional.mse_loss(m1(h, xs, g, vs, ea, qs), ts)
        loss.backward()
        o1.step()
        losses1.append(float(loss))
    gs = ab.GraphedTrainStep(m2, o2, (send, recv), B * N)
    losses2 = [float(gs.step(xs, vs, qs, ts)) for _ in range(4)]
    assert np.allclose(losses1, losses2, rtol=1e-5, atol=1e-8), (losses1, losses2)
    assert losses1[-1] < losses1[0]
    for (k, a), (_, b) in zip(m1.named_parameters(), m2.named_parameters()):
        assert rel_err(b.detach().cpu(), a.detach().cpu()) <= 1e-5, k


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_no_grad_forward_equals_grad_forward(precision):
    """Under torch.no_grad() the module takes the ping-pong (no save-for-backward) workspace; its output must be
    bit-identical to the grad-enabled forward, which keeps all eight per-layer pre-activations."""
    import aether_b200 as ab
    D, N, B, H = 2, 20, 6, 64
    torch.manual_seed(5)
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda", precision=precision)
    x, v, q = [t.cuda() for t in O.synth_state(B, N, D, seed=12)]
    send, recv = ab.fc_edges(N, B, "cuda")
    h, ea = ab.caller_prep(x, v, q.reshape(-1), send, recv)
    a = m(h, x, [send, recv], v, ea, q)
    assert a.requires_grad
    with torch.no_grad():
        b = m(h, x, [send, recv], v, ea, q)
    assert not b.requires_grad
    assert torch.equal(a.detach(), b)


def test_backward_edge_cases_raise_clearly():
    import aether_b200 as ab
    D, N, B, H = 2, 5, 2, 64
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda")
    x, v, q = [t.cuda() for t in O.synth_state(B, N, D, seed=2)]
    send, recv = ab.fc_edges(N, B, "cuda")
    h, ea = ab.caller_prep(x, v, q.reshape(-1), send, recv)
    out = m(h, x, [send, recv], v, ea, q)
    out.sum().backward(retain_graph=True)
    with pytest.raises(RuntimeError, match="already consumed"):
        out.sum().backward()
    with pytest.raises(RuntimeError, match="edge_attr_orig"):
        m(h, x, [send, recv], v, ea.clone().requires_grad_(True), q)
    # dropout_prob == 1: nn.Dropout zeroes everything (no 0/0)
    m1 = ab.Aether(2 * D, H, 1.0, D, device="cuda")
    m1.train()
    o1 = m1(h, x, [send, recv], v, ea, q)
    assert torch.isfinite(o1).all()


def test_out_of_range_edges_are_memory_safe_and_flagged():
    """The reference raises IndexError on an out-of-range edge; here the list is clamped (never an out-of-bounds read),
    the status word is set and validate() / AETHER_B200_VALIDATE=1 raises."""
    import aether_b200 as ab
    D, N, B, H = 2, 5, 2, 64
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda")
    x, v, q = [t.cuda() for t in O.synth_state(B, N, D, seed=2)]
    send, recv = ab.fc_edges(N, B, "cuda")
    bad_s, bad_r = send.clone(), recv.clone()
    bad_r[3] = B * N + 7
    bad_s[11] = -2
    g = ab.PreparedGraph(bad_s, bad_r, B * N)
    with pytest.raises(IndexError):
        g.validate()
    h, ea = ab.caller_prep(x, v, q.reshape(-1), send, recv)
    with torch.no_grad():
        out = m(h, x, g, v, ea, q)
    torch.cuda.synchronize()            # a fault would surface here
    assert out.shape == x.shape
    assert int(g.send.min()) >= 0 and int(g.send.max()) < B * N and int(g.recv.max()) < B * N
