"""Size-independent properties at BASELINE.json's full sizes (the oracle cannot run these in seconds):
config 4 (1024 graphs x 100 particles fully connected, H = 256, bf16 tensor-core path, train step) and
config 5 (100k particles, radius graph capped at 32, receiver-partitioned)."""
import pytest
import torch

from conftest import rel_err
from oracle import aether_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ab():
    import aether_b200
    return aether_b200


def _need(gb):
    free, _ = torch.cuda.mem_get_info()
    if free < gb * 2 ** 30:
        pytest.skip("needs %d GB of free device memory" % gb)


def _train_grads(ab, m, x, v, q, t, send, recv):
    for p in m.parameters():
        p.grad = None
    h, ea = ab.caller_prep(x, v, q.reshape(-1), send, recv)
    out = m(h, x, [send, recv], v, ea, q)
    loss = torch.nn.functional.mse_loss(out, t)
    loss.backward()
    return out.detach(), float(loss.detach()), [p.grad.clone() for p in m.parameters()]


def test_config4_full_size_properties(ab):
    _need(80)
    D, N, B, H = 2, 100, 1024, 256
    torch.manual_seed(1)
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda", precision="bf16")
    m.train()
    x, v, q = [t.cuda() for t in O.synth_state(B, N, D, seed=1)]
    t = x + 0.05
    send, recv = ab.fc_edges(N, B, "cuda")
    assert send.numel() == 10137600
    out, loss, grads = _train_grads(ab, m, x, v, q, t, send, recv)
    assert torch.isfinite(out).all() and all(torch.isfinite(g).all() for g in grads)
    # run-to-run determinism (no atomics anywhere on the path)
    out2, loss2, grads2 = _train_grads(ab, m, x, v, q, t, send, recv)
    assert torch.equal(out, out2) and loss == loss2 and all(torch.equal(a, b) for a, b in zip(grads, grads2))
    # batch independence: the first 8 graphs alone give the same rows (graphs share no edges)
    k = 8
    s8, r8 = ab.fc_edges(N, k, "cuda")
    out8, _, _ = _train_grads(ab, m, x[:k * N], v[:k * N], q[:k * N], t[:k * N], s8, r8)
    assert rel_err(out8.cpu(), out[:k * N].cpu()) <= 1e-6
    # linearity of the gradient in the batch: the mean loss over 1024 graphs = mean of the two halves' mean losses
    hb = B // 2
    sh, rh = ab.fc_edges(N, hb, "cuda")
    _, l0, g0 = _train_grads(ab, m, x[:hb * N], v[:hb * N], q[:hb * N], t[:hb * N], sh, rh)
    _, l1, g1 = _train_grads(ab, m, x[hb * N:], v[hb * N:], q[hb * N:], t[hb * N:], sh, rh)
    assert abs(0.5 * (l0 + l1) - loss) <= 1e-5 * abs(loss)
    for a, b0, b1, (name, _) in zip(grads, g0, g1, m.named_parameters()):
        if float(a.abs().max()) > 0:
            assert rel_err((0.5 * (b0 + b1)).cpu(), a.cpu()) <= 2e-3, name


def test_config5_full_size_partition(ab):
    _need(40)
    n, H, T = 100000, 64, 2
    g = torch.Generator().manual_seed(1)
    L = float(n) ** (1.0 / 3.0)
    x = (torch.rand(n, 3, generator=g) * L).cuda()
    v = torch.randn(n, 3, generator=g)
    v = (0.05 * v / v.norm(dim=-1, keepdim=True)).cuda()
    q = (torch.randint(0, 2, (n, 1), generator=g).float() * 2 - 1).cuda()
    send, recv = ab.radius_graph(x, 2.25, 32)
    E = send.numel()
    deg = torch.bincount(send, minlength=n)
    assert int(deg.max()) <= 32 and 3.0e6 < E <= 3.2e6
    d = (x[send] - x[recv]).norm(dim=1)
    assert float(d.max()) <= 2.25 * (1 + 1e-6) and bool((send != recv).all())
    # sender-major, nearest first within each sender
    assert bool((send[1:] >= send[:-1]).all())
    same = send[1:] == send[:-1]
    assert bool((d[1:][same] >= d[:-1][same]).all())
    torch.manual_seed(1)
    m = ab.Aether(6, H, 0.0, 3, device="cuda", precision="bf16")
    m.eval()
    whole = ab.PartitionedRollout(m, send, recv, n, T, parts=1).run(x, v, q).clone()
    parts = ab.PartitionedRollout(m, send, recv, n, T, parts=8).run(x, v, q)
    assert torch.isfinite(whole).all()
    assert rel_err(parts.cpu(), whole.cpu()) <= 1e-6
