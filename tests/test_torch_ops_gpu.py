"""The registered operator torch.ops.aether_b200.step on the GPU: same result as the module's forward, parity with the
oracle, capturable in a CUDA graph."""
import pytest
import torch

from conftest import rel_err
from oracle import aether_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-5), ("bf16", 2e-2)])
def test_registered_step_matches_module_and_oracle(precision, tol):
    import aether_b200 as ab
    import aether_b200.torch_ops as tops
    D, N, B, H = 3, 12, 5, 64
    torch.manual_seed(4)
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda", precision=precision)
    m.eval()
    p = {k: t.detach().cpu() for k, t in m.state_dict().items()}
    x, v, q = O.synth_state(B, N, D, seed=9)
    send, recv = O.fc_edges(N, B)
    _, ea = O.caller_prep(x, v, q, send, recv)
    xc, vc, qc, eac = x.cuda(), v.cuda(), q.cuda(), ea.cuda()
    g = ab.PreparedGraph(send.cuda(), recv.cuda(), B * N)
    with torch.no_grad():
        want = O.aether_step(p, x, (send, recv), v, ea, q)
        ref = m(None, xc, g, vc, eac, qc)
        got = tops.step_from_model(m, g, xc, vc, qc, eac)
    assert torch.equal(got, ref)                      # same library call underneath
    assert rel_err(got.cpu(), want) <= tol
    # capturable: no host reads, all work on the current stream
    static_out = torch.empty_like(got)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        static_out.copy_(tops.step_from_model(m, g, xc, vc, qc, eac))
    torch.cuda.current_stream().wait_stream(s)
    cg = torch.cuda.CUDAGraph()
    with torch.cuda.graph(cg):
        static_out.copy_(tops.step_from_model(m, g, xc, vc, qc, eac))
    static_out.zero_()
    cg.replay()
    torch.cuda.synchronize()
    assert torch.equal(static_out, ref)


def test_registered_training_pair_matches_autograd_function():
    """torch.ops.aether_b200.step_fwd / step_bwd (register_autograd) give the gradients of the module's default path, and
    Aether.forward routes through them when asked to."""
    import aether_b200 as ab
    import aether_b200.torch_ops as tops
    D, N, B, H = 2, 10, 4, 64
    torch.manual_seed(2)
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda")
    m.train()
    x, v, q = O.synth_state(B, N, D, seed=4)
    send, recv = O.fc_edges(N, B)
    _, ea = O.caller_prep(x, v, q, send, recv)
    g = ab.PreparedGraph(send.cuda(), recv.cuda(), B * N)
    tgt = (x + 0.1).cuda()

    def run(route):
        m.zero_grad(set_to_none=True)
        m.use_registered_ops = route
        xc = x.cuda().requires_grad_(True)
        vc = v.cuda().requires_grad_(True)
        out = m(None, xc, g, vc, ea.cuda(), q.cuda())
        torch.nn.functional.mse_loss(out, tgt).backward()
        return out.detach(), xc.grad, vc.grad, [p.grad.clone() for p in m.parameters()]

    a = run(False)
    b = run(True)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])
    assert all(torch.equal(p, r) for p, r in zip(a[3], b[3]))
    # fake kernels: shapes propagate without touching the device
    from torch._subclasses.fake_tensor import FakeTensorMode
    with FakeTensorMode(allow_non_fake_inputs=True) as mode:
        fx = mode.from_tensor(x.cuda())
        out, ws = torch.ops.aether_b200.step_fwd(fx, fx, mode.from_tensor(q.cuda().reshape(-1)), mode.from_tensor(ea.cuda()),
                                                 g.perm, g.rowptr, g.send, g.recv, g.sperm, g.colptr,
                                                 [p.detach() for p in m.param_list()], g.n_edges, H, m.flags())
        assert out.shape == x.shape and ws.dtype == torch.uint8 and ws.numel() > 0
