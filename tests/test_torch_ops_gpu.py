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
