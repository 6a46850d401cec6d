"""GPU property tests (passed on the B200 in the round-1 driver run).  Size-independent property at a size the oracle
does not finish in seconds:
LoCS on the GPU is equivariant to translations and to rotations (2D: any; 3D: about z) of the scene."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("D,precision,tol", [(2, "fp32", 2e-4), (3, "fp32", 2e-4), (2, "bf16", 5e-2)])
def test_locs_equivariance_large(D, precision, tol):
    import aether_b200 as ab
    N, B, H = 100, 64, 128
    torch.manual_seed(11)
    m = ab.LoCS(2 * D, H, 0.0, D, device="cuda", precision=precision)
    m.eval()
    x = torch.randn(B * N, D, device="cuda") * 3
    v = torch.randn(B * N, D, device="cuda")
    q = torch.zeros(B * N, 1, device="cuda")
    send, recv = ab.fc_edges(N, B, "cuda")
    g = ab.PreparedGraph(send, recv, B * N)
    a = 0.83
    R = torch.eye(D, device="cuda")
    R[0, 0], R[0, 1], R[1, 0], R[1, 1] = math.cos(a), -math.sin(a), math.sin(a), math.cos(a)
    t = torch.randn(D, device="cuda")
    with torch.no_grad():
        _, ea = ab.caller_prep(x, v, q, send, recv)
        out = m(None, x, g, v, ea)
        x2, v2 = (x @ R.T + t).contiguous(), (v @ R.T).contiguous()
        out2 = m(None, x2, g, v2, ea)
    # compare displacements (out - x): they rotate with the scene and are not swamped by the positions' magnitude
    d1, d2 = (out - x) @ R.T, out2 - x2
    # an edge whose angle feature sits within rounding of the +-pi branch cut may flip under the rotation (the network is
    # not periodic in it), so the bar is on all but 0.1 % of the nodes
    err = (d1 - d2).abs().amax(dim=1)
    assert float(torch.quantile(err, 0.999)) <= tol * float(d1.abs().max())
