"""GPU parity of the LoCS variant against the reference's golden vectors and the oracle."""
import numpy as np
import pytest
import torch

from conftest import golden_params, load_golden, rel_err
from oracle import aether_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ab():
    import aether_b200
    return aether_b200


@pytest.mark.parametrize("case", ["locs2d_h32", "locs3d_h64"])
def test_locs_golden(ab, case):
    g = load_golden(case)
    D, H = int(g["D"]), int(g["H"])
    m = ab.LoCS(2 * D, H, 0.0, D, device="cuda")
    m.load_state_dict({k: v.cuda() for k, v in golden_params(g).items()})       # the reference's state_dict loads as is
    m.train()
    x = torch.from_numpy(g["x"]).cuda().requires_grad_(True)
    v = torch.from_numpy(g["vel"]).cuda().requires_grad_(True)
    send, recv = torch.from_numpy(g["send"]).cuda(), torch.from_numpy(g["recv"]).cuda()
    ea = torch.from_numpy(g["edge_attr_orig"]).cuda()
    out = m(None, x, [send, recv], v, ea)
    assert rel_err(out.detach().cpu(), g["out"]) <= 1e-5
    loss = torch.nn.functional.mse_loss(out, torch.from_numpy(g["target"]).cuda())
    loss.backward()
    assert abs(float(loss) - float(g["loss"])) <= 1e-6 * max(1.0, abs(float(g["loss"])))
    assert rel_err(x.grad.cpu(), g["grad/x"]) <= 1e-4 and rel_err(v.grad.cpu(), g["grad/vel"]) <= 1e-4
    for k, p in m.named_parameters():
        if float(np.abs(g["grad/" + k]).max()) > 0:
            assert p.grad.shape == p.shape and rel_err(p.grad.cpu(), g["grad/" + k]) <= 1e-4, k


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-5), ("bf16", 2e-2)])
def test_locs_vs_oracle_fresh(ab, precision, tol):
    D, N, B, H = 3, 20, 6, 64
    torch.manual_seed(3)
    m = ab.LoCS(2 * D, H, 0.0, D, device="cuda", precision=precision)
    p = {k: t.detach().cpu() for k, t in m.state_dict().items()}
    x, v, q = O.synth_state(B, N, D, seed=8)
    send, recv = O.fc_edges(N, B)
    _, ea = O.caller_prep(x, v, q, send, recv)
    with torch.no_grad():
        want = O.locs_step(p, x, (send, recv), v, ea)
        got = m(None, x.cuda(), [send.cuda(), recv.cuda()], v.cuda(), ea.cuda())
    assert rel_err(got.cpu(), want) <= tol
