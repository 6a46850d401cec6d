"""DynamicFieldAether on the GPU (aether_b200/dynamic_field.py -> csrc/dynfield.cu + the step with an external field)
against fixtures recorded from the unmodified reference and against the oracle at a larger shape."""
import pytest
import torch

from conftest import golden_params, load_golden, rel_err

pytestmark = pytest.mark.gpu


def _run(g, precision):
    import aether_b200 as ab
    from aether_b200.dynamic_field import DynamicFieldAether
    D, H, N = int(g["D"]), int(g["H"]), int(g["N"])
    m = DynamicFieldAether(2 * D, H, 0.0, D, device="cuda", precision=precision)
    m.load_state_dict(golden_params(g))
    m.train()
    dev = torch.device("cuda")
    x = torch.from_numpy(g["x"]).to(dev).requires_grad_(True)
    vel = torch.from_numpy(g["vel"]).to(dev).requires_grad_(True)
    q = torch.from_numpy(g["charges"]).to(dev)
    ea = torch.from_numpy(g["edge_attr_orig"]).to(dev)
    send, recv = torch.from_numpy(g["send"]).to(dev), torch.from_numpy(g["recv"]).to(dev)
    out = m(None, x, [send, recv], vel, ea, q, N)
    loss = torch.nn.functional.mse_loss(out, torch.from_numpy(g["target"]).to(dev))
    loss.backward()
    return m, out, loss, x, vel


@pytest.mark.parametrize("name", ["dynfield2d_h32", "dynfield3d_h64"])
def test_dynfield_matches_reference_fp32(name):
    g = load_golden(name)
    m, out, loss, x, vel = _run(g, "fp32")
    assert rel_err(out.detach().cpu(), g["out"]) <= 1e-5
    assert abs(float(loss) - float(g["loss"])) <= 1e-6 * max(1.0, abs(float(g["loss"])))
    assert rel_err(x.grad.cpu(), g["grad/x"]) <= 1e-4 and rel_err(vel.grad.cpu(), g["grad/vel"]) <= 1e-4
    for k, p in m.named_parameters():
        ref = g["grad/" + k]
        if abs(ref).max() > 1e-9:          # (the gate's output bias has a zero gradient: soft-max shift invariance)
            assert rel_err(p.grad.cpu(), ref) <= 2e-4, k


def test_dynfield_bf16_and_no_grad():
    g = load_golden("dynfield3d_h64")
    m, out, loss, x, vel = _run(g, "bf16")
    assert rel_err(out.detach().cpu(), g["out"]) <= 2e-2
    for k, p in m.named_parameters():
        ref = g["grad/" + k]
        if abs(ref).max() > 1e-9:          # (the gate's output bias has a zero gradient: soft-max shift invariance)
            assert rel_err(p.grad.cpu(), ref) <= 2e-2, k
    with torch.no_grad():
        D, N = int(g["D"]), int(g["N"])
        dev = torch.device("cuda")
        args = [torch.from_numpy(g[k]).to(dev) for k in ("x", "vel", "edge_attr_orig", "charges", "send", "recv")]
        o2 = m(None, args[0], [args[4], args[5]], args[1], args[2], args[3], N)
    assert torch.equal(o2, out.detach())


def test_dynfield_vs_oracle_larger_graphs():
    """20 particles per graph, 16 graphs, H = 128: attention pooling over longer segments, FiLM per graph."""
    import aether_b200 as ab
    from aether_b200.dynamic_field import DynamicFieldAether
    from oracle import aether_oracle as O
    D, H, N, B = 3, 128, 20, 16
    torch.manual_seed(5)
    m = DynamicFieldAether(2 * D, H, 0.0, D, device="cuda")
    p = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    x, v, q = O.synth_state(B, N, D, seed=3)
    send, recv = O.fc_edges(N, B)
    _, ea = O.caller_prep(x, v, q, send, recv)
    dev = torch.device("cuda")
    out = m(None, x.to(dev), [send.to(dev), recv.to(dev)], v.to(dev), ea.to(dev), q.to(dev), N)
    torch.nn.functional.mse_loss(out, (x + 0.1).to(dev)).backward()
    want = O.dynamic_field_step(p, x, (send, recv), v, ea, q, N)
    torch.nn.functional.mse_loss(want, x + 0.1).backward()
    assert rel_err(out.detach().cpu(), want.detach()) <= 1e-5
    for k, prm in m.named_parameters():
        if float(p[k].grad.abs().max()) > 1e-9:
            assert rel_err(prm.grad.cpu(), p[k].grad) <= 3e-4, k
