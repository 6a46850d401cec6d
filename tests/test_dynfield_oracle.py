"""Oracle restatement of DynamicFieldAether (oracle/aether_oracle.py: latent_field_forward, dynamic_field_step) against
fixtures recorded from the unmodified reference (tests/golden/make_golden_dynfield.py; nn/state2state/
dynamic_field_aether.py:29-100, graph_pool.py:6-29, film.py:5-60), and the parameter holder's initialisation order."""
import pytest
import torch

from conftest import golden_params, load_golden, rel_err
from oracle import aether_oracle as O

CASES = ["dynfield2d_h32", "dynfield3d_h64"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference(name):
    g = load_golden(name)
    p = {k: v.clone().requires_grad_(True) for k, v in golden_params(g).items()}
    x = torch.from_numpy(g["x"]).requires_grad_(True)
    vel = torch.from_numpy(g["vel"]).requires_grad_(True)
    q, ea = torch.from_numpy(g["charges"]), torch.from_numpy(g["edge_attr_orig"])
    send, recv = torch.from_numpy(g["send"]), torch.from_numpy(g["recv"])
    N = int(g["N"])
    st = {}
    out = O.dynamic_field_step(p, x, (send, recv), vel, ea, q, N, stages=st)
    assert rel_err(st["field"].detach(), g["field"]) <= 1e-6
    assert rel_err(out.detach(), g["out"]) <= 1e-6
    loss = torch.nn.functional.mse_loss(out, torch.from_numpy(g["target"]))
    assert abs(float(loss) - float(g["loss"])) <= 1e-7
    loss.backward()
    assert rel_err(x.grad, g["grad/x"]) <= 1e-5 and rel_err(vel.grad, g["grad/vel"]) <= 1e-5
    for k, t in p.items():
        ref = g["grad/" + k]
        if abs(ref).max() < 1e-9:          # the gate's output bias: soft-max is shift invariant, its gradient is rounding noise
            assert float(t.grad.abs().max()) < 1e-9, k
        else:
            assert rel_err(t.grad, ref) <= 2e-5, k


@pytest.mark.parametrize("name", CASES)
def test_module_reproduces_reference_init_and_state_dict(name):
    """Same seed -> the reference's initial weights, key for key (incl. the second initialisation PyG's
    AttentionalAggregation gives its sub-networks); the holder is built on the CPU, no kernel runs."""
    from aether_b200.dynamic_field import DynamicFieldAether
    g = load_golden(name)
    torch.manual_seed(int(g["seed"]))
    m = DynamicFieldAether(2 * int(g["D"]), int(g["H"]), 0.0, int(g["D"]), device="cpu")
    ref = golden_params(g)
    sd = m.state_dict()
    assert list(sd.keys()) == list(ref.keys())
    for k in ref:
        assert torch.equal(sd[k], ref[k]), k
    assert len(m.param_list()) == 67
    with pytest.raises(RuntimeError):
        z = torch.zeros(5, int(g["D"]))
        m(None, z, [torch.zeros(0, dtype=torch.long)] * 2, z, torch.zeros(0, 2), torch.zeros(5, 1), 5)   # CPU tensors: no fallback
