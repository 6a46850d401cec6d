"""Host logic of the LoCS drop-in on CPU: the zero-field embedding of its parameters into the 47 Aether-shaped tensors
(aether_b200/locs.py) is checked by running the ORACLE's Aether step on the embedded tensors and comparing with the
golden LoCS vectors recorded from the unmodified reference; gradients flow back to the LoCS-shaped parameters."""
import numpy as np
import pytest
import torch

import aether_b200
from oracle import aether_oracle as O
from conftest import golden_params, load_golden, rel_err


@pytest.mark.parametrize("case", ["locs2d_h32", "locs3d_h64"])
def test_zero_field_embedding_reproduces_locs(case):
    g = load_golden(case)
    D = int(g["x"].shape[1])
    gp = golden_params(g)
    H = gp["gnn.layer_2.message_fn.0.weight"].shape[0]
    m = aether_b200.LoCS(2 * D, H, 0.0, D, device="cpu")
    assert list(m.state_dict().keys()) == list(gp.keys())          # the reference's names, in the reference's order
    m.load_state_dict(gp)
    names = [k for k, _ in O.param_shapes(D, H)]
    tensors = m._aether_params()
    assert len(tensors) == len(names)
    for (k, shape), t in zip(O.param_shapes(D, H), tensors):
        assert tuple(t.shape) == tuple(shape), k
    p = dict(zip(names, tensors))
    x = torch.from_numpy(g["x"])
    v = torch.from_numpy(g["vel"])
    send, recv = torch.from_numpy(g["send"]), torch.from_numpy(g["recv"])
    ea = torch.from_numpy(g["edge_attr_orig"])
    q = torch.zeros(x.shape[0], 1)
    out = O.aether_step(p, x, (send, recv), v, ea, q)
    assert float((out.detach() - torch.from_numpy(g["out"])).abs().max()) <= 2e-6
    loss = torch.nn.functional.mse_loss(out, torch.from_numpy(g["target"]))
    loss.backward()
    for k, t in m.named_parameters():
        want = g["grad/" + k]
        if float(np.abs(want).max()) > 0:
            assert t.grad is not None and rel_err(t.grad, want) <= 1e-4, k


def test_locs_rejects_cpu_forward():
    m = aether_b200.LoCS(4, 32, 0.0, 2, device="cpu")
    with pytest.raises(RuntimeError):
        m(None, torch.zeros(3, 2), (torch.zeros(0, dtype=torch.int64), torch.zeros(0, dtype=torch.int64)), torch.zeros(3, 2),
          torch.zeros(0, 2))
