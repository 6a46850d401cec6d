"""Domain properties of the oracle that do not depend on recorded vectors (they also hold for the reference):

* LoCS is equivariant to translations and to rotations of the scene (any rotation in 2D, rotations about z in 3D:
  nn/state2state/locs/locs.py:20-99 builds every feature in the receiver's velocity frame);
* Aether is equivariant to node relabelling;
* the receiver mean does not depend on the order of the edge list beyond fp32 summation order.
"""
import math

import numpy as np
import pytest
import torch

from oracle import aether_oracle as O


def _rot(D, seed):
    """2D: any rotation.  3D: a rotation about the z axis -- the frame is built from the velocity's spherical angles
    (nn/utils/geometry.py:37-73), which fixes the roll against the global z axis, so only azimuthal rotations commute."""
    g = torch.Generator().manual_seed(seed)
    a = float(torch.rand(1, generator=g)) * 2 * math.pi
    R = torch.eye(D, dtype=torch.float64)
    R[0, 0], R[0, 1], R[1, 0], R[1, 1] = math.cos(a), -math.sin(a), math.sin(a), math.cos(a)
    return R


def _locs_params(D, H, seed):
    """LoCS-shaped parameters: Aether's without the field net and the force columns (locs.py:102-139)."""
    p = O.init_params(D, H, seed)
    Od = D * (D - 1) // 2
    cols = list(range(3 * D + Od)) + list(range(4 * D + Od, 6 * D + Od)) + [7 * D + Od, 7 * D + Od + 1]
    out = {k: v for k, v in p.items() if not k.startswith("field_net.")}
    out["gnn.layer_1.message_fn.0.weight"] = p["gnn.layer_1.message_fn.0.weight"][:, cols].contiguous()
    out["gnn.layer_1.res.weight"] = p["gnn.layer_1.res.weight"][:, :2 * D].contiguous()
    return out


@pytest.mark.parametrize("D", [2, 3])
def test_locs_is_roto_translation_equivariant(D):
    N, B, H = 6, 3, 32
    p = {k: v.double() for k, v in _locs_params(D, H, 3).items()}
    x, v, q = O.synth_state(B, N, D, seed=4)
    x, v = x.double(), v.double()
    send, recv = O.fc_edges(N, B)
    _, ea = O.caller_prep(x.float(), v.float(), q, send, recv)
    ea = ea.double()
    R, t = _rot(D, 7), torch.randn(D, dtype=torch.float64)
    with torch.no_grad():
        out = O.locs_step(p, x, (send, recv), v, ea)
        out_r = O.locs_step(p, x @ R.T + t, (send, recv), v @ R.T, ea)
    assert float((out_r - (out @ R.T + t)).abs().max()) <= 1e-9 * max(1.0, float(out.abs().max()))


def test_aether_is_equivariant_to_node_relabelling():
    D, N, B, H = 3, 7, 2, 32
    p = O.init_params(D, H, 5)
    x, v, q = O.synth_state(B, N, D, seed=6)
    send, recv = O.knn_edges(x, 4)
    _, ea = O.caller_prep(x, v, q, send, recv)
    perm = torch.from_numpy(np.random.default_rng(0).permutation(B * N))
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(B * N)
    with torch.no_grad():
        out = O.aether_step(p, x, (send, recv), v, ea, q)
        out_p = O.aether_step(p, x[perm], (inv[send], inv[recv]), v[perm], ea, q[perm])
    assert float((out_p - out[perm]).abs().max()) <= 2e-6 * float(out.abs().max())


def test_edge_order_only_changes_summation_order():
    D, N, B, H = 2, 9, 2, 32
    p = O.init_params(D, H, 8)
    x, v, q = O.synth_state(B, N, D, seed=9)
    send, recv = O.fc_edges(N, B)
    _, ea = O.caller_prep(x, v, q, send, recv)
    sh = torch.from_numpy(np.random.default_rng(1).permutation(send.numel()))
    with torch.no_grad():
        a = O.aether_step(p, x, (send, recv), v, ea, q)
        b = O.aether_step(p, x, (send[sh], recv[sh]), v, ea[sh], q)
    assert float((a - b).abs().max()) <= 2e-6 * float(a.abs().max())
