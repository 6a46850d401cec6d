"""CPU oracle for the seq2seq decoder path of Aether.  TEST INFRASTRUCTURE ONLY.

A plain torch-CPU (fp32, eager) restatement of ``RecurrentDecoder.forward`` (nn/seq2seq/aether.py:589-654), the
``AugmentedLocalizer`` it calls (nn/utils/augmented_global_to_local.py:51-68, nn/utils/canonicalization.py:33-56,
111-172), the field predictor (nn/seq2seq/aether.py:71-87, nn/nn/fourier_feature_mapper.py:7-23) and the decoder side of
``predict_future`` (nn/seq2seq/aether.py:155-191) with the sampled edge types given as inputs.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import it; nothing under ``aether_b200/`` does.

Parity status: PINNED.  ``tests/golden/make_golden_seq2seq.py`` runs the unmodified reference model (real encoder, real
gumbel sampling under a seed) in the build container; ``tests/test_seq2seq_oracle.py`` checks this restatement against
those vectors.  The one third-party piece is ``torch_scatter.scatter(reduce="mean")`` as in ``aether_oracle.py``; on the
fully connected graphs of this path every receiver has N-1 incoming edges, so it is a plain mean.

Weights are a dict with the reference's ``state_dict`` keys of ``RecurrentDecoder`` (``msg_fc1.0.weight`` ...) and, for
the field, ``field_net.{0,2,4}.{weight,bias}`` + ``rff_B``.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch

Tensor = torch.Tensor
EPS_PHI = 1e-7  # nn/utils/geometry.py:63


def fc_edges_sender_major(n: int) -> Tuple[Tensor, Tensor]:
    """``torch.where(~torch.eye(n, dtype=bool))`` (nn/seq2seq/aether.py:548-549): row index = sender, column = receiver."""
    s, r = torch.where(~torch.eye(n, dtype=torch.bool))
    return s, r


def _rot2(a: Tensor) -> Tensor:
    """nn/utils/geometry.py:7-23 for ndim == 2; ``a`` has a trailing 1."""
    c, s = torch.cos(a), torch.sin(a)
    return torch.stack([torch.cat([c, -s], -1), torch.cat([s, c], -1)], -2)


def _rot3(theta: Tensor, phi: Tensor) -> Tensor:
    """nn/utils/geometry.py:24-34."""
    ct, st, cp, sp = torch.cos(theta), torch.sin(theta), torch.cos(phi), torch.sin(phi)
    return torch.stack([
        torch.cat([cp * ct, -st, sp * ct], -1),
        torch.cat([cp * st, ct, sp * st], -1),
        torch.cat([-sp, torch.zeros_like(ct), cp], -1)], -2)


def _spherical(x: Tensor):
    """nn/utils/geometry.py:37-67 with symmetric_theta=False, ndim == 3."""
    rho = torch.norm(x, p=2, dim=-1, keepdim=True)
    theta = torch.atan2(x[..., [1]], x[..., [0]])
    theta = theta + (theta < 0).type_as(theta) * (2 * math.pi)
    phi = torch.acos(torch.clamp(x[..., 2:] / (rho + EPS_PHI), min=-1.0, max=1.0))
    return rho, theta, phi


def _mv(R: Tensor, v: Tensor) -> Tensor:
    return (R @ v.unsqueeze(-1)).squeeze(-1)


def canonicalize_augmented(x: Tensor, D: int):
    """nn/utils/canonicalization.py:33-56: per-node frame from the velocity, velocity and field rotated into it."""
    if D == 3:
        vel, f = x[..., 3:6], x[..., 6:9]
        _, th, ph = _spherical(vel)
        Rinv = _rot3(th, ph)
        r = Rinv.transpose(-1, -2)
        canon = torch.cat([torch.zeros_like(x[..., :3]), _mv(r, vel), _mv(r, f)], -1)
    else:
        vel, f = x[..., 2:4], x[..., 4:6]
        Rinv = _rot2(torch.atan2(vel[..., [1]], vel[..., [0]]))
        canon = torch.zeros_like(x)
        canon[..., 2] = torch.norm(vel, dim=-1)
        canon[..., 4:6] = _mv(Rinv.transpose(-1, -2), f)
    return canon, Rinv


def pair_features(xj: Tensor, xi: Tensor, D: int) -> Tensor:
    """Sender j seen from receiver i.  2D: nn/utils/canonicalization.py:111-139; 3D: :142-172."""
    if D == 2:
        yaw = torch.atan2(xi[..., [3]], xi[..., [2]])
        r = _rot2(yaw).transpose(-1, -2)
        d = (torch.atan2(xj[..., 3], xj[..., 2]) - torch.atan2(xi[..., 3], xi[..., 2])).clone()   # geometry.py:117-128
        d[d >= math.pi] -= 2 * math.pi
        d[d < -math.pi] += 2 * math.pi
        dyaw = (d / math.pi).unsqueeze(-1)
        dp = xj[..., :2] - xi[..., :2]
        dist = torch.norm(dp, dim=-1, keepdim=True)
        dth = torch.atan2(dp[..., 1], dp[..., 0]).unsqueeze(-1) - yaw
        dth = dth + (dth <= -math.pi).type_as(dth) * (2 * math.pi)                                # geometry.py:108-113
        dth = dth - (dth > math.pi).type_as(dth) * (2 * math.pi)
        dth = dth / math.pi
        return torch.cat([_mv(r, dp), dyaw, dist, dth, _mv(r, xj[..., 2:4]), _mv(r, xj[..., 4:6])], -1)
    _, sy, sp = _spherical(xj[..., 3:6])
    _, ry, rp = _spherical(xi[..., 3:6])
    r = _rot3(ry, rp).transpose(-1, -2)
    dp = xj[..., :3] - xi[..., :3]
    dist = torch.norm(dp, p=2, dim=-1, keepdim=True)
    M = r @ _rot3(sy, sp).transpose(-1, -2)
    euler = torch.stack([torch.atan2(M[..., 1, 0], M[..., 0, 0]), torch.asin(-M[..., 2, 0]),
                         torch.atan2(M[..., 2, 1], M[..., 2, 2])], -1)                            # geometry.py:87-96, no /pi
    rp_ = _mv(r, dp)
    _, dth, dph = _spherical(rp_)
    return torch.cat([rp_, euler, dist, dth, dph, _mv(r, xj[..., 3:6]), _mv(r, xj[..., 6:9])], -1)


def augmented_localizer(x: Tensor, D: int):
    """nn/utils/augmented_global_to_local.py:51-68 for ``x[B, N, 3D]`` = [pos | vel | field].

    Returns ``rel_feat[B, N, 3D+F]``, ``Rinv[B, N, D, D]``, ``edge_attr[B, N(N-1), 2F+3D]`` with F = 4D + D(D-1)/2."""
    B, N, _ = x.shape
    canon, Rinv = canonicalize_augmented(x, D)
    origin = torch.zeros(3 * D, dtype=x.dtype)
    origin[D] = 1.0                                                  # :41: unit velocity along x, at the origin
    ext = torch.cat([x, origin.expand(B, 1, 3 * D)], 1)
    s, r = torch.where(~torch.eye(N + 1, N, dtype=torch.bool))       # :33-34 (origin node N only sends)
    attr = pair_features(ext[:, s], ext[:, r], D)
    E = N * (N - 1)
    origin_attr, attr = attr[:, E:], attr[:, :E]
    recv = r[:E]
    edge_attr = torch.cat([attr, canon[:, recv], origin_attr[:, recv]], -1)
    rel_feat = torch.cat([canon, origin_attr], -1)
    return rel_feat, Rinv, edge_attr


def _lin(w: Dict[str, Tensor], name: str, x: Tensor) -> Tensor:
    b = w.get(name + ".bias")
    y = x @ w[name + ".weight"].t()
    return y if b is None else y + b


def decoder_step(w: Dict[str, Tensor], inputs: Tensor, hidden: Tensor, edges: Tensor, field: Tensor,
                 D: int, n_types: int, skip_first: bool):
    """``RecurrentDecoder.forward`` in eval mode (nn/seq2seq/aether.py:589-654).

    inputs[B,N,2D], hidden[B,N,H], edges[B,N(N-1),n_types], field[B,N,D] -> (outputs[B,N,2D], hidden'[B,N,H])."""
    B, N, _ = inputs.shape
    send, recv = fc_edges_sender_major(N)
    pre = torch.cat([hidden[:, recv], hidden[:, send]], -1)                                   # :596-600
    start = 1 if skip_first else 0
    msgs = torch.zeros(B, send.numel(), hidden.shape[-1], dtype=inputs.dtype)
    for i in range(start, n_types):                                                           # :608-613
        m = torch.tanh(_lin(w, f"msg_fc1.{i}", pre))
        m = torch.tanh(_lin(w, f"msg_fc2.{i}", m))
        msgs = msgs + m * edges[:, :, i:i + 1]
    agg = torch.zeros(B, N, msgs.shape[-1], dtype=msgs.dtype).index_add_(1, recv, msgs) / (N - 1)   # :616-617
    rel_feat, Rinv, edge_attr = augmented_localizer(torch.cat([inputs, field], -1), D)        # :619-621
    pmsgs = torch.zeros_like(msgs)
    for i in range(start, n_types):                                                           # :625-632
        m = torch.relu(_lin(w, f"present_msg_fc1.{i}", edge_attr))
        m = torch.relu(_lin(w, f"present_msg_fc2.{i}", m))
        pmsgs = pmsgs + m * edges[:, :, i:i + 1]
    pagg = torch.zeros_like(agg).index_add_(1, recv, pmsgs) / (N - 1)                         # :634
    inp_r = _lin(w, "input_r", rel_feat) + _lin(w, "present_r", pagg)                         # :638-640
    inp_i = _lin(w, "input_i", rel_feat) + _lin(w, "present_i", pagg)
    inp_n = _lin(w, "input_n", rel_feat) + _lin(w, "present_n", pagg)
    r = torch.sigmoid(inp_r + _lin(w, "hidden_r", agg))                                       # :642-645
    i = torch.sigmoid(inp_i + _lin(w, "hidden_i", agg))
    n = torch.tanh(inp_n + r * _lin(w, "hidden_h", agg))
    hidden = (1 - i) * n + i * hidden
    p = torch.relu(_lin(w, "out_mlp.0", hidden))                                              # :648
    p = torch.relu(_lin(w, "out_mlp.3", p))
    p = _lin(w, "out_mlp.6", p)
    pg = torch.cat([torch.einsum("...ij,...j->...i", Rinv, c) for c in p.split(D, -1)], -1)   # local_to_global.py:12-13
    return inputs + pg, hidden


def predict_field(w: Dict[str, Tensor], x: Tensor, D: int) -> Tensor:
    """nn/seq2seq/aether.py:83-87 + nn/nn/fourier_feature_mapper.py:21-23."""
    proj = (2 * math.pi * x[..., :D]) @ w["rff_B"]
    h = torch.cat([torch.sin(proj), torch.cos(proj)], -1)
    h = torch.nn.functional.silu(_lin(w, "field_net.0", h))
    h = torch.nn.functional.silu(_lin(w, "field_net.2", h))
    return _lin(w, "field_net.4", h)


def predict_future(w: Dict[str, Tensor], inputs: Tensor, edges: Tensor, prediction_steps: int, D: int, n_types: int,
                   skip_first: bool, return_everything: bool = True) -> Tensor:
    """Decoder side of ``Aether.predict_future`` (nn/seq2seq/aether.py:155-191).  ``inputs[B,T,N,2D]`` is the burn-in,
    ``edges[B, T-1+steps, N(N-1), n_types]`` the edge types the encoder + sampler chose at every step."""
    B, T, N, _ = inputs.shape
    H = w["hidden_r.weight"].shape[0]
    hidden = torch.zeros(B, N, H, dtype=inputs.dtype)
    out = []
    for t in range(T - 1):                                                                    # :166-174
        f = predict_field(w, inputs[:, t], D)
        p, hidden = decoder_step(w, inputs[:, t], hidden, edges[:, t], f, D, n_types, skip_first)
        if return_everything:
            out.append(p)
    p = inputs[:, T - 1]
    for t in range(prediction_steps):                                                         # :176-184
        f = predict_field(w, p, D)
        p, hidden = decoder_step(w, p, hidden, edges[:, T - 1 + t], f, D, n_types, skip_first)
        out.append(p)
    return torch.stack(out, 1)


# --------------------------------------------------------------------------------------------------------------------
# Edge-type encoder, prior branch (nn/seq2seq/aether.py:249-438): what predict_future needs of it
# --------------------------------------------------------------------------------------------------------------------
def _bn_eval(w: Dict[str, Tensor], name: str, x: Tensor) -> Tensor:
    """nn.BatchNorm1d in eval mode on the flattened rows (nn/utils/model_utils.py:41-45)."""
    return torch.nn.functional.batch_norm(x.reshape(-1, x.shape[-1]), w[name + ".running_mean"], w[name + ".running_var"],
                                          w[name + ".weight"], w[name + ".bias"], False, 0.0, 1e-5).view(x.shape)


def _nri_mlp(w: Dict[str, Tensor], name: str, x: Tensor) -> Tensor:
    """RefNRIMLP.forward, dropout inactive (nn/utils/model_utils.py:15-54)."""
    x = torch.nn.functional.elu(_lin(w, name + ".model.0", x))
    x = torch.nn.functional.elu(_lin(w, name + ".model.3", x))
    return _bn_eval(w, name + ".bn", x) if (name + ".bn.weight") in w else x


def _fc_out(w: Dict[str, Tensor], name: str, x: Tensor) -> Tensor:
    """prior_fc_out: one Linear, or Linear/ELU pairs and a last Linear (nn/seq2seq/aether.py:289-299)."""
    if (name + ".weight") in w:
        return _lin(w, name, x)
    idx = sorted(int(k.split(".")[-2]) for k in w if k.startswith(name + ".") and k.endswith(".weight"))
    for i in idx[:-1]:
        x = torch.nn.functional.elu(_lin(w, "%s.%d" % (name, i), x))
    return _lin(w, "%s.%d" % (name, idx[-1]), x)


def encoder_embed(w: Dict[str, Tensor], states: Tensor, field: Tensor, D: int, pos_representation: str) -> Tensor:
    """The per-timestep part of Encoder.forward / single_step_forward (nn/seq2seq/aether.py:364-381, 400-413):
    states[B,N,2D], field[B,N,D] -> edge embeddings [B, N(N-1), hidden]."""
    B, N, _ = states.shape
    send, recv = fc_edges_sender_major(N)
    F = 4 * D + D * (D - 1) // 2
    rel_feat, _, edge_attr = augmented_localizer(torch.cat([states, field], -1), D)
    pos_idx = {(2, "cart"): [0, 1, 2], (2, "polar"): [2, 3, 4], (3, "cart"): [0, 1, 2, 3, 4, 5],
               (3, "polar"): [3, 4, 5, 6, 7, 8]}[(D, pos_representation)]      # augmented_global_to_local.py:18-24
    edge_pos = edge_attr[..., pos_idx]
    # AnisotropicEdgeFilter (nn/nn/anisotropic_filter.py:35-41): a per-edge [Fin, H] matrix from the positions
    a = torch.nn.functional.elu(_lin(w, "edge_filter.edge_filter.0", edge_pos))
    ew = _lin(w, "edge_filter.edge_filter.2", a)
    Fin = edge_attr.shape[-1]
    ew = ew.reshape(ew.shape[:-1] + (Fin, ew.shape[-1] // Fin))
    filt = (edge_attr.unsqueeze(-2) @ ew).squeeze(-2)
    res_x = _lin(w, "res1", rel_feat)
    incoming = torch.zeros(B, N, filt.shape[-1], dtype=filt.dtype).index_add_(1, recv, filt) / (N - 1)   # edge2node :351-360
    x = _nri_mlp(w, "mlp3", incoming + res_x)
    x = torch.cat([x[:, send], x[:, recv], filt], -1)                                                    # node2edge :344-349
    return _nri_mlp(w, "mlp4", x)


def lstm_cell(w: Dict[str, Tensor], name: str, x: Tensor, h: Tensor, c: Tensor):
    """One step of nn.LSTM (gate order i, f, g, o)."""
    g = x @ w[name + ".weight_ih_l0"].t() + w[name + ".bias_ih_l0"] + h @ w[name + ".weight_hh_l0"].t() + w[name + ".bias_hh_l0"]
    i, f, gg, o = g.chunk(4, -1)
    c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
    return torch.sigmoid(o) * torch.tanh(c), c


def encoder_step(w: Dict[str, Tensor], states: Tensor, field: Tensor, h: Tensor, c: Tensor, D: int,
                 pos_representation: str):
    """Encoder.single_step_forward (nn/seq2seq/aether.py:397-425): prior logits [B, N(N-1), types] and the new LSTM state
    (h, c: [B*N(N-1), rnn_hidden])."""
    x = encoder_embed(w, states, field, D, pos_representation)
    B, E, H = x.shape
    h, c = lstm_cell(w, "forward_rnn", x.reshape(B * E, H), h, c)
    return _fc_out(w, "prior_fc_out", h).view(B, E, -1), h, c


def encoder_prior(w: Dict[str, Tensor], inputs: Tensor, D: int, pos_representation: str):
    """The prior branch of Encoder.forward over a burn-in sequence (nn/seq2seq/aether.py:362-395): inputs[B,T,N,2D] ->
    prior logits [B,T,N(N-1),types] and the final LSTM state.  The field at every step is the field predictor's."""
    B, T, N, _ = inputs.shape
    R = w["forward_rnn.weight_hh_l0"].shape[1]
    h = torch.zeros(B * N * (N - 1), R, dtype=inputs.dtype)
    c = torch.zeros_like(h)
    out = []
    for t in range(T):
        lg, h, c = encoder_step(w, inputs[:, t], predict_field(w, inputs[:, t], D), h, c, D, pos_representation)
        out.append(lg)
    return torch.stack(out, 1), h, c
