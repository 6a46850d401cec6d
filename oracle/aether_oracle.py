"""CPU oracle for the Aether state-to-state step.  TEST INFRASTRUCTURE ONLY.

This file is a plain torch-CPU (fp32, eager) restatement of the algorithm of the
reference's hot path.  It exists so that the CUDA product path in
``aether_b200/`` can be checked against something that does not depend on it.
Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it.  Nothing under
``aether_b200/`` imports it, and the product path has no CPU fallback.

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the unmodified
reference (``/root/reference/nn/state2state/aether.py``) in the build container
and records inputs/outputs of every stage; ``tests/test_oracle_golden.py``
checks this restatement against those vectors (bit-exact for the integer path,
<= 2e-6 abs for fp32 stages, see the test).  The one third-party piece,
``torch_scatter.scatter(reduce="mean")`` (un-vendored, unpinned,
``nn/state2state/locs/locs.py:4,236``), is restated from its published
semantics: sum by index, divide by max(count, 1).

Each function cites the reference lines it follows (paths relative to
``/root/reference``).  Layout conventions are the reference's: nodes flattened
over the batch ``[Nn, .]``, edges as two int64 vectors ``send``/``recv``.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

Tensor = torch.Tensor
EPS_PHI = 1e-7  # nn/utils/geometry.py:63


# --------------------------------------------------------------------------
# Integer path: edge lists and CSR offsets (bit-exact contract)
# --------------------------------------------------------------------------

def fc_edges(n_nodes: int, batch: int = 1) -> Tuple[Tensor, Tensor]:
    """Fully connected, sender-major edge list, batched with per-graph offsets.

    experiments/lorentz/dataset4newton.py:54-61 (double loop, i != j) and
    :84-94 (offset ``n_nodes * b`` per graph, concatenated).  ``edges[0]`` are
    senders, ``edges[1]`` receivers (nn/state2state/locs/locs.py:231).
    """
    i = np.repeat(np.arange(n_nodes, dtype=np.int64), n_nodes)
    j = np.tile(np.arange(n_nodes, dtype=np.int64), n_nodes)
    keep = i != j
    i, j = i[keep], j[keep]
    off = (np.arange(batch, dtype=np.int64) * n_nodes)[:, None]
    send = (i[None, :] + off).reshape(-1)
    recv = (j[None, :] + off).reshape(-1)
    return torch.from_numpy(send), torch.from_numpy(recv)


def knn_edges(pos: Tensor, k: int = 10) -> Tuple[Tensor, Tensor]:
    """k-nearest-neighbour edges of ONE scene with all agents valid.

    experiments/ind/single_ind_data.py:186-217: pairwise distances, diagonal to
    +inf, ``topk(k=min(10, n-1), largest=False)``; sender = the query agent i,
    receivers = its k nearest, sender-major.  Tie order follows torch.topk.
    """
    n = pos.shape[0]
    if n <= 1:
        z = torch.zeros(0, dtype=torch.int64)
        return z, z.clone()
    d = torch.cdist(pos[:, :2], pos[:, :2])
    d[torch.arange(n), torch.arange(n)] = float("inf")
    kk = min(k, n - 1)
    _, nbr = d.topk(kk, dim=-1, largest=False)
    send = torch.arange(n, dtype=torch.int64).unsqueeze(1).repeat(1, kk).reshape(-1)
    recv = nbr.reshape(-1).to(torch.int64)
    return send, recv


def knn_edges_batched(pos: Tensor, sizes: Sequence[int], k: int = 10) -> Tuple[Tensor, Tensor]:
    """Concatenate per-scene kNN lists with node offsets (scenes are ragged)."""
    s_all, r_all, off = [], [], 0
    for n in sizes:
        s, r = knn_edges(pos[off:off + n], k)
        s_all.append(s + off)
        r_all.append(r + off)
        off += n
    return torch.cat(s_all), torch.cat(r_all)


def radius_knn_edges(pos: Tensor, radius: float, k: int = 32) -> Tuple[Tensor, Tensor]:
    """Radius graph capped at the k nearest, brute force (derived oracle for BASELINE config 5).

    Same conventions as ``knn_edges`` above (experiments/ind/single_ind_data.py:186-217: sender = the query
    particle, receivers = its nearest neighbours in ascending distance, sender-major) with two changes the large
    scene needs: only neighbours with |x_i - x_j|^2 <= radius^2 count, and ties are broken toward the lower index
    (torch.topk leaves tie order unspecified).  Squared distances are accumulated left to right in float32
    (dx*dx + dy*dy [+ dz*dz], each operation rounded), the arithmetic the CUDA builder reproduces bit for bit.
    """
    p = pos.numpy().astype(np.float32)
    n, D = p.shape
    r2 = np.float32(radius) * np.float32(radius)
    send, recv = [], []
    step = max(1, min(n, (1 << 24) // max(n, 1)))
    ar = np.arange(n)
    for b in range(0, n, step):
        q = p[b:b + step]
        dx = q[:, None, 0] - p[None, :, 0]
        dy = q[:, None, 1] - p[None, :, 1]
        d2 = dx * dx + dy * dy
        if D == 3:
            dz = q[:, None, 2] - p[None, :, 2]
            d2 = d2 + dz * dz
        d2 = d2.astype(np.float32)
        ok = d2 <= r2
        ok[np.arange(q.shape[0]), np.arange(b, b + q.shape[0])] = False
        for i in range(q.shape[0]):
            cand = ar[ok[i]]
            order = np.lexsort((cand, d2[i, cand]))[:k]        # by distance, then index
            recv.append(cand[order])
            send.append(np.full(order.shape[0], b + i, dtype=np.int64))
    if not send:
        z = torch.zeros(0, dtype=torch.int64)
        return z, z.clone()
    return torch.from_numpy(np.concatenate(send).astype(np.int64)), torch.from_numpy(np.concatenate(recv).astype(np.int64))


def receiver_csr(send: Tensor, recv: Tensor, n_nodes: int):
    """Receiver-sorted view of an edge list (derived oracle).

    perm = stable argsort of recv; rowptr = [0, cumsum(bincount(recv))].
    Returns (perm, rowptr, send_sorted, recv_sorted) as int64 numpy arrays.
    """
    s = send.numpy().astype(np.int64)
    r = recv.numpy().astype(np.int64)
    perm = np.argsort(r, kind="stable").astype(np.int64)
    cnt = np.bincount(r, minlength=n_nodes).astype(np.int64)
    rowptr = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int64)
    return perm, rowptr, s[perm], r[perm]


def sender_csr(send_sorted: np.ndarray, n_nodes: int):
    """Sender-grouped view of the receiver-sorted list (used by backward)."""
    perm = np.argsort(send_sorted, kind="stable").astype(np.int64)
    cnt = np.bincount(send_sorted, minlength=n_nodes).astype(np.int64)
    colptr = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int64)
    return perm, colptr


# --------------------------------------------------------------------------
# Geometry
# --------------------------------------------------------------------------

def heading_angles(v: Tensor, symmetric: bool) -> List[Tensor]:
    """Spherical angles of a D-vector, each ``[..., 1]``.

    nn/utils/geometry.py:37-66.  theta = atan2(y, x), moved to [0, 2pi) unless
    ``symmetric`` (:51-53).  3D adds phi = acos(clamp(z / (|v| + 1e-7))) (:63-64).
    """
    th = torch.atan2(v[..., 1:2], v[..., 0:1])
    if not symmetric:
        th = th + (th < 0).to(th.dtype) * (2 * np.pi)
    if v.shape[-1] == 2:
        return [th]
    rho = torch.norm(v, p=2, dim=-1, keepdim=True)
    ph = torch.acos(torch.clamp(v[..., 2:] / (rho + EPS_PHI), min=-1.0, max=1.0))
    return [th, ph]


def heading_frame(v: Tensor) -> Tensor:
    """Rotation matrix whose heading axis is the velocity direction, ``[...,D,D]``.

    nn/utils/geometry.py:69-73 -> :7-34.  2D: [[c,-s],[s,c]].  3D:
    [[cp*ct, -st, sp*ct], [cp*st, ct, sp*st], [-sp, 0, cp]].
    """
    ang = heading_angles(v, symmetric=False)
    ct, st = torch.cos(ang[0]), torch.sin(ang[0])
    if v.shape[-1] == 2:
        return torch.stack([torch.cat([ct, -st], -1), torch.cat([st, ct], -1)], -2)
    cp, sp = torch.cos(ang[1]), torch.sin(ang[1])
    rows = [
        torch.cat([cp * ct, -st, sp * ct], -1),
        torch.cat([cp * st, ct, sp * st], -1),
        torch.cat([-sp, torch.zeros_like(ct), cp], -1),
    ]
    return torch.stack(rows, -2)


def apply_rot(R: Tensor, x: Tensor) -> Tensor:
    """nn/utils/geometry.py:104-105."""
    return torch.einsum("...ij,...j->...i", R, x)


def frame_to_euler(M: Tensor, num_dims: int) -> Tensor:
    """Euler angles / pi.  nn/utils/geometry.py:76-101 (ZYX in 3D)."""
    if num_dims == 2:
        e = torch.atan2(M[..., 1, 0:1], M[..., 0, 0:1])
    else:
        e = torch.stack(
            [
                torch.atan2(M[..., 1, 0], M[..., 0, 0]),
                torch.asin(-M[..., 2, 0]),
                torch.atan2(M[..., 2, 1], M[..., 2, 2]),
            ],
            -1,
        )
    return e / np.pi


# --------------------------------------------------------------------------
# Stages of the step
# --------------------------------------------------------------------------

def _lin(p: Dict[str, Tensor], key: str, x: Tensor) -> Tensor:
    return torch.nn.functional.linear(x, p[key + ".weight"], p[key + ".bias"])


def field_forward(p: Dict[str, Tensor], x: Tensor, vel: Tensor, charges: Tensor) -> Tensor:
    """Neural field queried at each particle.  nn/state2state/aether.py:123-134.

    Embedding row = (charge + 1) truncated to integer (:125), concatenated after
    [pos, vel]; MLP (2D+16)->32->32->D with SiLU between (:113-119).
    """
    idx = (charges + 1).long().reshape(-1)
    emb = p["field_net.class_embedding.weight"][idx]
    z = torch.cat([x, vel, emb], -1)
    z = torch.nn.functional.silu(_lin(p, "field_net.net.0", z))
    z = torch.nn.functional.silu(_lin(p, "field_net.net.2", z))
    return _lin(p, "field_net.net.4", z)


def localize(x: Tensor, vel: Tensor, f: Tensor, send: Tensor, recv: Tensor):
    """Local-frame node and edge features.  nn/state2state/aether.py:33-100.

    Returns rel_feat [Nn,3D] = [0, R^T v, R^T f] (:38-48), R [Nn,D,D] and
    edge_attr [E,7D+O] with column order
    [R_i^T (p_j-p_i) | euler(R_i^T R_j)/pi | |p_j-p_i| | theta(,phi) of the
    rotated offset (theta symmetric) | R_i^T v_j | R_i^T f_j | rel_feat[recv]]
    (:81-99).  Frames are computed once per node here and gathered, which is
    bit-identical to the reference's per-edge recomputation (:56,64).
    """
    D = x.shape[-1]
    R = heading_frame(vel)
    Rt = R.transpose(-1, -2)
    rel_feat = torch.cat([torch.zeros_like(vel), apply_rot(Rt, vel), apply_rot(Rt, f)], -1)

    Rt_i = Rt[recv]
    d = x[send] - x[recv]
    d_loc = apply_rot(Rt_i, d)
    euler = frame_to_euler(Rt_i @ R[send], D)
    dist = torch.norm(d, p=2, dim=-1, keepdim=True)
    sph = torch.cat(heading_angles(d_loc, symmetric=True), -1)
    v_loc = apply_rot(Rt_i, vel[send])
    f_loc = apply_rot(Rt_i, f[send])
    edge_attr = torch.cat([d_loc, euler, dist, sph, v_loc, f_loc, rel_feat[recv]], -1)
    return rel_feat, R, edge_attr


def scatter_mean(src: Tensor, index: Tensor, n_rows: int) -> Tensor:
    """torch_scatter.scatter(src, index, dim=0, reduce='mean') restated.

    Call site nn/state2state/locs/locs.py:236-238.  Published semantics: sum
    rows by index, divide by the in-degree clamped to >= 1.  The reference sizes
    the output as max(index)+1; the oracle always uses ``n_rows`` = Nn (superset,
    SURVEY.md 7.2 last bullet).
    """
    out = torch.zeros(n_rows, src.shape[1], dtype=src.dtype, device=src.device).index_add_(0, index, src)
    cnt = torch.zeros(n_rows, dtype=src.dtype, device=src.device).index_add_(0, index, torch.ones_like(index, dtype=src.dtype))
    return out / cnt.clamp(min=1).unsqueeze(1)


def gnn_layer(p: Dict[str, Tensor], name: str, x: Tensor, e: Tensor, send: Tensor, recv: Tensor, first: bool):
    """One message-passing layer.  nn/state2state/locs/locs.py:227-243."""
    silu = torch.nn.functional.silu
    if not first:
        e = torch.cat([x[send], x[recv], e], -1)                          # :232-233
    e = silu(_lin(p, name + ".message_fn.2", silu(_lin(p, name + ".message_fn.0", e))))  # :235
    agg = scatter_mean(e, recv, x.shape[0])                                # :236-238
    x = (_lin(p, name + ".res", x) if first else x) + agg                  # :240
    x = x + _lin(p, name + ".update_fn.2", silu(_lin(p, name + ".update_fn.0", x)))      # :241
    return x, e


def aether_step(
    p: Dict[str, Tensor],
    x: Tensor,
    edges: Tuple[Tensor, Tensor],
    vel: Tensor,
    edge_attr_orig: Tensor,
    charges: Tensor,
    dropout_prob: float = 0.0,
    training: bool = False,
    stages: Optional[dict] = None,
) -> Tensor:
    """Aether.forward.  nn/state2state/aether.py:169-186 (``h`` is unused there)."""
    send, recv = edges
    f = field_forward(p, x, vel, charges)                                   # :171-173
    rel_feat, R, edge_attr = localize(x, vel, f, send, recv)                # :176
    e = torch.cat([edge_attr, edge_attr_orig], -1)                          # :177
    if stages is not None:
        stages.update(field=f, rel_feat=rel_feat, R=R, edge_attr=e)
    h = rel_feat
    for l in range(1, 5):                                                   # locs.py:187-190
        h, e = gnn_layer(p, "gnn.layer_%d" % l, h, e, send, recv, first=(l == 1))
        if stages is not None:
            stages["x%d" % l] = h
            stages["e%d" % l] = e
    silu = torch.nn.functional.silu
    drop = lambda t: torch.nn.functional.dropout(t, dropout_prob, training)
    h = drop(silu(_lin(p, "gnn.out_mlp.0", h)))                             # locs.py:160-168
    h = drop(silu(_lin(p, "gnn.out_mlp.3", h)))
    pred = _lin(p, "gnn.out_mlp.6", h)
    pred_g = apply_rot(R, pred)                                             # local_to_global.py:12-13
    out = x + pred_g                                                        # aether.py:185
    if stages is not None:
        stages.update(pred=pred, out=out)
    return out


def latent_field_forward(p: Dict[str, Tensor], x: Tensor, vel: Tensor, charges: Tensor, num_nodes: int,
                         stages: Optional[dict] = None) -> Tensor:
    """LatentFieldNetwork.forward.  nn/state2state/dynamic_field_aether.py:29-47.

    Graph summary (graph_pool.py:6-29): torch_geometric's AttentionalAggregation is an un-vendored dependency
    (requirements of the reference, README.md:31; not installed here) -- parity UNPINNED at that boundary.  Its published
    semantics are restated: gate = gate_nn(x); alpha = softmax of the gate over the nodes of a graph
    (exp(g - max) / (sum + 1e-16)); summary = sum_n alpha_n * nn(x_n).  FiLM (film.py:40-60): (1 + gamma) y + beta with
    [gamma | beta] = modulator(z); FilmedNetwork (film.py:24-37).
    """
    silu = torch.nn.functional.silu
    inp = torch.cat([x, vel], -1)                                            # dynamic_field_aether.py:84
    pre = "field_net.summary_net.summary_net."
    gate = _lin(p, pre + "gate_nn.2", silu(_lin(p, pre + "gate_nn.0", inp)))  # [Nn, 1]
    val = _lin(p, pre + "nn.2", silu(_lin(p, pre + "nn.0", inp)))             # [Nn, 32]
    g = gate.reshape(-1, num_nodes, 1)
    a = torch.exp(g - g.max(dim=1, keepdim=True).values)
    alpha = a / (a.sum(dim=1, keepdim=True) + 1e-16)
    z = (alpha * val.reshape(-1, num_nodes, val.shape[-1])).sum(dim=1)       # [B, 32]
    zb = z.repeat_interleave(num_nodes, dim=0)                               # :33-39

    def film(name, y):
        m = _lin(p, name + ".modulator.4", silu(_lin(p, name + ".modulator.2", silu(_lin(p, name + ".modulator.0", zb)))))
        gamma, beta = torch.chunk(m, 2, dim=-1)
        return (1.0 + gamma) * y + beta

    idx = (charges + 1).long().reshape(-1)
    emb = p["field_net.class_embedding.weight"][idx]
    y = _lin(p, "field_net.wrapper.linear_1", torch.cat([x, vel, emb], -1))  # :41-42
    y = silu(film("field_net.wrapper.film_1", y))
    y = _lin(p, "field_net.wrapper.linear_2", y)
    y = silu(film("field_net.wrapper.film_2", y))
    f = _lin(p, "field_net.wrapper.linear_3", y)
    if stages is not None:
        stages.update(summary=z, alpha=alpha.reshape(-1), field=f)
    return f


def dynamic_field_step(p: Dict[str, Tensor], x: Tensor, edges: Tuple[Tensor, Tensor], vel: Tensor, edge_attr_orig: Tensor,
                       charges: Tensor, num_nodes: int, stages: Optional[dict] = None) -> Tensor:
    """DynamicFieldAether.forward.  nn/state2state/dynamic_field_aether.py:79-100: Aether's step with the field predicted
    by the latent field network."""
    send, recv = edges
    f = latent_field_forward(p, x, vel, charges, num_nodes, stages)
    rel_feat, R, edge_attr = localize(x, vel, f, send, recv)
    e = torch.cat([edge_attr, edge_attr_orig], -1)
    h = rel_feat
    for l in range(1, 5):
        h, e = gnn_layer(p, "gnn.layer_%d" % l, h, e, send, recv, first=(l == 1))
    silu = torch.nn.functional.silu
    h = silu(_lin(p, "gnn.out_mlp.0", h))
    h = silu(_lin(p, "gnn.out_mlp.3", h))
    pred = _lin(p, "gnn.out_mlp.6", h)
    out = x + apply_rot(R, pred)
    if stages is not None:
        stages.update(pred=pred, out=out)
    return out


def localize_locs(x: Tensor, vel: Tensor, send: Tensor, recv: Tensor):
    """The base Localizer of LoCS (no field columns).  nn/state2state/locs/locs.py:37-99.

    rel_feat [Nn,2D] = [0, R^T v] (:37-46); edge_attr [E,5D+O] =
    [R_i^T (p_j-p_i) | euler(R_i^T R_j)/pi | |p_j-p_i| | theta(,phi) | R_i^T v_j | rel_feat[recv]] (:48-99).
    """
    D = x.shape[-1]
    R = heading_frame(vel)
    Rt = R.transpose(-1, -2)
    rel_feat = torch.cat([torch.zeros_like(vel), apply_rot(Rt, vel)], -1)
    Rt_i = Rt[recv]
    d = x[send] - x[recv]
    d_loc = apply_rot(Rt_i, d)
    euler = frame_to_euler(Rt_i @ R[send], D)
    dist = torch.norm(d, p=2, dim=-1, keepdim=True)
    sph = torch.cat(heading_angles(d_loc, symmetric=True), -1)
    v_loc = apply_rot(Rt_i, vel[send])
    edge_attr = torch.cat([d_loc, euler, dist, sph, v_loc, rel_feat[recv]], -1)
    return rel_feat, R, edge_attr


def locs_step(p: Dict[str, Tensor], x: Tensor, edges: Tuple[Tensor, Tensor], vel: Tensor, edge_attr_orig: Tensor,
              stages: Optional[dict] = None) -> Tensor:
    """LoCS.forward.  nn/state2state/locs/locs.py:124-139: the same GNN on the field-free local features."""
    send, recv = edges
    rel_feat, R, edge_attr = localize_locs(x, vel, send, recv)
    e = torch.cat([edge_attr, edge_attr_orig], -1)
    if stages is not None:
        stages.update(rel_feat=rel_feat, R=R, edge_attr=e)
    h = rel_feat
    for l in range(1, 5):
        h, e = gnn_layer(p, "gnn.layer_%d" % l, h, e, send, recv, first=(l == 1))
    silu = torch.nn.functional.silu
    h = silu(_lin(p, "gnn.out_mlp.0", h))
    h = silu(_lin(p, "gnn.out_mlp.3", h))
    pred = _lin(p, "gnn.out_mlp.6", h)
    return x + apply_rot(R, pred)


def caller_prep(x: Tensor, vel: Tensor, charges: Tensor, send: Tensor, recv: Tensor):
    """Per-step features the caller builds.  experiments/lorentz/main.py:243-246.

    h = |v|;  edge_attr_orig = [q_i q_j, |x_i - x_j|] over (rows, cols) = (send, recv).
    (The charge product is stored by the dataset, dataset4newton.py:53-58; the
    synthetic harness computes it from the per-node charges.)
    """
    h = torch.sqrt(torch.sum(vel ** 2, dim=1)).unsqueeze(1)
    qq = charges[send] * charges[recv]
    dist = torch.sqrt(torch.sum((x[send] - x[recv]) ** 2, 1)).unsqueeze(1)
    return h, torch.cat([qq.reshape(-1, 1), dist], 1)


def rollout(p, x0: Tensor, v0: Tensor, charges: Tensor, edges_fn, steps: int) -> Tensor:
    """Harness-level rollout: x_{t+1} = step(x_t, v_t), v_{t+1} = x_{t+1} - x_t.

    ``edges_fn(x)`` returns the edge list for the current positions (static FC
    lists ignore x; kNN lists are rebuilt every step).  Returns [steps, Nn, D].
    """
    xs, x, v = [], x0, v0
    for _ in range(steps):
        send, recv = edges_fn(x)
        _, ea = caller_prep(x, v, charges, send, recv)
        xn = aether_step(p, x, (send, recv), v, ea, charges)
        v = xn - x
        x = xn
        xs.append(x)
    return torch.stack(xs)


# --------------------------------------------------------------------------
# Parameters and synthetic inputs
# --------------------------------------------------------------------------

def param_shapes(num_dims: int, hidden: int) -> List[Tuple[str, Tuple[int, ...]]]:
    """state_dict keys and shapes in the reference's construction order.

    nn/state2state/aether.py:143-159 builds GNN first (locs.py:160-181: out_mlp,
    layer_1..4 each message_fn, res, update_fn) then the field network
    (aether.py:113-121: net, class_embedding).
    """
    D, H = num_dims, hidden
    O = D * (D - 1) // 2
    fe = 7 * D + O + 2
    out: List[Tuple[str, Tuple[int, ...]]] = []

    def lin(name, i, o):
        out.append((name + ".weight", (o, i)))
        out.append((name + ".bias", (o,)))

    lin("gnn.out_mlp.0", H, H)
    lin("gnn.out_mlp.3", H, H)
    lin("gnn.out_mlp.6", H, D)
    for l in range(1, 5):
        n = "gnn.layer_%d" % l
        lin(n + ".message_fn.0", fe if l == 1 else 3 * H, H)
        lin(n + ".message_fn.2", H, H)
        if l == 1:
            lin(n + ".res", 3 * D, H)
        lin(n + ".update_fn.0", H, 2 * H)
        lin(n + ".update_fn.2", 2 * H, H)
    lin("field_net.net.0", 2 * D + 16, 32)
    lin("field_net.net.2", 32, 32)
    lin("field_net.net.4", 32, D)
    out.append(("field_net.class_embedding.weight", (3, 16)))
    return out


def init_params(num_dims: int, hidden: int, seed: int = 1) -> Dict[str, Tensor]:
    """Default PyTorch inits drawn in the reference's construction order, so a
    given seed yields the reference's initial weights (checked in the golden test)."""
    torch.manual_seed(seed)
    p: Dict[str, Tensor] = {}
    shapes = dict(param_shapes(num_dims, hidden))
    for key, shp in param_shapes(num_dims, hidden):
        if key.endswith(".bias"):
            continue
        if key == "field_net.class_embedding.weight":
            p[key] = torch.empty(shp).normal_(0, 1)                         # nn.Embedding default
            continue
        w = torch.empty(shp)
        torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))                   # nn.Linear.reset_parameters
        p[key] = w
        bkey = key[:-7] + ".bias"
        bound = 1.0 / math.sqrt(shp[1])
        p[bkey] = torch.empty(shapes[bkey]).uniform_(-bound, bound)
    return p


def synth_state(n_graphs: int, sizes, num_dims: int, seed: int = 1, speed_norm: bool = False):
    """Synthetic particles following the reference simulators' distributions.

    positions N(0, (N/5)^(2/3)) (experiments/lorentz/dataset/synthetic_sim.py:469),
    velocities random direction with norm 0.5 (:553-555), charges uniform +-1
    (:476).  ``sizes`` is an int (all graphs equal) or a per-graph list.
    """
    g = torch.Generator().manual_seed(seed)
    if isinstance(sizes, int):
        sizes = [sizes] * n_graphs
    xs, vs = [], []
    for n in sizes:
        sigma = (n / 5.0) ** (1.0 / 3.0)
        xs.append(torch.randn(n, num_dims, generator=g) * sigma)
        v = torch.randn(n, num_dims, generator=g)
        vs.append(0.5 * v / torch.norm(v, dim=-1, keepdim=True))
    x, v = torch.cat(xs), torch.cat(vs)
    q = (torch.randint(0, 2, (x.shape[0], 1), generator=g).float() * 2 - 1)
    if speed_norm:
        x = x / 0.5
        v = v / 0.5
    return x, v, q
