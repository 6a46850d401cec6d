"""Edge lists on the device: the reference's builders and the receiver-sorted CSR.

Reference: experiments/lorentz/dataset4newton.py:54-61,84-94 (fully connected, sender-major),
experiments/ind/single_ind_data.py:186-217 (kNN), experiments/lorentz/main.py:211-212 (the
per-batch edge handling).  The reference scatters unsorted inside torch_scatter
(nn/state2state/locs/locs.py:236-238); here the list is sorted by receiver once and the CSR is
reused by every layer, forward and backward.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence, Tuple

import torch

from . import _lib


def _stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class PreparedGraph:
    """Receiver-sorted view of an edge list (device int32 arrays, see AetherGraph in the header)."""

    __slots__ = ("n_nodes", "n_edges", "perm", "rowptr", "send", "recv", "sperm", "colptr", "status", "_c", "_key")

    def __init__(self, send: torch.Tensor, recv: torch.Tensor, n_nodes: int):
        lib = _lib.load()
        if send.device.type != "cuda":
            raise RuntimeError("aether_b200 runs on CUDA devices only (no CPU fallback)")
        send = send.to(torch.int64).contiguous()
        recv = recv.to(device=send.device, dtype=torch.int64).contiguous()
        E = int(send.numel())
        dev = send.device
        i32 = dict(dtype=torch.int32, device=dev)
        self.n_nodes, self.n_edges = int(n_nodes), E
        self.perm = torch.empty(max(E, 1), **i32)
        self.send = torch.empty(max(E, 1), **i32)
        self.recv = torch.empty(max(E, 1), **i32)
        self.sperm = torch.empty(max(E, 1), **i32)
        self.rowptr = torch.empty(n_nodes + 1, **i32)
        self.colptr = torch.empty(n_nodes + 1, **i32)
        self.status = torch.zeros(1, **i32)
        nbytes = lib.aether_graph_scratch_bytes(n_nodes, E)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        self._c = _lib.AetherGraphC(self.perm.data_ptr(), self.rowptr.data_ptr(), self.send.data_ptr(),
                                    self.recv.data_ptr(), self.sperm.data_ptr(), self.colptr.data_ptr())
        with torch.cuda.device(dev):
            _lib.check(lib.aether_graph_prepare(_lib.ptr(send), _lib.ptr(recv), E, n_nodes, ctypes.byref(self._c),
                                                _lib.ptr(self.status), _lib.ptr(scratch), nbytes, _stream()),
                       "aether_graph_prepare")
        self._key = None
        # Out-of-range indices never cause out-of-bounds reads (graph.cu clamps them and sets the status word), but
        # the reference raises on them; reading the status word synchronises, so the check is opt-in.
        if os.environ.get("AETHER_B200_VALIDATE", "0") == "1":
            self.validate()

    def c_struct(self) -> "_lib.AetherGraphC":
        return self._c

    def validate(self) -> None:
        """Host read of the status word (synchronises): raises on out-of-range indices."""
        if int(self.status.item()) != 0:
            raise IndexError("aether_b200: edge index out of range [0, %d)" % self.n_nodes)


def fc_edges(n_nodes: int, batch: int, device) -> Tuple[torch.Tensor, torch.Tensor]:
    """Fully connected sender-major edge list with per-graph offsets (dataset4newton.py:54-61,84-94)."""
    lib = _lib.load()
    E = n_nodes * (n_nodes - 1) * batch
    send = torch.empty(E, dtype=torch.int64, device=device)
    recv = torch.empty(E, dtype=torch.int64, device=device)
    with torch.cuda.device(send.device):
        _lib.check(lib.aether_fc_edges(n_nodes, batch, _lib.ptr(send), _lib.ptr(recv), _stream()), "aether_fc_edges")
    return send, recv


class KnnPlan:
    """Static part of a ragged kNN batch: scene offsets and edge offsets (sizes only)."""

    def __init__(self, sizes: Sequence[int], k: int, device):
        self.sizes = [int(s) for s in sizes]
        self.k = int(k)
        node_off, edge_off = [0], [0]
        for n in self.sizes:
            if n > 1024:
                raise ValueError("aether_knn_edges supports scenes of at most 1024 agents")
            node_off.append(node_off[-1] + n)
            edge_off.append(edge_off[-1] + n * min(self.k, max(n - 1, 0)))
        self.n_nodes, self.n_edges = node_off[-1], edge_off[-1]
        self.node_off = torch.tensor(node_off, dtype=torch.int32, device=device)
        self.edge_off = torch.tensor(edge_off, dtype=torch.int64, device=device)


def knn_edges(pos: torch.Tensor, plan: KnnPlan, out: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
    """k nearest neighbours per agent, sender = query agent (single_ind_data.py:186-217)."""
    lib = _lib.load()
    pos = pos.contiguous()
    if out is None:
        send = torch.empty(plan.n_edges, dtype=torch.int64, device=pos.device)
        recv = torch.empty(plan.n_edges, dtype=torch.int64, device=pos.device)
    else:
        send, recv = out
    with torch.cuda.device(pos.device):
        _lib.check(lib.aether_knn_edges(_lib.ptr(pos), pos.shape[1], _lib.ptr(plan.node_off), _lib.ptr(plan.edge_off),
                                        len(plan.sizes), plan.k, _lib.ptr(send), _lib.ptr(recv), _stream()),
                   "aether_knn_edges")
    return send, recv


def caller_prep(x, vel, charges, send, recv):
    """h = |v| and edge_attr = [q_i q_j, |x_i - x_j|] (experiments/lorentz/main.py:243-246)."""
    lib = _lib.load()
    Nn, D = x.shape
    E = send.numel()
    h = torch.empty(Nn, 1, dtype=torch.float32, device=x.device)
    ea = torch.empty(E, 2, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.aether_caller_prep(_lib.ptr(x.contiguous()), _lib.ptr(vel.contiguous()),
                                          _lib.ptr(charges.contiguous()), D, _lib.ptr(send), _lib.ptr(recv), E, Nn,
                                          _lib.ptr(h), _lib.ptr(ea), _stream()), "aether_caller_prep")
    return h, ea


def radius_graph(pos: torch.Tensor, radius: float, k: int = 32) -> Tuple[torch.Tensor, torch.Tensor]:
    """Radius graph capped at the ``k`` nearest neighbours of every sender (k <= 32), sender-major, nearest first
    (ties toward the lower index), like get_knn_graph_info (experiments/ind/single_ind_data.py:186-217) but built
    with a uniform cell list so that 100k-particle scenes work (BASELINE config 5).  Returns int64 (send, recv).

    Construction synchronises once (the edge count is read back to size the output)."""
    lib = _lib.load()
    if pos.device.type != "cuda":
        raise RuntimeError("aether_b200 runs on CUDA devices only (no CPU fallback)")
    pos = pos.contiguous().float()
    n, D = int(pos.shape[0]), int(pos.shape[1])
    if D not in (2, 3) or not 1 <= k <= 32:
        raise ValueError("radius_graph needs 2-D or 3-D positions and 1 <= k <= 32")
    dev = pos.device
    if n == 0:
        z = torch.zeros(0, dtype=torch.int64, device=dev)
        return z, z.clone()
    cell = float(radius) * 1.001
    lo = pos.min(0).values.cpu()
    hi = pos.max(0).values.cpu()
    dims = [max(1, int((float(hi[a]) - float(lo[a])) / cell) + 1) for a in range(D)]
    while True:                                   # keep the cell table small relative to the particle count
        cells = 1
        for d in dims:
            cells *= d
        if cells <= max(8 * n, 4096):
            break
        cell *= 2.0
        dims = [max(1, int((float(hi[a]) - float(lo[a])) / cell) + 1) for a in range(D)]
    c_lo = (ctypes.c_float * D)(*[float(v) for v in lo])
    c_dim = (ctypes.c_int32 * D)(*dims)
    nbytes = lib.aether_radius_graph_scratch_bytes(n, D, c_dim)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    nbr = torch.empty(n, k, dtype=torch.int32, device=dev)
    off = torch.empty(n + 1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.aether_radius_graph_neighbors(_lib.ptr(pos), pos.stride(0), D, n, float(radius), cell, k, c_lo, c_dim,
                                                     _lib.ptr(nbr), _lib.ptr(off), _lib.ptr(scratch), nbytes, _stream()),
                   "aether_radius_graph_neighbors")
        E = int(off[-1].item())
        send = torch.empty(E, dtype=torch.int64, device=dev)
        recv = torch.empty(E, dtype=torch.int64, device=dev)
        _lib.check(lib.aether_radius_graph_fill(_lib.ptr(nbr), _lib.ptr(off), n, k, _lib.ptr(send), _lib.ptr(recv),
                                                _stream()), "aether_radius_graph_fill")
    return send, recv
