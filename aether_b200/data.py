"""Loader for the reference's n-body ``.npy`` files (experiments/lorentz/dataset4newton.py:7-94).

``NBodyDataset`` keeps the reference's constructor, file naming (:16-26,34-38), frame choice (30 -> 40, :31-32,76-79),
``__getitem__`` / ``__len__`` / ``get_n_nodes`` / ``get_edges`` contract, so ``torch.utils.data.DataLoader`` and the train
loop of experiments/lorentz/main.py:197-312 run on it unchanged.  What differs is how it gets there:

* the off-diagonal edge attributes are taken with one boolean mask over the ``[S, N, N]`` interaction array instead of an
  N^2 Python loop that appends array slices (:54-60) -- same values, same (row-major, i != j) order;
* ``get_edges(batch_size, n_nodes, device=...)`` can build the batched list on the GPU (``aether_graph``'s fc kernel,
  bit-identical to the concatenation loop of :84-94, tests/test_step_gpu.py) instead of B host tensors + ``torch.cat``;
* ``device_batches`` yields what the train loop derives per batch (flattened state, edge list, ``|v|`` node feature,
  ``[q_i q_j, |x_i - x_j|]`` edge attribute: main.py:205-246) from pinned host memory, the derived tensors by the
  library's caller-prep kernel.
"""
from __future__ import annotations

import os
from typing import Iterator, List, Optional

import numpy as np
import torch


def file_suffix(partition: str, data_mode: str) -> str:
    """dataset4newton.py:16-26."""
    if data_mode == "small":
        return partition + "_charged5_initvel1small"
    if data_mode in ("static", "dynamic"):
        return partition + "_%s5_initvel1%s" % (data_mode, data_mode)
    if data_mode == "small_20body":
        return partition + "_charged20_initvel1%s" % data_mode
    return partition + "_%s20_initvel1%s" % (data_mode[:-7], data_mode)


class NBodyDataset:
    def __init__(self, partition: str = "train", max_samples=1e8, data_root: Optional[str] = None, data_mode: str = "small"):
        self.partition = partition
        self.sufix = file_suffix(partition, data_mode)          # (sic) the reference's attribute name
        self.data_root = data_root
        self.max_samples = int(max_samples)
        self.frame_0, self.frame_T = 30, 40
        self.data, self.edges = self.load()

    # ---- files -> tensors ------------------------------------------------------------------------------------
    def load(self):
        rd = lambda name: np.load(os.path.join(self.data_root, "%s_%s.npy" % (name, self.sufix)))
        return self.preprocess(rd("loc"), rd("vel"), rd("edges"), rd("charges"))

    def preprocess(self, loc, vel, edges, charges):
        """[S, T, D, N] trajectories -> ([S, T, N, D] loc, vel, [S, N(N-1), 1] edge attributes, charges), [rows, cols]."""
        S = min(self.max_samples, loc.shape[0])
        loc_t = torch.from_numpy(np.ascontiguousarray(loc[:S], dtype=np.float32)).transpose(2, 3)
        vel_t = torch.from_numpy(np.ascontiguousarray(vel[:S], dtype=np.float32)).transpose(2, 3)
        n = loc_t.size(2)
        off = ~np.eye(n, dtype=bool)
        rows, cols = np.nonzero(off)                              # row-major over (i, j), i != j: the reference's loop order
        inter = np.asarray(edges, dtype=np.float32)               # all samples, like the reference (only loc/vel/charges are cut)
        edge_attr = torch.from_numpy(inter[:, off]).unsqueeze(2) if n > 1 else torch.empty(loc_t.shape[0], 0, 1)
        q = torch.from_numpy(np.ascontiguousarray(charges[:S], dtype=np.float32))
        return (loc_t, vel_t, edge_attr, q), [rows.tolist(), cols.tolist()]

    def set_max_samples(self, max_samples):
        self.max_samples = int(max_samples)
        self.data, self.edges = self.load()

    def get_n_nodes(self) -> int:
        """The reference returns ``self.data[0].size(1)`` (:73-74), which on the [S, T, N, D] tensor is the number of
        time steps; only experiments/lorentz/dataloader.py:10 reads it.  Kept as is for drop-in behaviour --
        ``n_particles`` is the particle count."""
        return self.data[0].size(1)

    @property
    def n_particles(self) -> int:
        return self.data[0].size(2)

    def __getitem__(self, i):
        loc, vel, edge_attr, charges = self.data
        return loc[i, self.frame_0], vel[i, self.frame_0], edge_attr[i], charges[i], loc[i, self.frame_T]

    def __len__(self) -> int:
        return len(self.data[0])

    def get_edges(self, batch_size: int, n_nodes: int, device=None) -> List[torch.Tensor]:
        """[rows, cols] of ``batch_size`` graphs, graph b offset by ``n_nodes * b`` (:84-94)."""
        if device is not None and torch.device(device).type == "cuda" and n_nodes == self.n_particles:
            from .graph import fc_edges
            s, r = fc_edges(n_nodes, batch_size, device)
            return [s, r]
        r0 = torch.as_tensor(self.edges[0], dtype=torch.long)
        c0 = torch.as_tensor(self.edges[1], dtype=torch.long)
        if batch_size == 1:
            return [r0, c0]
        shift = (torch.arange(batch_size, dtype=torch.long) * n_nodes).unsqueeze(1)
        return [(r0.unsqueeze(0) + shift).reshape(-1), (c0.unsqueeze(0) + shift).reshape(-1)]

    # ---- the per-batch preparation of the train loop, on the device -------------------------------------------
    def device_batches(self, batch_size: int, device, shuffle: bool = False, drop_last: bool = False,
                       generator: Optional[torch.Generator] = None) -> Iterator[dict]:
        """Batches as the train loop consumes them (main.py:205-246): ``x, vel, charges, target`` flattened over the
        batch, ``edges`` (device), node feature ``h = |v|`` and ``edge_attr = [stored interaction, |x_i - x_j|]``."""
        from .graph import caller_prep
        loc, vel, edge_attr, charges = self.data
        S, n = len(self), self.n_particles
        order = torch.randperm(S, generator=generator) if shuffle else torch.arange(S)
        stage, copied = None, None
        for b0 in range(0, S, batch_size):
            idx = order[b0:b0 + batch_size]
            if drop_last and idx.numel() < batch_size:
                break
            B = int(idx.numel())
            host = [loc[idx, self.frame_0].reshape(B * n, -1), vel[idx, self.frame_0].reshape(B * n, -1),
                    charges[idx].reshape(B * n, -1), loc[idx, self.frame_T].reshape(B * n, -1),
                    edge_attr[idx].reshape(B * n * (n - 1), -1)]
            if stage is None or stage[0].shape[0] != B * n:
                stage = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in host]
            if copied is not None:
                copied.synchronize()                                   # the previous batch's H2D copies have read the stage
            for s, t in zip(stage, host):
                s.copy_(t)
            x, v, q, tgt, inter = [s.to(device, non_blocking=True) for s in stage]
            copied = torch.cuda.Event()
            copied.record()
            send, recv = self.get_edges(B, n, device=device)
            h, ea = caller_prep(x, v, q.reshape(-1), send, recv)       # [q_i q_j, |x_i - x_j|] from the per-node charges
            ea = torch.cat([inter, ea[:, 1:2]], 1)                     # the stored interaction strength replaces the product
            yield {"x": x, "vel": v, "charges": q, "target": tgt, "edges": [send, recv], "h": h, "edge_attr": ea,
                   "batch_size": B, "n_nodes": n}
