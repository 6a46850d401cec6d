"""Sharding of the step across GPUs: one process per GPU, torch.distributed for the plumbing.

Two cases, as the reference's data allows:

* **Batch of trajectories** -- graphs of a batch share no edges (experiments/lorentz/dataset4newton.py:89-93
  offsets each graph by ``n_nodes * b``), so rank r takes a contiguous block of graphs and builds its edge
  list with local node numbering.  No data-path collective; training adds ONE all-reduce of the flat
  gradient (the backward of ``aether_b200.ops`` returns all 47 gradients as views of one buffer, so the
  collective runs in place, no concatenate / copy-back).
* **Single large scene** -- receiver partition: rank r owns a contiguous block of nodes and every edge
  whose receiver it owns (a contiguous slice of the receiver-sorted edge array, cut at ``rowptr``
  boundaries), so aggregation, node MLP and output are local; the only exchange is an all-gather of the
  per-node sender term before each of layers 2-4 and of the new positions after the step.

Everything here is host logic and works with the ``gloo`` backend on CPU tensors (tests) as well as ``nccl``.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch


def shard_range(n_units: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block [begin, end) of ``n_units`` for ``rank``; the first ``n_units % world`` ranks hold one more."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    q, r = divmod(int(n_units), world)
    begin = rank * q + min(rank, r)
    return begin, begin + q + (1 if rank < r else 0)


def shard_batch_edges(send: torch.Tensor, recv: torch.Tensor, nodes_per_graph: int, n_graphs: int, world: int,
                      rank: int):
    """Slice a batched edge list (graphs concatenated with per-graph node offsets, equal-size graphs whose edges
    are stored graph by graph) down to the graphs of ``rank`` and renumber nodes locally.

    Returns (send_local, recv_local, graph_begin, graph_end).  For fully connected batches the result is
    bit-identical to building ``fc_edges(nodes_per_graph, graph_end - graph_begin)`` directly.
    """
    g0, g1 = shard_range(n_graphs, world, rank)
    E = int(send.numel())
    if n_graphs <= 0 or E % n_graphs:
        raise ValueError("edge count is not a multiple of the number of graphs")
    per = E // n_graphs
    off = g0 * nodes_per_graph
    return send[g0 * per:g1 * per] - off, recv[g0 * per:g1 * per] - off, g0, g1


def flat_grad(params: Sequence[torch.Tensor]) -> Optional[torch.Tensor]:
    """The single 1-D buffer all ``p.grad`` are views of, in parameter order with gaps of at most 3 pad
    elements (16-byte aligned views), or None if the gradients are laid out any other way."""
    base = None
    off = 0
    for p in params:
        g = p.grad
        if g is None or g._base is None or not g.is_contiguous():
            return None
        if base is None:
            base = g._base
            if base.dim() != 1:
                return None
        rel = g.storage_offset() - base.storage_offset()
        if g._base is not base or not off <= rel <= off + 3:
            return None
        off = rel + g.numel()
    if base is None or not off <= base.numel() <= off + 3:
        return None
    return base


def allreduce_grads(params: Sequence[torch.Tensor], world: int, group=None) -> int:
    """Average gradients over ranks (MSELoss is a mean, experiments/lorentz/main.py:86,289: with equal shards the
    mean of per-rank means is the global mean).  Returns the number of collectives issued (1)."""
    if world <= 1:
        return 0
    import torch.distributed as dist
    flat = flat_grad(params)
    if flat is not None:
        dist.all_reduce(flat, group=group)
        flat.div_(world)
        return 1
    flat = torch.cat([p.grad.reshape(-1) for p in params])
    dist.all_reduce(flat, group=group)
    flat.div_(world)
    off = 0
    for p in params:
        n = p.numel()
        p.grad.copy_(flat[off:off + n].view_as(p.grad))
        off += n
    return 1


class ReceiverPartition:
    """Receiver partition of ONE scene over ``world`` ranks.

    ``rowptr`` is the receiver CSR of the whole scene ([Nn+1], host or device tensor).  Nodes are split into
    ``world`` equal blocks of ``block = ceil(Nn / world)`` rows (the last block may be short; exchange
    buffers are padded to ``world * block`` rows so that an in-place all-gather has equal counts).
    """

    def __init__(self, rowptr: torch.Tensor, n_nodes: int, world: int, rank: int):
        if rowptr.numel() != n_nodes + 1:
            raise ValueError("rowptr must have n_nodes + 1 entries")
        self.n_nodes, self.world, self.rank = int(n_nodes), int(world), int(rank)
        self.block = (self.n_nodes + world - 1) // world
        self.padded_nodes = self.block * world
        self.node_begin = min(rank * self.block, self.n_nodes)
        self.node_end = min(self.node_begin + self.block, self.n_nodes)
        cut = rowptr[[self.node_begin, self.node_end]].to("cpu", torch.int64)   # one host read, at plan time
        self.edge_begin, self.edge_end = int(cut[0]), int(cut[1])

    @property
    def n_own(self) -> int:
        return self.node_end - self.node_begin

    @property
    def n_local_edges(self) -> int:
        return self.edge_end - self.edge_begin

    def local_rowptr(self, rowptr: torch.Tensor) -> torch.Tensor:
        """CSR of the owned receivers over the local edge slice ([n_own + 1], rebased to 0)."""
        return (rowptr[self.node_begin:self.node_end + 1] - rowptr[self.node_begin]).contiguous()

    def all_gather_rows(self, buf: torch.Tensor, group=None) -> None:
        """In-place all-gather of ``buf`` ([padded_nodes, C]): every rank has written its own block of rows."""
        if self.world <= 1:
            return
        import torch.distributed as dist
        if buf.shape[0] != self.padded_nodes:
            raise ValueError("exchange buffer must have world * block rows")
        own = buf[self.rank * self.block:(self.rank + 1) * self.block]
        if buf.device.type == "cuda":
            dist.all_gather_into_tensor(buf, own, group=group)
        else:                                    # gloo: list form (no in-place aliasing guarantees)
            parts = [torch.empty_like(own) for _ in range(self.world)]
            dist.all_gather(parts, own.clone(), group=group)
            for r, p in enumerate(parts):
                buf[r * self.block:(r + 1) * self.block].copy_(p)
