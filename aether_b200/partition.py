"""Receiver-partitioned inference of ONE large scene across GPUs (BASELINE config 5).

Rank r owns a contiguous block of nodes and every edge whose receiver it owns, so the segmented mean, the node
MLPs and the output are local.  What crosses ranks is the sender term ``P_s = x W_s^T`` of the node-factorised
first Linear of layers 2-4 (one all-gather of ``[Nn, H]`` per layer, bf16 on the bf16 path) and the new
positions (one all-gather of ``[Nn, D]`` per step).  The compute between two exchanges is one C-ABI call
(``aether_part_forward_phase``).  The exchange is either torch.distributed's in-place all-gather (NCCL over NVLink,
``exchange="nccl"``) or the library's own (``exchange="peer"``, csrc/peer.cu): the exchange buffers live in symmetric
memory (``torch.distributed._symmetric_memory``: every rank maps every peer's copy), and one kernel per rank stores its
block into all peers' copies over NVLink, raises a flag there and waits for the peers' flags -- no NCCL call on the
data path.

``parts > 1`` with ``world == 1`` runs the partitions one after another on one GPU over a shared exchange
buffer -- the same kernels and the same slicing, used by the single-GPU parity test.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from . import _lib
from .dist import ReceiverPartition
from .graph import PreparedGraph


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class _LocalPart:
    """Edge slice, CSR and workspace of one partition."""

    def __init__(self, lib, model, full: PreparedGraph, part: ReceiverPartition, flags: int):
        dev = full.send.device
        Nn = full.n_nodes
        e0, e1 = part.edge_begin, part.edge_end
        self.part = part
        self.E = e1 - e0
        self.send = full.send[e0:e1].contiguous() if self.E else torch.zeros(1, dtype=torch.int32, device=dev)
        self.recv = full.recv[e0:e1].contiguous() if self.E else torch.zeros(1, dtype=torch.int32, device=dev)
        self.perm = torch.arange(max(self.E, 1), dtype=torch.int32, device=dev)
        self.rowptr = (full.rowptr.to(torch.int64) - e0).clamp_(0, self.E).to(torch.int32).contiguous()
        self.send64 = self.send.to(torch.int64)
        self.recv64 = self.recv.to(torch.int64)
        self.ea = torch.empty(max(self.E, 1), 2, dtype=torch.float32, device=dev)
        self.c_graph = _lib.AetherGraphC(self.perm.data_ptr(), self.rowptr.data_ptr(), self.send.data_ptr(),
                                         self.recv.data_ptr(), 0, 0)
        self.dims = _lib.AetherDims(model.num_dims, model.hidden_size, Nn, self.E)
        nbytes = lib.aether_workspace_bytes(ctypes.byref(self.dims), flags)
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)


class PartitionedRollout:
    """Inference rollout of one scene with a static edge list, receiver-partitioned.

    ``run(x0, v0, charges)`` returns positions ``[steps, Nn, D]`` (identical on every rank)."""

    def __init__(self, model, send: torch.Tensor, recv: torch.Tensor, n_nodes: int, steps: int, world: int = 1,
                 rank: int = 0, parts: Optional[int] = None, group=None, cuda_graph: bool = False,
                 exchange: str = "nccl"):
        self.lib = _lib.load()
        self.model, self.steps, self.world, self.rank, self.group = model, int(steps), int(world), int(rank), group
        self.D, self.H, self.Nn = model.num_dims, model.hidden_size, int(n_nodes)
        dev = next(model.parameters()).device
        self.dev = dev
        self.flags = model.flags()
        full = PreparedGraph(send.to(dev), recv.to(dev), self.Nn)
        full.validate()
        n_parts = world if world > 1 else int(parts or 1)
        ranks = [rank] if world > 1 else list(range(n_parts))
        self.local: List[_LocalPart] = [
            _LocalPart(self.lib, model, full, ReceiverPartition(full.rowptr, self.Nn, n_parts, r), self.flags) for r in ranks]
        self.padded = self.local[0].part.padded_nodes
        xdtype = torch.bfloat16 if (self.flags & _lib.F_BF16 and self.H in (64, 128, 256)) else torch.float32
        f32 = dict(dtype=torch.float32, device=dev)
        if exchange not in ("nccl", "peer"):
            raise ValueError("exchange must be 'nccl' or 'peer'")
        self.exchange = exchange if world > 1 else "nccl"
        # two exchange buffers: a phase gathers sender terms from one and writes the next layer's into the other
        if self.exchange == "peer":
            self._alloc_symmetric(xdtype)
        else:
            self.xchg = [torch.zeros(self.padded, self.H, dtype=xdtype, device=dev) for _ in range(2)]
            self.xbuf = [torch.zeros(self.padded, self.D, **f32) for _ in range(2)]   # positions ping-pong (gathered)
        self.xin = torch.zeros(self.padded, self.D, **f32)                          # static inputs of a run
        self.vin = torch.zeros(self.padded, self.D, **f32)
        self.vbuf = [torch.zeros(self.padded, self.D, **f32) for _ in range(2)]
        self.q = torch.zeros(self.padded, **f32)
        self.traj = torch.empty(self.steps, self.Nn, self.D, **f32)
        self.edges_total = full.n_edges
        self.launch_exchanges = 0
        # the whole T-step loop, all-gathers included, as one CUDA graph (NCCL collectives are capturable; every rank
        # captures and replays the same sequence).  With world > 1, delete the object (and its graph) before
        # torch.distributed.destroy_process_group(): tearing the communicator down under a live captured graph hung
        # in our 8-GPU run.
        self._use_graph, self._g, self.launches_per_replay = bool(cuda_graph), None, 0

    def _alloc_symmetric(self, xdtype):
        """Exchange buffers + flag row in ONE symmetric allocation; peer base pointers for aether_peer_allgather."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        dev = self.dev
        group = self.group if self.group is not None else dist.group.WORLD
        al = lambda n: (n + 255) // 256 * 256
        xb = al(self.padded * self.H * (2 if xdtype == torch.bfloat16 else 4))
        pb = al(self.padded * self.D * 4)
        self._off = {"xchg0": 0, "xchg1": xb, "xbuf0": 2 * xb, "xbuf1": 2 * xb + pb, "flags": 2 * xb + 2 * pb}
        total = self._off["flags"] + 256
        self._symm = symm_mem.empty(total, dtype=torch.uint8, device=dev)
        self._symm.zero_()
        try:
            hdl = symm_mem.rendezvous(self._symm, group)
        except TypeError:
            hdl = symm_mem.rendezvous(self._symm, group.group_name)
        self._symm_hdl = hdl
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        if len(ptrs) != self.world or ptrs[self.rank] != self._symm.data_ptr():
            raise RuntimeError("aether_b200: symmetric memory rendezvous returned unexpected peer pointers")
        self._peer_base = (ctypes.c_void_p * self.world)(*ptrs)

        def view(name, dtype, cols):
            o = self._off[name]
            n = self.padded * cols * (2 if dtype == torch.bfloat16 else 4)
            return self._symm[o:o + n].view(dtype).view(self.padded, cols)

        self.xchg = [view("xchg0", xdtype, self.H), view("xchg1", xdtype, self.H)]
        self.xbuf = [view("xbuf0", torch.float32, self.D), view("xbuf1", torch.float32, self.D)]
        self._peer_state = torch.zeros(4, dtype=torch.int32, device=dev)
        self._name_of = {self.xchg[0].data_ptr(): "xchg0", self.xchg[1].data_ptr(): "xchg1",
                         self.xbuf[0].data_ptr(): "xbuf0", self.xbuf[1].data_ptr(): "xbuf1"}
        torch.cuda.synchronize(dev)
        dist.barrier(group=self.group)          # every copy and flag row is zero before anybody pushes

    def peer_error(self) -> bool:
        """True when a wait of the peer exchange gave up (a rank never arrived); synchronises."""
        return self.exchange == "peer" and bool(int(self._peer_state[2].item()))

    def _gather(self, buf: torch.Tensor):
        if self.world <= 1:
            return
        if self.exchange == "peer":
            block = self.local[0].part.block
            row_bytes = buf.shape[1] * buf.element_size()
            _lib.check(self.lib.aether_peer_allgather(self._peer_base, self.world, self.rank,
                                                      self._off[self._name_of[buf.data_ptr()]], block * row_bytes,
                                                      self._off["flags"], _lib.ptr(self._peer_state), _stream()),
                       "aether_peer_allgather")
        else:
            self.local[0].part.all_gather_rows(buf, self.group)
        self.launch_exchanges += 1

    def _step(self, x, v, xn, vn):
        lib = self.lib
        params = [p.detach() for p in self.model.param_list()]
        parr = _lib.ptr_array(params)
        for lp in self.local:    # caller prep on the local (receiver-sorted) edges: [q_i q_j, |x_i - x_j|]
            _lib.check(lib.aether_caller_prep(_lib.ptr(x), _lib.ptr(v), _lib.ptr(self.q), self.D, _lib.ptr(lp.send64),
                                              _lib.ptr(lp.recv64), lp.E, self.Nn, None, _lib.ptr(lp.ea), _stream()),
                       "aether_caller_prep")
        for phase in range(4):
            for lp in self.local:
                _lib.check(lib.aether_part_forward_phase(ctypes.byref(lp.dims), parr, ctypes.byref(lp.c_graph), _lib.ptr(x),
                                                         _lib.ptr(v), _lib.ptr(self.q), _lib.ptr(lp.ea),
                                                         lp.part.node_begin, lp.part.node_end, phase,
                                                         _lib.ptr(self.xchg[(phase + 1) & 1]),
                                                         _lib.ptr(self.xchg[phase & 1]), _lib.ptr(xn), _lib.ptr(lp.ws), lp.ws.numel(), self.flags,
                                                         _stream()), "aether_part_forward_phase")
            self._gather(self.xchg[phase & 1] if phase < 3 else xn)
        torch.sub(xn, x, out=vn)

    def _loop(self):
        x, v = self.xin, self.vin
        for t in range(self.steps):
            xn, vn = self.xbuf[t & 1], self.vbuf[t & 1]
            self._step(x, v, xn, vn)
            self.traj[t].copy_(xn[:self.Nn])
            x, v = xn, vn

    def release(self) -> None:
        """Drop the captured graph (it references the communicator's kernels): call before the process group is
        destroyed; the next ``run`` captures again."""
        if self._g is not None:
            torch.cuda.synchronize(self.dev)
            self._g = None

    @torch.no_grad()
    def run(self, x0, v0, charges) -> torch.Tensor:
        with torch.cuda.device(self.dev):
            self.xin[:self.Nn].copy_(x0, non_blocking=True)
            self.vin[:self.Nn].copy_(v0, non_blocking=True)
            self.q[:self.Nn].copy_(charges.reshape(-1), non_blocking=True)
            if not self._use_graph:
                self._loop()
                return self.traj
            if self._g is None:
                self._loop()                                   # warm-up outside capture (lazy init, NCCL channels)
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                n0 = self.lib.aether_launch_count()
                with torch.cuda.graph(g):
                    self._loop()
                self.launches_per_replay = int(self.lib.aether_launch_count() - n0)
                self._g = g
            self._g.replay()
            return self.traj
