"""Multi-step rollouts of the state-to-state step (harness-level).

One step t -> t+1, as the reference's caller would run it (experiments/lorentz/main.py:243-247):
    a_t = [q_i q_j, |x_i - x_j|] over the edge list;  x_{t+1} = Aether(|v_t|, x_t, edges_t, v_t, a_t, q)
    v_{t+1} = x_{t+1} - x_t;  dynamic-edge scenes rebuild edges_{t+1} = knn(x_{t+1}).
Each step is one C-ABI call (aether_rollout_step); the whole T-step loop can be captured in a CUDA
graph (``Rollout(..., cuda_graph=True)``) because no call synchronises or reads device data on the host.

Batches of small scenes (hidden size 64, at most 64 particles per scene, edge state within one SM's shared memory:
BASELINE configs 1-3) take the single-launch path instead: ``aether_small_rollout`` runs all T steps of every scene in
one kernel, one CTA per scene (csrc/small.cu).  It is chosen automatically when the scene sizes are known -- always
for kNN plans, for static edge lists when ``sizes`` is passed -- and ``persistent=False`` keeps the multi-launch path.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch

from . import _lib
from .graph import KnnPlan, PreparedGraph, knn_edges


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class Rollout:
    """Inference rollout with static buffers.

    ``edges``: (send, recv) int64 device tensors for a static edge list, or None with ``knn_plan`` for
    per-step kNN edges.  ``run(x0, v0, charges)`` returns positions ``[steps, Nn, D]``.
    """

    def __init__(self, model, n_nodes: int, steps: int, edges=None, knn_plan: Optional[KnnPlan] = None,
                 cuda_graph: bool = False, sizes: Optional[Sequence[int]] = None, persistent: Optional[bool] = None):
        self.lib = _lib.load()
        self.model, self.steps, self.plan = model, int(steps), knn_plan
        self.D, self.H = model.num_dims, model.hidden_size
        dev = next(model.parameters()).device
        self.dev = dev
        self.Nn = int(n_nodes)
        if (edges is None) == (knn_plan is None):
            raise ValueError("pass exactly one of edges / knn_plan")
        if edges is not None:
            self.send = edges[0].to(dev, torch.int64).contiguous()
            self.recv = edges[1].to(dev, torch.int64).contiguous()
            self.graph = PreparedGraph(self.send, self.recv, self.Nn)
            self.E = self.graph.n_edges
        else:
            self.E = knn_plan.n_edges
            self.send = torch.empty(self.E, dtype=torch.int64, device=dev)
            self.recv = torch.empty(self.E, dtype=torch.int64, device=dev)
            self.graph = None
        f32 = dict(dtype=torch.float32, device=dev)
        self.x = torch.empty(self.Nn, self.D, **f32)
        self.v = torch.empty(self.Nn, self.D, **f32)
        self.q = torch.empty(self.Nn, **f32)
        self.traj = torch.empty(self.steps, self.Nn, self.D, **f32)
        self.vbuf = [torch.empty(self.Nn, self.D, **f32) for _ in range(2)]
        self.ea = torch.empty(max(self.E, 1), 2, **f32)
        self.flags = model.flags()
        self.dims = _lib.AetherDims(self.D, self.H, self.Nn, self.E)
        nbytes = self.lib.aether_workspace_bytes(ctypes.byref(self.dims), self.flags)
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        self._graph_exec = None
        self._use_graph = cuda_graph
        self.launches_per_replay = 0
        self.persistent = False
        if persistent is not False:
            why = self._plan_persistent(sizes)
            if why and persistent:
                raise ValueError("Rollout(persistent=True): " + why)

    # ---- single-launch path ------------------------------------------------------------------------------------------
    def _plan_persistent(self, sizes) -> Optional[str]:
        """Scene offsets for aether_small_rollout; returns why the path does not apply, or None when it does."""
        from .aether import Aether
        if type(self.model) is not Aether:
            return "only the Aether step has a single-launch rollout"
        if self.plan is not None:
            sizes = self.plan.sizes
        if sizes is None:
            return "scene sizes unknown (pass sizes=[particles per scene])"
        sizes = [int(s) for s in sizes]
        if sum(sizes) != self.Nn or any(s < 0 for s in sizes):
            return "sizes do not add up to the number of particles"
        dev = self.dev
        node_off = torch.zeros(len(sizes) + 1, dtype=torch.int64)
        node_off[1:] = torch.tensor(sizes, dtype=torch.int64).cumsum(0) if sizes else 0
        if self.plan is not None:
            edge_off = self.plan.edge_off.to(torch.int64).cpu()
            k = self.plan.k
        else:
            # the edges of a scene must be contiguous in the list and stay inside the scene (one host check, at construction)
            bounds = node_off[1:].to(dev)
            ss = torch.bucketize(self.send, bounds, right=True)
            sr = torch.bucketize(self.recv, bounds, right=True)
            if self.E and (bool((ss != sr).any()) or bool((ss[1:] < ss[:-1]).any()) or int(self.send.min()) < 0 or
                           int(self.recv.min()) < 0 or int(ss.max()) >= len(sizes)):
                return "edges are not grouped by scene"
            edge_off = torch.zeros(len(sizes) + 1, dtype=torch.int64)
            if self.E:
                edge_off[1:] = torch.bincount(ss, minlength=len(sizes)).cumsum(0).cpu()
            k = 0
        per_edges = (edge_off[1:] - edge_off[:-1])
        self._max_nodes = max(max(sizes), 1) if sizes else 1
        self._max_edges = int(per_edges.max()) if len(sizes) else 0
        if not self.lib.aether_small_rollout_supported(self.D, self.H, self._max_nodes, self._max_edges, self.flags):
            return "shape not supported by the single-launch kernel (hidden 64, <= 64 particles, edge state in shared memory)"
        self._n_scenes = len(sizes)
        self._node_off = node_off.to(torch.int32).to(dev)
        self._edge_off = edge_off.to(dev)
        self._knn_k = int(k)
        # largest scenes first: the hardware hands CTAs to SMs in index order, so the long ones start early
        work = per_edges.to(torch.float64) + 0.25 * torch.tensor(sizes, dtype=torch.float64)
        self._order = torch.argsort(work, descending=True, stable=True).to(torch.int32).to(dev)
        self._scratch = torch.empty(self.lib.aether_small_rollout_scratch_bytes(), dtype=torch.uint8, device=dev)
        self.last_edges = None
        self.persistent = True
        return None

    def _run_persistent(self, dump_edges: bool = False):
        params = [p.detach() for p in self.model.param_list()]
        parr = _lib.ptr_array(params)
        so = ro = None
        if dump_edges and self.plan is not None:
            so, ro = torch.empty_like(self.send), torch.empty_like(self.recv)
            self.last_edges = (so, ro)
        static = self.plan is None
        rc = self.lib.aether_small_rollout(ctypes.byref(self.dims), parr, self._n_scenes, _lib.ptr(self._node_off),
                                           _lib.ptr(self._edge_off), _lib.ptr(self._order),
                                           _lib.ptr(self.send if static else None), _lib.ptr(self.recv if static else None),
                                           self._knn_k, self._max_nodes, self._max_edges, _lib.ptr(self.x), _lib.ptr(self.v),
                                           _lib.ptr(self.q), self.steps, _lib.ptr(self.traj), _lib.ptr(so), _lib.ptr(ro),
                                           _lib.ptr(self._scratch), self._scratch.numel(), self.flags, _stream())
        if rc == -3:            # parameter tensors not 16-byte aligned (views into a flat buffer): multi-launch path
            return False
        _lib.check(rc, "aether_small_rollout")
        return True

    def _steps(self):
        params = [p.detach() for p in self.model.param_list()]
        parr = _lib.ptr_array(params)
        x, v = self.x, self.v
        for t in range(self.steps):
            if self.plan is not None:
                knn_edges(x, self.plan, out=(self.send, self.recv))
                graph = PreparedGraph(self.send, self.recv, self.Nn)
            else:
                graph = self.graph
            xn, vn = self.traj[t], self.vbuf[t & 1]
            _lib.check(self.lib.aether_rollout_step(ctypes.byref(self.dims), parr, ctypes.byref(graph.c_struct()),
                                                    _lib.ptr(self.send), _lib.ptr(self.recv), _lib.ptr(x), _lib.ptr(v),
                                                    _lib.ptr(self.q), _lib.ptr(xn), _lib.ptr(vn), _lib.ptr(self.ea),
                                                    _lib.ptr(self.ws), self.ws.numel(), self.flags, _stream()),
                       "aether_rollout_step")
            self._keep = getattr(self, "_keep", [])
            self._keep.append(graph)          # CSR buffers must outlive the captured graph
            x, v = xn, vn

    @torch.no_grad()
    def run(self, x0, v0, charges, dump_edges: bool = False) -> torch.Tensor:
        with torch.cuda.device(self.dev):
            self.x.copy_(x0, non_blocking=True)
            self.v.copy_(v0, non_blocking=True)
            self.q.copy_(charges.reshape(-1), non_blocking=True)
            if self.persistent:
                if self._run_persistent(dump_edges):  # weight pack + ONE rollout launch
                    return self.traj
                self.persistent = False
            if not self._use_graph:
                self._keep = []
                self._steps()
                return self.traj
            if self._graph_exec is None:
                self._keep = []
                self._steps()                      # warm-up outside capture (allocations, lazy init)
                torch.cuda.synchronize()
                self._keep = []
                g = torch.cuda.CUDAGraph()
                n0 = self.lib.aether_launch_count()
                with torch.cuda.graph(g):
                    self._steps()
                self.launches_per_replay = int(self.lib.aether_launch_count() - n0)   # kernels one replay launches
                self._graph_exec = g
            self._graph_exec.replay()
            return self.traj


def rollout(model, x0, v0, charges, steps: int, edges=None, knn_plan: Optional[KnnPlan] = None,
            cuda_graph: bool = False, sizes: Optional[Sequence[int]] = None, persistent: Optional[bool] = None) -> torch.Tensor:
    r = Rollout(model, x0.shape[0], steps, edges=edges, knn_plan=knn_plan, cuda_graph=cuda_graph, sizes=sizes,
                persistent=persistent)
    return r.run(x0, v0, charges).clone()
