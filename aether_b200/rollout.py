"""Multi-step rollouts of the state-to-state step (harness-level).

One step t -> t+1, as the reference's caller would run it (experiments/lorentz/main.py:243-247):
    a_t = [q_i q_j, |x_i - x_j|] over the edge list;  x_{t+1} = Aether(|v_t|, x_t, edges_t, v_t, a_t, q)
    v_{t+1} = x_{t+1} - x_t;  dynamic-edge scenes rebuild edges_{t+1} = knn(x_{t+1}).
Each step is one C-ABI call (aether_rollout_step); the whole T-step loop can be captured in a CUDA
graph (``Rollout(..., cuda_graph=True)``) because no call synchronises or reads device data on the host.
"""
from __future__ import annotations

import ctypes
from typing import Optional

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
                 cuda_graph: bool = False):
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
    def run(self, x0, v0, charges) -> torch.Tensor:
        with torch.cuda.device(self.dev):
            self.x.copy_(x0, non_blocking=True)
            self.v.copy_(v0, non_blocking=True)
            self.q.copy_(charges.reshape(-1), non_blocking=True)
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
            cuda_graph: bool = False) -> torch.Tensor:
    r = Rollout(model, x0.shape[0], steps, edges=edges, knn_plan=knn_plan, cuda_graph=cuda_graph)
    return r.run(x0, v0, charges).clone()
