"""Whole-train-step CUDA graph for shapes where launch latency, not bandwidth, sets the pace.

One training step of the reference's caller (experiments/lorentz/main.py:243-247,289-292):
    edge_attr = [q_i q_j, |x_i - x_j|];  out = model(|v|, x, edges, v, edge_attr, q);  loss = MSE(out, target)
    loss.backward();  optimizer.step()
is a few hundred kernel launches of libaether_b200.so plus the loss and the fused AdamW.  For the 5- and 20-particle
configurations (a few thousand edges) every kernel is shorter than its launch, so the step is captured once and replayed:
no call of the library synchronises or reads device data on the host, and all buffers come from torch's graph-private pool.
"""
from __future__ import annotations

import torch

from .graph import PreparedGraph, caller_prep


class GraphedTrainStep:
    """``step(x, vel, charges, target) -> loss`` (a device scalar; read it when you need it).

    ``edges``: (send, recv) int64 device tensors, static across steps.  The optimizer must be capturable
    (``aether_b200.FusedAdamW``, or ``torch.optim.AdamW(..., fused=True, capturable=True)``).  ``allreduce``: optional callable run after backward
    (data-parallel gradient averaging; NCCL collectives are capturable)."""

    def __init__(self, model, optimizer, edges, n_nodes: int, allreduce=None, warmup: int = 3):
        self.model, self.opt, self.allreduce = model, optimizer, allreduce
        dev = next(model.parameters()).device
        D = model.num_dims
        self.send, self.recv = edges[0].to(dev), edges[1].to(dev)
        self.graph_csr = PreparedGraph(self.send, self.recv, n_nodes)
        self.graph_csr.validate()
        f32 = dict(dtype=torch.float32, device=dev)
        self.x = torch.zeros(n_nodes, D, **f32)
        self.v = torch.zeros(n_nodes, D, **f32)
        self.q = torch.zeros(n_nodes, 1, **f32)
        self.t = torch.zeros(n_nodes, D, **f32)
        self.loss = torch.zeros((), **f32)
        self._warmup, self._g = int(warmup), None
        self.launches_per_replay = 0

    def _body(self):
        self.opt.zero_grad(set_to_none=True)
        h, ea = caller_prep(self.x, self.v, self.q.reshape(-1), self.send, self.recv)
        out = self.model(h, self.x, self.graph_csr, self.v, ea, self.q)
        loss = torch.nn.functional.mse_loss(out, self.t)
        loss.backward()
        if self.allreduce is not None:
            self.allreduce()
        self.opt.step()
        self.loss.copy_(loss.detach())

    def _capture(self):
        import copy
        # the warm-up iterations below take optimizer steps: snapshot parameters and optimizer state, restore after
        model_state = copy.deepcopy(self.model.state_dict())
        opt_state = copy.deepcopy(self.opt.state_dict())
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(self._warmup):        # lazy initialisation (optimizer state, function attributes) outside capture
                self._body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.model.load_state_dict(model_state)          # in place: parameter addresses are unchanged
        if opt_state["state"]:
            self.opt.load_state_dict(opt_state)          # before capture: the captured step sees the final state tensors
        else:                                            # a fresh optimizer: keep the lazily created state, zeroed in
            for st in self.opt.state.values():           # place (re-creating it inside the capture would bake the
                for v in st.values():                    # zero-fill into every replay)
                    if torch.is_tensor(v):
                        v.zero_()
            for group in self.opt.param_groups:          # FusedAdamW keeps its step count per group, on the device
                if torch.is_tensor(group.get("_step_dev")):
                    group["_step_dev"].zero_()
        torch.cuda.synchronize()
        from . import _lib
        lib = _lib.load()
        g = torch.cuda.CUDAGraph()
        n0 = lib.aether_launch_count()
        with torch.cuda.graph(g):
            self._body()
        self.launches_per_replay = int(lib.aether_launch_count() - n0)   # library kernels one replay launches
        self._g = g

    def step(self, x, vel, charges, target) -> torch.Tensor:
        self.x.copy_(x, non_blocking=True)
        self.v.copy_(vel, non_blocking=True)
        self.q.copy_(charges.reshape(-1, 1), non_blocking=True)
        self.t.copy_(target, non_blocking=True)
        if self._g is None:
            self._capture()
        self._g.replay()
        return self.loss
