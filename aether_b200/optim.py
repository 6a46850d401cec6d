"""Fused AdamW for the drop-in module.

The reference's caller builds ``torch.optim.AdamW(model.parameters(), lr=5e-4, weight_decay=1e-12)``
(experiments/lorentz/main.py:164) and calls ``optimizer.step()`` after ``loss.backward()`` (:291-292).  ``FusedAdamW``
keeps that constructor / ``step`` / ``zero_grad`` / ``state_dict`` surface and runs the whole update -- all 47 parameter
tensors -- as ONE kernel of ``libaether_b200.so`` (``aether_adamw_step``) plus a one-thread step-count tick, instead of
torch's multi-tensor passes.  The step count lives on the device, so the update can be captured in a CUDA graph
(``aether_b200.train.GraphedTrainStep``).  Same arithmetic as ``torch.optim.AdamW`` (not amsgrad / maximize).
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-2):
        if lr < 0 or eps < 0 or not 0 <= betas[0] < 1 or not 0 <= betas[1] < 1 or weight_decay < 0:
            raise ValueError("invalid AdamW hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._tables = {}

    def _group_state(self, gi, group):
        """Flat moment buffers + pointer tables of one parameter group (built on first use, rebuilt when the set of
        parameters that carry a gradient changes)."""
        ps = [p for p in group["params"] if p.grad is not None]
        key = tuple((p.data_ptr(), p.grad.data_ptr(), self.state[p]["exp_avg"].data_ptr() if "exp_avg" in self.state[p] else 0)
                    for p in ps)
        tab = self._tables.get(gi)
        if tab is not None and tab["key"] == key and torch.is_tensor(group.get("_step_dev")):
            return tab
        for p in ps:
            if p.device.type != "cuda" or p.dtype != torch.float32 or not p.is_contiguous() or not p.grad.is_contiguous():
                raise RuntimeError("FusedAdamW: contiguous fp32 CUDA parameters and gradients only (no CPU fallback)")
            if p.grad.is_sparse:
                raise RuntimeError("FusedAdamW does not support sparse gradients")
        if len(ps) > 64:
            raise RuntimeError("FusedAdamW: at most 64 tensors per parameter group")
        dev = ps[0].device
        for p in ps:                                   # moments are per-parameter tensors (state_dict compatible)
            st = self.state[p]
            if "exp_avg" not in st:
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        if not torch.is_tensor(group.get("_step_dev")):
            group["_step_dev"] = torch.zeros(1, dtype=torch.float32, device=dev)
        key = tuple((p.data_ptr(), p.grad.data_ptr(), self.state[p]["exp_avg"].data_ptr()) for p in ps)
        n = len(ps)
        arr = lambda ts: (ctypes.c_void_p * n)(*[t.data_ptr() for t in ts])
        tab = {"key": key, "n": n, "p": arr(ps), "g": arr([p.grad for p in ps]),
               "m": arr([self.state[p]["exp_avg"] for p in ps]), "v": arr([self.state[p]["exp_avg_sq"] for p in ps]),
               "numel": (ctypes.c_int64 * n)(*[p.numel() for p in ps]), "dev": dev}
        self._tables[gi] = tab
        return tab

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        for gi, group in enumerate(self.param_groups):
            if not any(p.grad is not None for p in group["params"]):
                continue
            tab = self._group_state(gi, group)
            b1, b2 = group["betas"]
            with torch.cuda.device(tab["dev"]):
                st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
                _lib.check(lib.aether_adamw_step(tab["n"], tab["p"], tab["g"], tab["m"], tab["v"], tab["numel"],
                                                 _lib.ptr(group["_step_dev"]), float(group["lr"]), float(b1), float(b2),
                                                 float(group["eps"]), float(group["weight_decay"]), st), "aether_adamw_step")
        return loss

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._tables = {}                              # moments and step counter are new tensors now
        for group in self.param_groups:
            s = group.get("_step_dev")
            if torch.is_tensor(s) and group["params"]:
                group["_step_dev"] = s.to(group["params"][0].device, torch.float32)

    def step_count(self, group: int = 0) -> int:
        """Number of updates taken (reads the device counter)."""
        s = self.param_groups[group].get("_step_dev")
        return int(s.item()) if torch.is_tensor(s) else 0
