"""The step as a differentiable op: thin autograd wrapper over the C ABI.

forward  -> aether_step_forward   (Aether.forward, nn/state2state/aether.py:169-186)
backward -> aether_step_backward  (loss.backward(), experiments/lorentz/main.py:291)

All buffers (workspace, scratch, gradients) are torch allocations handed to the library as raw
pointers; launches go to torch's current stream, so the calls can be captured in a CUDA graph.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch
from torch.autograd.function import once_differentiable

from . import _lib
from .graph import PreparedGraph


def _stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _dims(D: int, H: int, graph: PreparedGraph) -> "_lib.AetherDims":
    return _lib.AetherDims(D, H, graph.n_nodes, graph.n_edges)


def step_forward_raw(params: Sequence[torch.Tensor], graph: PreparedGraph, x, vel, charges, edge_attr_orig, D: int,
                     H: int, flags: int, dropout_mask: Optional[torch.Tensor] = None,
                     workspace: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None):
    """One forward step on raw tensors.  Returns (out, workspace)."""
    lib = _lib.load()
    if x.device.type != "cuda":
        raise RuntimeError("aether_b200 runs on CUDA devices only (no CPU fallback)")
    if len(params) != _lib.NUM_PARAMS:
        raise ValueError("expected %d parameter tensors" % _lib.NUM_PARAMS)
    dims = _dims(D, H, graph)
    if x.shape[0] != graph.n_nodes or edge_attr_orig.shape[0] != graph.n_edges:
        raise ValueError("node/edge counts do not match the prepared graph")
    nbytes = lib.aether_workspace_bytes(ctypes.byref(dims), flags)
    if workspace is None or workspace.numel() < nbytes:
        workspace = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    if out is None:
        out = torch.empty(graph.n_nodes, D, dtype=torch.float32, device=x.device)
    parr = _lib.ptr_array(params)
    with torch.cuda.device(x.device):
        _lib.check(lib.aether_step_forward(ctypes.byref(dims), parr, ctypes.byref(graph.c_struct()), _lib.ptr(x),
                                           _lib.ptr(vel), _lib.ptr(charges), _lib.ptr(edge_attr_orig),
                                           _lib.ptr(dropout_mask), _lib.ptr(out), _lib.ptr(workspace),
                                           workspace.numel(), flags, _stream()), "aether_step_forward")
    return out, workspace


class AetherStepFn(torch.autograd.Function):
    """Use ``step_apply``: it decides whether the forward keeps what the backward needs."""

    @staticmethod
    def forward(ctx, x, vel, charges, edge_attr_orig, graph, D, H, flags, dropout_mask, save, *params):
        x, vel, charges, edge_attr_orig = _f32c(x), _f32c(vel), _f32c(charges), _f32c(edge_attr_orig)
        params = [_f32c(p) for p in params]
        # grad mode is always off in here and needs_input_grad stays True under torch.no_grad() (the parameters still
        # require grad), so the caller passes `save`: without it the forward ping-pongs 3 edge buffers instead of 8
        need = bool(save) and any(ctx.needs_input_grad)
        if need and (ctx.needs_input_grad[2] or ctx.needs_input_grad[3]):
            raise RuntimeError("aether_b200: gradients w.r.t. charges / edge_attr_orig are not implemented; detach "
                               "them (the reference's caller passes constants, experiments/lorentz/main.py:243-247)")
        f = (int(flags) & ~_lib.F_SAVE) | (_lib.F_SAVE if need else 0)
        out, ws = step_forward_raw(params, graph, x, vel, charges, edge_attr_orig, D, H, f, dropout_mask)
        if need:
            ctx.save_for_backward(x, vel, charges, edge_attr_orig, *params)
            ctx.ws, ctx.graph, ctx.dims, ctx.flags, ctx.dmask = ws, graph, (D, H), f, dropout_mask
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_out):
        lib = _lib.load()
        if getattr(ctx, "ws", None) is None:
            raise RuntimeError("aether_b200: the step's saved workspace was already consumed; a second backward "
                               "through the same forward (retain_graph=True) is not supported -- run forward again")
        x, vel, charges, ea, *params = ctx.saved_tensors
        D, H = ctx.dims
        graph = ctx.graph
        dims = _dims(D, H, graph)
        grad_out = _f32c(grad_out)
        # all 47 gradients are views of ONE buffer (state_dict order): the data-parallel all-reduce runs on it
        # in place (aether_b200.dist.allreduce_grads).  Views start 16-byte aligned (float4 stores); the pad words stay zero.
        offs, tot = [], 0
        for p in params:
            offs.append(tot)
            tot += (p.numel() + 3) // 4 * 4
        flat = torch.zeros(tot, dtype=torch.float32, device=x.device)
        grads = [flat[o:o + p.numel()].view_as(p) for o, p in zip(offs, params)]
        gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        gv = torch.empty_like(vel) if ctx.needs_input_grad[1] else None
        nbytes = lib.aether_backward_scratch_bytes(ctypes.byref(dims), ctx.flags)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(lib.aether_step_backward(ctypes.byref(dims), _lib.ptr_array(params),
                                                ctypes.byref(graph.c_struct()), _lib.ptr(x), _lib.ptr(vel),
                                                _lib.ptr(charges), _lib.ptr(ea), _lib.ptr(ctx.dmask),
                                                _lib.ptr(grad_out), _lib.ptr_array(grads), _lib.ptr(gx), _lib.ptr(gv),
                                                _lib.ptr(ctx.ws), ctx.ws.numel(), _lib.ptr(scratch), nbytes,
                                                ctx.flags, _stream()), "aether_step_backward")
        ctx.ws = None
        return (gx, gv, None, None, None, None, None, None, None, None, *grads)


def step_apply(x, vel, charges, edge_attr_orig, graph, D, H, flags, dropout_mask, params):
    """The differentiable step.  Keeps the per-layer pre-activations only when a backward can follow."""
    save = torch.is_grad_enabled() and (x.requires_grad or vel.requires_grad or any(p.requires_grad for p in params))
    return AetherStepFn.apply(x, vel, charges, edge_attr_orig, graph, D, H, flags, dropout_mask, save, *params)


def dropout_masks(training: bool, p: float, n_nodes: int, hidden: int, device):
    """The two inverted-dropout masks of out_mlp (locs.py:163,166) as one [2, Nn, H] tensor, or None."""
    if not training or p <= 0.0:
        return None
    keep = 1.0 - p
    if keep <= 0.0:
        return torch.zeros(2, n_nodes, hidden, device=device)        # nn.Dropout(p=1) zeroes everything
    return torch.bernoulli(torch.full((2, n_nodes, hidden), keep, device=device)) / keep


def debug_stage(ws: torch.Tensor, graph: PreparedGraph, D: int, H: int, flags: int, stage: str) -> torch.Tensor:
    """Copy a named intermediate out of a forward workspace (tests only)."""
    lib = _lib.load()
    dims = _dims(D, H, graph)
    Nn, E = graph.n_nodes, graph.n_edges
    O = D * (D - 1) // 2
    shapes = {"field": (Nn, D), "rel_feat": (Nn, 3 * D), "R": (Nn, D, D), "pred": (Nn, D),
              "edge_attr": (E, 7 * D + O + 2)}
    for l in "1234":
        shapes["e" + l] = (E, H)
        shapes["x" + l] = (Nn, H)
    out = torch.empty(shapes[stage], dtype=torch.float32, device=ws.device)
    with torch.cuda.device(ws.device):
        n = lib.aether_debug_stage(ctypes.byref(dims), _lib.ptr(ws), flags, stage.encode(), _lib.ptr(out), out.numel(),
                                   _stream())
    if n < 0:
        _lib.check(int(n), "aether_debug_stage")
    return out
