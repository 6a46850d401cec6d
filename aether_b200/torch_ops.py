"""``torch.ops.aether_b200.step``: the inference step as a registered operator (SURVEY.md 8b boundary row).

``aether_b200.Aether`` by default goes through ``ops.AetherStepFn`` (a ``torch.autograd.Function`` over the same C ABI);
this module is the ``torch.library`` form for callers that want operators with fake (meta) kernels -- export, FakeTensor
shape propagation, custom schedulers: ``aether_b200::step`` (inference), and the differentiable pair
``aether_b200::step_fwd -> (out, workspace)`` / ``aether_b200::step_bwd`` wired with ``register_autograd``.  With
``AETHER_B200_TORCH_OPS=1`` (or ``model.use_registered_ops = True``) ``Aether.forward`` routes through the pair whenever it
trains without dropout.  It is not imported by the package otherwise; ``import aether_b200.torch_ops`` registers it.  There is no CPU kernel: the op is registered for CUDA only, like the rest of the product path.

    out = torch.ops.aether_b200.step(x, vel, charges, edge_attr_orig, perm, rowptr, send, recv, sperm, colptr,
                                     params, n_edges, hidden, flags)

replaces ``Aether.forward`` (nn/state2state/aether.py:169-186) without gradient tracking; the six int32 arrays are the
members of ``AetherGraph`` (include/aether_b200.h) as produced by ``aether_graph_prepare`` / ``PreparedGraph``.
"""
from __future__ import annotations

from typing import List, Tuple

import torch

from . import _lib
from .ops import step_forward_raw


class _GraphView:
    """The fields ``step_forward_raw`` reads from a ``PreparedGraph``, over caller-provided arrays."""

    __slots__ = ("n_nodes", "n_edges", "_c", "_keep")

    def __init__(self, perm, rowptr, send, recv, sperm, colptr, n_edges: int):
        arrs = (perm, rowptr, send, recv, sperm, colptr)
        for a in arrs:
            if a.dtype != torch.int32 or not a.is_contiguous() or a.device.type != "cuda":
                raise ValueError("graph arrays must be contiguous int32 CUDA tensors")
        self.n_nodes, self.n_edges = int(rowptr.numel()) - 1, int(n_edges)
        if min(perm.numel(), send.numel(), recv.numel(), sperm.numel()) < self.n_edges or colptr.numel() != rowptr.numel():
            raise ValueError("graph arrays are shorter than n_edges / n_nodes + 1")
        self._keep = arrs
        self._c = _lib.AetherGraphC(*[a.data_ptr() for a in arrs])

    def c_struct(self):
        return self._c


@torch.library.custom_op("aether_b200::step", mutates_args=(), device_types="cuda")
def step(x: torch.Tensor, vel: torch.Tensor, charges: torch.Tensor, edge_attr_orig: torch.Tensor, perm: torch.Tensor,
         rowptr: torch.Tensor, send: torch.Tensor, recv: torch.Tensor, sperm: torch.Tensor, colptr: torch.Tensor,
         params: List[torch.Tensor], n_edges: int, hidden: int, flags: int) -> torch.Tensor:
    graph = _GraphView(perm, rowptr, send, recv, sperm, colptr, n_edges)
    f32 = lambda t: t.detach().to(torch.float32).contiguous()
    out, _ = step_forward_raw([f32(p) for p in params], graph, f32(x), f32(vel), f32(charges), f32(edge_attr_orig),
                              int(x.shape[1]), int(hidden), int(flags) & ~_lib.F_SAVE)
    return out


@step.register_fake
def _(x, vel, charges, edge_attr_orig, perm, rowptr, send, recv, sperm, colptr, params, n_edges, hidden, flags):
    torch._check(x.dim() == 2 and x.shape[1] in (2, 3), lambda: "x must be [Nn, 2|3]")
    torch._check(vel.shape == x.shape, lambda: "vel must have the shape of x")
    torch._check(rowptr.shape[0] == x.shape[0] + 1, lambda: "rowptr must have Nn + 1 entries")
    torch._check(edge_attr_orig.shape[0] == n_edges, lambda: "edge_attr_orig must have n_edges rows")
    torch._check(len(params) == _lib.NUM_PARAMS, lambda: "expected %d parameter tensors" % _lib.NUM_PARAMS)
    return x.new_empty(x.shape, dtype=torch.float32)


# ---- the differentiable pair: step_fwd (keeps its workspace) / step_bwd, wired together with register_autograd ----------

def _ws_bytes(D: int, H: int, n_nodes: int, n_edges: int, flags: int) -> int:
    import ctypes
    dims = _lib.AetherDims(D, H, n_nodes, n_edges)
    return int(_lib.load().aether_workspace_bytes(ctypes.byref(dims), flags))


@torch.library.custom_op("aether_b200::step_fwd", mutates_args=(), device_types="cuda")
def step_fwd(x: torch.Tensor, vel: torch.Tensor, charges: torch.Tensor, edge_attr_orig: torch.Tensor, perm: torch.Tensor,
             rowptr: torch.Tensor, send: torch.Tensor, recv: torch.Tensor, sperm: torch.Tensor, colptr: torch.Tensor,
             params: List[torch.Tensor], n_edges: int, hidden: int, flags: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """(out, workspace): the training forward (activations kept: AETHER_F_SAVE_FOR_BACKWARD is set here)."""
    graph = _GraphView(perm, rowptr, send, recv, sperm, colptr, n_edges)
    f32 = lambda t: t.detach().to(torch.float32).contiguous()
    out, ws = step_forward_raw([f32(p) for p in params], graph, f32(x), f32(vel), f32(charges), f32(edge_attr_orig),
                               int(x.shape[1]), int(hidden), int(flags) | _lib.F_SAVE)
    return out, ws


@step_fwd.register_fake
def _(x, vel, charges, edge_attr_orig, perm, rowptr, send, recv, sperm, colptr, params, n_edges, hidden, flags):
    torch._check(x.dim() == 2 and x.shape[1] in (2, 3), lambda: "x must be [Nn, 2|3]")
    torch._check(len(params) == _lib.NUM_PARAMS, lambda: "expected %d parameter tensors" % _lib.NUM_PARAMS)
    nbytes = _ws_bytes(int(x.shape[1]), int(hidden), int(x.shape[0]), int(n_edges), int(flags) | _lib.F_SAVE)
    return x.new_empty(x.shape, dtype=torch.float32), x.new_empty((nbytes,), dtype=torch.uint8)


@torch.library.custom_op("aether_b200::step_bwd", mutates_args=(), device_types="cuda")
def step_bwd(grad_out: torch.Tensor, workspace: torch.Tensor, x: torch.Tensor, vel: torch.Tensor, charges: torch.Tensor,
             edge_attr_orig: torch.Tensor, perm: torch.Tensor, rowptr: torch.Tensor, send: torch.Tensor, recv: torch.Tensor,
             sperm: torch.Tensor, colptr: torch.Tensor, params: List[torch.Tensor], n_edges: int, hidden: int,
             flags: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(flat parameter gradients [sum of 16-byte-padded sizes], d/dx, d/dvel) from the workspace step_fwd left."""
    import ctypes
    lib = _lib.load()
    graph = _GraphView(perm, rowptr, send, recv, sperm, colptr, n_edges)
    f32 = lambda t: t.detach().to(torch.float32).contiguous()
    ps = [f32(p) for p in params]
    x, vel, charges, ea, go = f32(x), f32(vel), f32(charges), f32(edge_attr_orig), f32(grad_out)
    D, f = int(x.shape[1]), int(flags) | _lib.F_SAVE
    dims = _lib.AetherDims(D, int(hidden), graph.n_nodes, graph.n_edges)
    offs, tot = [], 0
    for p in ps:
        offs.append(tot)
        tot += (p.numel() + 3) // 4 * 4
    flat = torch.zeros(tot, dtype=torch.float32, device=x.device)
    grads = [flat[o:o + p.numel()] for o, p in zip(offs, ps)]
    gx, gv = torch.empty_like(x), torch.empty_like(vel)
    nbytes = lib.aether_backward_scratch_bytes(ctypes.byref(dims), f)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    with torch.cuda.device(x.device):
        _lib.check(lib.aether_step_backward(ctypes.byref(dims), _lib.ptr_array(ps), ctypes.byref(graph.c_struct()),
                                            _lib.ptr(x), _lib.ptr(vel), _lib.ptr(charges), _lib.ptr(ea), _lib.ptr(None),
                                            _lib.ptr(go), _lib.ptr_array(grads), _lib.ptr(gx), _lib.ptr(gv),
                                            _lib.ptr(workspace), workspace.numel(), _lib.ptr(scratch), nbytes, f, st),
                   "aether_step_backward")
    return flat, gx, gv


@step_bwd.register_fake
def _(grad_out, workspace, x, vel, charges, edge_attr_orig, perm, rowptr, send, recv, sperm, colptr, params, n_edges, hidden,
      flags):
    tot = sum((p.numel() + 3) // 4 * 4 for p in params)
    return x.new_empty((tot,), dtype=torch.float32), x.new_empty(x.shape, dtype=torch.float32), x.new_empty(x.shape, dtype=torch.float32)


def _setup_context(ctx, inputs, output):
    x, vel, charges, ea, perm, rowptr, send, recv, sperm, colptr, params, n_edges, hidden, flags = inputs
    _, ws = output
    ctx.save_for_backward(ws, x, vel, charges, ea, perm, rowptr, send, recv, sperm, colptr, *params)
    ctx.meta = (n_edges, hidden, flags)


def _backward(ctx, grad_out, grad_ws):
    ws, x, vel, charges, ea, perm, rowptr, send, recv, sperm, colptr, *params = ctx.saved_tensors
    n_edges, hidden, flags = ctx.meta
    flat, gx, gv = torch.ops.aether_b200.step_bwd(grad_out, ws, x, vel, charges, ea, perm, rowptr, send, recv, sperm, colptr,
                                                  list(params), n_edges, hidden, flags)
    grads, off = [], 0
    for p in params:
        grads.append(flat[off:off + p.numel()].view_as(p))
        off += (p.numel() + 3) // 4 * 4
    return gx, gv, None, None, None, None, None, None, None, None, grads, None, None, None


step_fwd.register_autograd(_backward, setup_context=_setup_context)


def train_step_from_model(model, graph, x, vel, charges, edge_attr_orig) -> torch.Tensor:
    """The differentiable operator call for an ``aether_b200.Aether`` module (dropout-free: ``dropout_prob == 0``)."""
    out, _ = torch.ops.aether_b200.step_fwd(x, vel, charges.reshape(-1), edge_attr_orig, graph.perm, graph.rowptr, graph.send,
                                            graph.recv, graph.sperm, graph.colptr, list(model.param_list()), graph.n_edges,
                                            model.hidden_size, model.flags())
    return out


def step_from_model(model, graph, x, vel, charges, edge_attr_orig) -> torch.Tensor:
    """Convenience: the operator call for an ``aether_b200.Aether`` module and a ``PreparedGraph``."""
    return torch.ops.aether_b200.step(x, vel, charges, edge_attr_orig, graph.perm, graph.rowptr, graph.send, graph.recv,
                                      graph.sperm, graph.colptr, [p.detach() for p in model.param_list()],
                                      graph.n_edges, model.hidden_size, model.flags())
