"""``torch.ops.aether_b200.step``: the inference step as a registered operator (SURVEY.md 8b boundary row).

``aether_b200.Aether`` itself goes through ``ops.AetherStepFn`` (a ``torch.autograd.Function`` over the same C ABI);
this module is the ``torch.library`` form for callers that want an operator with a fake (meta) kernel -- export,
FakeTensor shape propagation, custom schedulers.  It is not imported by the package; ``import aether_b200.torch_ops``
registers it.  There is no CPU kernel: the op is registered for CUDA only, like the rest of the product path.

    out = torch.ops.aether_b200.step(x, vel, charges, edge_attr_orig, perm, rowptr, send, recv, sperm, colptr,
                                     params, n_edges, hidden, flags)

replaces ``Aether.forward`` (nn/state2state/aether.py:169-186) without gradient tracking; the six int32 arrays are the
members of ``AetherGraph`` (include/aether_b200.h) as produced by ``aether_graph_prepare`` / ``PreparedGraph``.
"""
from __future__ import annotations

from typing import List

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


def step_from_model(model, graph, x, vel, charges, edge_attr_orig) -> torch.Tensor:
    """Convenience: the operator call for an ``aether_b200.Aether`` module and a ``PreparedGraph``."""
    return torch.ops.aether_b200.step(x, vel, charges, edge_attr_orig, graph.perm, graph.rowptr, graph.send, graph.recv,
                                      graph.sperm, graph.colptr, [p.detach() for p in model.param_list()],
                                      graph.n_edges, model.hidden_size, model.flags())
