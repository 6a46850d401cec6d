"""Drop-in ``DynamicFieldAether`` (nn/state2state/dynamic_field_aether.py:50-100 of the reference).

Same constructor, ``forward(h, x, edges, vel, edge_attr_orig, charges, num_nodes)`` signature, sub-module / parameter
names and default initialisation order as the reference (``gnn`` first, then ``field_net``: summary gate / value
networks, the FiLM-modulated field MLP, the class embedding), so a seed reproduces the reference's weights and
``state_dict``s are interchangeable.  The sub-modules only hold parameters.  ``forward`` runs

    latent field network   aether_dynfield_forward   (csrc/dynfield.cu: attention pooling + FiLM kernels, fp32 GEMMs)
    step with that field   aether_step_forward        (AETHER_F_EXTERNAL_FIELD; the same kernels as ``Aether``)

and the backward chains ``aether_step_backward`` -> ``aether_dynfield_backward``.  No ATen or CPU fallback.
"""
from __future__ import annotations

import ctypes
import os
from typing import List

import numpy as np
import torch
from torch import nn
from torch.autograd.function import once_differentiable

from . import _lib
from .aether import Aether, AetherLocalizer, Globalizer, GNN, _warn_bf16_fallback, check_charges
from .ops import _dims, _f32c, _stream, dropout_masks

F_EXTERNAL_FIELD = 4
NUM_FIELD_PARAMS = 27
NUM_GNN_PARAMS = 40


class AttentionalAggregation(nn.Module):
    """Parameter holder with PyG's attribute names (torch_geometric.nn.aggr.AttentionalAggregation: ``gate_nn``, ``nn``)."""

    def __init__(self, gate_nn, nn_):
        super().__init__()
        self.gate_nn = gate_nn
        self.nn = nn_
        # PyG's constructor ends with reset_parameters(): every Linear of gate_nn, then of nn, is initialised a second
        # time -- reproduced so that a seed yields the reference's weights
        for seq in (self.gate_nn, self.nn):
            for m in seq:
                if hasattr(m, "reset_parameters"):
                    m.reset_parameters()


class GraphSummary(nn.Module):
    """graph_pool.py:6-29."""

    def __init__(self, input_size, hidden_size):
        super().__init__()
        self.summary_net = AttentionalAggregation(
            nn.Sequential(nn.Linear(input_size, hidden_size), nn.SiLU(), nn.Linear(hidden_size, 1)),
            nn.Sequential(nn.Linear(input_size, hidden_size), nn.SiLU(), nn.Linear(hidden_size, hidden_size)))


class FiLM(nn.Module):
    """film.py:40-60."""

    def __init__(self, x_size, z_size, hidden_size):
        super().__init__()
        self.modulator = nn.Sequential(nn.Linear(z_size, hidden_size), nn.SiLU(), nn.Linear(hidden_size, hidden_size),
                                       nn.SiLU(), nn.Linear(hidden_size, 2 * x_size))


class FilmedNetwork(nn.Module):
    """film.py:5-37 (construction order: the three Linears, then the two FiLM blocks)."""

    def __init__(self, x_size, z_size, hidden_size, out_size):
        super().__init__()
        self.linear_1 = nn.Linear(x_size, hidden_size)
        self.linear_2 = nn.Linear(hidden_size, hidden_size)
        self.linear_3 = nn.Linear(hidden_size, out_size)
        self.film_1 = FiLM(hidden_size, z_size, hidden_size)
        self.film_2 = FiLM(hidden_size, z_size, hidden_size)
        self.act_1 = nn.SiLU()
        self.act_2 = nn.SiLU()


class LatentFieldNetwork(nn.Module):
    """dynamic_field_aether.py:11-47."""

    def __init__(self, num_dims, hidden_size, class_embedding_dim):
        super().__init__()
        self.num_dims = num_dims
        self.summary_dim = hidden_size
        self.summary_net = GraphSummary(2 * num_dims, self.summary_dim)
        self.wrapper = FilmedNetwork(2 * num_dims + class_embedding_dim, self.summary_dim, hidden_size, num_dims)
        self.class_embedding = nn.Embedding(3, class_embedding_dim)

    @staticmethod
    def charge_to_index(charges):
        return (charges + 1).long()


class _DynStepFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, vel, charges, edge_attr_orig, graph, D, H, flags, num_nodes, dropout_mask, save, *params):
        lib = _lib.load()
        x, vel, charges, edge_attr_orig = _f32c(x), _f32c(vel), _f32c(charges), _f32c(edge_attr_orig)
        params = [_f32c(p) for p in params]
        gnn_p, fld_p = params[:NUM_GNN_PARAMS], params[NUM_GNN_PARAMS:]
        need = bool(save) and any(ctx.needs_input_grad)
        if need and (ctx.needs_input_grad[2] or ctx.needs_input_grad[3]):
            raise RuntimeError("aether_b200: gradients w.r.t. charges / edge_attr_orig are not implemented; detach them")
        Nn, dev = x.shape[0], x.device
        f = (int(flags) & ~_lib.F_SAVE) | (_lib.F_SAVE if need else 0) | F_EXTERNAL_FIELD
        dims = _dims(D, H, graph)
        field = torch.empty(Nn, D, dtype=torch.float32, device=dev)
        fws = torch.empty(lib.aether_dynfield_workspace_bytes(D, Nn, num_nodes), dtype=torch.uint8, device=dev)
        ws = torch.empty(lib.aether_workspace_bytes(ctypes.byref(dims), f), dtype=torch.uint8, device=dev)
        out = torch.empty(Nn, D, dtype=torch.float32, device=dev)
        # parameter table of the step: slots 0..39 the GNN, slot 40 the field, 41..46 unused
        table = list(gnn_p) + [field] * (_lib.NUM_PARAMS - NUM_GNN_PARAMS)
        with torch.cuda.device(dev):
            _lib.check(lib.aether_dynfield_forward(D, Nn, num_nodes, _lib.ptr_array(fld_p), _lib.ptr(x), _lib.ptr(vel),
                                                   _lib.ptr(charges), _lib.ptr(field), _lib.ptr(fws), fws.numel(), _stream()),
                       "aether_dynfield_forward")
            _lib.check(lib.aether_step_forward(ctypes.byref(dims), _lib.ptr_array(table), ctypes.byref(graph.c_struct()),
                                               _lib.ptr(x), _lib.ptr(vel), _lib.ptr(charges), _lib.ptr(edge_attr_orig),
                                               _lib.ptr(dropout_mask), _lib.ptr(out), _lib.ptr(ws), ws.numel(), f, _stream()),
                       "aether_step_forward")
        if need:
            ctx.save_for_backward(x, vel, charges, edge_attr_orig, field, *params)
            ctx.ws, ctx.fws, ctx.graph, ctx.cfg, ctx.dmask = ws, fws, graph, (D, H, f, num_nodes), dropout_mask
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_out):
        lib = _lib.load()
        if getattr(ctx, "ws", None) is None:
            raise RuntimeError("aether_b200: the step's saved workspace was already consumed (run forward again)")
        x, vel, charges, ea, field, *params = ctx.saved_tensors
        gnn_p, fld_p = params[:NUM_GNN_PARAMS], params[NUM_GNN_PARAMS:]
        D, H, f, num_nodes = ctx.cfg
        graph = ctx.graph
        dims = _dims(D, H, graph)
        Nn, dev = x.shape[0], x.device
        grad_out = _f32c(grad_out)
        offs, tot = [], 0
        for p in params:
            offs.append(tot)
            tot += (p.numel() + 3) // 4 * 4
        flat = torch.zeros(tot, dtype=torch.float32, device=dev)       # one buffer: the data-parallel all-reduce runs in place
        grads = [flat[o:o + p.numel()].view_as(p) for o, p in zip(offs, params)]
        gfield = torch.empty(Nn, D, dtype=torch.float32, device=dev)
        want_in = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        gx = torch.empty_like(x) if want_in else None
        gv = torch.empty_like(vel) if want_in else None
        scratch = torch.empty(lib.aether_backward_scratch_bytes(ctypes.byref(dims), f), dtype=torch.uint8, device=dev)
        fscr = torch.empty(lib.aether_dynfield_scratch_bytes(D, Nn, num_nodes), dtype=torch.uint8, device=dev)
        ptable = list(gnn_p) + [field] * (_lib.NUM_PARAMS - NUM_GNN_PARAMS)
        gtable = list(grads[:NUM_GNN_PARAMS]) + [gfield] * (_lib.NUM_PARAMS - NUM_GNN_PARAMS)
        with torch.cuda.device(dev):
            _lib.check(lib.aether_step_backward(ctypes.byref(dims), _lib.ptr_array(ptable), ctypes.byref(graph.c_struct()),
                                                _lib.ptr(x), _lib.ptr(vel), _lib.ptr(charges), _lib.ptr(ea),
                                                _lib.ptr(ctx.dmask), _lib.ptr(grad_out), _lib.ptr_array(gtable), _lib.ptr(gx),
                                                _lib.ptr(gv), _lib.ptr(ctx.ws), ctx.ws.numel(), _lib.ptr(scratch),
                                                scratch.numel(), f, _stream()), "aether_step_backward")
            _lib.check(lib.aether_dynfield_backward(D, Nn, num_nodes, _lib.ptr_array(fld_p), _lib.ptr(charges),
                                                    _lib.ptr(gfield), _lib.ptr_array(grads[NUM_GNN_PARAMS:]), _lib.ptr(gx),
                                                    _lib.ptr(gv), _lib.ptr(ctx.fws), ctx.fws.numel(), _lib.ptr(fscr),
                                                    fscr.numel(), _stream()), "aether_dynfield_backward")
        ctx.ws = ctx.fws = None
        return (gx if ctx.needs_input_grad[0] else None, gv if ctx.needs_input_grad[1] else None, None, None, None, None,
                None, None, None, None, None, *grads)


class DynamicFieldAether(nn.Module):
    """``DynamicFieldAether(input_size, hidden_size, dropout_prob, num_dims, device="cuda")``;
    ``forward(h, x, edges, vel, edge_attr_orig, charges, num_nodes)``.  ``precision`` as for ``Aether``."""

    def __init__(self, input_size, hidden_size, dropout_prob, num_dims, device="cuda", *, precision=None):
        super().__init__()
        if num_dims not in (2, 3) or input_size != 2 * num_dims:
            raise ValueError("aether_b200 supports num_dims in {2,3} with input_size == 2*num_dims")
        if hidden_size % 4 or hidden_size < 4:
            raise ValueError("aether_b200: hidden_size must be a positive multiple of 4 (rows of every [.,H] tensor are read "
                             "and written as 16-byte vectors); the reference has no such restriction")
        if hidden_size == 3 * num_dims:
            raise ValueError("aether_b200: hidden_size == 3*num_dims makes layer_1.res an nn.Identity in the reference "
                             "(locs.py:214-218: input_size == hidden_size), i.e. a different parameter set; not supported")
        _warn_bf16_fallback(precision or os.environ.get("AETHER_B200_PRECISION", "fp32"), hidden_size)
        self.gnn = GNN(input_size, hidden_size, dropout_prob, num_dims, additional_features=num_dims)
        self.localizer = AetherLocalizer(num_dims)
        self.globalizer = Globalizer(num_dims)
        self.num_dims = num_dims
        self.hidden_size = hidden_size
        self.dropout_prob = float(dropout_prob)
        self.field_net = LatentFieldNetwork(num_dims, 32, 16)
        self.precision = precision or os.environ.get("AETHER_B200_PRECISION", "fp32")
        if self.precision not in ("fp32", "bf16"):
            raise ValueError("precision must be 'fp32' or 'bf16'")
        self._graph_cache = None
        self.to(device)
        self.params = self.__str__()

    __str__ = Aether.__str__
    __getstate__ = Aether.__getstate__
    prepare_graph = Aether.prepare_graph
    flags = Aether.flags

    def param_list(self) -> List[torch.Tensor]:
        """40 GNN parameters then the 27 of ``field_net``, in state_dict order."""
        ps = [p for _, p in self.named_parameters()]
        if len(ps) != NUM_GNN_PARAMS + NUM_FIELD_PARAMS:
            raise RuntimeError("unexpected parameter count %d" % len(ps))
        return ps

    def forward(self, h, x, edges, vel, edge_attr_orig, charges, num_nodes):
        if x.device.type != "cuda":
            raise RuntimeError("aether_b200.DynamicFieldAether runs on CUDA devices only (no CPU fallback)")
        D, H = self.num_dims, self.hidden_size
        Nn = x.shape[0]
        num_nodes = int(num_nodes)
        if num_nodes <= 0 or Nn % num_nodes:
            raise ValueError("the number of nodes must be a multiple of num_nodes (the reference reshapes to "
                             "[-1, num_nodes, .], dynamic_field_aether.py:32)")
        graph = self.prepare_graph(edges, Nn)
        mask = dropout_masks(self.training, self.dropout_prob, Nn, H, x.device)
        params = self.param_list()
        if os.environ.get("AETHER_B200_VALIDATE", "0") == "1":
            check_charges(charges.reshape(-1))
        save = torch.is_grad_enabled() and (x.requires_grad or vel.requires_grad or any(p.requires_grad for p in params))
        return _DynStepFn.apply(x, vel, charges.reshape(-1), edge_attr_orig, graph, D, H, self.flags(), num_nodes, mask,
                                save, *params)
