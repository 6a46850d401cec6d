"""Drop-in ``Aether`` module for the reference's training and evaluation scripts.

Same constructor and forward signature as nn/state2state/aether.py:142-186 of the reference,
same sub-module / parameter names, shapes and default initialisation (so a seed reproduces the
reference's initial weights and ``state_dict``s are interchangeable), but ``forward`` runs the
whole step through libaether_b200.so (hand-written sm_100a CUDA) -- there is no ATen or CPU
fallback.  The sub-modules below only hold parameters; they are never called.
"""
from __future__ import annotations

import os
from typing import List

import numpy as np
import torch
from torch import nn

from . import _lib
from .graph import PreparedGraph
from .ops import dropout_masks, step_apply


_warned_bf16 = set()


def _warn_bf16_fallback(precision, hidden_size):
    """bf16 requested at a hidden size the tcgen05 kernels do not cover: the step silently ran on the fp32 CUDA-core kernels
    (about 18x slower at H = 256-class sizes); say so once per size."""
    if precision == "bf16" and hidden_size not in (64, 128, 256) and hidden_size not in _warned_bf16:
        import warnings
        _warned_bf16.add(hidden_size)
        warnings.warn("aether_b200: precision='bf16' is implemented for hidden_size in {64, 128, 256}; hidden_size=%d runs on "
                      "the fp32 CUDA-core kernels (much slower, fp32 accuracy)" % hidden_size, RuntimeWarning, stacklevel=3)


def check_charges(charges):
    """The reference indexes ``nn.Embedding(3, .)`` with ``(charges + 1).long()`` (aether.py:123-125) and raises on a class
    outside 0..2; the kernels clamp the index for memory safety.  With AETHER_B200_VALIDATE=1 the modules check it here (a
    host read: synchronises)."""
    idx = (charges + 1).long()
    if idx.numel() and (int(idx.min()) < 0 or int(idx.max()) > 2):
        raise IndexError("aether_b200: charge class (charges + 1).long() outside 0..2")


class GNNLayer(nn.Module):
    """Parameter holder of one message-passing layer (nn/state2state/locs/locs.py:197-225)."""

    def __init__(self, input_size, hidden_size, only_edge_attr=False, num_edge_features=0):
        super().__init__()
        self.only_edge_attr = only_edge_attr
        num_edge_features = num_edge_features if only_edge_attr else 3 * hidden_size
        self.message_fn = nn.Sequential(nn.Linear(num_edge_features, hidden_size), nn.SiLU(),
                                        nn.Linear(hidden_size, hidden_size), nn.SiLU())
        self.res = nn.Linear(input_size, hidden_size) if input_size != hidden_size else nn.Identity()
        self.update_fn = nn.Sequential(nn.Linear(hidden_size, 2 * hidden_size), nn.SiLU(),
                                       nn.Linear(2 * hidden_size, hidden_size))


class GNN(nn.Module):
    """Parameter holder of the 4-layer network + output MLP (locs.py:142-181)."""

    def __init__(self, input_size, hidden_size, dropout_prob, num_dims, additional_features=0, out_size=0):
        super().__init__()
        out_size = input_size // 2 if out_size == 0 else out_size
        num_orientations = num_dims * (num_dims - 1) // 2
        self.num_relative_features = input_size + num_dims + num_orientations
        self.out_mlp = nn.Sequential(nn.Linear(hidden_size, hidden_size), nn.SiLU(), nn.Dropout(p=dropout_prob),
                                     nn.Linear(hidden_size, hidden_size), nn.SiLU(), nn.Dropout(p=dropout_prob),
                                     nn.Linear(hidden_size, out_size))
        self.layer_1 = GNNLayer(input_size + additional_features, hidden_size, only_edge_attr=True,
                                num_edge_features=self.num_relative_features + input_size + 2 + 2 * additional_features)
        self.layer_2 = GNNLayer(hidden_size, hidden_size)
        self.layer_3 = GNNLayer(hidden_size, hidden_size)
        self.layer_4 = GNNLayer(hidden_size, hidden_size)


class AetherLocalizer(nn.Module):
    """Marker for the local-frame transform (aether.py:14-100); the math lives in csrc/geom.cu."""

    def __init__(self, num_dims):
        super().__init__()
        self.num_dims = num_dims
        self.num_orientations = num_dims * (num_dims - 1) // 2


class Globalizer(nn.Module):
    """Marker for the rotation back to the global frame (nn/utils/local_to_global.py:7-13)."""

    def __init__(self, num_dims):
        super().__init__()
        self.num_dims = num_dims


class FieldNetwork(nn.Module):
    """Parameter holder of the neural field (aether.py:108-134)."""

    def __init__(self, num_dims, hidden_size, class_embedding_dim):
        super().__init__()
        self.num_dims = num_dims
        self.net = nn.Sequential(nn.Linear(2 * num_dims + class_embedding_dim, hidden_size), nn.SiLU(),
                                 nn.Linear(hidden_size, hidden_size), nn.SiLU(), nn.Linear(hidden_size, num_dims))
        self.class_embedding = nn.Embedding(3, class_embedding_dim)

    @staticmethod
    def charge_to_index(charges):
        return (charges + 1).long()


class Aether(nn.Module):
    """``Aether(input_size, hidden_size, dropout_prob, num_dims, device="cuda")``.

    ``forward(h, x, edges, vel, edge_attr_orig, charges) -> [Nn, num_dims]``; ``h`` is unused, as
    in the reference.  Optional keyword ``precision``: "fp32" (default, CUDA-core GEMMs, matches the
    reference to ~1e-6) or "bf16" (tcgen05 edge-MLP GEMMs, fp32 accumulate).  The environment
    variable AETHER_B200_PRECISION overrides the default.
    """

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
        self.field_net = FieldNetwork(num_dims, 32, 16)
        self.precision = precision or os.environ.get("AETHER_B200_PRECISION", "fp32")
        if self.precision not in ("fp32", "bf16"):
            raise ValueError("precision must be 'fp32' or 'bf16'")
        self._graph_cache = None
        self.to(device)
        self.params = self.__str__()

    def __str__(self):
        n = sum(int(np.prod(p.size())) for p in self.parameters() if p.requires_grad)
        print("Network Size", n)
        return str(n)

    def __getstate__(self):
        d = self.__dict__.copy()
        d["_graph_cache"] = None          # device CSR is rebuilt on first use after unpickling
        return d

    # -- helpers -----------------------------------------------------------------------
    def param_list(self) -> List[torch.Tensor]:
        """The 47 parameters in state_dict order (the order the C ABI expects)."""
        ps = [p for _, p in self.named_parameters()]
        if len(ps) != _lib.NUM_PARAMS:
            raise RuntimeError("unexpected parameter count %d" % len(ps))
        return ps

    def prepare_graph(self, edges, n_nodes: int) -> PreparedGraph:
        """Receiver-sorted CSR of ``edges``; reused when the same tensors are passed again."""
        if isinstance(edges, PreparedGraph):
            return edges
        send, recv = edges[0], edges[1]
        key = (id(send), id(recv), send._version, recv._version, send.data_ptr(), recv.data_ptr(), int(send.numel()),
               int(n_nodes))
        c = self._graph_cache
        if c is not None and c[0] == key and c[1]() is send and c[2]() is recv:
            return c[3]
        import weakref
        g = PreparedGraph(send, recv, n_nodes)
        try:
            self._graph_cache = (key, weakref.ref(send), weakref.ref(recv), g)
        except TypeError:
            self._graph_cache = None
        return g

    def flags(self) -> int:
        return _lib.F_BF16 if self.precision == "bf16" else 0

    def forward(self, h, x, edges, vel, edge_attr_orig, charges):
        if x.device.type != "cuda":
            raise RuntimeError("aether_b200.Aether runs on CUDA devices only (no CPU fallback)")
        D, H = self.num_dims, self.hidden_size
        Nn = x.shape[0]
        graph = self.prepare_graph(edges, Nn)
        mask = dropout_masks(self.training, self.dropout_prob, Nn, H, x.device)
        charges = charges.reshape(-1)
        if os.environ.get("AETHER_B200_VALIDATE", "0") == "1":
            check_charges(charges)
        if mask is None and torch.is_grad_enabled() and (getattr(self, "use_registered_ops", False) or
                                                          os.environ.get("AETHER_B200_TORCH_OPS", "0") == "1"):
            from . import torch_ops          # the same step as registered torch.library operators (step_fwd / step_bwd)
            return torch_ops.train_step_from_model(self, graph, x, vel, charges, edge_attr_orig)
        return step_apply(x, vel, charges, edge_attr_orig, graph, D, H, self.flags(), mask, self.param_list())
