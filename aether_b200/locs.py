"""Drop-in ``LoCS`` (nn/state2state/locs/locs.py:102-139 of the reference): the same graph network on the base
Localizer's features, i.e. Aether without the neural field.

It runs through the SAME library entry points as ``Aether``.  The field-free step is the Aether step with a field that is
identically zero: the force columns of the edge features and of ``rel_feat`` vanish, so LoCS's narrower first-layer
weights are embedded into Aether-shaped ones (zero columns at the force positions) and the field network is given zero
weights (its output is then exactly 0).  The embedding is a differentiable torch index op, so autograd hands the
gradients of the real columns back to the LoCS parameters; parameter names and shapes are the reference's, so
``state_dict``s are interchangeable.
"""
from __future__ import annotations

import os

import numpy as np
import torch
from torch import nn

from . import _lib
from .aether import GNN, Globalizer
from .graph import PreparedGraph
from .ops import dropout_masks, step_apply


class Localizer(nn.Module):
    """Marker for the base local-frame transform (locs.py:20-99); the math lives in csrc/geom.cu."""

    def __init__(self, num_dims: int = 2):
        super().__init__()
        self.num_dims = num_dims
        self.num_orientations = num_dims * (num_dims - 1) // 2
        self.num_relative_features = 3 * num_dims + self.num_orientations


class LoCS(nn.Module):
    """``LoCS(input_size, hidden_size, dropout_prob, num_dims, device="cuda")``;
    ``forward(h, x, edges, vel, edge_attr_orig) -> [Nn, num_dims]``."""

    def __init__(self, input_size, hidden_size, dropout_prob, num_dims, device="cuda", *, precision=None):
        super().__init__()
        if num_dims not in (2, 3) or input_size != 2 * num_dims:
            raise ValueError("aether_b200 supports num_dims in {2,3} with input_size == 2*num_dims")
        if hidden_size % 4 or hidden_size < 4 or hidden_size in (2 * num_dims, 3 * num_dims):
            raise ValueError("hidden_size must be a multiple of 4")
        self.gnn = GNN(input_size, hidden_size, dropout_prob, num_dims, additional_features=0, out_size=0)
        self.localizer = Localizer(num_dims)
        self.globalizer = Globalizer(num_dims)
        self.num_dims, self.hidden_size, self.dropout_prob = num_dims, hidden_size, float(dropout_prob)
        self.precision = precision or os.environ.get("AETHER_B200_PRECISION", "fp32")
        D, O = num_dims, num_dims * (num_dims - 1) // 2
        # LoCS feature column -> Aether feature column (aether.py:81-99 inserts R^T f_j after R^T v_j and R^T f_i last)
        cols = list(range(3 * D + O)) + list(range(4 * D + O, 6 * D + O)) + [7 * D + O, 7 * D + O + 1]
        self.register_buffer("_feat_cols", torch.tensor(cols, dtype=torch.int64), persistent=False)
        self._graph_cache = None
        self.to(device)
        self.params = self.__str__()

    def __str__(self):
        n = sum(int(np.prod(p.size())) for p in self.parameters() if p.requires_grad)
        print("Network Size", n)
        return str(n)

    def flags(self) -> int:
        return _lib.F_BF16 if self.precision == "bf16" else 0

    def _aether_params(self):
        """The 47 Aether-ordered tensors: LoCS parameters, first-layer weights widened with zero columns, zero field net."""
        D, H = self.num_dims, self.hidden_size
        O = D * (D - 1) // 2
        dev = self._feat_cols.device
        out = []
        for name, p in self.named_parameters():
            if name == "gnn.layer_1.message_fn.0.weight":
                p = torch.zeros(H, 7 * D + O + 2, device=dev, dtype=p.dtype).index_copy(1, self._feat_cols, p)
            elif name == "gnn.layer_1.res.weight":
                p = torch.cat([p, torch.zeros(H, D, device=dev, dtype=p.dtype)], 1)
            out.append(p)
        z = lambda *s: torch.zeros(*s, device=dev)
        out += [z(32, 2 * D + 16), z(32), z(32, 32), z(32), z(D, 32), z(D), z(3, 16)]      # field_net.*: output == 0
        if len(out) != _lib.NUM_PARAMS:
            raise RuntimeError("unexpected parameter count %d" % len(out))
        return out

    def forward(self, h, x, edges, vel, edge_attr_orig):
        if x.device.type != "cuda":
            raise RuntimeError("aether_b200.LoCS runs on CUDA devices only (no CPU fallback)")
        Nn = x.shape[0]
        graph = edges if isinstance(edges, PreparedGraph) else PreparedGraph(edges[0], edges[1], Nn)
        mask = dropout_masks(self.training, self.dropout_prob, Nn, self.hidden_size, x.device)
        charges = torch.zeros(Nn, device=x.device)
        return step_apply(x, vel, charges, edge_attr_orig, graph, self.num_dims, self.hidden_size, self.flags(), mask,
                          self._aether_params())
