"""aether_b200 -- B200 (sm_100a) implementation of Aether's state-to-state step.

Public surface (mirrors the reference's nn/state2state/aether.py):
    Aether              drop-in nn.Module (same constructor / forward signature / state_dict)
    LoCS                the field-free variant (nn/state2state/locs/locs.py), same kernels
    DynamicFieldAether  the attention-pooled / FiLM-modulated field variant (nn/state2state/dynamic_field_aether.py)
    NBodyDataset        loader of the reference's n-body .npy files (experiments/lorentz/dataset4newton.py)
    PreparedGraph       receiver-sorted CSR of an edge list (device)
    fc_edges, knn_edges, KnnPlan, caller_prep   the reference's edge builders / caller prep on device
    Rollout, rollout    multi-step rollouts (optionally CUDA-graph captured)
    radius_graph        cell-list radius / kNN graph for large single scenes
    FusedAdamW          torch.optim.AdamW semantics, all parameter tensors in one launch (csrc/optim.cu)
    PartitionedRollout  receiver-partitioned rollout of one large scene across GPUs (aether_b200.dist for the plan)
Everything computes through aether_b200/lib/libaether_b200.so; there is no CPU or ATen fallback.
"""
from .aether import Aether, AetherLocalizer, FieldNetwork, Globalizer, GNN, GNNLayer  # noqa: F401
from .locs import LoCS  # noqa: F401
from .dynamic_field import DynamicFieldAether  # noqa: F401
from .data import NBodyDataset  # noqa: F401
from .graph import KnnPlan, PreparedGraph, caller_prep, fc_edges, knn_edges, radius_graph  # noqa: F401
from .partition import PartitionedRollout  # noqa: F401
from .train import GraphedTrainStep  # noqa: F401
from .optim import FusedAdamW  # noqa: F401
from .rollout import Rollout, rollout  # noqa: F401

__all__ = ["Aether", "LoCS", "DynamicFieldAether", "NBodyDataset", "PreparedGraph", "KnnPlan", "fc_edges", "knn_edges", "caller_prep", "Rollout", "rollout", "radius_graph",
           "PartitionedRollout", "GraphedTrainStep", "FusedAdamW"]
