"""Synthetic workloads of the shapes BASELINE.json names (inputs only; no model code).

Distributions follow the reference simulators: positions N(0, (N/5)^(2/3))
(experiments/lorentz/dataset/synthetic_sim.py:469), velocities of norm 0.5 in a random direction
(:553-555), charges uniform in {-1, +1} (:476).  bench.py's product arm draws its inputs here so that it
imports nothing from oracle/; tests/test_bench_cpu.py checks that the oracle's generator draws the same numbers.
"""
from __future__ import annotations

from typing import Sequence, Union

import torch


def synth_state(n_graphs: int, sizes: Union[int, Sequence[int]], num_dims: int, seed: int = 1):
    """(x [Nn,D], v [Nn,D], q [Nn,1]) on the CPU, graphs concatenated."""
    g = torch.Generator().manual_seed(seed)
    if isinstance(sizes, int):
        sizes = [sizes] * n_graphs
    xs, vs = [], []
    for n in sizes:
        xs.append(torch.randn(n, num_dims, generator=g) * (n / 5.0) ** (1.0 / 3.0))
        d = torch.randn(n, num_dims, generator=g)
        vs.append(0.5 * d / torch.norm(d, dim=-1, keepdim=True))
    x, v = torch.cat(xs), torch.cat(vs)
    q = torch.randint(0, 2, (x.shape[0], 1), generator=g).float() * 2 - 1
    return x, v, q


def ind_style_scenes(sizes: Sequence[int], num_dims: int, seed: int = 1):
    """inD-style ragged scenes (config 3): positions in the unit square, speeds U(-0.05, 0.05) per axis
    (speed-normalised tracks), charges +-1."""
    g = torch.Generator().manual_seed(seed)
    n = sum(sizes)
    x = torch.rand(n, num_dims, generator=g)
    v = (torch.rand(n, num_dims, generator=g) - 0.5) * 0.1
    q = torch.randint(0, 2, (n, 1), generator=g).float() * 2 - 1
    return x, v, q


def unit_density_scene(n: int, seed: int = 1):
    """One 3-D scene of n particles at unit density in a cube, speeds 0.05, charges +-1 (config 5)."""
    g = torch.Generator().manual_seed(seed)
    L = float(n) ** (1.0 / 3.0)
    x = torch.rand(n, 3, generator=g) * L
    v = torch.randn(n, 3, generator=g)
    v = 0.05 * v / v.norm(dim=-1, keepdim=True)
    q = torch.randint(0, 2, (n, 1), generator=g).float() * 2 - 1
    return x, v, q
