"""aether_b200.data.NBodyDataset against a loop restatement of the reference's loader
(experiments/lorentz/dataset4newton.py:34-94) on synthetic .npy files."""
import os

import numpy as np
import pytest
import torch

from aether_b200.data import NBodyDataset, file_suffix


def _write(root, suffix, S=6, T=49, D=3, N=5, seed=0):
    rng = np.random.default_rng(seed)
    arrs = {"loc": rng.standard_normal((S, T, D, N)).astype(np.float32), "vel": rng.standard_normal((S, T, D, N)).astype(np.float32),
            "edges": rng.standard_normal((S, N, N)).astype(np.float32), "charges": rng.choice([-1.0, 1.0], (S, N, 1)).astype(np.float32)}
    for k, v in arrs.items():
        np.save(os.path.join(root, "%s_%s.npy" % (k, suffix)), v)
    return arrs


def _reference_loops(a, max_samples):
    """dataset4newton.py:43-66, restated with its loops."""
    loc = torch.Tensor(a["loc"]).transpose(2, 3)[:max_samples]
    vel = torch.Tensor(a["vel"]).transpose(2, 3)[:max_samples]
    n = loc.size(2)
    ea, rows, cols = [], [], []
    for i in range(n):
        for j in range(n):
            if i != j:
                ea.append(a["edges"][:, i, j])
                rows.append(i)
                cols.append(j)
    ea = torch.Tensor(np.array(ea)).transpose(0, 1).unsqueeze(2)
    return loc, vel, ea, torch.Tensor(a["charges"][:max_samples]), [rows, cols]


@pytest.mark.parametrize("mode,partition", [("small", "train"), ("dynamic", "valid"), ("small_20body", "test"), ("dynamic_20body", "train")])
def test_file_names(mode, partition):
    want = {"small": "_charged5_initvel1small", "dynamic": "_dynamic5_initvel1dynamic", "small_20body": "_charged20_initvel1small_20body",
            "dynamic_20body": "_dynamic20_initvel1dynamic_20body"}[mode]
    assert file_suffix(partition, mode) == partition + want


def test_loader_matches_reference_semantics(tmp_path):
    a = _write(str(tmp_path), file_suffix("train", "small"))
    ds = NBodyDataset("train", max_samples=4, data_root=str(tmp_path), data_mode="small")
    loc, vel, ea, q, edges = _reference_loops(a, 4)
    assert len(ds) == 4 and ds.n_particles == 5 and ds.get_n_nodes() == loc.size(1)
    assert ds.edges == edges
    for i in range(len(ds)):
        got = ds[i]
        want = (loc[i][30], vel[i][30], ea[i], q[i], loc[i][40])
        for g, w in zip(got, want):
            assert torch.equal(g, w)
    # batched edge lists: graph b offset by n_nodes * b (:84-94)
    r0, c0 = torch.LongTensor(edges[0]), torch.LongTensor(edges[1])
    for B in (1, 3):
        s, r = ds.get_edges(B, 5)
        assert torch.equal(s, torch.cat([r0 + 5 * b for b in range(B)])) and torch.equal(r, torch.cat([c0 + 5 * b for b in range(B)]))
    # a DataLoader collates it like the reference's
    dl = torch.utils.data.DataLoader(ds, batch_size=2, shuffle=False)
    batch = next(iter(dl))
    assert batch[0].shape == (2, 5, 3) and batch[2].shape == (2, 20, 1) and batch[3].shape == (2, 5, 1)
