"""The reference's train loop (experiments/lorentz/main.py:197-312, branches 'aether' :240-245 and 'film_aether' :246-251)
restated step for step -- DataLoader over the n-body .npy loader, edge list from the dataset per batch and ``.to(device)``,
``nodes`` / ``edge_attr`` prepared as the script does, ``AdamW(lr=5e-4, weight_decay=1e-12)`` (:164), ``loss.item()`` per batch,
``torch.save(model)`` at the end -- driving the drop-in modules through the aliases ``aether_b200.dropin`` installs, for three
epochs on the GPU.  The same loop runs on the CPU with the oracle; per-batch losses must agree to 1e-4."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _files(root, S=24, T=49, D=3, N=5, seed=0):
    from aether_b200.data import file_suffix
    rng = np.random.default_rng(seed)
    loc = np.cumsum(0.02 * rng.standard_normal((S, T, D, N)), axis=1).astype(np.float32)
    vel = (0.3 * rng.standard_normal((S, T, D, N))).astype(np.float32)
    q = rng.choice([-1.0, 1.0], (S, N, 1)).astype(np.float32)
    inter = (q @ q.transpose(0, 2, 1)).astype(np.float32)
    for k, v in {"loc": loc, "vel": vel, "edges": inter, "charges": q}.items():
        np.save(os.path.join(root, "%s_%s.npy" % (k, file_suffix("train", "small"))), v)


def _loop(model_fn, loader, device, epochs, film):
    """main.py:197-312 for --model aether / film_aether; returns the per-batch losses."""
    model, params = model_fn()
    optimizer = torch.optim.AdamW(params, lr=5e-4, weight_decay=1e-12)
    loss_mse = torch.nn.MSELoss()
    losses = []
    for _ in range(epochs):
        for data in loader:
            batch_size, n_nodes, _ = data[0].size()
            data = [d.to(device) for d in data]
            data = [d.flatten(0, 1) for d in data]
            loc, vel, edge_attr, charges, loc_end = data
            edges = loader.dataset.get_edges(batch_size, n_nodes)
            edges = [edges[0].to(device), edges[1].to(device)]
            optimizer.zero_grad()
            nodes = torch.sqrt(torch.sum(vel ** 2, dim=1)).unsqueeze(1).detach()
            rows, cols = edges
            loc_dist = torch.sqrt(torch.sum((loc[rows] - loc[cols]) ** 2, 1)).unsqueeze(1)
            edge_attr = torch.cat([edge_attr, loc_dist], 1).detach()
            if film:
                loc_pred = model(nodes, loc.detach(), edges, vel, edge_attr, charges, n_nodes)
            else:
                loc_pred = model(nodes, loc.detach(), edges, vel, edge_attr, charges)
            loss = loss_mse(loc_pred, loc_end)
            loss.backward()
            optimizer.step()
            losses.append(loss.item())
    return losses, model


@pytest.mark.parametrize("film", [False, True])
def test_reference_train_loop_on_the_dropin(tmp_path, film):
    import aether_b200.dropin
    aether_b200.dropin.install()
    from nn.state2state.aether import Aether                                  # the names the reference script imports
    from nn.state2state.dynamic_field_aether import DynamicFieldAether
    from aether_b200.data import NBodyDataset
    from oracle import aether_oracle as O
    _files(str(tmp_path))
    ds = NBodyDataset("train", data_root=str(tmp_path), data_mode="small")
    loader = torch.utils.data.DataLoader(ds, batch_size=8, shuffle=False, drop_last=True)
    D, H = 3, 64

    def gpu_model():
        torch.manual_seed(11)
        m = (DynamicFieldAether if film else Aether)(input_size=2 * D, hidden_size=H, dropout_prob=0.0, num_dims=D, device="cuda")
        return m, m.parameters()

    got, model = _loop(gpu_model, loader, torch.device("cuda"), 3, film)

    # the same loop on the CPU with the oracle (same initial weights: the holders reproduce the reference's init order)
    p0 = {k: v.detach().cpu().clone() for k, v in gpu_model()[0].state_dict().items()}

    class OracleModel:
        def __init__(self):
            self.p = {k: v.clone().requires_grad_(True) for k, v in p0.items()}

        def __call__(self, nodes, x, edges, vel, edge_attr, charges, n_nodes=None):
            if film:
                return O.dynamic_field_step(self.p, x, (edges[0], edges[1]), vel, edge_attr, charges, n_nodes)
            return O.aether_step(self.p, x, (edges[0], edges[1]), vel, edge_attr, charges)

    def cpu_model():
        m = OracleModel()
        return m, list(m.p.values())

    want, _ = _loop(cpu_model, loader, torch.device("cpu"), 3, film)
    assert len(got) == len(want) == 9
    for a, b in zip(got, want):
        assert abs(a - b) <= 1e-4 * max(abs(b), 1e-3), (got, want)
    assert got[-1] < got[0]                                                  # it trains
    path = os.path.join(str(tmp_path), "model.pt")
    torch.save(model, path)                                                  # main.py saves the whole module
    again = torch.load(path, weights_only=False)
    assert all(torch.equal(a, b) for a, b in zip(again.state_dict().values(), model.state_dict().values()))
