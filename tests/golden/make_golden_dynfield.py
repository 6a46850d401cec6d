"""Golden vectors of DynamicFieldAether from the UNMODIFIED reference (nn/state2state/dynamic_field_aether.py).

Run (only where /root/reference exists):
    python tests/golden/make_golden_dynfield.py

Two dependencies of the reference are neither vendored nor installed here and get stand-ins: ``torch_scatter.scatter`` (as
in make_golden.py) and ``torch_geometric.nn.aggr.AttentionalAggregation`` (graph_pool.py:3).  The latter restates PyG's
published behaviour -- constructor stores ``gate_nn`` / ``nn`` and calls ``reset_parameters()`` on both; forward computes
``softmax(gate_nn(x), index)`` per segment (``exp(g - max) / (sum + 1e-16)``) and sums ``gate * nn(x)`` per segment -- so the
fixtures are pinned to the reference's own code everywhere except inside that aggregation ("parity unpinned" at that
boundary, stated in DESIGN.md).  Only inputs and outputs are recorded.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402


def _install_pyg_shim():
    class AttentionalAggregation(torch.nn.Module):
        def __init__(self, gate_nn, nn=None):
            super().__init__()
            self.gate_nn = gate_nn
            self.nn = nn
            self.reset_parameters()

        def reset_parameters(self):
            def reset(v):
                if hasattr(v, "reset_parameters"):
                    v.reset_parameters()
                else:
                    for c in v.children() if hasattr(v, "children") else []:
                        reset(c)
            reset(self.gate_nn)
            reset(self.nn)

        def forward(self, x, index=None, ptr=None, dim_size=None, dim=-2):
            gate = self.gate_nn(x)
            x = self.nn(x) if self.nn is not None else x
            n = int(index.max()) + 1 if dim_size is None else dim_size
            idx = index.view(-1, 1)
            mx = torch.full((n, gate.shape[1]), float("-inf"), dtype=gate.dtype).scatter_reduce(
                0, idx.expand_as(gate), gate, reduce="amax", include_self=True)
            e = (gate - mx[index]).exp()
            den = torch.zeros(n, gate.shape[1], dtype=gate.dtype).scatter_add_(0, idx.expand_as(e), e) + 1e-16
            a = e / den[index]
            return torch.zeros(n, x.shape[1], dtype=x.dtype).scatter_add_(0, idx.expand_as(x), a * x)

    tg = types.ModuleType("torch_geometric")
    tgn = types.ModuleType("torch_geometric.nn")
    tga = types.ModuleType("torch_geometric.nn.aggr")
    tga.AttentionalAggregation = AttentionalAggregation
    tg.nn = tgn
    tgn.aggr = tga
    sys.modules.update({"torch_geometric": tg, "torch_geometric.nn": tgn, "torch_geometric.nn.aggr": tga})


def run_case(name, D, H, B, N, seed=1):
    from nn.state2state.dynamic_field_aether import DynamicFieldAether
    torch.manual_seed(seed)
    model = DynamicFieldAether(input_size=2 * D, hidden_size=H, dropout_prob=0.0, num_dims=D, device="cpu")
    model.train()
    x, vel, q = mg.synth(B, N, D, seed)
    send, recv = mg.fc_edge_list(N, B)
    out = {"D": D, "H": H, "B": B, "N": N, "seed": seed, "x": x, "vel": vel, "charges": q, "send": send, "recv": recv}
    for k, v in model.state_dict().items():
        out["param/" + k] = v.clone()
    h = torch.sqrt(torch.sum(vel ** 2, dim=1)).unsqueeze(1)
    dist = torch.sqrt(torch.sum((x[send] - x[recv]) ** 2, 1)).unsqueeze(1)
    ea = torch.cat([(q[send] * q[recv]).reshape(-1, 1), dist], 1)
    out["edge_attr_orig"] = ea
    xg = x.clone().requires_grad_(True)
    vg = vel.clone().requires_grad_(True)
    out["field"] = model.field_net(torch.cat([xg, vg], -1), q, N)
    y = model(h, xg, [send, recv], vg, ea, q, N)
    out["out"] = y
    g = torch.Generator().manual_seed(seed + 100)
    target = x + 0.1 * torch.randn(x.shape, generator=g)
    loss = torch.nn.functional.mse_loss(y, target)
    loss.backward()
    out.update(target=target, loss=loss)
    out["grad/x"] = xg.grad
    out["grad/vel"] = vg.grad
    for k, p in model.named_parameters():
        out["grad/" + k] = p.grad if p.grad is not None else torch.zeros_like(p)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **mg._np(out))
    print("wrote", name, float(loss))


def main():
    mg._install_shims()
    _install_pyg_shim()
    run_case("dynfield2d_h32", 2, 32, 3, 5)
    run_case("dynfield3d_h64", 3, 64, 2, 7, seed=2)


if __name__ == "__main__":
    main()
