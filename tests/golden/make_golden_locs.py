"""Golden vectors for the LoCS variant (nn/state2state/locs/locs.py:102-139), recorded from the UNMODIFIED
reference exactly like make_golden.py does for Aether (same torch_scatter stand-in, nothing copied).

    python tests/golden/make_golden_locs.py          # only where /root/reference exists
"""
import os

import numpy as np
import torch

import make_golden as G

HERE = os.path.dirname(os.path.abspath(__file__))


def run_case(name, D, H, x, vel, q, send, recv, seed=1, rollout_steps=3):
    from nn.state2state.locs.locs import LoCS
    torch.manual_seed(seed)
    model = LoCS(input_size=2 * D, hidden_size=H, dropout_prob=0.0, num_dims=D, device="cpu")
    model.train()
    out = {"D": D, "H": H, "seed": seed, "x": x, "vel": vel, "charges": q, "send": send, "recv": recv}
    for k, v in model.state_dict().items():
        out["param/" + k] = v.clone()

    def prep(x_, v_):
        h = torch.sqrt(torch.sum(v_ ** 2, dim=1)).unsqueeze(1)
        dist = torch.sqrt(torch.sum((x_[send] - x_[recv]) ** 2, 1)).unsqueeze(1)
        return h, torch.cat([(q[send] * q[recv]).reshape(-1, 1), dist], 1)

    h, ea = prep(x, vel)
    out["edge_attr_orig"] = ea
    xg = x.clone().requires_grad_(True)
    vg = vel.clone().requires_grad_(True)
    rel_feat, R, edge_attr = model.localizer(torch.cat([xg, vg], -1), [send, recv])
    out.update(rel_feat=rel_feat, R=R, edge_attr=torch.cat([edge_attr, ea], -1))
    y = model(h, xg, [send, recv], vg, ea)
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
    xs, xc, vc = [], x, vel
    with torch.no_grad():
        for _ in range(rollout_steps):
            hh, eaa = prep(xc, vc)
            xn = model(hh, xc, [send, recv], vc, eaa)
            vc = xn - xc
            xc = xn
            xs.append(xc)
    out["rollout"] = torch.stack(xs)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **G._np(out))
    print("wrote", name, float(loss))


def main():
    G._install_shims()
    for name, D, H, B, N in [("locs2d_h32", 2, 32, 3, 5), ("locs3d_h64", 3, 64, 2, 6)]:
        x, v, q = G.synth(B, N, D, seed=2)
        s, r = G.fc_edge_list(N, B)
        run_case(name, D, H, x, v, q, s, r)


if __name__ == "__main__":
    main()
