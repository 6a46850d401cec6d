"""Generate golden vectors by running the UNMODIFIED reference in the build container.

Run (only where /root/reference exists):
    python tests/golden/make_golden.py

It imports ``nn.state2state.aether.Aether`` from /root/reference with a small
``torch_scatter`` stand-in injected into ``sys.modules`` (torch_scatter is not
installed and not vendored; the stand-in implements scatter sum/mean with
``scatter_add_`` + count clamp, i.e. the reference's own
``unsorted_segment_mean``, nn/state2state/gnn/gcl.py:203-210).  No reference
source is copied; only inputs and outputs are recorded, as ``.npz`` files next
to this script.  The GPU box has no /root/reference: tests read the .npz only.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("AETHER_REFERENCE", "/root/reference")


def _install_shims():
    def scatter(src, index, dim=-1, out=None, dim_size=None, reduce="sum"):
        if dim < 0:
            dim += src.dim()
        if dim_size is None:
            dim_size = int(index.max()) + 1 if index.numel() else 0
        shape = list(src.shape)
        shape[dim] = dim_size
        view = [1] * src.dim()
        view[dim] = -1
        res = torch.zeros(shape, dtype=src.dtype).scatter_add_(dim, index.view(view).expand_as(src), src)
        if reduce in ("sum", "add"):
            return res
        cnt = torch.zeros(dim_size, dtype=src.dtype).scatter_add_(
            0, index, torch.ones_like(index, dtype=src.dtype)).clamp_(min=1)
        return res / cnt.view(view)

    m = types.ModuleType("torch_scatter")
    m.scatter = scatter
    sys.modules["torch_scatter"] = m
    try:
        import PIL  # noqa: F401
    except Exception:  # single_ind_data imports PIL.Image at module level only
        pil = types.ModuleType("PIL")
        pil.Image = types.ModuleType("PIL.Image")
        sys.modules["PIL"] = pil
        sys.modules["PIL.Image"] = pil.Image
    sys.path.insert(0, REF)


def _np(d):
    return {k: (v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in d.items()}


def fc_edge_list(n, b):
    """Edge list through the reference's own builders (dataset4newton.py:54-61, 84-94)."""
    from experiments.lorentz.dataset4newton import NBodyDataset
    ds = object.__new__(NBodyDataset)
    ds.max_samples = 4
    loc = np.zeros((2, 3, 2, n), np.float32)
    vel = np.zeros((2, 3, 2, n), np.float32)
    e = np.zeros((2, n, n), np.float32)
    q = np.zeros((2, n, 1), np.float32)
    _, _, _, edges, _ = ds.preprocess(loc, vel, e, q)
    ds.edges = edges
    s, r = ds.get_edges(b, n)
    return s, r


def run_case(name, D, H, x, vel, q, send, recv, seed=1, with_grad=True, rollout_steps=0):
    from nn.state2state.aether import Aether
    torch.manual_seed(seed)
    model = Aether(input_size=2 * D, hidden_size=H, dropout_prob=0.0, num_dims=D, device="cpu")
    model.train()
    out = {"D": D, "H": H, "seed": seed, "x": x, "vel": vel, "charges": q, "send": send, "recv": recv}
    for k, v in model.state_dict().items():
        out["param/" + k] = v.clone()

    def prep(x_, v_):
        h = torch.sqrt(torch.sum(v_ ** 2, dim=1)).unsqueeze(1)
        dist = torch.sqrt(torch.sum((x_[send] - x_[recv]) ** 2, 1)).unsqueeze(1)
        qq = (q[send] * q[recv]).reshape(-1, 1)
        return h, torch.cat([qq, dist], 1)

    h, ea = prep(x, vel)
    out["edge_attr_orig"] = ea

    # Stage outputs, captured by calling the reference's own sub-modules in order.
    xg = x.clone().requires_grad_(with_grad)
    vg = vel.clone().requires_grad_(with_grad)
    inputs = torch.cat([xg, vg], -1)
    f = model.field_net(inputs, q)
    rel_feat, R, edge_attr = model.localizer(torch.cat([inputs, f], -1), [send, recv])
    e = torch.cat([edge_attr, ea], -1)
    out.update(field=f, rel_feat=rel_feat, R=R, edge_attr=e)
    hx = rel_feat
    for l in range(1, 5):
        hx, e = getattr(model.gnn, "layer_%d" % l)(hx, e, [send, recv])
        out["x%d" % l] = hx
        out["e%d" % l] = e
    pred = model.gnn.out_mlp(hx)
    out["pred"] = pred

    y = model(h, xg, [send, recv], vg, ea, q)
    out["out"] = y
    if with_grad:
        g = torch.Generator().manual_seed(seed + 100)
        target = x + 0.1 * torch.randn(x.shape, generator=g)
        loss = torch.nn.functional.mse_loss(y, target)
        loss.backward()
        out["target"] = target
        out["loss"] = loss
        out["grad/x"] = xg.grad
        out["grad/vel"] = vg.grad
        for k, p in model.named_parameters():
            out["grad/" + k] = p.grad if p.grad is not None else torch.zeros_like(p)
    if rollout_steps:
        xs, xc, vc = [], x, vel
        with torch.no_grad():
            for _ in range(rollout_steps):
                hh, eaa = prep(xc, vc)
                xn = model(hh, xc, [send, recv], vc, eaa, q)
                vc = xn - xc
                xc = xn
                xs.append(xc)
        out["rollout"] = torch.stack(xs)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **_np(out))
    print("wrote", name, "loss" if with_grad else "", float(out.get("loss", 0)))


def synth(B, N, D, seed):
    g = torch.Generator().manual_seed(seed)
    sigma = (N / 5.0) ** (1.0 / 3.0)
    x = torch.randn(B * N, D, generator=g) * sigma
    v = torch.randn(B * N, D, generator=g)
    v = 0.5 * v / torch.norm(v, dim=-1, keepdim=True)
    q = torch.randint(0, 2, (B * N, 1), generator=g).float() * 2 - 1
    return x, v, q


def main():
    _install_shims()
    from nn.state2state.aether import AetherLocalizer

    # ---- integer path -----------------------------------------------------
    ints = {}
    for n, b in [(5, 1), (5, 3), (4, 2), (2, 2), (20, 2)]:
        s, r = fc_edge_list(n, b)
        ints["fc_%d_%d/send" % (n, b)] = s.numpy()
        ints["fc_%d_%d/recv" % (n, b)] = r.numpy()
    from experiments.ind.single_ind_data import get_knn_graph_info
    g = torch.Generator().manual_seed(7)
    for n in [2, 7, 11, 25, 40]:
        pos = torch.rand(n, 4, generator=g)
        s, r = get_knn_graph_info(pos, torch.ones(n), n, use_edge2node=False)
        ints["knn_%d/pos" % n] = pos.numpy()
        ints["knn_%d/send" % n] = s.numpy()
        ints["knn_%d/recv" % n] = r.numpy()
    np.savez_compressed(os.path.join(HERE, "integer_path.npz"), **ints)
    print("wrote integer_path")

    # ---- weight-free localizer KATs incl. edge cases ----------------------
    kat = {}
    send = torch.tensor([0, 0, 1, 1, 2, 2])
    recv = torch.tensor([1, 2, 0, 2, 0, 1])
    nodes2 = torch.tensor([[0, 0, 1, 0, .5, -.25], [1, 2, 0, 2, -1, 0], [-3, .5, -1, -1, .25, .75]])
    nodes3 = torch.tensor([[0, 0, 0, 1, 0, 0, .5, -.25, .125], [1, 2, -1, 0, 2, 1, -1, 0, .5],
                           [-3, .5, 2, -1, -1, -.5, .25, .75, -.5]])
    # edge cases: zero velocity, +-z heading, coincident particles, anti-parallel headings
    edge2 = torch.tensor([[0, 0, 0, 0, 1, 1], [0, 0, 1, 0, 0, 1], [2, 1, -1, 0, .3, .2]])
    edge3 = torch.tensor([[0, 0, 0, 0, 0, 0, 1, 1, 1], [0, 0, 0, 0, 0, 2, 0, 1, 0],
                          [1, 1, 1, 0, 0, -2, .3, .2, .1]])
    anti3 = torch.tensor([[0, 0, 0, 1, 1, 0, 1, 0, 0], [1, 0, 0, -1, -1, 0, 0, 1, 0],
                          [0, 1, 0, 0, 0, 0, 0, 0, 1]])
    for nm, D, nodes in [("kat2d", 2, nodes2), ("kat3d", 3, nodes3), ("edge2d", 2, edge2),
                         ("edge3d", 3, edge3), ("anti3d", 3, anti3)]:
        rel, R, ea = AetherLocalizer(D)(nodes.float(), [send, recv])
        kat[nm + "/nodes"] = nodes.float().numpy()
        kat[nm + "/rel_feat"] = rel.numpy()
        kat[nm + "/R"] = R.numpy()
        kat[nm + "/edge_attr"] = ea.numpy()
    kat["send"] = send.numpy()
    kat["recv"] = recv.numpy()
    np.savez_compressed(os.path.join(HERE, "localizer_kat.npz"), **kat)
    print("wrote localizer_kat")

    # ---- full-step cases ---------------------------------------------------
    for name, D, H, B, N in [("step2d_h32", 2, 32, 3, 5), ("step3d_h32", 3, 32, 3, 5),
                             ("step3d_h64", 3, 64, 2, 5), ("step2d_h64_n12", 2, 64, 2, 12)]:
        x, v, q = synth(B, N, D, seed=1)
        s, r = fc_edge_list(N, B)
        run_case(name, D, H, x, v, q, s, r, rollout_steps=3)

    # ragged kNN scenes (inD-like): positions U(0,1)^2, in-degree varies, some nodes isolated
    g = torch.Generator().manual_seed(3)
    sizes = [12, 3, 17]
    xs, ss, rs, off = [], [], [], 0
    for n in sizes:
        pos = torch.rand(n, 2, generator=g)
        s, r = get_knn_graph_info(pos, torch.ones(n), n, use_edge2node=False)
        xs.append(pos)
        ss.append(s + off)
        rs.append(r + off)
        off += n
    x = torch.cat(xs)
    v = torch.rand(off, 2, generator=g) - 0.5
    q = torch.randint(0, 2, (off, 1), generator=g).float() * 2 - 1
    run_case("step2d_knn_h32", 2, 32, x, v, q, torch.cat(ss), torch.cat(rs))

    # unsorted edges + an isolated (in-degree 0) node that is not the last one
    x, v, q = synth(1, 6, 3, seed=5)
    s = torch.tensor([5, 0, 3, 1, 4, 0, 2, 5, 1, 3])
    r = torch.tensor([0, 5, 1, 3, 0, 4, 5, 3, 0, 5])   # node 2 never receives
    run_case("step3d_unsorted_h32", 3, 32, x, v, q, s, r)


if __name__ == "__main__":
    main()
