"""GPU parity tests: the CUDA path through the C ABI against the oracle / golden vectors.

Bars (BASELINE.json north_star): integer path bit-exact; fp32 path rel <= 1e-5 of the tensor scale
per stage and for the step output; gradients rel <= 1e-4 of the tensor scale (fp32 sums over edges
in a different association than ATen's).  Every comparison is also appended to
gpurun_out/parity_report.json so one GPU run shows all stages, not only the first failure.
"""
import json
import os

import numpy as np
import pytest
import torch

from conftest import ROOT, golden_params, load_golden, rel_err
from oracle import aether_oracle as O

pytestmark = pytest.mark.gpu

STEP_CASES = ["step2d_h32", "step3d_h32", "step3d_h64", "step2d_h64_n12", "step2d_knn_h32", "step3d_unsorted_h32"]
FWD_TOL = 1e-5
GRAD_TOL = 1e-4
REPORT = os.path.join(ROOT, "gpurun_out", "parity_report.json")


def report(name, **kv):
    os.makedirs(os.path.dirname(REPORT), exist_ok=True)
    try:
        data = json.load(open(REPORT))
    except Exception:
        data = {}
    data[name] = kv
    json.dump(data, open(REPORT, "w"), indent=1, sort_keys=True)


@pytest.fixture(scope="module")
def ab():
    import aether_b200
    return aether_b200


def dev(a):
    return torch.as_tensor(a).cuda()


def build_model(ab, g, precision="fp32"):
    D, H = int(g["D"]), int(g["H"])
    m = ab.Aether(2 * D, H, 0.0, D, device="cuda", precision=precision)
    m.load_state_dict({k: v for k, v in golden_params(g).items()})
    return m


# ---- integer path ---------------------------------------------------------------------------------

def _check_graph(ab, send, recv, n_nodes):
    g = ab.PreparedGraph(send.cuda(), recv.cuda(), n_nodes)
    g.validate()
    perm, rowptr, ss, rs = O.receiver_csr(send, recv, n_nodes)
    E = send.numel()
    assert np.array_equal(g.rowptr.cpu().numpy(), rowptr.astype(np.int32))
    assert np.array_equal(g.perm.cpu().numpy()[:E], perm.astype(np.int32))
    assert np.array_equal(g.send.cpu().numpy()[:E], ss.astype(np.int32))
    assert np.array_equal(g.recv.cpu().numpy()[:E], rs.astype(np.int32))
    sp, colptr = O.sender_csr(ss, n_nodes)
    assert np.array_equal(g.colptr.cpu().numpy(), colptr.astype(np.int32))
    assert np.array_equal(g.sperm.cpu().numpy()[:E], sp.astype(np.int32))


def test_graph_prepare_bit_exact(ab):
    gi = load_golden("integer_path")
    for n, b in [(5, 1), (5, 3), (4, 2), (2, 2), (20, 2)]:
        s, r = torch.from_numpy(gi["fc_%d_%d/send" % (n, b)]), torch.from_numpy(gi["fc_%d_%d/recv" % (n, b)])
        _check_graph(ab, s, r, n * b)
    for n in [2, 7, 11, 25, 40]:
        _check_graph(ab, torch.from_numpy(gi["knn_%d/send" % n]), torch.from_numpy(gi["knn_%d/recv" % n]), n)
    # unsorted with an isolated node, empty list, a hub node above the warp-sort threshold, random multigraph
    _check_graph(ab, torch.tensor([5, 0, 3, 1, 4, 0, 2, 5, 1, 3]), torch.tensor([0, 5, 1, 3, 0, 4, 5, 3, 0, 5]), 6)
    _check_graph(ab, torch.zeros(0, dtype=torch.int64), torch.zeros(0, dtype=torch.int64), 4)
    gen = torch.Generator().manual_seed(0)
    s = torch.randint(0, 3000, (5000,), generator=gen)
    r = torch.where(torch.rand(5000, generator=gen) < 0.6, torch.tensor(7), torch.randint(0, 3000, (5000,), generator=gen))
    _check_graph(ab, s, r, 3000)
    s, r = O.fc_edges(100, 16)
    _check_graph(ab, s, r, 1600)


def test_graph_prepare_rejects_bad_index(ab):
    g = ab.PreparedGraph(torch.tensor([0, 1]).cuda(), torch.tensor([1, 9]).cuda(), 3)
    with pytest.raises(IndexError):
        g.validate()


def test_edge_builders_bit_exact(ab):
    gi = load_golden("integer_path")
    for n, b in [(5, 1), (5, 3), (4, 2), (2, 2), (20, 2)]:
        s, r = ab.fc_edges(n, b, "cuda")
        assert s.dtype == torch.int64
        assert np.array_equal(s.cpu().numpy(), gi["fc_%d_%d/send" % (n, b)])
        assert np.array_equal(r.cpu().numpy(), gi["fc_%d_%d/recv" % (n, b)])
    s, r = ab.fc_edges(100, 64, "cuda")
    so, ro = O.fc_edges(100, 64)
    assert torch.equal(s.cpu(), so) and torch.equal(r.cpu(), ro)
    sizes = [2, 7, 11, 25, 40]
    pos = torch.cat([torch.from_numpy(gi["knn_%d/pos" % n]) for n in sizes]).cuda()
    plan = ab.KnnPlan(sizes, 10, "cuda")
    s, r = ab.knn_edges(pos, plan)
    off = 0
    ws = np.concatenate([gi["knn_%d/send" % n] + o for n, o in zip(sizes, np.cumsum([0] + sizes[:-1]))])
    wr = np.concatenate([gi["knn_%d/recv" % n] + o for n, o in zip(sizes, np.cumsum([0] + sizes[:-1]))])
    assert np.array_equal(s.cpu().numpy(), ws)
    assert np.array_equal(r.cpu().numpy(), wr)


def test_caller_prep(ab):
    g = load_golden("step3d_h64")
    h, ea = ab.caller_prep(dev(g["x"]), dev(g["vel"]), dev(g["charges"]).reshape(-1), dev(g["send"]), dev(g["recv"]))
    assert rel_err(ea.cpu(), g["edge_attr_orig"]) <= 1e-6
    assert rel_err(h.cpu(), np.linalg.norm(g["vel"], axis=1, keepdims=True)) <= 1e-6


# ---- the step, stage by stage ---------------------------------------------------------------------

@pytest.mark.parametrize("case", STEP_CASES)
def test_step_stages(ab, case):
    from aether_b200 import _lib, ops
    g = load_golden(case)
    m = build_model(ab, g)
    D, H = int(g["D"]), int(g["H"])
    Nn = g["x"].shape[0]
    graph = ab.PreparedGraph(dev(g["send"]), dev(g["recv"]), Nn)
    params = [p.detach() for p in m.param_list()]
    out, ws = ops.step_forward_raw(params, graph, dev(g["x"]), dev(g["vel"]), dev(g["charges"]).reshape(-1),
                                   dev(g["edge_attr_orig"]), D, H, _lib.F_SAVE)
    perm = graph.perm.cpu().numpy()[: graph.n_edges]
    errs = {}
    for k in ["field", "rel_feat", "R", "edge_attr", "e1", "e2", "e3", "e4", "x1", "x2", "x3", "x4", "pred"]:
        got = ops.debug_stage(ws, graph, D, H, _lib.F_SAVE, k).cpu()
        want = g[k]
        if k == "edge_attr" or k[0] == "e":
            want = want[perm]
        want = want.reshape(got.shape) if got.numel() == want.size else want
        errs[k] = rel_err(got[: want.shape[0]], want)
    errs["out"] = rel_err(out.cpu(), g["out"])
    report("stages/" + case, **errs)
    bad = {k: v for k, v in errs.items() if not v <= FWD_TOL}
    assert not bad, bad


@pytest.mark.parametrize("case", STEP_CASES)
def test_step_grads(ab, case):
    g = load_golden(case)
    m = build_model(ab, g)
    m.train()
    x = dev(g["x"]).requires_grad_(True)
    v = dev(g["vel"]).requires_grad_(True)
    h = torch.zeros(x.shape[0], 1, device="cuda")
    out = m(h, x, [dev(g["send"]), dev(g["recv"])], v, dev(g["edge_attr_orig"]), dev(g["charges"]))
    loss = torch.nn.functional.mse_loss(out, dev(g["target"]))
    loss.backward()
    errs = {"loss": abs(float(loss) - float(g["loss"])) / max(abs(float(g["loss"])), 1e-30),
            "x": rel_err(x.grad.cpu(), g["grad/x"]), "vel": rel_err(v.grad.cpu(), g["grad/vel"])}
    for k, p in m.named_parameters():
        want = g["grad/" + k]
        if np.abs(want).max() == 0:
            errs[k] = float(p.grad.abs().max())
        else:
            errs[k] = rel_err(p.grad.cpu(), want)
    report("grads/" + case, **errs)
    bad = {k: e for k, e in errs.items() if not e <= GRAD_TOL}
    assert not bad, bad


@pytest.mark.parametrize("case", STEP_CASES[:4])
def test_rollout_golden(ab, case):
    g = load_golden(case)
    m = build_model(ab, g).eval()
    edges = (dev(g["send"]), dev(g["recv"]))
    T = g["rollout"].shape[0]
    xs = ab.rollout(m, dev(g["x"]), dev(g["vel"]), dev(g["charges"]), T, edges=edges)
    xg = ab.rollout(m, dev(g["x"]), dev(g["vel"]), dev(g["charges"]), T, edges=edges, cuda_graph=True)
    e1, e2 = rel_err(xs.cpu(), g["rollout"]), rel_err(xg.cpu(), g["rollout"])
    report("rollout/" + case, eager=e1, graph=e2)
    assert e1 <= 2e-5 and e2 <= 2e-5
    assert torch.equal(xs, xg)       # same kernels, same order: bitwise identical
