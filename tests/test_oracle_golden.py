"""The oracle restatement is pinned to vectors recorded from the unmodified reference.

Bars: integer path bit-exact; fp32 stages within 2e-6 of the tensor scale
(same torch build generated both sides; the slack covers op-fusion order only).
"""
import numpy as np
import pytest
import torch

from conftest import golden_params, load_golden, rel_err
from oracle import aether_oracle as O

STEP_CASES = ["step2d_h32", "step3d_h32", "step3d_h64", "step2d_h64_n12", "step2d_knn_h32", "step3d_unsorted_h32"]
FP_TOL = 2e-6


def test_fc_edges_bit_exact():
    g = load_golden("integer_path")
    for n, b in [(5, 1), (5, 3), (4, 2), (2, 2), (20, 2)]:
        s, r = O.fc_edges(n, b)
        assert s.dtype == torch.int64 and r.dtype == torch.int64
        assert np.array_equal(s.numpy(), g["fc_%d_%d/send" % (n, b)])
        assert np.array_equal(r.numpy(), g["fc_%d_%d/recv" % (n, b)])


def test_knn_edges_bit_exact():
    g = load_golden("integer_path")
    for n in [2, 7, 11, 25, 40]:
        s, r = O.knn_edges(torch.from_numpy(g["knn_%d/pos" % n]))
        assert np.array_equal(s.numpy(), g["knn_%d/send" % n])
        assert np.array_equal(r.numpy(), g["knn_%d/recv" % n])


def test_receiver_csr_properties():
    s, r = O.fc_edges(5, 3)
    perm, rowptr, ss, rs = O.receiver_csr(s, r, 15)
    assert np.all(np.diff(rs) >= 0) and rowptr[0] == 0 and rowptr[-1] == len(rs)
    assert np.array_equal(np.sort(perm), np.arange(len(rs)))
    for n in range(15):
        seg = perm[rowptr[n]:rowptr[n + 1]]
        assert np.all(r.numpy()[seg] == n) and np.all(np.diff(seg) > 0)   # stable
    ps, colptr = O.sender_csr(ss, 15)
    assert np.all(np.diff(ss[ps]) >= 0) and colptr[-1] == len(ss)


@pytest.mark.parametrize("case", ["kat2d", "kat3d", "edge2d", "edge3d", "anti3d"])
def test_localizer_kat(case):
    g = load_golden("localizer_kat")
    nodes = torch.from_numpy(g[case + "/nodes"])
    D = nodes.shape[1] // 3
    send, recv = torch.from_numpy(g["send"]), torch.from_numpy(g["recv"])
    rel, R, ea = O.localize(nodes[:, :D], nodes[:, D:2 * D], nodes[:, 2 * D:], send, recv)
    for got, key in [(rel, "rel_feat"), (R, "R"), (ea, "edge_attr")]:
        want = g[case + "/" + key]
        assert np.array_equal(np.isnan(got.numpy()), np.isnan(want))
        assert np.allclose(got.numpy(), want, rtol=0, atol=FP_TOL, equal_nan=True), key


def test_survey_literals():
    """The literal KATs quoted in SURVEY.md 8c."""
    g = load_golden("localizer_kat")
    e0 = g["kat2d/edge_attr"][0]
    want = [-2.0, 1.00000012, -0.5, 2.23606801, 2.67794490, -4.37113883e-08, -1.0, -0.25000003, -0.5,
            0, 0, 2.0, -8.74227766e-08, 4.37113883e-08, 1.0]
    assert np.allclose(e0, want, rtol=0, atol=1e-7)


@pytest.mark.parametrize("case", STEP_CASES)
def test_init_matches_reference(case):
    g = load_golden(case)
    p = O.init_params(int(g["D"]), int(g["H"]), seed=int(g["seed"]))
    ref = golden_params(g)
    assert list(p.keys()) == [k for k, _ in O.param_shapes(int(g["D"]), int(g["H"]))] or set(p) == set(ref)
    for k in ref:
        assert torch.equal(p[k], ref[k]), k


@pytest.mark.parametrize("case", STEP_CASES)
def test_step_stages(case):
    g = load_golden(case)
    p = golden_params(g)
    t = lambda k: torch.from_numpy(g[k])
    st = {}
    out = O.aether_step(p, t("x"), (t("send"), t("recv")), t("vel"), t("edge_attr_orig"), t("charges"), stages=st)
    h, ea = O.caller_prep(t("x"), t("vel"), t("charges"), t("send"), t("recv"))
    assert rel_err(ea, g["edge_attr_orig"]) <= FP_TOL
    n_ref_rows = g["x1"].shape[0]          # reference scatter output has max(recv)+1 rows
    for k in ["field", "rel_feat", "R", "edge_attr", "e1", "e2", "e3", "e4", "x1", "x2", "x3", "x4", "pred", "out"]:
        assert rel_err(st[k][: g[k].shape[0]], g[k]) <= FP_TOL, k
    assert n_ref_rows == t("x").shape[0]
    assert rel_err(out, g["out"]) <= FP_TOL


@pytest.mark.parametrize("case", STEP_CASES)
def test_step_grads(case):
    g = load_golden(case)
    p = {k: v.clone().requires_grad_(True) for k, v in golden_params(g).items()}
    t = lambda k: torch.from_numpy(g[k])
    x = t("x").requires_grad_(True)
    v = t("vel").requires_grad_(True)
    out = O.aether_step(p, x, (t("send"), t("recv")), v, t("edge_attr_orig"), t("charges"))
    loss = torch.nn.functional.mse_loss(out, t("target"))
    loss.backward()
    assert abs(float(loss) - float(g["loss"])) <= 1e-7
    assert rel_err(x.grad, g["grad/x"]) <= 1e-5
    assert rel_err(v.grad, g["grad/vel"]) <= 1e-5
    for k in p:
        got = p[k].grad if p[k].grad is not None else torch.zeros_like(p[k])
        assert rel_err(got, g["grad/" + k]) <= 1e-5, k


@pytest.mark.parametrize("case", STEP_CASES[:4])
def test_rollout(case):
    g = load_golden(case)
    p = golden_params(g)
    t = lambda k: torch.from_numpy(g[k])
    edges = (t("send"), t("recv"))
    with torch.no_grad():
        xs = O.rollout(p, t("x"), t("vel"), t("charges"), lambda x: edges, g["rollout"].shape[0])
    assert rel_err(xs, g["rollout"]) <= 1e-5


def test_radius_knn_oracle_matches_knn_and_properties():
    """The large-scene builder's oracle: with an unbounded radius it is the reference's kNN list (tie-free inputs),
    and with a finite radius every edge is within the radius, nearest first, at most k per sender."""
    g = torch.Generator().manual_seed(3)
    pos = torch.rand(60, 2, generator=g)
    s0, r0 = O.knn_edges(pos, 10)
    s1, r1 = O.radius_knn_edges(pos, 1e3, 10)
    assert torch.equal(s0, s1) and torch.equal(r0, r1)
    pos3 = torch.randn(400, 3, generator=g)
    s, r = O.radius_knn_edges(pos3, 0.9, 8)
    d = (pos3[s] - pos3[r]).norm(dim=1)
    assert float(d.max()) <= 0.9 * (1 + 1e-6) and bool((s != r).all())
    deg = torch.bincount(s, minlength=400)
    assert int(deg.max()) <= 8
    for i in (0, 17, 399):
        di = d[s == i]
        assert bool((di[1:] >= di[:-1]).all())
    full = (torch.cdist(pos3, pos3) <= 0.9).sum(1) - 1
    assert torch.equal(deg, full.clamp(max=8))


@pytest.mark.parametrize("case", ["locs2d_h32", "locs3d_h64"])
def test_locs_oracle_vs_reference(case):
    """LoCS (nn/state2state/locs/locs.py:102-139): stages, output, gradients and a 3-step rollout recorded from the
    unmodified reference (tests/golden/make_golden_locs.py)."""
    g = load_golden(case)
    p = {k: v.clone().requires_grad_(True) for k, v in golden_params(g).items()}
    x = torch.from_numpy(g["x"]).requires_grad_(True)
    v = torch.from_numpy(g["vel"]).requires_grad_(True)
    send, recv = torch.from_numpy(g["send"]), torch.from_numpy(g["recv"])
    ea = torch.from_numpy(g["edge_attr_orig"])
    st = {}
    out = O.locs_step(p, x, (send, recv), v, ea, stages=st)
    for k in ("rel_feat", "R", "edge_attr"):
        assert float((st[k].detach() - torch.from_numpy(g[k])).abs().max()) <= 2e-6, k
    assert float((out.detach() - torch.from_numpy(g["out"])).abs().max()) <= 2e-6
    loss = torch.nn.functional.mse_loss(out, torch.from_numpy(g["target"]))
    loss.backward()
    assert abs(float(loss.detach()) - float(g["loss"])) <= 1e-7
    assert rel_err(x.grad, g["grad/x"]) <= 1e-4 and rel_err(v.grad, g["grad/vel"]) <= 1e-4
    for k, t in p.items():
        if float(np.abs(g["grad/" + k]).max()) > 0:
            assert rel_err(t.grad, g["grad/" + k]) <= 1e-4, k
    q = torch.from_numpy(g["charges"])
    xs, xc, vc = [], x.detach(), v.detach()
    pd = {k: t.detach() for k, t in p.items()}
    for _ in range(g["rollout"].shape[0]):
        _, eaa = O.caller_prep(xc, vc, q, send, recv)
        xn = O.locs_step(pd, xc, (send, recv), vc, eaa)
        vc, xc = xn - xc, xn
        xs.append(xc)
    assert rel_err(torch.stack(xs), g["rollout"]) <= 1e-5
