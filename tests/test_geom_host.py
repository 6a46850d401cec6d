"""CPU check of the geometry code the CUDA kernels run.

aether_b200/csrc/geom.cu and geom_bwd.cu keep their per-element math in __host__ __device__
bodies; tests/host_geom_check.cu loops over them on the host.  Here the forward features are
compared with the oracle and the hand-derived backward with autograd through the oracle.
"""
import ctypes
import os

import numpy as np
import pytest
import torch

from conftest import load_golden, rel_err
from oracle import aether_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host():
    from aether_b200 import build
    try:
        so = build.build_hostcheck()
    except Exception as e:  # pragma: no cover
        pytest.skip("nvcc not available: %s" % e)
    return ctypes.CDLL(so)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def run_forward(host, D, x, vel, f, send, recv, orig):
    Nn, E = x.shape[0], send.shape[0]
    fe = 7 * D + D * (D - 1) // 2 + 2
    ld = (fe + 3) // 4 * 4
    R = np.zeros((Nn, D * D), np.float32)
    x0 = np.zeros((Nn, 3 * D), np.float32)
    feat = np.zeros((E, ld), np.float32)
    perm = np.arange(E, dtype=np.int32)
    host.hostcheck_geom_forward(D, ctypes.c_int64(Nn), ctypes.c_int64(E), _p(send), _p(recv), _p(perm), _p(x), _p(vel),
                                _p(f), _p(orig), _p(R), _p(x0), _p(feat))
    return R, x0, feat[:, :fe]


@pytest.mark.parametrize("case", ["kat2d", "kat3d", "edge2d", "edge3d", "anti3d"])
def test_forward_kat(host, case):
    g = load_golden("localizer_kat")
    nodes = g[case + "/nodes"]
    D = nodes.shape[1] // 3
    send, recv = g["send"].astype(np.int32), g["recv"].astype(np.int32)
    x, v, f = [np.ascontiguousarray(nodes[:, i * D:(i + 1) * D]) for i in range(3)]
    orig = np.zeros((send.shape[0], 2), np.float32)
    R, x0, feat = run_forward(host, D, x, v, f, send, recv, orig)
    want = g[case + "/edge_attr"]
    got = feat[:, : want.shape[1]]
    # branch-cut features (+-pi) may flip sign under 1-ulp libm differences: compare on the circle
    diff = np.abs(got - want)
    wrap = np.minimum(diff, np.abs(diff - 2 * np.pi))
    wrap = np.minimum(wrap, np.abs(diff - 2.0))          # euler features are divided by pi
    assert np.nanmax(wrap) <= 5e-6, np.nanmax(wrap)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.allclose(R.reshape(-1, D, D), g[case + "/R"], atol=2e-6)
    assert np.allclose(x0, g[case + "/rel_feat"], atol=2e-6)


@pytest.mark.parametrize("D,N,B,seed", [(2, 5, 3, 0), (3, 5, 3, 1), (3, 7, 2, 2), (2, 9, 2, 3)])
def test_forward_and_backward_random(host, D, N, B, seed):
    x, v, q = O.synth_state(B, N, D, seed=seed)
    g = torch.Generator().manual_seed(seed + 50)
    f = torch.randn(x.shape, generator=g)
    send, recv = O.fc_edges(N, B)
    Nn, E = x.shape[0], send.shape[0]
    orig = torch.randn(E, 2, generator=g)
    pred = torch.randn(Nn, D, generator=g)

    xr, vr, fr = [t.clone().requires_grad_(True) for t in (x, v, f)]
    rel, R, ea = O.localize(xr, vr, fr, send, recv)
    out = xr + O.apply_rot(R, pred)
    gfeat = torch.randn(ea.shape, generator=g)
    gx0 = torch.randn(rel.shape, generator=g)
    gout = torch.randn(out.shape, generator=g)
    (ea * gfeat).sum().add((rel * gx0).sum()).add((out * gout).sum()).backward()

    s32, r32 = send.numpy().astype(np.int32), recv.numpy().astype(np.int32)
    Rh, x0h, feath = run_forward(host, D, x.numpy(), v.numpy(), f.numpy(), s32, r32, orig.numpy())
    assert rel_err(feath[:, : ea.shape[1]], ea.detach()) <= 2e-6
    assert rel_err(x0h, rel.detach()) <= 2e-6

    fe = ea.shape[1] + 2
    ld = (fe + 3) // 4 * 4
    gfeat_p = np.zeros((E, ld), np.float32)
    gfeat_p[:, : ea.shape[1]] = gfeat.numpy()
    GS = 3 * D + D * D
    z = lambda *s: np.zeros(s, np.float32)
    gsend, grecv, ssend, srecv, gRg = z(E, GS), z(E, GS), z(Nn, GS), z(Nn, GS), z(Nn, D * D)
    gpred, gf, gx, gv = z(Nn, D), z(Nn, D), z(Nn, D), z(Nn, D)
    host.hostcheck_geom_backward(D, ctypes.c_int64(Nn), ctypes.c_int64(E), _p(s32), _p(r32), _p(x.numpy()),
                                 _p(v.numpy()), _p(f.numpy()), _p(Rh), _p(pred.numpy()), _p(gfeat_p),
                                 _p(gx0.numpy()), _p(gout.numpy()), _p(gsend), _p(grecv), _p(ssend), _p(srecv),
                                 _p(gRg), _p(gpred), _p(gf), _p(gx), _p(gv))
    assert rel_err(gf, fr.grad) <= 1e-5
    assert rel_err(gx, xr.grad) <= 1e-5
    assert rel_err(gv, vr.grad) <= 2e-5
    want_gpred = torch.einsum("nij,ni->nj", R.detach(), gout)
    assert rel_err(gpred, want_gpred) <= 1e-5
