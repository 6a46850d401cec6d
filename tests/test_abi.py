"""The C-ABI library builds, loads and exports every symbol include/aether_b200.h declares.
No compute calls here (no GPU): only size queries and argument validation."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def lib():
    from aether_b200 import _lib
    return _lib.load()


def header_functions():
    src = open(os.path.join(ROOT, "include", "aether_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(aether_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    from aether_b200 import _lib
    names = header_functions()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_lib.SIGNATURES) == names          # the ctypes table covers the header one to one


def test_version_and_errors(lib):
    assert lib.aether_abi_version() == 1
    assert lib.aether_error_string(0) == b"success"
    assert b"workspace" in lib.aether_error_string(-2)


def test_size_queries(lib):
    from aether_b200 import _lib
    d = _lib.AetherDims(3, 64, 640, 2560)
    inf = lib.aether_workspace_bytes(ctypes.byref(d), 0)
    sav = lib.aether_workspace_bytes(ctypes.byref(d), _lib.F_SAVE)
    assert 0 < inf < sav
    assert sav >= 8 * 2560 * 64 * 4                  # four layers keep two [E,H] fp32 pre-activations
    assert lib.aether_backward_scratch_bytes(ctypes.byref(d), _lib.F_SAVE) > 2 * 2560 * 64 * 4
    assert lib.aether_graph_scratch_bytes(640, 2560) >= 2560 * 4
    bad = _lib.AetherDims(4, 64, 10, 10)
    assert lib.aether_workspace_bytes(ctypes.byref(bad), 0) == 0


def test_small_rollout_shape_query(lib):
    """Host-side logic of the single-launch rollout (no GPU work): which shapes fit one SM's shared memory."""
    ok = lib.aether_small_rollout_supported
    assert ok(2, 64, 5, 20, 0) == 1 and ok(3, 64, 40, 400, 0) == 1 and ok(2, 64, 21, 420, 0) == 1
    assert ok(2, 32, 5, 20, 0) == 0 and ok(2, 128, 5, 20, 0) == 0           # built for hidden 64
    assert ok(2, 64, 65, 100, 0) == 0                                     # more than 64 particles per scene
    assert ok(2, 64, 40, 1560, 0) == 0 and ok(2, 64, 40, 1560, 2) == 0 and ok(2, 64, 30, 870, 2) == 1                                  # fully connected 40: edge state beyond 227 KB
    assert ok(4, 64, 5, 20, 0) == 0 and ok(2, 64, 0, 0, 0) == 0
    d = __import__("aether_b200")._lib.AetherDims(2, 32, 10, 40)
    arr = (ctypes.c_void_p * 47)()
    one = ctypes.c_void_p(256)
    rc = lib.aether_small_rollout(ctypes.byref(d), arr, 2, one, one, None, one, one, 0, 5, 20, one, one, one, 3, one, None, None,
                                  one, 1 << 20, 0, None)
    assert rc == -3                                                    # AETHER_E_UNSUPPORTED, before any launch
    rc = lib.aether_small_rollout(ctypes.byref(d), arr, 2, one, one, None, one, None, 0, 5, 20, one, one, one, 3, one, None, None,
                                  one, 1 << 20, 0, None)
    assert rc == -1                                                    # send without recv
    d64 = __import__("aether_b200")._lib.AetherDims(2, 64, 10, 40)
    rc = lib.aether_small_rollout(ctypes.byref(d64), arr, 2, one, one, None, one, one, 0, 5, 20, one, one, one, 3, one, None, None,
                                  one, 16, 0, None)
    assert rc == -2                                                    # scratch too small
    assert lib.aether_small_rollout_scratch_bytes() == 32 * 64 * 64 * 4


def test_bad_arguments_return_codes(lib):
    from aether_b200 import _lib
    bad = _lib.AetherDims(4, 64, 10, 10)
    rc = lib.aether_step_forward(ctypes.byref(bad), None, None, None, None, None, None, None, None, None, 0, 0, None)
    assert rc == -1
    d = _lib.AetherDims(2, 64, 10, 10)
    g = _lib.AetherGraphC()
    arr = (ctypes.c_void_p * 47)()
    rc = lib.aether_step_forward(ctypes.byref(d), arr, ctypes.byref(g), None, None, None, None, None,
                                 ctypes.c_void_p(256), None, 0, 0, None)
    assert rc == -2                                  # workspace missing


def test_module_state_dict_matches_reference_layout():
    import torch
    import aether_b200
    from oracle import aether_oracle as O
    for D, H in [(2, 32), (3, 64)]:
        torch.manual_seed(1)
        m = aether_b200.Aether(2 * D, H, 0.0, D, device="cpu")
        assert [k for k, _ in O.param_shapes(D, H)] == list(m.state_dict().keys())
        p = O.init_params(D, H, seed=1)
        assert all(torch.equal(p[k], v) for k, v in m.state_dict().items())
    with pytest.raises(RuntimeError):
        m(None, torch.zeros(4, 3), [torch.zeros(0, dtype=torch.long)] * 2, torch.zeros(4, 3), torch.zeros(0, 2),
          torch.zeros(4, 1))


def test_dropin_alias_and_pickle_on_cpu():
    """The reference's import line (experiments/lorentz/main.py:16) resolves to this package after install(); the module
    pickles like the reference's (main.py:187) and `print(model)` keeps the reference's side effect (aether.py:163-167)."""
    import importlib
    import io
    import sys

    import torch

    import aether_b200
    import aether_b200.dropin as dropin
    saved = {k: sys.modules.get(k) for k in ("nn.state2state.aether", "nn.state2state.locs.locs")}
    try:
        dropin.install()
        mod = importlib.import_module("nn.state2state.aether")
        assert mod.Aether is aether_b200.Aether
        assert importlib.import_module("nn.state2state.locs.locs").LoCS is aether_b200.LoCS
        with pytest.raises(NotImplementedError):
            mod.ParallelAether(6, 64, 0.0, 3)
        m = mod.Aether(input_size=6, hidden_size=64, dropout_prob=0.0, num_dims=3, device="cpu")
        assert m.params == "132822" and str(m) == "132822"            # probed parameter count (D=3, H=64)
        buf = io.BytesIO()
        torch.save(m, buf)
        buf.seek(0)
        m2 = torch.load(buf, weights_only=False)
        assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), m2.state_dict().values()))
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def test_header_is_plain_c_and_c_caller_links(lib, tmp_path):
    """include/aether_b200.h compiles as C99 and a C program links against the library (examples/abi_sizes.c)."""
    import subprocess
    inc = os.path.join(ROOT, "include")
    libdir = os.path.join(ROOT, "aether_b200", "lib")
    r = subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-x", "c",
                        os.path.join(inc, "aether_b200.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    exe = str(tmp_path / "abi_sizes")
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", inc, os.path.join(ROOT, "examples", "abi_sizes.c"),
                        "-L", libdir, "-laether_b200", "-Wl,-rpath," + libdir, "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe, "2", "256", "102400", "10137600"], capture_output=True, text=True)
    assert r.returncode == 0 and "abi 1" in r.stdout, r.stdout + r.stderr
    from aether_b200 import _lib
    d = _lib.AetherDims(2, 256, 102400, 10137600)
    assert ("training fp32 %d " % lib.aether_workspace_bytes(ctypes.byref(d), _lib.F_SAVE)) in r.stdout
    sd = _lib.AetherSeqDims(2, 5, 512, 2, 0, 512, 128, 256, 0, 64)
    assert ("seq2seq decoder + encoder workspace %d" % lib.aether_seq_workspace_bytes(ctypes.byref(sd))) in r.stdout
    assert subprocess.run([exe, "4", "64", "1", "1"], capture_output=True).returncode == 1      # D = 4 is rejected
