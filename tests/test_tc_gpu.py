"""GPU tests of the tcgen05 building blocks (tc_gemm.cu, tc_wgrad.cu) through the C ABI test hooks.

Reference: the same product formed in fp32 by torch from the bf16-rounded operands (so only the
accumulation order and the tanh-based SiLU differ).  Tolerance 2e-2 of the tensor scale is the bf16-path
bar of BASELINE.json; the plain-GEMM cases are checked much tighter (5e-3, bf16 output rounding).
"""
import ctypes

import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from aether_b200 import _lib
    return _lib.load(build_if_missing=False)


def P(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def run_gemm(lib, A, W, N, K, *, n_valid=None, k_valid=None, transpose=0, kmod=0, silu=0, bias=None, gat=None,
             idx_a=None, idx_b=None, goff_b=0, dsilu=None, f32=False, bf=True):
    """Returns (bf16 output or None, fp32 output or None); the two epilogues are separate launches."""
    M = A.shape[0]
    n_valid = N if n_valid is None else n_valid
    k_valid = K if k_valid is None else k_valid
    scratch = torch.empty(N * K * 2, dtype=torch.uint8, device="cuda")
    out_bf = out_f = None
    if bf and N >= 64:
        out_bf = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device="cuda")
        rc = lib.aether_dbg_tc_gemm(P(A), M, K, A.stride(0), P(W), W.stride(0), N, n_valid, k_valid, transpose, kmod, silu,
                                    P(bias), P(gat), 0 if gat is None else gat.stride(0), P(idx_a), P(idx_b), goff_b,
                                    P(dsilu), P(out_bf), None, P(scratch), stream())
        assert rc == 0, rc
    if f32:
        out_f = torch.full((M, N), float("nan"), dtype=torch.float32, device="cuda")
        rc = lib.aether_dbg_tc_gemm(P(A), M, K, A.stride(0), P(W), W.stride(0), N, n_valid, k_valid, transpose, kmod, silu,
                                    P(bias), None, 0, None, None, 0, None, None, P(out_f), P(scratch), stream())
        assert rc == 0, rc
    torch.cuda.synchronize()
    return out_bf, out_f


@pytest.mark.parametrize("K,N", [(64, 64), (64, 256), (128, 128), (256, 256), (256, 32), (64, 32), (128, 32), (64, 128)])
@pytest.mark.parametrize("M", [128, 1000, 40000])
def test_plain_gemm(lib, K, N, M):
    torch.manual_seed(K * 7 + N + M)
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = torch.randn(N, K, device="cuda") / K ** 0.5
    bias = torch.randn(N, device="cuda")
    out_bf, out_f = run_gemm(lib, A, W, N, K, bias=bias, f32=True)
    ref = A.float() @ W.bfloat16().float().t() + bias
    assert rel_err(out_f, ref) < 1e-4
    if N >= 64:
        assert rel_err(out_bf.float(), ref) < 5e-3


def test_transposed_weight_and_padding(lib):
    torch.manual_seed(3)
    M, K, N = 3000, 256, 32
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = torch.randn(K, 17, device="cuda")            # dgrad form: B[n][k] = W[k][n], n < 17
    _, out_f = run_gemm(lib, A, W, N, K, n_valid=17, transpose=1, f32=True)
    ref = A.float() @ W.bfloat16().float()
    assert rel_err(out_f[:, :17], ref) < 1e-4
    # hi/lo operand of layer 1: K = 64 = [hi(32) | lo(32)], weights repeated with k % 32
    feat = torch.randn(M, 17, device="cuda") * 3
    hi = feat.bfloat16()
    lo = (feat - hi.float()).bfloat16()
    A2 = torch.zeros(M, 64, device="cuda", dtype=torch.bfloat16)
    A2[:, :17] = hi
    A2[:, 32:49] = lo
    W1 = torch.randn(256, 17, device="cuda")
    _, out2 = run_gemm(lib, A2, W1, 256, 64, k_valid=17, kmod=32, f32=True)
    ref2 = feat @ W1.bfloat16().float().t()
    assert rel_err(out2, ref2) < 1e-4


@pytest.mark.parametrize("M,K,N", [(5000, 256, 256), (777, 128, 128), (130, 64, 64), (4000, 256, 128)])
def test_silu_on_load_gather_and_dsilu(lib, M, K, N):
    torch.manual_seed(5)
    Nn = 300
    pre = torch.randn(M, K, device="cuda").bfloat16() * 2
    W = torch.randn(N, K, device="cuda") / K ** 0.5
    bias = torch.randn(N, device="cuda")
    gat = torch.randn(Nn, 2 * N, device="cuda").bfloat16()
    ia = torch.randint(0, Nn, (M,), device="cuda", dtype=torch.int32)
    ib = torch.randint(0, Nn, (M,), device="cuda", dtype=torch.int32)
    a = torch.nn.functional.silu(pre.float()).bfloat16().float()
    _, out_f = run_gemm(lib, pre, W, N, K, silu=1, bias=bias, f32=True, bf=False)
    ref0 = a @ W.bfloat16().float().t() + bias
    assert rel_err(out_f, ref0) < 5e-3               # tanh.approx SiLU before bf16 rounding
    out_bf, _ = run_gemm(lib, pre, W, N, K, silu=1, bias=bias, gat=gat, idx_a=ia, idx_b=ib, goff_b=N)
    ref = ref0 + gat[ia.long(), :N].float() + gat[ib.long(), N:].float()
    assert rel_err(out_bf.float(), ref) < 1e-2
    # backward epilogue: (A W) * silu'(z)
    z = torch.randn(M, N, device="cuda").bfloat16() * 2
    o2, _ = run_gemm(lib, pre, W, N, K, dsilu=z)
    zf = z.float()
    s = torch.sigmoid(zf)
    ref2 = (pre.float() @ W.bfloat16().float().t()) * (s * (1 + zf * (1 - s)))
    assert rel_err(o2.float(), ref2) < 1e-2


@pytest.mark.parametrize("MG,NA", [(256, 256), (128, 128), (64, 64), (256, 64), (128, 64)])
@pytest.mark.parametrize("E", [64, 1000, 70001])
def test_wgrad(lib, MG, NA, E):
    torch.manual_seed(MG + NA + E)
    G = torch.randn(E, MG, device="cuda").bfloat16()
    ACT = torch.randn(E, NA, device="cuda").bfloat16()
    nf = lib.aether_dbg_tc_wgrad_partial_floats(MG, NA)
    partial = torch.empty(nf, dtype=torch.float32, device="cuda")
    for silu in (0, 1):
        dW = torch.full((MG, NA), float("nan"), device="cuda")
        db = torch.full((MG,), float("nan"), device="cuda") if silu else None
        rc = lib.aether_dbg_tc_wgrad(P(G), MG, P(ACT), NA, silu, E, P(dW), NA, 0, P(db), P(partial), nf, stream())
        assert rc == 0, rc
        torch.cuda.synchronize()
        if db is not None:
            assert rel_err(db, G.float().sum(0)) < 1e-4
        act = torch.nn.functional.silu(ACT.float()).bfloat16().float() if silu else ACT.float()
        ref = G.float().t() @ act
        assert rel_err(dW, ref) < (5e-3 if silu else 1e-4)
    if NA == 64:
        dW = torch.full((MG, 17), float("nan"), device="cuda")
        rc = lib.aether_dbg_tc_wgrad(P(G), MG, P(ACT), NA, 0, E, P(dW), 17, 1, None, P(partial), nf, stream())
        assert rc == 0
        torch.cuda.synchronize()
        ref = G.float().t() @ ACT.float()
        assert rel_err(dW, ref[:, :17] + ref[:, 32:49]) < 1e-4


def test_gemm_throughput_report(lib):
    """Not a pass/fail perf gate: records the kernel's rate for the notes (gpurun_out/tc_rates.txt)."""
    import os
    M, K, N = 128 * 148 * 40, 256, 256
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = torch.randn(N, K, device="cuda") / 16
    lines = []
    for silu in (0, 1):
        run_gemm(lib, A, W, N, K, silu=silu)
        gat = torch.randn(3000, 2 * N, device="cuda").bfloat16()
        ia = torch.randint(0, 3000, (M,), device="cuda", dtype=torch.int32)
        out_bf = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
        scratch = torch.empty(N * K * 2, dtype=torch.uint8, device="cuda")
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            lib.aether_dbg_tc_gemm(P(A), M, K, K, P(W), K, N, N, K, 0, 0, silu, None, None, 0, None, None, 0, None,
                                   P(out_bf), None, P(scratch), stream())
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        lines.append("tc_gemm M=%d K=%d N=%d silu=%d: %.3f ms, %.1f TFLOP/s, %.0f GB/s" %
                     (M, K, N, silu, ms, 2.0 * M * K * N / ms / 1e9, M * (K + N) * 2 / ms / 1e6))
        a.record()
        for _ in range(5):
            lib.aether_dbg_tc_gemm(P(A), M, K, K, P(W), K, N, N, K, 0, 0, silu, None, P(gat), 2 * N, P(ia), None, N, P(A),
                                   P(out_bf), None, P(scratch), stream())
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        lines.append("tc_gemm + gather + dsilu epilogue silu=%d: %.3f ms, %.0f GB/s (in+aux+out)" %
                     (silu, ms, M * (K + 2 * N) * 2 / ms / 1e6))
    E = M
    G = torch.randn(E, 256, device="cuda").bfloat16()
    nf = lib.aether_dbg_tc_wgrad_partial_floats(256, 256)
    partial = torch.empty(nf, dtype=torch.float32, device="cuda")
    dW = torch.empty(256, 256, device="cuda")
    for silu in (0, 1):
        lib.aether_dbg_tc_wgrad(P(G), 256, P(A), 256, silu, E, P(dW), 256, 0, P(dW), P(partial), nf, stream())
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            lib.aether_dbg_tc_wgrad(P(G), 256, P(A), 256, silu, E, P(dW), 256, 0, P(dW), P(partial), nf, stream())
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        lines.append("tc_wgrad E=%d 256x256 silu=%d: %.3f ms, %.1f TFLOP/s, %.0f GB/s" %
                     (E, silu, ms, 2.0 * E * 65536 / ms / 1e9, E * 512 * 2 / ms / 1e6))
    os.makedirs("gpurun_out", exist_ok=True)
    open("gpurun_out/tc_rates.txt", "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))
