"""GPU tests of the chained edge-MLP kernel (csrc/tc_chain.cu: GEMM -> SiLU -> GEMM on a CTA pair, intermediate in
tensor memory) through the C ABI test hook.

Two references: (1) the two separate tcgen05 GEMM launches it replaces -- same operands, same rounding points and the
same K order, so the results must be BIT-IDENTICAL; (2) the same products formed in fp32 by torch from the
bf16-rounded operands, tolerance 2e-2 of the tensor scale (the bf16-path bar of BASELINE.json).
Reference op: GNNLayer.message_fn, nn/state2state/locs/locs.py:206-212,232-235.
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


def chain(lib, A, W1, W2, b1, b2, silu, gat=None, send=None, recv=None, goff=0, save=True, kmod=0, k1_valid=None):
    M, K1 = A.shape
    H = W2.shape[0]
    scratch = torch.empty((K1 + H) * H * 2, dtype=torch.uint8, device="cuda")
    out1 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda") if save else None
    out2 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.aether_dbg_tc_chain(P(A), M, K1, A.stride(0), P(W1), W1.stride(0), K1 if k1_valid is None else k1_valid, kmod,
                                 P(W2), H, silu, P(b1), P(b2), P(gat), 0 if gat is None else gat.stride(0), P(send), P(recv),
                                 goff, P(out1), P(out2), P(scratch), stream())
    assert rc == 0, rc
    torch.cuda.synchronize()
    return out1, out2


def two_gemms(lib, A, W1, W2, b1, b2, silu, gat=None, send=None, recv=None, goff=0, kmod=0, k1_valid=None):
    M, K1 = A.shape
    H = W2.shape[0]
    scratch = torch.empty(max(K1, H) * H * 2, dtype=torch.uint8, device="cuda")
    out1 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    out2 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.aether_dbg_tc_gemm(P(A), M, K1, A.stride(0), P(W1), W1.stride(0), H, H, K1 if k1_valid is None else k1_valid, 0,
                                kmod, silu, P(b1), P(gat), 0 if gat is None else gat.stride(0), P(send), P(recv), goff, None,
                                P(out1), None, P(scratch), stream())
    assert rc == 0, rc
    rc = lib.aether_dbg_tc_gemm(P(out1), M, H, H, P(W2), H, H, H, H, 0, 0, 1, P(b2), None, 0, None, None, 0, None,
                                P(out2), None, P(scratch), stream())
    assert rc == 0, rc
    torch.cuda.synchronize()
    return out1, out2


def torch_ref(A, W1, W2, b1, b2, silu, gat=None, send=None, recv=None, goff=0):
    a = A.float()
    if silu:
        a = torch.nn.functional.silu(a).bfloat16().float()
    p1 = a @ W1.bfloat16().float().t() + b1
    if gat is not None:
        H = W2.shape[0]
        g = gat.float()
        p1 = p1 + g[send.long(), :H] + g.reshape(-1)[(recv.long() * gat.stride(0) + goff)[:, None] + torch.arange(H, device="cuda")]
    p1b = p1.bfloat16()
    p2 = torch.nn.functional.silu(p1b.float()).bfloat16().float() @ W2.bfloat16().float().t() + b2
    return p1, p2


def make(M, K1, H, seed, gather):
    torch.manual_seed(seed)
    A = torch.randn(M, K1, device="cuda").bfloat16()
    W1 = torch.randn(H, K1, device="cuda") / K1 ** 0.5
    W2 = torch.randn(H, H, device="cuda") / H ** 0.5
    b1 = torch.randn(H, device="cuda") * 0.5
    b2 = torch.randn(H, device="cuda") * 0.5
    gat = send = recv = None
    if gather:
        Nn = max(M // 9, 1) + 1
        gat = (torch.randn(Nn, 2 * H, device="cuda") * 0.5).bfloat16()     # planes P_s | P_r side by side
        recv = (torch.arange(M, device="cuda", dtype=torch.int32) // 9)
        send = torch.randint(0, Nn, (M,), device="cuda", dtype=torch.int32)
    return A, W1, W2, b1, b2, gat, send, recv


@pytest.mark.parametrize("H", [64, 128, 256])
@pytest.mark.parametrize("M", [256, 100, 1000, 37888, 74 * 256 * 3 + 129])
def test_chain_layers_2_to_4(lib, H, M):
    """silu(A) We^T + b1 + P_s[send] + P_r[recv] -> silu -> W2^T + b2 (K1 = H), training (pre1 stored) and inference."""
    A, W1, W2, b1, b2, gat, send, recv = make(M, H, H, 11 + H + M, True)
    o1, o2 = chain(lib, A, W1, W2, b1, b2, 1, gat, send, recv, H)
    r1, r2 = torch_ref(A, W1, W2, b1, b2, 1, gat, send, recv, H)
    assert rel_err(o1.float(), r1) < 1e-2 and rel_err(o2.float(), r2) < 2e-2
    u1, u2 = two_gemms(lib, A, W1, W2, b1, b2, 1, gat, send, recv, H)
    assert torch.equal(o1, u1), "pre1 differs from the separate GEMM"
    assert torch.equal(o2, u2), "pre2 differs from the separate GEMMs"
    _, i2 = chain(lib, A, W1, W2, b1, b2, 1, gat, send, recv, H, save=False)
    assert torch.equal(i2, o2), "inference (intermediate not stored) differs from training mode"


@pytest.mark.parametrize("H", [64, 128, 256])
@pytest.mark.parametrize("M", [128, 5000, 40001])
def test_chain_layer_1(lib, H, M):
    """feat[hi|lo] W1^T + b1 -> silu -> W2^T + b2 (K1 = 64: 17 feature columns repeated for the hi and lo halves)."""
    torch.manual_seed(5 + H + M)
    feat = torch.randn(M, 17, device="cuda") * 3
    hi = feat.bfloat16()
    lo = (feat - hi.float()).bfloat16()
    A = torch.zeros(M, 64, device="cuda", dtype=torch.bfloat16)
    A[:, :17], A[:, 32:49] = hi, lo
    W1 = torch.randn(H, 17, device="cuda") / 4
    W2 = torch.randn(H, H, device="cuda") / H ** 0.5
    b1, b2 = torch.randn(H, device="cuda"), torch.randn(H, device="cuda")
    o1, o2 = chain(lib, A, W1, W2, b1, b2, 0, kmod=32, k1_valid=17)
    u1, u2 = two_gemms(lib, A, W1, W2, b1, b2, 0, kmod=32, k1_valid=17)
    assert torch.equal(o1, u1) and torch.equal(o2, u2)
    p1 = (hi.float() + lo.float()) @ W1.bfloat16().float().t() + b1
    p2 = torch.nn.functional.silu(p1.bfloat16().float()).bfloat16().float() @ W2.bfloat16().float().t() + b2
    assert rel_err(o1.float(), p1) < 1e-2 and rel_err(o2.float(), p2) < 2e-2


def test_chain_is_deterministic(lib):
    A, W1, W2, b1, b2, gat, send, recv = make(50000, 256, 256, 3, True)
    a = chain(lib, A, W1, W2, b1, b2, 1, gat, send, recv, 256)
    b = chain(lib, A, W1, W2, b1, b2, 1, gat, send, recv, 256)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


@pytest.mark.parametrize("H", [64, 128, 256])
@pytest.mark.parametrize("pattern", ["fc", "random", "mixed"])
def test_chain_box_gather(lib, H, pattern):
    """Sender term fed by TMA boxes (tiles whose senders span < 128 node rows) with the direct-load fallback for the rest:
    bit-identical to the register-gather variant on fully connected batches (every tile inside a graph or at a graph
    boundary), random senders (every tile falls back) and a mix."""
    N, B = 100, 12
    M = N * (N - 1) * B
    torch.manual_seed(17 + H)
    A = torch.randn(M, H, device="cuda").bfloat16()
    W1 = torch.randn(H, H, device="cuda") / H ** 0.5
    W2 = torch.randn(H, H, device="cuda") / H ** 0.5
    b2 = torch.randn(H, device="cuda") * 0.5
    Nn = N * B
    gat = (torch.randn(Nn, 2 * H, device="cuda") * 0.5).bfloat16()
    r = torch.arange(N, device="cuda").repeat_interleave(N - 1)                       # receiver-sorted fully connected graph
    s_ = torch.cat([torch.cat([torch.arange(0, i), torch.arange(i + 1, N)]) for i in range(N)]).cuda()
    recv = torch.cat([r + g * N for g in range(B)]).int()
    send = torch.cat([s_ + g * N for g in range(B)]).int()
    if pattern == "random":
        send = torch.randint(0, Nn, (M,), device="cuda", dtype=torch.int32)
    elif pattern == "mixed":
        half = M // 2
        send = torch.cat([send[:half], torch.randint(0, Nn, (M - half,), device="cuda", dtype=torch.int32)])
    scratch = torch.empty(2 * H * H * 2, dtype=torch.uint8, device="cuda")

    def run(box):
        o1 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
        o2 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
        rc = lib.aether_dbg_tc_chain_ex(P(A), M, H, H, P(W1), H, H, 0, P(W2), H, 1, None, P(b2), P(gat), 2 * H, P(send), P(recv), H,
                                        P(o1), P(o2), P(scratch), Nn, box, stream())
        assert rc == 0, rc
        torch.cuda.synchronize()
        return o1, o2

    a1, a2 = run(0)
    b1, b2_ = run(1)
    assert torch.equal(a1, b1) and torch.equal(a2, b2_)
