"""GPU tests of the chained backward kernel (csrc/tc_chain_bwd.cu) through the C ABI test hook: bit-identical to the two
separate tcgen05 data-gradient launches it replaces (same operands, rounding points and K order) and within 2e-2 of the
fp32 products formed by torch.  Reference op: autograd of GNNLayer.message_fn, nn/state2state/locs/locs.py:206-212,232-235."""
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


def dsilu(z):
    s = torch.sigmoid(z)
    return s * (1 + z * (1 - s))


@pytest.mark.parametrize("H", [64, 128, 256])
@pytest.mark.parametrize("M", [256, 100, 1000, 37888, 74 * 256 * 3 + 129])
@pytest.mark.parametrize("gather", [True, False])
def test_chain_bwd(lib, H, M, gather):
    torch.manual_seed(7 + H + M)
    A = (torch.randn(M, H, device="cuda") * 0.5).bfloat16()
    W1 = torch.randn(H, H, device="cuda") / H ** 0.5
    W2 = torch.randn(H, H, device="cuda") / H ** 0.5
    z1 = (torch.randn(M, H, device="cuda") * 2).bfloat16()
    z2 = (torch.randn(M, H, device="cuda") * 2).bfloat16()
    gat = recv = None
    if gather:
        Nn = max(M // 9, 1) + 1
        gat = (torch.randn(Nn, H, device="cuda") * 0.5).bfloat16()
        recv = torch.arange(M, device="cuda", dtype=torch.int32) // 9
    scratch = torch.empty(2 * H * H * 2, dtype=torch.uint8, device="cuda")
    o1 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    o2 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.aether_dbg_tc_chain_bwd(P(A), M, H, P(W1), P(W2), P(gat), H, P(recv), P(z1), P(z2), P(o1), P(o2), P(scratch), stream())
    assert rc == 0, rc
    torch.cuda.synchronize()
    # the two separate launches
    u1 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    u2 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.aether_dbg_tc_gemm(P(A), M, H, H, P(W1), H, H, H, H, 1, 0, 0, None, P(gat), H, P(recv), None, 0, P(z1), P(u1), None,
                                P(scratch), stream())
    assert rc == 0, rc
    rc = lib.aether_dbg_tc_gemm(P(u1), M, H, H, P(W2), H, H, H, H, 1, 0, 0, None, None, 0, None, None, 0, P(z2), P(u2), None,
                                P(scratch), stream())
    assert rc == 0, rc
    torch.cuda.synchronize()
    assert torch.equal(o1, u1), "g2 differs from the separate GEMM"
    assert torch.equal(o2, u2), "g1 differs from the separate GEMMs"
    # fp32 products of the bf16-rounded operands
    p1 = A.float() @ W1.bfloat16().float()
    if gather:
        p1 = p1 + gat.float()[recv.long()]
    p1 = p1 * dsilu(z1.float())
    p2 = (p1.bfloat16().float() @ W2.bfloat16().float()) * dsilu(z2.float())
    assert rel_err(o1.float(), p1) < 2e-2 and rel_err(o2.float(), p2) < 2e-2


@pytest.mark.parametrize("H", [64, 128, 256])
@pytest.mark.parametrize("M", [100, 5000, 74 * 256 * 2 + 77])
def test_chain_bwd_top_layer(lib, H, M):
    """No first GEMM: out1 = gyd[recv] * silu'(z1) (the mean's backward), out2 = (out1 W2) * silu'(z2)."""
    torch.manual_seed(3 + H + M)
    W2 = torch.randn(H, H, device="cuda") / H ** 0.5
    z1 = (torch.randn(M, H, device="cuda") * 2).bfloat16()
    z2 = (torch.randn(M, H, device="cuda") * 2).bfloat16()
    Nn = max(M // 9, 1) + 1
    gat = (torch.randn(Nn, H, device="cuda") * 0.5).bfloat16()
    recv = torch.arange(M, device="cuda", dtype=torch.int32) // 9
    scratch = torch.empty(2 * H * H * 2, dtype=torch.uint8, device="cuda")
    o1 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    o2 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.aether_dbg_tc_chain_bwd(None, M, H, None, P(W2), P(gat), H, P(recv), P(z1), P(z2), P(o1), P(o2), P(scratch), stream())
    assert rc == 0, rc
    torch.cuda.synchronize()
    p1 = gat.float()[recv.long()] * dsilu(z1.float())
    assert rel_err(o1.float(), p1) < 2e-2
    u2 = torch.full((M, H), float("nan"), dtype=torch.bfloat16, device="cuda")
    rc = lib.aether_dbg_tc_gemm(P(o1), M, H, H, P(W2), H, H, H, H, 1, 0, 0, None, None, 0, None, None, 0, P(z2), P(u2), None,
                                P(scratch), stream())
    assert rc == 0, rc
    torch.cuda.synchronize()
    assert torch.equal(o2, u2), "g1 differs from the separate GEMM on the same g2"
