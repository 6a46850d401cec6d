#!/usr/bin/env python
"""Micro-benchmark of the chained edge-MLP kernel (tc_chain.cu) against the two separate GEMM launches it replaces,
config-4 row pattern (99-edge receiver segments, contiguous senders), H from the environment (default 256)."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aether_b200 import _lib
lib = _lib.load(build_if_missing=False)
P = lambda t: ctypes.c_void_p(0 if t is None else t.data_ptr())
stream = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
H = int(os.environ.get("H", 256))
M = int(os.environ.get("M", 128 * 148 * 64))
A = torch.randn(M, H, device="cuda").bfloat16()
W1 = torch.randn(H, H, device="cuda") / 16
W2 = torch.randn(H, H, device="cuda") / 16
b1 = torch.randn(H, device="cuda")
b2 = torch.randn(H, device="cuda")
o1 = torch.empty(M, H, dtype=torch.bfloat16, device="cuda")
o2 = torch.empty(M, H, dtype=torch.bfloat16, device="cuda")
scratch = torch.empty(2 * H * H * 2, dtype=torch.uint8, device="cuda")
Nn = M // 99 + 1
gat = torch.randn(Nn, 2 * H, device="cuda").bfloat16()
recv = torch.arange(M, device="cuda", dtype=torch.int32) // 99
send = ((torch.arange(M, device="cuda", dtype=torch.int32) % 99) + (recv // 100) * 100).clamp(max=Nn - 1)


send_rand = torch.randint(0, Nn, (M,), device="cuda", dtype=torch.int32)


def chain(save, gather, snd=None, bias=True):
    snd = send if snd is None else snd
    return lib.aether_dbg_tc_chain(P(A), M, H, H, P(W1), H, H, 0, P(W2), H, 1, P(b1 if bias else None), P(b2), P(gat if gather else None), 2 * H,
                                   P(snd if gather else None), P(recv if gather else None), H, P(o1 if save else None), P(o2),
                                   P(scratch), stream())


def chain_box(save, snd):
    return lib.aether_dbg_tc_chain_ex(P(A), M, H, H, P(W1), H, H, 0, P(W2), H, 1, None, P(b2), P(gat), 2 * H, P(snd), P(recv), H,
                                      P(o1 if save else None), P(o2), P(scratch), Nn, 1, stream())


def two(gather):
    rc = lib.aether_dbg_tc_gemm(P(A), M, H, H, P(W1), H, H, H, H, 0, 0, 1, P(b1), P(gat if gather else None), 2 * H,
                                P(send if gather else None), P(recv if gather else None), H, None, P(o1), None, P(scratch), stream())
    rc |= lib.aether_dbg_tc_gemm(P(o1), M, H, H, P(W2), H, H, H, H, 0, 0, 1, P(b2), None, 0, None, None, 0, None, P(o2), None,
                                 P(scratch), stream())
    return rc


def timeit(name, fn, passes):
    assert fn() == 0
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(4):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / 5)
    flops = 2 * 2.0 * M * H * H
    print("%-44s %.3f ms  %.2f us per 128-row tile per SM  %.0f GB/s (%d passes)  %.0f TFLOP/s" %
          (name, best, best * 1e3 / (M / 128 / 148), M * H * 2 * passes / best / 1e6, passes, flops / best / 1e9), flush=True)


MODES = {"two_g": ("two GEMMs, gather (incl. 2 weight packs)", lambda: two(True), 4),
         "two": ("two GEMMs, no gather", lambda: two(False), 4),
         "train_g": ("chain train  (pre1 stored), gather", lambda: chain(True, True), 3),
         "train": ("chain train  (pre1 stored), no gather", lambda: chain(True, False), 3),
         "infer_g": ("chain infer  (pre1 not stored), gather", lambda: chain(False, True), 2),
         "train_g_bcast": ("chain train, gather with send := recv (warp-uniform rows)", lambda: chain(True, True, recv), 3),
         "train_g_rand": ("chain train, gather with random senders", lambda: chain(True, True, send_rand), 3),
         "train_g_nobias": ("chain train, gather, bias folded into P_r", lambda: chain(True, True, None, False), 3),
         "train_box": ("chain train, sender term by TMA boxes (contiguous senders)", lambda: chain_box(True, send), 3),
         "infer_box": ("chain infer, sender term by TMA boxes (contiguous senders)", lambda: chain_box(False, send), 2),
         "train_box_rand": ("chain train, box variant on random senders (all tiles fall back)", lambda: chain_box(True, send_rand), 3),
         "infer": ("chain infer  (pre1 not stored), no gather", lambda: chain(False, False), 2)}
for m in os.environ.get("MODES", ",".join(MODES)).split(","):
    timeit(*MODES[m])
