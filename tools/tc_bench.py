#!/usr/bin/env python
"""Micro-benchmark of the tcgen05 GEMM kernel in the shapes the train step uses (H=256)."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aether_b200 import _lib
lib = _lib.load(build_if_missing=False)
P = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
stream = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
H = int(os.environ.get("H", 256))
M = 128 * 148 * 64
A = torch.randn(M, H, device="cuda").bfloat16()
Z = torch.randn(M, H, device="cuda").bfloat16()
W = torch.randn(H, H, device="cuda") / 16
out = torch.empty(M, H, dtype=torch.bfloat16, device="cuda")
scratch = torch.empty(H * H * 2, dtype=torch.uint8, device="cuda")
Nn = M // 99 + 1
gat = torch.randn(Nn, 2 * H, device="cuda").bfloat16()
recv = torch.arange(M, device="cuda", dtype=torch.int32) // 99
send = ((torch.arange(M, device="cuda", dtype=torch.int32) % 99) + (recv // 100) * 100).clamp(max=Nn - 1)

def run(name, silu, gat_, ia, ib, z, passes, N=None):
    # N < H: the kernel a column-split CTA pair would run (K = H inputs, N output columns, weights N*K*2 bytes in smem)
    N = N or H
    Wn = W[:N].contiguous()
    zn = None if z is None else (z if N == H else z[:, :N].contiguous())
    outn = out if N == H else torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    def call():
        return lib.aether_dbg_tc_gemm(P(A), M, H, H, P(Wn), H, N, N, H, 0, 0, silu, None, P(gat_), 2 * H, P(ia), P(ib), H, P(zn),
                                      P(outn), None, P(scratch), stream())
    assert call() == 0
    best = 1e9
    for _ in range(4):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            call()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / 5)
    ms = best
    print("%-34s dbg=%s %.3f ms  %.2f us/tile  %.0f GB/s" % (name, os.environ.get("AETHER_TC_DBG", "0"), ms,
          ms * 1e3 / (M / 128 / 148), M * 2 * (H + N * (passes - 1)) / ms / 1e6), flush=True)

for dbg in os.environ.get("DBGS", "0").split(","):
    os.environ["AETHER_TC_DBG"] = dbg
    if os.environ.get("ONLY_B"):
        run("B1: dsilu", 0, None, None, None, Z, 3)
        run("B2: gather recv + dsilu", 0, gat, recv, None, Z, 3)
        continue
    run("plain", 0, None, None, None, None, 2)
    run("silu-on-load", 1, None, None, None, None, 2)
    run("B1: dsilu", 0, None, None, None, Z, 3)
    run("B2: gather recv + dsilu", 0, gat, recv, None, Z, 3)
    run("fwdA: silu + gather send,recv", 1, gat, send, recv, None, 2)
    run("fwdA': silu + gather recv,recv", 1, gat, recv, recv, None, 2)
    run("fwdA'': silu + gather send only", 1, gat, send, None, None, 2)
    if os.environ.get("HALF"):        # half-width kernels: twice their us/tile is what a column-split CTA pair would cost
        run("plain N=H/2", 0, None, None, None, None, 2, H // 2)
        run("B1 N=H/2", 0, None, None, None, Z, 3, H // 2)
        run("B2 N=H/2", 0, gat, recv, None, Z, 3, H // 2)
        run("fwdA N=H/2", 1, gat, send, recv, None, 2, H // 2)

# node-level shape: fp32 output (+ residual through the debug hook is not exposed; plain + bias only)
Mn = 102400
An = torch.randn(Mn, H, device="cuda").bfloat16()
outf = torch.empty(Mn, H, dtype=torch.float32, device="cuda")
bias = torch.randn(H, device="cuda")
def call_node():
    return lib.aether_dbg_tc_gemm(P(An), Mn, H, H, P(W), H, H, H, H, 0, 0, 0, P(bias), None, 0, None, None, 0, None,
                                  None, P(outf), P(scratch), stream())
assert call_node() == 0
best = 1e9
for _ in range(4):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        call_node()
    b.record()
    torch.cuda.synchronize()
    best = min(best, a.elapsed_time(b) / 10)
print("node GEMM fp32 out M=%d: %.1f us (incl. weight pack), %.0f GB/s" % (Mn, best * 1e3, Mn * H * 6 / best / 1e6))
