#!/usr/bin/env python
"""Does cp.async.bulk.tensor.2d.tile::gather4 fill an operand slot like a box load?  (run on the GPU box)"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aether_b200 import _lib
lib = _lib.load(build_if_missing=False)
P = lambda t: ctypes.c_void_p(t.data_ptr())
torch.manual_seed(0)
Nn, H = 1000, 256
tab = torch.randn(Nn, H, device="cuda").bfloat16()
idx = torch.randint(0, Nn, (128,), device="cuda", dtype=torch.int32)
for box_rows in (1, 4, 128):
    for col0 in (0, 64, 192):
        raw = torch.zeros(16384, dtype=torch.uint8, device="cuda")
        rc = lib.aether_dbg_gather4_probe(P(tab), Nn, H, H, box_rows, P(idx), col0, P(raw), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        try:
            torch.cuda.synchronize()
        except Exception as e:
            print("box_rows", box_rows, "col0", col0, "rc", rc, "CUDA error", str(e)[:80]); sys.exit(0)
        got = raw.view(torch.bfloat16).view(128, 64)          # raw rows of 128 bytes
        r = torch.arange(128, device="cuda")
        exp = torch.empty(128, 64, dtype=torch.bfloat16, device="cuda")
        rows = tab[idx.long(), col0:col0 + 64]                 # [128, 64]
        for j in range(8):                                     # 16-byte piece j of row r sits at piece (j ^ (r % 8))
            pj = (j ^ (r % 8))
            for rr in range(128):
                exp[rr, int(pj[rr]) * 8:int(pj[rr]) * 8 + 8] = rows[rr, j * 8:j * 8 + 8]
        print("box_rows", box_rows, "col0", col0, "rc", rc, "match" if torch.equal(got, exp) else "MISMATCH", flush=True)
