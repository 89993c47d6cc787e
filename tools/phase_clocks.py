#!/usr/bin/env python
"""Per-phase cycle counters of the build pass (block 0 of the last build launch = level 0).
Needs the library built with `make -C tsqr_gpu_b200/csrc PHASES=1` (rebuild without it afterwards)."""
import ctypes, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import tsqr_gpu_b200 as tq
from tsqr_gpu_b200._lib import lib

m, n, mode = 1 << int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22, 64, "fp32_tc_cor"
a = torch.empty((n, m), device="cuda").uniform_(-1, 1).t()
q, r = tq.qr(a, mode)
q, r = tq.qr(a, mode)
torch.cuda.synchronize()
out = (ctypes.c_ulonglong * 14)()
lib.tsqr_b200_debug_phases.argtypes = [ctypes.POINTER(ctypes.c_ulonglong)]
rc = lib.tsqr_b200_debug_phases(out)
names = ["wait loads + ns staging", "S mma", "raw spill", "combine", "T_p rows", "solve", "rn staging", "Q mma (+next loads issue)", "Q store",
         "  solve: accumulate", "  solve: ldtm + shuffles", "  solve: T_p apply", "  solve: stores + exchange + sttm", "-"]
tot = sum(out[:9])
print("rc", rc, "total cycles", tot)
for nm, v in zip(names, out):
    print(f"{nm:28s} {v:12d} {100.0 * v / max(tot, 1):5.1f}%")
