#!/usr/bin/env python
"""Phase clocks (SM cycles) of the single-CTA kernels of the fast path: prints what tsqr_b200_pc_status reports with
TSQR_B200_PC_PHASES=1, for a few shapes."""
import os, sys, torch
os.environ["TSQR_B200_PC_PHASES"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tsqr_gpu_b200 as tq
for (m, n) in [(1 << 24, 64), (1 << 20, 32), (1 << 22, 64)]:
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    a = tq.to_colmajor(torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t())
    buf = tq.Buffer().allocate(m, n, False)
    for _ in range(2):
        q, r = tq.qr(a, "fp32_tc_cor", False, buf)
    torch.cuda.synchronize()
    print(m, n, tq.pc_status(buf.ws, m, n, "fp32_tc_cor"), flush=True)
