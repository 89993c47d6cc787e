#!/usr/bin/env python
"""BASELINE config C2 (TSQR 2^20 x 32 fp32_tc_cor) a few times: the ncu launch-list target for the small-problem overheads."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tsqr_gpu_b200 as tq
m, n = 1 << 20, 32
g = torch.Generator(device="cuda"); g.manual_seed(2)
a = tq.to_colmajor(torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t())
buf = tq.Buffer().allocate(m, n, False)
for _ in range(4):
    q, r = tq.qr(a, "fp32_tc_cor", False, buf)
torch.cuda.synchronize()
print("ok")
