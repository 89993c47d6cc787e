#!/usr/bin/env python
"""One BlockQR factorisation at BASELINE config C5 scaled by argv[1] rows (default 2^22 x 512, tf32_tc_cor + reorth):
the ncu target for the projection kernels tc_tn_kernel / tc_nn_kernel."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tsqr_gpu_b200 as tq
m = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 22)
n = 512
g = torch.Generator(device="cuda"); g.manual_seed(5)
a = tq.to_colmajor(torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t())
for _ in range(2):
    q, r = tq.qr(a.clone(), "tf32_tc_cor", True)
torch.cuda.synchronize()
print("ok", float(q[0, 0]))
