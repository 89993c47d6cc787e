#!/usr/bin/env python
"""BASELINE config C5 (BlockQR 2^22 x 512, tf32_tc_cor + reorth) timed with CUDA events, with the on-device fp64 metrics.
argv: [log2 rows] [mode]"""
import sys, os, json, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tsqr_gpu_b200 as tq
m = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 22)
mode = sys.argv[2] if len(sys.argv) > 2 else "tf32_tc_cor"
n = 512
g = torch.Generator(device="cuda"); g.manual_seed(5)
dt = tq.io_dtype(mode)
a0 = tq.to_colmajor(torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t().to(dt))
buf = tq.Buffer().allocate(m, n, True)
q = tq.colmajor_empty(m, n, dt)
r = torch.zeros((n, n), dtype=dt, device="cuda").t()
a = a0.clone()
tq.qr(a, mode, True, buf, q, r)
torch.cuda.synchronize()
ts = []
for _ in range(3):
    a.copy_(a0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); tq.qr(a, mode, True, buf, q, r); e1.record(); e1.synchronize()
    ts.append(e0.elapsed_time(e1))
orth = float(tq.orthogonality(q, mode)); res = float(tq.residual(a0, q, r, mode))
print(json.dumps({"config": "C5", "m": m, "n": n, "mode": mode, "reorth": True, "ms": min(ts), "ms_all": ts,
                  "tma": os.environ.get("TSQR_B200_TMA", "1"), "orthogonality": orth, "residual": res}))
