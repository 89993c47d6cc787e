#!/usr/bin/env python
"""Orthogonality / residual of one mode as a function of the row count (fp64 metrics on the device, the
definitions of reference src/validation.cu:43-80 and src/test.cu:146-165). One JSON line per size."""
import json, os, sys
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import tsqr_gpu_b200 as tq  # noqa: E402
from fullsize_parity import metrics  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "fp32_tc_cor"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
for lg in (16, 18, 20, 22, 24):
    m = 1 << lg
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    a = torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t().to(tq.io_dtype(mode))
    a = tq.to_colmajor(a)
    buf = tq.Buffer().allocate(m, n, True)
    q, r = tq.tsqr(a, mode, buffer=buf)
    torch.cuda.synchronize()
    st, est = tq.pc_status(buf, m, n, mode)
    orth, res = metrics(q, r, a)
    # diagonal and off-diagonal parts of Q^T Q - I separately: a bias shows up on the diagonal
    gram = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    for i in range(0, m, 1 << 20):
        qs = q[i:i + (1 << 20)].double(); gram += qs.t() @ qs
    d = (gram.diagonal() - 1.0)
    off = gram - torch.diag(gram.diagonal())
    print(json.dumps({"mode": mode, "m_log2": lg, "n": n, "status": st, "est": est, "orthogonality": orth, "residual": res,
                      "diag_mean": float(d.mean()), "diag_rms": float((d ** 2).mean().sqrt()),
                      "offdiag_rms": float((off ** 2).sum().div(n * (n - 1)).sqrt())}), flush=True)
    del a, q, buf
    torch.cuda.empty_cache()
