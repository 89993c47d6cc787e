#!/usr/bin/env python
"""which Gram entries differ between repeated pass-1 runs (debug aid)"""
import sys, os, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tsqr_gpu_b200 as tq  # noqa: E402
mode, n, m = "fp32_tc_cor", 64, 1 << 24
g = torch.Generator(device="cuda"); g.manual_seed(7)
a = tq.to_colmajor(torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t())
rpre = torch.eye(64, device="cuda") * 20.0 + torch.triu(torch.rand(64, 64, device="cuda"), 1)
c = torch.eye(64, device="cuda")
_, ref = tq.pc_probe(a, mode, rpre, c, use_tma=False)
for rep in range(8):
    _, gs = tq.pc_probe(a, mode, rpre, c, use_tma=True)
    torch.cuda.synchronize()
    d = (gs != ref)
    hh, lh = d[:64], d[64:]
    rows_all = [i for i in range(64) if bool(hh[i].all())]
    cols_all = [j for j in range(64) if bool(hh[:, j].all())]
    lrows = [i for i in range(64) if bool(lh[i].all())]
    lcols = [j for j in range(64) if bool(lh[:, j].all())]
    print(json.dumps({"rep": rep, "ndiff": int(d.sum()), "hh_full_rows": rows_all, "hh_full_cols": cols_all, "lh_full_rows": lrows, "lh_full_cols": lcols,
                      "maxdiff": float((gs - ref).abs().max())}), flush=True)
