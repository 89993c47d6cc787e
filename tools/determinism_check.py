#!/usr/bin/env python
"""Same input, repeated runs: are R, Q and the pass-1 Gram sums bit-identical? (debug aid)"""
import sys, os, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import tsqr_gpu_b200 as tq  # noqa: E402

mode, n = "fp32_tc_cor", 64
for lg in [int(x) for x in sys.argv[1:]] or [22, 24]:
    m = 1 << lg
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    a = tq.to_colmajor(torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1, generator=g).t())
    rpre = torch.eye(64, device="cuda") * 20.0 + torch.triu(torch.rand(64, 64, device="cuda"), 1)
    c = torch.eye(64, device="cuda")
    gs, qs, rs, q2s = [], [], [], []
    for rep in range(4):
        q, gsum = tq.pc_probe(a, mode, rpre, c, use_tma=(rep % 2 == 0))
        torch.cuda.synchronize()
        gs.append(gsum.clone()); qs.append(q.clone())
        q2, r2 = tq.tsqr(a, mode)
        torch.cuda.synchronize()
        rs.append(r2.clone()); q2s.append(q2.clone())
    out = {"m_log2": lg}
    for name, lst in (("gsum", gs), ("probe_q", qs), ("r", rs), ("q", q2s)):
        diffs = [float((lst[0].double() - x.double()).abs().max()) for x in lst[1:]]
        nz = [int((lst[0] != x).sum()) for x in lst[1:]]
        out[name + "_maxdiff"] = diffs; out[name + "_ndiff"] = nz
    # exact Gram of the probe's Z (= probe_q since c = I), fp64
    z = qs[0]
    gram = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    for i in range(0, m, 1 << 20):
        zz = z[i:i + (1 << 20)].double(); gram += zz.t() @ zz
    for rep in range(4):
        hh, lh = gs[rep][:64], gs[rep][64:]
        gg = 0.5 * (hh + hh.t()) + lh + lh.t()
        rel = (gg - gram) / gram.diagonal().mean()
        out[f"gram_err_rep{rep}"] = {"diag_mean": float(rel.diagonal().mean()), "offdiag_rms": float(((rel - torch.diag(rel.diagonal())) ** 2).sum().div(n * (n - 1)).sqrt())}
    print(json.dumps(out), flush=True)
    del a, gs, qs, rs, q2s
    torch.cuda.empty_cache()
