#!/usr/bin/env python
"""repeatability of tsqr (out of place and in place) over shapes (debug aid)"""
import sys, os, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tsqr_gpu_b200 as tq
for (m, n) in [(20000, 32), (1 << 17, 32), (20000, 64), (20096, 32), (16384, 32), (40000, 16), (70001, 48)]:
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    a = tq.to_colmajor(torch.empty((n, m), device="cuda").normal_(generator=g).t())
    outs = []
    for rep in range(4):
        buf = tq.Buffer().allocate(m, n, True)
        if rep % 2 == 0:
            q, r = tq.tsqr(a, "fp32_tc_cor", buffer=buf)
        else:
            b = a.clone(); q, r = tq.tsqr(b, "fp32_tc_cor", buffer=buf, q=b)
        torch.cuda.synchronize()
        st = tq.pc_status(buf, m, n, "fp32_tc_cor")
        outs.append((q.clone(), r.clone(), st))
    print(json.dumps({"m": m, "n": n, "status": [o[2][0] for o in outs], "est": [round(o[2][1], 4) for o in outs],
                      "q_ndiff_vs0": [int((outs[0][0] != o[0]).sum()) for o in outs[1:]],
                      "r_ndiff_vs0": [int((outs[0][1] != o[1]).sum()) for o in outs[1:]]}), flush=True)
