#!/usr/bin/env python
"""Hottest SASS instructions (warp stall samples) and the opcode mix of every kernel in an ncu report.
  python tools/ncu_hot.py gpurun_out/x.ncu-rep [top]"""
import collections, csv, io, subprocess, sys
top = int(sys.argv[2]) if len(sys.argv) > 2 else 24
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
kern, cur = [], None
for r in csv.reader(io.StringIO(out)):
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; kern.append(cur); continue
    if r and r[0] == "Address":
        cur["hdr"] = r; continue
    if cur is not None and r:
        cur["rows"].append(r)
seen = set()
for k in kern:
    if k["name"] in seen:
        continue
    seen.add(k["name"])
    h = k["hdr"]; si = h.index("Warp Stall Sampling (All Samples)"); ii = h.index("Instructions Executed"); so = h.index("Source")
    tot = sum(int(r[si]) for r in k["rows"]) or 1; toti = sum(int(r[ii]) for r in k["rows"]) or 1
    print("=====", k["name"][:90], "samples", tot, "warp-instructions", toti)
    for r in sorted(k["rows"], key=lambda r: -int(r[si]))[:top]:
        print(f"{int(r[si]) * 100 / tot:5.1f}%  exec {int(r[ii]):>9d}  {r[so].strip()[:100]}")
    hist, hs = collections.Counter(), collections.Counter()
    for r in k["rows"]:
        w = r[so].strip().split()
        op = (w[1] if w[0].startswith("@") else w[0]).split(".")[0]
        hist[op] += int(r[ii]); hs[op] += int(r[si])
    print("  opcode mix (executed %, stall %):", [(o, round(c * 100 / toti, 1), round(hs[o] * 100 / tot, 1)) for o, c in hist.most_common(20)])
