#!/usr/bin/env python
"""Join an ncu source-page CSV (SASS, per instruction samples) with nvdisasm -g line info so that
stall samples and executed instructions can be read per CUDA source line.

  ncu -i prof.ncu-rep --page source --csv > src.csv
  python tools/ncu_lines.py src.csv build/tc_pass.o <kernel-substring> [kernel-occurrence]
"""
import collections, csv, re, subprocess, sys, tempfile, os, glob

def sass_lines(obj, kern_sub):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, stdout=subprocess.DEVNULL)
    cub = glob.glob(os.path.join(d, "*.cubin"))[0]
    out = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout.splitlines()
    res, inside, cur = [], False, None
    for ln in out:
        if ln.startswith("//---") and ".text." in ln:
            inside = kern_sub in ln
            cur = None
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            res.append((cur, m.group(2).strip()))
    return res

def main():
    src_csv, obj, kern_sub = sys.argv[1:4]
    occ = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    rows = list(csv.reader(open(src_csv)))
    hdr = rows[1]
    iS, iE, iSm = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    kerns, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = []
            kerns.append((r[1], cur))
            continue
        if len(r) == len(hdr) and r[iE].isdigit():
            cur.append(r)
    sel = [k for k in kerns if kern_sub.split("ILi")[0].split("_kernel")[0][-6:] in k[0] or True]
    name, data = [k for k in kerns if kern_sub_matches(k[0], kern_sub)][occ]
    sl = sass_lines(obj, kern_sub)
    if len(sl) != len(data):
        print(f"warning: {len(sl)} disassembled vs {len(data)} profiled instructions", file=sys.stderr)
    per = collections.defaultdict(lambda: [0, 0])
    tot_s = sum(int(r[iSm]) for r in data) or 1
    tot_e = sum(int(r[iE]) for r in data) or 1
    for (loc, _), r in zip(sl, data):
        per[loc][0] += int(r[iSm])
        per[loc][1] += int(r[iE])
    print(f"{name[:90]}  samples {tot_s} warp-instr {tot_e}")
    srcs = {}
    for loc, (s, e) in sorted(per.items(), key=lambda x: -x[1][0])[:45]:
        text = ""
        if loc:
            f = loc[0]
            if f not in srcs:
                cand = glob.glob(os.path.join(os.path.dirname(os.path.abspath(obj)), "..", f))
                srcs[f] = open(cand[0]).read().splitlines() if cand else []
            if 0 < loc[1] <= len(srcs[f]):
                text = srcs[f][loc[1] - 1].strip()[:90]
        print(f"{100*s/tot_s:5.1f}% smp {100*e/tot_e:5.1f}% ins  {loc}  {text}")

def kern_sub_matches(ncu_name, mangled_sub):
    # mangled: tc_reduce_kernelILi64ELb1EffE / cs_reduce_kernelILi64ELi4EffE -> base name + first two template args
    base = re.search(r"([a-z]+_[a-z]+_kernel)", mangled_sub).group(1)
    m = re.search(r"ILi(\d+)EL([bi])(\d+)", mangled_sub)
    if m is None:   # kernels without the <int, bool|int, ...> head: match on the base name only
        return base in ncu_name
    second = f"(bool){m.group(3)}" if m.group(2) == "b" else f"(int){m.group(3)}"
    return base in ncu_name and f"(int){m.group(1)}, {second}" in ncu_name

if __name__ == "__main__":
    main()
