#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel: joins the SASS page of an ncu report (executed instruction counts
per SASS address) with the line table nvdisasm prints for the same kernel in the built library.

    python tools/ncu_lines.py gpurun_out/x.ncu-rep k_draw [library.so] [launches]
"""
import csv, os, re, subprocess, sys, tempfile
from collections import Counter

rep, kern = sys.argv[1], sys.argv[2]
lib = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "xash3d-fwgs_b200", "libref_cuda.so")
lib = os.path.abspath(lib)
launches = float(sys.argv[4]) if len(sys.argv) > 4 else None
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern], capture_output=True, text=True).stdout
rows = list(csv.reader(sass.splitlines()))
hdr = rows[1]
iA, iI, iT, iS = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Source")
iSmp = hdr.index("# Samples")
insts = [(int(r[iA], 16), int(r[iI]), int(r[iT]), r[iS].strip(), int(r[iSmp] or 0)) for r in rows[2:] if len(r) > iT and r[iA].startswith("0x")]
base = insts[0][0]
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=d, capture_output=True)
    cub = [f for f in os.listdir(d) if f.endswith(".cubin")]
    text = ""
    for c in cub:
        text += subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, c)], capture_output=True, text=True).stdout
line_of, cur, inside = {}, None, False
for ln in text.splitlines():
    if ln.startswith("//---------------------"):
        inside = (".text." in ln) and (kern in ln)
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+\S", ln)
    if m and cur:
        line_of[int(m.group(1), 16)] = cur
per, thr, smp = Counter(), Counter(), Counter()
tot = 0
for a, n, t, s, sm in insts:
    key = line_of.get(a - base, ("?", 0))
    per[key] += n; thr[key] += t; smp[key] += sm; tot += n
scale = launches or 1.0
print(f"kernel {kern}: {tot / scale:.0f} warp instructions per launch ({len(insts)} SASS instructions)")
src_cache = {}
def src(f, l):
    for root in ("xash3d-fwgs_b200/csrc", "."):
        p = os.path.join(root, f)
        if os.path.exists(p):
            if p not in src_cache: src_cache[p] = open(p).read().splitlines()
            return src_cache[p][l - 1].strip()[:90] if 0 < l <= len(src_cache[p]) else ""
    return ""
totsmp = sum(smp.values()) or 1
for (f, l), n in per.most_common(45):
    print(f"{100 * n / tot:5.1f}% inst {100 * smp[(f, l)] / totsmp:5.1f}% stall-samples  thr/inst {thr[(f, l)] / max(n, 1):4.1f}  {f}:{l}  {src(f, l)}")
