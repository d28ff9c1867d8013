#!/usr/bin/env python
"""Attribute the warp-stall samples / executed instructions of one kernel in an `ncu --set full` report to CUDA
source lines. ncu's CSV source page is SASS only, so the line table comes from `nvdisasm -g` of the cubin inside
the shared library (built with -lineinfo); instruction i of the report is offset 16*i of the function.
usage: ncu_lines.py report.ncu-rep kernel_substring [lib.so] [top]"""
import csv, io, os, re, subprocess, sys, tempfile

rep, kname = sys.argv[1], sys.argv[2]
lib = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(__file__), "..", "rslqr_b200", "librslqr.so")
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cub = [f for f in os.listdir(tmp) if "sm_100" in f][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.splitlines()
lines, cur, inside = {}, None, False
for ln in dis:
    if ln.startswith(".text."):
        inside = kname in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        lines[int(m.group(1), 16)] = (cur, m.group(2).strip())

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = src.split('"Kernel Name",')
for blk in blocks[1:]:
    rows = list(csv.reader(io.StringIO('"Kernel Name",' + blk)))
    if kname not in rows[0][1]:
        continue
    hh = rows[1]
    si, ie = hh.index("# Samples"), hh.index("Instructions Executed")
    ci = {k: hh.index(k) for k in hh if k.startswith("stall_") and "Not Issued" not in k}
    wi, wid = hh.index("L1 Wavefronts Shared"), hh.index("L1 Wavefronts Shared Ideal")
    agg, tot, totins, totw = {}, 0, 0, 0
    for i, r in enumerate(rows[2:]):
        try:
            n, ins = int(r[si]), int(r[ie])
        except Exception:
            continue
        key = lines.get(16 * i, (None, ""))[0]
        a = agg.setdefault(key, [0, 0, {}, 0, 0])
        a[0] += n; a[1] += ins
        a[3] += int(r[wi] or 0); a[4] += int(r[wid] or 0); totw += int(r[wi] or 0)
        for k, c in ci.items():
            v = int(r[c] or 0)
            if v:
                a[2][k[6:]] = a[2].get(k[6:], 0) + v
        tot += n; totins += ins
    print(f"## {rows[0][1][:80]}: {tot} samples, {totins/1e6:.1f}M warp instructions, {totw/1e6:.1f}M shared wavefronts")
    order = int(os.environ.get("NCU_LINES_SORT", "0"))  # 0 samples, 1 instructions, 3 shared wavefronts
    for key, (n, ins, st, w, wideal) in sorted(agg.items(), key=lambda x: -x[1][order])[:top]:
        s = sorted(st.items(), key=lambda x: -x[1])[:3]
        print(f"  {str(key):34s} {100*n/tot:5.1f}% smp {ins/1e6:7.2f}M inst {w/1e6:7.2f}M smem-wf (ideal {wideal/1e6:6.2f})  {[(k, round(100*v/tot,1)) for k, v in s]}")
