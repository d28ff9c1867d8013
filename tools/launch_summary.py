#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list, for the launches of the first
chunk / step (up to the second launch of the first kernel name).
usage: launch_summary.py launches.csv [first_kernel_substring]"""
import collections, csv, sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hdr]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
scale = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0}
seq = [(r[ki].split("(")[0].replace("void ", ""), float(r[vi].replace(",", "")) * scale[r[ui]]) for r in rows[hdr + 1:] if len(r) > vi]
first = sys.argv[2] if len(sys.argv) > 2 else seq[0][0]
cnt, sel = 0, []
for k, v in seq:
    if first in k:
        cnt += 1
    if cnt == 1:
        sel.append((k, v))
tot, n = collections.OrderedDict(), collections.Counter()
for k, v in sel:
    tot[k] = tot.get(k, 0.0) + v
    n[k] += 1
T = sum(tot.values())
print("| kernel | launches | total (ms) | share |\n|---|---|---|---|")
for k, v in tot.items():
    print(f"| `{k}` | {n[k]} | {v:.3f} | {100 * v / T:.1f} % |")
print(f"| sum | {sum(n.values())} | {T:.3f} | |")
