#!/usr/bin/env python
"""Compact markdown table of an `ncu --set full` report: one row per captured launch.
usage: ncu_table.py report.ncu-rep [name_filter]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
filt = sys.argv[2] if len(sys.argv) > 2 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units = rows[0], rows[1]
cols = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("Block Size", "block"), ("gpu__time_duration.sum", "time"),
        ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("launch__registers_per_thread", "regs"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64 pipe %"),
        ("sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed", "fp64 tensor %"),
        ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "LSU wavefronts %")]
idx = [(h.index(k), n) for k, n in cols if k in h]
print("| " + " | ".join(n + (f" ({units[i]})" if units[i] and n in ("time", "dram rd", "dram wr") else "") for i, n in idx) + " |")
print("|" + "---|" * len(idx))
for r in rows[2:]:
    if filt and filt not in r[h.index("Kernel Name")]:
        continue
    out = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = "`" + v.split("(")[0].replace("void ", "") + "`"
        else:
            try:
                v = f"{float(v.replace(',', '')):.4g}"
            except ValueError:
                pass
        out.append(v)
    print("| " + " | ".join(out) + " |")
