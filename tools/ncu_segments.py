#!/usr/bin/env python
"""Summarise an `ncu --set full --import-source on` report per kernel: headline metrics and the warp-stall
samples segmented by barrier (BAR.SYNC) so that a phase-structured kernel can be read phase by phase.
usage: ncu_segments.py report.ncu-rep"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__ops_path_tensor_src_fp64.sum", "sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum"]
for r in rows[2:]:
    print("==", {k: r[h.index(k)] for k in keys if k in h})
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = src.split('"Kernel Name",')
for blk in blocks[1:]:
    lines = list(csv.reader(io.StringIO('"Kernel Name",' + blk)))
    print("\n##", lines[0][1][:90])
    hh = lines[1]; si = hh.index('# Samples'); so = hh.index('Source'); ie = hh.index('Instructions Executed')
    ci = {k: hh.index(k) for k in hh if k.startswith('stall_') and 'Not Issued' not in k}
    data = []
    for r in lines[2:]:
        try: data.append((int(r[si]), r[so], int(r[ie]), {k: int(r[i] or 0) for k, i in ci.items()}))
        except Exception: pass
    tot = sum(d[0] for d in data) or 1
    start = 0
    def flush(i):
        global start
        seg = data[start:i]
        n = sum(d[0] for d in seg); ins = sum(d[2] for d in seg)
        if n == 0 and ins == 0: start = i; return
        st = {}; kinds = {}
        for d in seg:
            for k, v in d[3].items(): st[k] = st.get(k, 0) + v
            t = d[1].split()
            op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
            kinds[op] = kinds.get(op, 0) + d[2]
        top = [(k[6:], round(100 * v / tot, 1)) for k, v in sorted(st.items(), key=lambda x: -x[1])[:5]]
        topk = [(k, round(v / 1e6, 1)) for k, v in sorted(kinds.items(), key=lambda x: -x[1])[:7]]
        print(f"  sass {start}-{i}: {100*n/tot:5.1f}% samples, {ins/1e6:7.1f}M inst | stalls% {top} | Minst {topk}")
        start = i
    for i, d in enumerate(data):
        if 'BAR.SYNC' in d[1] or 'EXIT' in d[1]: flush(i + 1)
    flush(len(data))
