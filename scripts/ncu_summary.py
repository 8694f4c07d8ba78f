#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key throughput metrics, stall reasons, hot source lines."""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
 'sm__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed',
 'sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum',
 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
 'sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active',
 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','launch__registers_per_thread',
 'launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','launch__occupancy_limit_warps','launch__shared_mem_per_block_dynamic','launch__shared_mem_per_block_static',
 'launch__grid_size','launch__block_size','sm__cycles_elapsed.max','l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum','lts__t_sectors_srcunit_tex_op_read.sum']
for r in rows[2:]:
    print("=====", r[hdr.index('Kernel Name')][:90])
    for k in keys:
        if k in hdr: print(f"  {k:78s} {r[hdr.index(k)]:>18s} {units[hdr.index(k)]}")
    st = []
    for i, h in enumerate(hdr):
        if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio'):
            try: st.append((float(r[i].replace(',', '')), h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')))
            except: pass
    print("  stall reasons (warps per issue):", ", ".join(f"{n}={v:.2f}" for v, n in sorted(st, reverse=True)[:8]))
if len(sys.argv) > 2:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = None
    for i, r in enumerate(rows):
        if 'Source' in r and any('Samples' in c for c in r): h = r; start = i; break
    if h:
        si = h.index('Source'); 
        smp = [i for i, c in enumerate(h) if c.startswith('# Samples') or c == 'Warp Stall Sampling (All Samples)' or 'Sampling (All' in c]
        ex = [i for i, c in enumerate(h) if 'Instructions Executed' == c]
        col = smp[0] if smp else None
        data = []
        for r in rows[start+1:]:
            if len(r) != len(h): continue
            try: v = float(r[col])
            except: continue
            data.append((v, r[0] if r[0] else '', r[si][:110]))
        tot = sum(d[0] for d in data) or 1
        print("  hot source lines (sampling share):")
        for v, ln, s in sorted(data, reverse=True)[:int(sys.argv[2])]:
            print(f"   {100*v/tot:5.1f}%  L{ln:>4s} {s}")
