#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics, stall totals, hottest SASS lines.
usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [n_top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__occupancy_limit', 'launch__grid_size',
        'sm__throughput.avg.pct', 'l1tex__throughput.avg.pct', 'lts__throughput.avg.pct', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum',
        'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio', 'sass__inst_executed_local',
        'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum', 'l1tex__t_sector_pipe_lsu_mem_local', 'smsp__cycles_active.avg',
        'sm__cycles_elapsed.max', 'smsp__warps_eligible.avg.per_cycle_active', 'smsp__issue_inst0']
print("kernel:", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
for h, u, v in zip(hdr, units, vals):
    if any(w in h for w in want) and '.per_second' not in h and 'pct_of_peak_sustained_elapsed' not in h.replace('throughput.avg.pct_of_peak_sustained_elapsed', ''):
        print(f"{h} [{u}] = {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
d = list(csv.reader(io.StringIO(src)))
h = d[1]; d = d[2:]; ix = {x: i for i, x in enumerate(h)}
print("\nsass lines", len(d), "samples", sum(int(r[ix['# Samples']]) for r in d))
print("# warp stall samples (all)")
for st in ['stall_long_sb', 'stall_wait', 'stall_selected', 'stall_not_selected', 'stall_branch_resolving', 'stall_short_sb', 'stall_math', 'stall_no_inst',
           'stall_dispatch', 'stall_mio', 'stall_lg', 'stall_barrier', 'stall_sleeping', 'stall_tex', 'stall_imc', 'stall_membar', 'stall_drain', 'stall_misc']:
    if st in ix: print(" ", st, sum(int(r[ix[st]]) for r in d))
print(f"# top {ntop} SASS by samples: samples, warp-instr, avg threads, sass")
for r in sorted(d, key=lambda r: -int(r[ix['# Samples']]))[:ntop]:
    print(r[ix['# Samples']], r[ix['Instructions Executed']], r[ix['Avg. Threads Executed']], r[1].strip())
