#!/usr/bin/env python
"""Turns `ncu --set full` captures of the search kernels (strand_filter_kernel + search_reads_kernel, one launch each,
captured with tools/kernel_probe.py as the workload) into
  profiles/<name>_ncu_full.txt            the raw metrics bench.py's roofline cites, stall totals, hottest SASS lines
  profiles/search_kernel_traffic.json     one entry per workload (genome, k): DRAM bytes per read, issue statistics
Read here, no GPU needed.

  python tools/ncu_traffic.py --version "<kernel version text>" --head <git sha> \
      gpurun_out/r2_search_cfg2.ncu-rep:5000000:31:1000000 gpurun_out/r2_search_1gbp_v2.ncu-rep:1000000000:31:4000000
"""
import argparse
import csv
import io
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__warps_eligible.avg.per_cycle_active", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]


def to_bytes(v, unit):
    return float(v) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]


def to_ms(v, unit):
    return float(v) * {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(unit, {"nsecond": 1e-6, "usecond": 1e-3, "msecond": 1, "second": 1e3}.get(unit, 1))


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def source_rows(rep, skip):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(skip), "--launch-count", "1"],
                         capture_output=True, text=True).stdout
    d = list(csv.reader(io.StringIO(out)))
    h = d[1]
    ix = {x: i for i, x in enumerate(h)}
    rows = [r for r in d[2:] if len(r) == len(h) and r[ix["# Samples"]].isdigit()]
    return h, ix, rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--version", required=True)
    ap.add_argument("--head", default="")
    ap.add_argument("--probe-dram-gbs", type=float, default=3900.0,
                    help="DRAM GB/s of data-independent random 32-byte reads (profiles/r1_probe_random_sectors_ncu.txt)")
    ap.add_argument("captures", nargs="+", help="file.ncu-rep:genome:k:reads_in_launch")
    a = ap.parse_args()
    entries = []
    for spec in a.captures:
        rep, genome, k, reads = spec.split(":")
        genome, k, reads = int(genome), int(k), int(reads)
        hdr, units, rows = raw_rows(rep)
        idx = {h: i for i, h in enumerate(hdr)}
        name = os.path.splitext(os.path.basename(rep))[0]
        txt = os.path.join(ROOT, "profiles", name + "_ncu_full.txt")
        entry = {"genome": genome, "k": k, "reads_in_captured_launch": reads, "version": a.version, "head": a.head,
                 "source": f"ncu --set full --clock-control none, profiles/{name}_ncu_full.txt", "kernels": [],
                 "probe_dram_gbs_at_32B": a.probe_dram_gbs}
        with open(txt, "w") as f:
            f.write(f"# {rep}: genome {genome}, k {k}, {reads} reads in the captured launch; kernel version: {a.version}; head {a.head}\n")
            total_dram = 0.0
            for li, r in enumerate(rows):
                kname = r[idx["Kernel Name"]]
                f.write(f"\n== launch {li}: {kname}\n")
                for w in WANT:
                    if w in idx:
                        f.write(f"{w} [{units[idx[w]]}] = {r[idx[w]]}\n")
                rd = to_bytes(r[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]])
                wr = to_bytes(r[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
                ms = to_ms(r[idx["gpu__time_duration.sum"]], units[idx["gpu__time_duration.sum"]])
                total_dram += rd + wr
                kd = {"kernel": kname.split("(")[0], "ms_under_ncu": ms, "dram_bytes_read": rd, "dram_bytes_written": wr}
                entry["kernels"].append(kd)
                h, ix, src = source_rows(rep, li)
                tot = sum(int(x[ix["# Samples"]]) for x in src)
                f.write(f"sass lines {len(src)}, samples {tot}\n# warp stall samples (all)\n")
                stalls = {}
                for st in [c for c in h if c.startswith("stall_") and "Not" not in c]:
                    stalls[st] = sum(int(x[ix[st]]) for x in src)
                    f.write(f"  {st} {stalls[st]}\n")
                if "search_reads_kernel" in kname:
                    entry.update(issue_active_pct=float(r[idx["smsp__issue_active.avg.pct_of_peak_sustained_active"]]),
                                 lanes_per_instruction=float(r[idx["smsp__thread_inst_executed_per_inst_executed.ratio"]]),
                                 l1_hit_pct=float(r[idx["l1tex__t_sector_hit_rate.pct"]]), l2_hit_pct=float(r[idx["lts__t_sector_hit_rate.pct"]]),
                                 stall_long_sb_share=stalls.get("stall_long_sb", 0) / max(1, tot),
                                 warps_active_pct=float(r[idx["sm__warps_active.avg.pct_of_peak_sustained_active"]]))
                f.write("# top 30 SASS by samples: samples, warp-instr, avg threads, stall_long_sb, sass\n")
                seen = set()
                for x in sorted(src, key=lambda x: -int(x[ix["# Samples"]])):
                    key = (x[0], x[1])
                    if key in seen:
                        continue
                    seen.add(key)
                    f.write(f"{x[ix['# Samples']]} {x[ix['Instructions Executed']]} {x[ix['Avg. Threads Executed']]} {x[ix['stall_long_sb']]} {x[1].strip()}\n")
                    if len(seen) >= 30:
                        break
        entry["dram_bytes_per_launch"] = total_dram
        entry["dram_bytes_per_read"] = total_dram / reads
        entry["limiter"] = ("instruction issue under divergence; the index is L2-resident" if genome < 50_000_000 else
                            "memory latency (long-scoreboard stalls) of dependent random sectors from HBM")
        entries.append(entry)
    out = os.path.join(ROOT, "profiles", "search_kernel_traffic.json")
    with open(out, "w") as f:
        json.dump(entries, f, indent=1)
    print(json.dumps(entries, indent=1)[:3000])


if __name__ == "__main__":
    main()
