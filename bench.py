#!/usr/bin/env python
"""search-fmin throughput on B200: k-mers localized per second (BASELINE.json metric, config #2:
5 Mbp genome unitigs, k=31, t=1 rarest, 10M simulated 150 bp reads at 1 % error per GPU).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--reads R] [--genome G]

One step = one pass of the hot path over one batch of R reads (both strands, merged).  `value` is measured
with the batch already in HBM (fin_search_device, CUDA events inside the library, on the stream the kernels
run on); `e2e` goes through fin_search_batch32 with pinned HOST buffers, copies included.  Under torchrun every
rank searches its own R reads against a replicated index (weak scaling, no data-path collective).
--impl reference times the reference's own CPU implementation (oracle/_ref/finito_ref, one process per host
core over disjoint read shards) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))

METRIC = "search-fmin k-mers localized/sec"
UNIT = "k-mers/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reads", type=int, default=10_000_000, help="reads per GPU per step")
    ap.add_argument("--genome", type=int, default=5_000_000)
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--cpu-sample-reads", type=int, default=0, help="reads of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-build-arm", action="store_true", help="skip the build-fmin bp/s arm (kernel experiments)")
    ap.add_argument("--no-hbm-arm", action="store_true", help="skip the second roofline entry (index far beyond the L2)")
    ap.add_argument("--hbm-genome", type=int, default=1_000_000_000)
    ap.add_argument("--hbm-reads", type=int, default=4_000_000)
    return ap.parse_args()


def dataset(args, rank=0):
    from fin_datasets import make_dataset

    if args.genome >= 50_000_000:
        return device_built_dataset(args, rank)
    # planted repeats as in SURVEY.md §8(d) config 2: 50 families x 5 copies x 2 kb at 1 % divergence per 5 Mbp
    return make_dataset("bench", genome_len=args.genome, k=args.k, n_reads=args.reads, read_len=args.read_len,
                        seed=20260102 + rank, families=max(1, args.genome // 100_000))


def device_built_dataset(args, rank):
    """Large inputs (configs[3] / [4]): the genome comes from tools/fin_datagen bigenome, the disjoint SPSS ("flipped unitigs") and
    its plain-matrix SBWT from the device pipeline fin_spss_build -> fin_sbwt_build (seconds instead of the minutes the CPU
    generator needs); the reads are simulated by tools/fin_datagen from the same genome.  Rank 0's genome is shared by all
    ranks (one index), every rank draws its own reads."""
    import finito_b200 as fb

    work = os.environ.get("FIN_WORK") or ("/dev/shm/finwork" if os.path.isdir("/dev/shm") else os.path.join(ROOT, "_work"))
    d = os.path.join(work, f"benchbig_g{args.genome}_k{args.k}")
    os.makedirs(d, exist_ok=True)
    genome, spss, sbwt = (os.path.join(d, f) for f in ("genome.fna", "unitigs.fna", "index.sbwt"))
    reads_bin = os.path.join(d, f"reads_r{rank}_{args.reads}x{args.read_len}.bin")
    datagen = os.path.join(ROOT, "tools", "fin_datagen")
    done = os.path.join(d, ".done")
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if rank == 0 and not os.path.exists(done):
        subprocess.run([datagen, "bigenome", genome, str(args.genome), "20260104", "1", "0.30", str(max(2000, args.genome // 50_000))],
                       check=True, capture_output=True)
        fb.spss_build(genome, args.k, spss, flip=True, seed=7, device=local_rank)
        fb.sbwt_build(spss, args.k, sbwt, 8, device=local_rank)
        open(done, "w").close()
    while not os.path.exists(done):  # the other ranks of the node wait for rank 0's files
        time.sleep(0.2)
    if not os.path.exists(reads_bin):  # several ranks may want rank 0's reads at once: write aside, rename
        tmp = f"{reads_bin}.{os.getpid()}.bin"
        subprocess.run([datagen, "reads", genome, str(args.reads), str(args.read_len), "0.01", str(9 + rank), tmp], check=True, capture_output=True)
        os.replace(tmp, reads_bin)
    seqs = np.fromfile(reads_bin, dtype=np.uint8)
    offs = np.arange(args.reads + 1, dtype=np.uint64) * np.uint64(args.read_len)
    return {"dir": d, "genome": genome, "unitigs": spss, "sbwt": sbwt, "reads_bin": reads_bin, "reads": (seqs, offs), "k": args.k,
            "n_reads": args.reads, "read_len": args.read_len}


def workload_name(args):
    which = "configs[1]" if (args.genome, args.k) == (5_000_000, 31) else "configs[3]" if args.k == 63 else "configs[4]" if args.genome >= 1_000_000_000 else "other size"
    return (f"synthetic {args.genome / 1e6:g} Mbp genome with planted repeats, flipped unitigs, k={args.k}, t=1 rarest, "
            f"{args.reads} simulated {args.read_len} bp reads at 1% error per GPU (BASELINE.json {which})")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons during the timed region."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device = device
        self.rows = []
        self.stop_flag = threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                r = subprocess.run(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5)
                if r.returncode == 0 and r.stdout.strip():
                    self.rows.append([x.strip() for x in r.stdout.strip().split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) > 2 + i and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def algorithmic_bytes(acc, n_bases_in, n_kmers, hit_bytes=8):
    """SURVEY.md §8(d) sector model: one 32-byte sector per semantically required random access."""
    return (32 * (acc["n_rank"] + acc["n_lcs_reads"] + acc["n_ustart_reads"] + 3 * acc["n_anchor"] + acc["n_walk_words"])
            + 0.25 * n_bases_in + hit_bytes * n_kmers)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    return 6650.0, "B200_PROFILING.md fallback (of fallback)"


def profiled_traffic():
    """dram bytes per launch of the search kernel from the committed ncu capture, if there is one."""
    p = os.path.join(ROOT, "profiles", "search_kernel_traffic.json")
    if os.path.exists(p):
        return json.load(open(p))  # one capture, or a list of captures (one per workload)
    return None


# ---------------------------------------------------------------------------------------------------------
# reference CPU implementation
# ---------------------------------------------------------------------------------------------------------
def reference_binary():
    p = os.path.join(ROOT, "oracle", "_ref", "finito_ref")
    return p if os.path.exists(p) else None


def cpu_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return os.cpu_count() or 1


def write_fasta_shards(seqs, read_len, n_reads, shards, out_dir):
    paths = []
    per = (n_reads + shards - 1) // shards
    arr = seqs[: n_reads * read_len].reshape(n_reads, read_len)
    for s in range(shards):
        a, b = s * per, min(n_reads, (s + 1) * per)
        if a >= b:
            break
        p = os.path.join(out_dir, f"cpu_shard{s}.fna")
        rows = arr[a:b]
        rec = np.empty((b - a, read_len + 3), dtype=np.uint8)
        rec[:, 0] = ord(">")
        rec[:, 1] = ord("\n")
        rec[:, 2:2 + read_len] = rows
        rec[:, -1] = ord("\n")
        rec.tofile(p)
        paths.append((p, b - a))
    return paths


def run_reference_sample(index_prefix, seqs, read_len, k, n_reads, threads, work_dir):
    """P reference processes over P disjoint shards; returns (k-mers/s over the wall time of the slowest, detail)."""
    ref = reference_binary()
    shards = write_fasta_shards(seqs, read_len, n_reads, threads, work_dir)
    t0 = time.perf_counter()
    procs = [subprocess.Popen([ref, "search-fmin", "-i", index_prefix, "-q", p, "-o", p + ".out"], stdout=subprocess.DEVNULL,
                              stderr=subprocess.PIPE, text=True) for p, _ in shards]
    logs = [pr.communicate()[1] for pr in procs]
    wall = time.perf_counter() - t0
    for pr in procs:
        if pr.returncode != 0:
            raise RuntimeError("reference search-fmin failed: " + logs[0][-500:])
    kmers = sum(n for _, n in shards) * (read_len - k + 1)
    us = [float(m.group(1)) for log in logs for m in re.finditer(r"us/query: ([0-9.]+) \(excluding", log)]
    for p, _ in shards:
        for f in (p, p + ".out"):
            if os.path.exists(f):
                os.remove(f)
    for junk in (index_prefix + ".stats", index_prefix + "stats.txt"):
        if os.path.exists(junk):
            os.remove(junk)
    return kmers / wall, {"wall_s": wall, "kmers": kmers, "processes": len(shards),
                          "ref_us_per_kmer_excl_io": float(np.mean(us)) if us else None}


def oracle_port_sample(index_prefix, seqs, offs, n_reads):
    import helpers as H

    orc = H.Oracle(index_prefix)
    t0 = time.perf_counter()
    _, c = orc.search_reads(seqs[: int(offs[n_reads])], offs[: n_reads + 1])
    wall = time.perf_counter() - t0
    return int(c[3]) / wall, {"wall_s": wall, "kmers": int(c[3]), "processes": 1}


def cpu_baseline(index_prefix, ds, args, all_threads, sample_reads=None):
    seqs, offs = ds["reads"]
    threads = cpu_threads() if all_threads else 1
    per_core_rate = 1.0e6  # k-mers/s, survey figure; only sizes the sample
    if not sample_reads:
        sample_reads = int(min(args.reads, max(2000, 15 * per_core_rate * threads / (args.read_len - args.k + 1))))
    if reference_binary():
        v, detail = run_reference_sample(index_prefix, seqs, args.read_len, args.k, sample_reads, threads, ds["dir"])
        kind = "reference"
    else:
        v, detail = oracle_port_sample(index_prefix, seqs, offs, min(sample_reads, 100_000))
        kind, threads = "port", 1
    return {"value": v, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"first {sample_reads} reads of the workload ({detail['kmers']} k-mer positions) in {detail['processes']} "
                      f"process(es), wall {detail['wall_s']:.1f} s incl. index load and text output",
            "detail": detail}


def unitig_bases(ds):
    n = 0
    with open(ds["unitigs"], "rb") as f:
        for line in f:
            if not line.startswith(b">"):
                n += len(line.strip())
    return n


def reference_build(ds, out_prefix):
    """build-fmin of the unmodified reference binary (LCS on 8 OpenMP threads, the rest on one core), wall time."""
    ref = reference_binary()
    if not ref:
        return None
    t0 = time.perf_counter()
    r = subprocess.run([ref, "build-fmin", "-o", out_prefix, "-i", ds["sbwt"], "-u", ds["unitigs"]], capture_output=True, text=True)
    wall = time.perf_counter() - t0
    if r.returncode != 0:
        return None
    bp = unitig_bases(ds)
    return {"value": bp / wall, "unit": "bp/s", "wall_s": wall, "unitig_bp": bp, "kind": "reference",
            "threads": "8 OpenMP threads for the LCS, 1 core for the rest (as shipped)"}


def build_arm(fb, ds, rank, world, local_rank, steps, barrier):
    """BASELINE.json's second metric: build-fmin unitig bp/sec = unitig bases / wall time of the whole build-fmin call
    (SBWT + unitig file parsing, LCS, permutation, stream, merge, the seven index files written) through the C ABI.
    N > 1: fin_builder_* sharded over the ranks with one NCCL all-to-all (finito_b200/distributed.py)."""
    import torch

    bp = unitig_bases(ds)
    out = os.path.join(ds["dir"], "bench_build")
    walls, stages = [], None
    big = bp >= 1_000_000_000  # human-scale inputs: one timed build, no warm-up build (each takes tens of seconds)
    if big:
        steps = 1
    for i in range((0 if big else 1) + steps):  # one untimed warm-up build
        barrier()
        t0 = time.perf_counter()
        if world == 1:
            fb.build_fmin(ds["sbwt"], ds["unitigs"], out, device=local_rank)
            stages = fb.build_timing()
        else:
            from finito_b200.distributed import build_fmin_sharded

            b = fb.ShardedBuilder(ds["sbwt"], ds["unitigs"], out, rank, world, device=local_rank)
            stages = build_fmin_sharded(b)
            b.close()
        barrier()
        if i or big:
            walls.append(time.perf_counter() - t0)
    wall = float(np.mean(walls))
    parity = None
    if world > 1:
        import torch.distributed as dist

        t = torch.tensor([wall], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall = float(t.item())
        # parity of the NCCL path on this very hardware: the seven files the sharded build wrote against a single-GPU
        # fin_build of the same inputs (itself pinned byte for byte to the reference's files by tests/)
        if rank == 0:
            single = os.path.join(ds["dir"], "bench_build_single")
            fb.build_fmin(ds["sbwt"], ds["unitigs"], single, device=local_rank)
            sha = {suf: (file_sha256(out + suf), file_sha256(single + suf)) for suf in INDEX_SUFFIXES}
            parity = {"parity_ok": all(a == b for a, b in sha.values()),
                      "sha256_sharded": {suf: a[:16] for suf, (a, _) in sha.items()},
                      "sha256_single_gpu": {suf: b[:16] for suf, (_, b) in sha.items()},
                      "compared": f"{len(sha)} index files of the {world}-rank NCCL build vs fin_build on one GPU, same inputs"}
        barrier()
    res = {"metric": "build-fmin unitig bp/sec", "value": bp / wall, "unit": "bp/s", "wall_s": wall, "unitig_bp": bp, "builds_timed": steps,
           "includes": "file parsing, LCS (device), permutation, stream, merge, index files written", "stages": stages}
    if parity:
        res.update(parity)
    return res


INDEX_SUFFIXES = [".O.sdsl", ".FBV.sdsl", ".packed_unitigs.sdsl", ".unitig_endpoints.sdsl", ".Ustart.sdsl", ".LCS.sdsl", ".sbwt"]


def file_sha256(path):
    import hashlib

    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


def sharded_search_parity(fb, ix, ds, rank, world, n_reads=200_000):
    """N > 1 only: the first n_reads of rank 0's workload searched (a) sharded over all ranks through
    finito_b200.distributed.search_sharded and (b) on rank 0's GPU alone; hits and the four counters must be equal."""
    import hashlib

    from finito_b200.distributed import search_sharded

    seqs, offs = ds["reads"]
    n = min(n_reads, len(offs) - 1)
    seqs, offs = seqs[: int(offs[n])], offs[: n + 1]
    all_hits, counts, _ = search_sharded(ix, seqs, offs, hit_bits=32)
    if rank != 0:
        return None
    want, c = ix.search_batch((seqs, offs), hit_bits=32)
    wc = [c.found_fwd, c.found_rc, c.positive, c.n_queries]
    return {"parity_ok": bool(np.array_equal(all_hits, want) and counts == wc), "reads": int(n), "counters_sharded": counts,
            "counters_single_gpu": wc, "sha256_hits_sharded": hashlib.sha256(np.ascontiguousarray(all_hits).tobytes()).hexdigest()[:16],
            "sha256_hits_single_gpu": hashlib.sha256(np.ascontiguousarray(want).tobytes()).hexdigest()[:16]}


def bind_to_gpu_numa_node(device):
    """One process per GPU: run on the cores next to that GPU (NVML's ideal CPU set) so that the pinned host buffers
    of the end-to-end arm are allocated on the GPU's own NUMA node.  Best effort; FIN_BENCH_NO_AFFINITY=1 turns it off."""
    if os.environ.get("FIN_BENCH_NO_AFFINITY") == "1":
        return None
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus and len(cpus) < n_cpu:
            os.sched_setaffinity(0, cpus)
            return f"{len(cpus)} of {n_cpu} cpus"
    except Exception:
        pass
    return None


# ---------------------------------------------------------------------------------------------------------
def search_arms(fb, ix, seqs, offs, k, read_len, n_reads, warmup, steps, barrier, with_e2e=True):
    """Device-resident dense + runs arms (CUDA events inside the library) and the two end-to-end calls (pinned host buffers)
    on one workload.  Returns times, counts and the run statistics."""
    import torch

    n_bases = int(offs[n_reads])
    kmers = n_reads * (read_len - k + 1)
    d_seqs = torch.from_numpy(seqs[:n_bases]).cuda()
    d_offs = torch.from_numpy(offs[: n_reads + 1].astype(np.int64)).cuda()
    d_out = torch.empty((kmers, 2), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    res = {"kmers": kmers, "n_bases": n_bases}
    # ---- dense table in HBM (fin_search_device)
    for _ in range(warmup):
        ix.search_device(d_seqs.data_ptr(), d_offs.data_ptr(), n_reads, n_bases, kmers, d_out.data_ptr(), 4)
    barrier()
    launches0 = fb.launch_count()
    dense_ms, dense_kernel_ms = [], []
    t0 = time.perf_counter()
    for _ in range(steps):
        counts, ms = ix.search_device(d_seqs.data_ptr(), d_offs.data_ptr(), n_reads, n_bases, kmers, d_out.data_ptr(), 4)
        dense_ms.append(ms)
        dense_kernel_ms.append(ix.last_timing()["search_kernel_ms"])
    barrier()
    res.update(wall_dense_s=time.perf_counter() - t0, launches_dense=fb.launch_count() - launches0, counts=counts,
               dense_ms=dense_ms, dense_kernel_ms=dense_kernel_ms)
    # ---- run-length records in HBM (fin_search_device_runs)
    runs_cap = 16 * n_reads
    d_ro = torch.empty(n_reads + 1, dtype=torch.int64, device="cuda")
    d_runs = torch.empty((runs_cap, 4), dtype=torch.int32, device="cuda")
    for _ in range(warmup):
        ix.search_device_runs(d_seqs.data_ptr(), d_offs.data_ptr(), n_reads, n_bases, d_ro.data_ptr(), d_runs.data_ptr(), runs_cap)
    barrier()
    launches0 = fb.launch_count()
    runs_ms, runs_kernel_ms = [], []
    for _ in range(steps):
        n_runs, counts_runs, ms = ix.search_device_runs(d_seqs.data_ptr(), d_offs.data_ptr(), n_reads, n_bases, d_ro.data_ptr(),
                                                        d_runs.data_ptr(), runs_cap)
        runs_ms.append(ms)
        runs_kernel_ms.append(ix.last_timing()["search_kernel_ms"])
    barrier()
    assert counts_runs == counts
    res.update(launches_runs=fb.launch_count() - launches0, runs_ms=runs_ms, runs_kernel_ms=runs_kernel_ms, n_runs=n_runs)
    # ---- access counters of the counting kernel variant on a sample (algorithmic bytes of the roofline)
    sample = min(n_reads, 1_000_000)
    ix.set_counting(True)
    ix.search_device(d_seqs.data_ptr(), d_offs.data_ptr(), sample, sample * read_len, sample * (read_len - k + 1), d_out.data_ptr(), 4)
    res["access"] = {k_: v * (n_reads / sample) for k_, v in ix.access_counters().items()}
    ix.set_counting(False)
    del d_runs, d_ro
    if not with_e2e:
        return res
    # ---- end to end, answers as runs: pinned host buffers through fin_search_batch_runs
    h_seqs = torch.empty(n_bases, dtype=torch.uint8).pin_memory()
    h_seqs.numpy()[:] = seqs[:n_bases]
    h_seqs_np = h_seqs.numpy()
    offs_n = np.ascontiguousarray(offs[: n_reads + 1])
    h_runs = torch.empty((int(n_runs * 1.05) + 1024, 4), dtype=torch.int32).pin_memory()
    h_ro = torch.empty(n_reads + 1, dtype=torch.int64).pin_memory()
    h_runs_np = h_runs.numpy().view(np.uint32).reshape(-1).view(fb.RUN_DTYPE)
    h_ro_np = h_ro.numpy().view(np.uint64)
    for _ in range(min(warmup, 2)):
        ix.search_batch_runs((h_seqs_np, offs_n), runs=h_runs_np, run_offsets=h_ro_np)
    barrier()
    e2e_runs = []
    for _ in range(steps):
        t1 = time.perf_counter()
        _, runs_e, c_e = ix.search_batch_runs((h_seqs_np, offs_n), runs=h_runs_np, run_offsets=h_ro_np)
        e2e_runs.append(time.perf_counter() - t1)
    barrier()
    assert c_e == counts and len(runs_e) == n_runs
    # what a caller that wants the dense table pays on the host (fin_expand_runs on all host threads), on a sample
    sample = min(n_reads, 1_000_000)
    t1 = time.perf_counter()
    fb.expand_runs(offs_n[: sample + 1], k, h_ro_np[: sample + 1], h_runs_np, hit_bits=32, threads=cpu_threads())
    res["expand_ms_per_step"] = (time.perf_counter() - t1) / sample * n_reads * 1e3
    # ---- end to end, dense table: fin_search_batch32 (8 bytes per k-mer over PCIe)
    h_out = torch.empty((kmers, 2), dtype=torch.int32).pin_memory()
    h_out_np = h_out.numpy()
    for _ in range(min(warmup, 2)):
        ix.search_batch((h_seqs_np, offs_n), hit_bits=32, out=h_out_np)
    barrier()
    e2e_dense = []
    for _ in range(steps):
        t1 = time.perf_counter()
        _, c_d = ix.search_batch((h_seqs_np, offs_n), hit_bits=32, out=h_out_np)
        e2e_dense.append(time.perf_counter() - t1)
    barrier()
    assert c_d == counts
    link = []
    for _ in range(2):  # what the host link gives for the dense result buffer in one plain copy
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        link.append(h_out.numel() * 4 / (time.perf_counter() - t1) / 1e9)
    res.update(e2e_runs_s=e2e_runs, e2e_dense_s=e2e_dense, d2h_link_gbs=max(link))
    return res


def algorithmic_bytes_breakdown(acc, n_bases_in, out_bytes, reference=False):
    """SURVEY.md §8(d) sector model (+ the strand prefilter's table lookups, one sector each).  The access counters tally what
    the REFERENCE's loops touch; `reference=False` leaves out the rank rows and LCS values of the drops / retried extensions
    that this library's closed forms never issue (n_rank_skipped, n_lcs_skipped): the bytes THIS algorithm needs."""
    n_rank = acc["n_rank"] - (0 if reference else acc.get("n_rank_skipped", 0))
    n_lcs = acc["n_lcs_reads"] - (0 if reference else acc.get("n_lcs_skipped", 0))
    parts = {"rank_rows": 32 * n_rank, "lcs_words": 32 * n_lcs, "ustart": 32 * acc["n_ustart_reads"],
             "anchor_lookups": 96 * acc["n_anchor"], "walk_words": 32 * acc["n_walk_words"],
             "prefilter_lookups": 32 * acc.get("n_filter_lookups", 0), "packed_reads": 0.25 * n_bases_in, "answers": out_bytes}
    return float(sum(parts.values())), parts


def committed_capture(genome, k):
    """DRAM traffic and issue statistics of the search kernel from the committed ncu capture of this workload (a profiler
    cannot run inside the timed bench): profiles/search_kernel_traffic.json, one entry per (genome, k)."""
    t = profiled_traffic()
    if isinstance(t, dict):
        t = [t]
    return next((x for x in t or [] if int(x.get("genome", 5_000_000)) == genome and int(x.get("k", 31)) == k), None)


def roofline_entry(fb, res, info, args_k, read_len, n_reads, genome, world, peak, peak_src, random_peaks):
    kmers = res["kmers"]
    kernel_s = float(np.mean(res["dense_kernel_ms"])) / 1e3
    total, parts = algorithmic_bytes_breakdown(res["access"], res["n_bases"], 8 * kmers)
    ref_total, _ = algorithmic_bytes_breakdown(res["access"], res["n_bases"], 8 * kmers, reference=True)
    achieved = total / kernel_s / 1e9
    cap = committed_capture(genome, args_k)
    l2_bytes = 126e6
    resident = info["device_bytes"] < l2_bytes
    entry = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
             "traffic": cap["dram_bytes_per_read"] * n_reads if cap else None,
             "kernel": "strand_filter_kernel + search_reads_kernel (dense)", "kernel_ms": kernel_s * 1e3,
             "kernel_share_of_step": float(np.sum(res["dense_kernel_ms"]) / np.sum(res["dense_ms"])),
             "algorithmic_bytes_per_launch": total, "algorithmic_bytes_per_kmer": total / kmers,
             "algorithmic_bytes_breakdown_per_kmer": {k_: v / kmers for k_, v in parts.items()},
             "algorithmic_bytes_note": "sector model of SURVEY.md section 8(d) over the accesses this kernel's algorithm needs: the drops and retried "
                                       "extensions that the closed-form shrink / contraction skip are not billed",
             "reference_algorithm_bytes_per_kmer": ref_total / kmers, "frac_reference_algorithm": ref_total / kernel_s / 1e9 / peak,
             "peak_source": peak_src, "index_device_bytes": info["device_bytes"],
             "regime": ("index L2-resident: the sector model bills every L1/L2 hit as HBM traffic, so `frac` is NOT a DRAM utilisation; "
                        "the kernel is limited by instruction issue under divergence (see `limiter`)") if resident else
                       f"index {info['device_bytes'] / l2_bytes:.0f}x the 126 MB L2: random sectors come from HBM"}
    if cap:
        dram_gbs = cap["dram_bytes_per_read"] * n_reads / kernel_s / 1e9
        entry.update(traffic_source=f"{cap['source']} (kernel version {cap['version']}), {cap['reads_in_captured_launch']}-read launch scaled per read",
                     dram_gbs=dram_gbs, frac_of_copy_peak=dram_gbs / peak, limiter=cap.get("limiter"),
                     issue_active_pct=cap.get("issue_active_pct"), lanes_per_instruction=cap.get("lanes_per_instruction"),
                     stall_long_scoreboard_share=cap.get("stall_long_sb_share"), l1_hit_pct=cap.get("l1_hit_pct"), l2_hit_pct=cap.get("l2_hit_pct"))
        if random_peaks and cap.get("probe_dram_gbs_at_32B"):
            entry["frac_of_random_access_ceiling"] = dram_gbs / cap["probe_dram_gbs_at_32B"]
            entry["random_access_ceiling_gbs"] = cap["probe_dram_gbs_at_32B"]
    return entry


def hbm_regime_arm(fb, args, local_rank, barrier, peak, peak_src, random_peaks):
    """Second roofline entry: the same kernels on an index far beyond the L2 (1 Gbp genome, ~3.8 GB on the device), built
    in this run by the device pipeline fin_spss_build -> fin_sbwt_build -> fin_build (tools/fin_datagen only makes the genome
    and the reads).  No end-to-end arm, no CPU baseline: this entry is about the kernel."""
    G, n_reads = args.hbm_genome, args.hbm_reads
    work = os.environ.get("FIN_WORK") or ("/dev/shm/finwork" if os.path.isdir("/dev/shm") else os.path.join(ROOT, "_work"))
    d = os.path.join(work, f"bench_hbm_g{G}_k{args.k}")
    os.makedirs(d, exist_ok=True)
    genome, spss, sbwt, prefix, reads_bin = (os.path.join(d, f) for f in ("genome.fna", "spss.fna", "input.sbwt", "fmin", "reads.bin"))
    datagen = os.path.join(ROOT, "tools", "fin_datagen")
    t0 = time.perf_counter()
    timings = {}
    if not os.path.exists(prefix + ".FBV.sdsl"):
        subprocess.run([datagen, "bigenome", genome, str(G), "20260104", "1", "0.45", "2000"], check=True, capture_output=True)
        timings["spss"] = fb.spss_build(genome, args.k, spss, flip=True, seed=7, device=local_rank)
        timings["sbwt_build"] = fb.sbwt_build(spss, args.k, sbwt, 8, device=local_rank)
        t1 = time.perf_counter()
        fb.build_fmin(sbwt, spss, prefix, device=local_rank)
        timings["build_fmin_s"] = time.perf_counter() - t1
        timings["build_fmin_bp_per_s"] = timings["spss"]["bases"] / timings["build_fmin_s"]
        timings["build_fmin_stages"] = fb.build_timing()
    if not os.path.exists(reads_bin):
        subprocess.run([datagen, "reads", genome, str(n_reads), str(args.read_len), "0.01", "9", reads_bin], check=True, capture_output=True)
    seqs = np.fromfile(reads_bin, dtype=np.uint8)
    offs = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(args.read_len)
    ix = fb.FinimizerIndex.load(prefix, device=local_rank)
    info = ix.info
    timings["prepare_s"] = time.perf_counter() - t0
    res = search_arms(fb, ix, seqs, offs, args.k, args.read_len, n_reads, 2, max(2, min(args.steps, 3)), barrier, with_e2e=False)
    entry = roofline_entry(fb, res, info, args.k, args.read_len, n_reads, G, 1, peak, peak_src, random_peaks)
    kmers = res["kmers"]
    entry.update(workload=f"synthetic {G / 1e6:g} Mbp genome (45 % repeats), disjoint SPSS of its canonical k-mers, k={args.k}, {n_reads} simulated "
                          f"{args.read_len} bp reads at 1% error",
                 value=kmers / np.mean(res["dense_ms"]) * 1e3, runs_value=kmers / np.mean(res["runs_ms"]) * 1e3, unit_value=UNIT,
                 found_fraction=res["counts"].positive / res["counts"].n_queries, runs_per_read=res["n_runs"] / n_reads,
                 index_nodes=info["n_nodes"], pipeline=timings)
    ix.close()
    for f in os.listdir(d):
        os.remove(os.path.join(d, f))
    os.rmdir(d)
    return entry


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return
        reference_arm(args)
        return

    import torch
    import torch.distributed as dist

    import finito_b200 as fb

    if not torch.cuda.is_available() or fb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    affinity = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    ds = dataset(args, rank)
    prefix = os.path.join(ds["dir"], "fmin")  # not "index": <prefix>.sbwt would overwrite the input index.sbwt
    shared_index = args.genome >= 50_000_000  # device-built datasets: one index for all ranks, built by rank 0
    built = prefix + ".built"
    build_s = None
    if not os.path.exists(built) and (rank == 0 or not shared_index):
        t0 = time.perf_counter()
        fb.build_fmin(ds["sbwt"], ds["unitigs"], prefix, device=local_rank)
        build_s = time.perf_counter() - t0
        open(built, "w").close()
    while not os.path.exists(built):
        time.sleep(0.2)
    ix = fb.FinimizerIndex.load(prefix, device=local_rank)
    info = ix.info
    seqs, offs = ds["reads"]
    n_reads = args.reads
    n_bases = int(offs[-1])
    kmers_per_step = n_reads * (args.read_len - args.k + 1)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    sampler.start()
    res = search_arms(fb, ix, seqs, offs, args.k, args.read_len, n_reads, args.warmup, args.steps, barrier)
    sampler.stop_flag.set()
    sampler.join(timeout=3)
    clocks = sampler.summary()
    counts = res["counts"]
    dev_total_ms, runs_total_ms = float(sum(res["dense_ms"])), float(sum(res["runs_ms"]))
    e2e_total, e2e_dense_total = float(sum(res["e2e_runs_s"])), float(sum(res["e2e_dense_s"]))

    # ---- build-fmin arm (every rank builds from rank 0's inputs when sharded)
    build = None
    if not args.no_build_arm:
        build = build_arm(fb, ds if world == 1 else dataset(args, 0), rank, world, local_rank, min(args.steps, 3), barrier)

    # ---- N > 1: the sharded search against the single-GPU search on the same reads (every rank uses rank 0's reads)
    search_parity = None
    if world > 1:
        ds0 = dataset(args, 0)
        pfx0 = os.path.join(ds0["dir"], "fmin")
        if rank == 0 and not shared_index and not os.path.exists(pfx0 + ".FBV.sdsl"):
            fb.build_fmin(ds0["sbwt"], ds0["unitigs"], pfx0, device=local_rank)
        barrier()
        same = rank == 0 or shared_index  # device-built datasets: every rank already holds rank 0's index
        ix0 = ix if same else fb.FinimizerIndex.load(pfx0, device=local_rank)
        search_parity = sharded_search_parity(fb, ix0, ds0, rank, world)
        if not same:
            ix0.close()

    # ---- max over ranks
    wall_dev = res["wall_dense_s"]
    if world > 1:
        t = torch.tensor([dev_total_ms, e2e_total, wall_dev, runs_total_ms, e2e_dense_total], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_total_ms, e2e_total, wall_dev, runs_total_ms, e2e_dense_total = (float(x) for x in t.cpu())
    total_kmers = world * kmers_per_step * args.steps
    value = total_kmers / (dev_total_ms / 1e3)

    if rank == 0:
        peak, peak_src = measured_peak()
        try:
            random_peaks = {b: fb.probe_random_sectors(local_rank, 8 << 30, 3, b) for b in (32, 64, 128)}
        except fb.FinitoError:
            random_peaks = None
        roofline = roofline_entry(fb, res, info, args.k, args.read_len, n_reads, args.genome, world, peak, peak_src, random_peaks)
        roofline["random_access_probe_gbs_by_bytes"] = random_peaks
        roofline["random_access_probe_note"] = ("fin_probe_random_sectors, measured in this run: REQUESTED bytes/s of data-independent random 32/64/128-byte reads "
                                                "over an 8 GiB table; the DRAM moves ~4x the requested bytes at 32 B (profiles/r1_probe_random_sectors_ncu.txt), "
                                                "that DRAM rate is the random-access ceiling `frac_of_random_access_ceiling` refers to")
        hbm = None
        if world == 1 and not args.no_hbm_arm:
            try:
                hbm = hbm_regime_arm(fb, args, local_rank, barrier, peak, peak_src, random_peaks)
            except Exception as e:  # the headline line must not depend on the side measurement
                hbm = {"unavailable": repr(e)[:300]}
        roofline["hbm_regime"] = hbm
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu = cpu_baseline(prefix, ds, args, all_threads=False, sample_reads=args.cpu_sample_reads)
            allc = cpu_baseline(prefix, ds, args, all_threads=True, sample_reads=0)
            cpu["all_cores"] = {k_: allc[k_] for k_ in ("value", "unit", "cores", "sample")}
            cpu["ref_us_per_kmer_excl_io"] = cpu["detail"].get("ref_us_per_kmer_excl_io")
            cpu["note"] = ("value = k-mer positions answered / wall time of the reference process (incl. index load and text output); "
                           "ref_us_per_kmer_excl_io is the reference's own 'us/query (excluding I/O)' log line; all_cores = one reference process per host core")
            if build:
                build["cpu_baseline"] = reference_build(ds, os.path.join(ds["dir"], "bench_build_ref"))
        n_runs = res["n_runs"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": workload_name(args), "k": args.k, "t": 1, "type": "rarest", "reads_per_gpu_per_step": n_reads,
                       "kmers_per_gpu_per_step": kmers_per_step, "found_fraction": counts.positive / counts.n_queries,
                       "index_nodes": info["n_nodes"], "index_device_bytes": info["device_bytes"], "parallelism": f"reads sharded x{world}, index replicated, no collective", "cpu_affinity": affinity,
                       "l2": "per-step inputs+outputs (1.5 GB reads, 9.6 GB hits at the default size) exceed the 126 MB L2; the index itself is L2-resident by design of this config",
                       "value_is": "fin_search_device: the dense (unitig, offset) table written to HBM, inputs resident in HBM",
                       "index_build_s": build_s},
            "e2e": {"value": total_kmers / e2e_total, "unit": UNIT, "h2d_bytes_per_step": int(n_bases + 8 * (n_reads + 1)),
                    "d2h_bytes_per_step": int(16 * n_runs + 8 * (n_reads + 1)), "ms_per_step": e2e_total / args.steps * 1e3,
                    "api": "fin_search_batch_runs: pinned host buffers in, the complete answer out as run-length records (fin_run, 16 B per run; "
                           "fin_expand_runs turns them into the dense table bit for bit)",
                    "runs_per_read": n_runs / n_reads, "expand_on_host_ms_per_step": res["expand_ms_per_step"], "expand_threads": cpu_threads(),
                    "dense": {"value": total_kmers / e2e_dense_total, "unit": UNIT, "ms_per_step": e2e_dense_total / args.steps * 1e3,
                              "d2h_bytes_per_step": int(kmers_per_step * 8), "api": "fin_search_batch32 (dense table, 8 B per k-mer over PCIe)",
                              "d2h_link_gbs": res["d2h_link_gbs"], "d2h_floor_ms_per_step": kmers_per_step * 8 / res["d2h_link_gbs"] / 1e6}},
            "runs": {"value": total_kmers / (runs_total_ms / 1e3), "unit": UNIT, "ms_per_step": runs_total_ms / args.steps,
                     "search_kernel_ms": float(np.mean(res["runs_kernel_ms"])), "api": "fin_search_device_runs (records sorted by read on the device)"},
            "gpu_launches": int(res["launches_dense"]), "gpu_launches_runs_arm": int(res["launches_runs"]), "clocks": clocks, "roofline": roofline,
            "cpu_baseline": cpu, "build": build, "search_parity": search_parity, "wall_s_device_arm": wall_dev,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def reference_arm(args):
    """The reference's own CPU implementation on this box's host cores; each step = a bounded sample."""
    ds = dataset(args, 0)
    prefix = os.path.join(ds["dir"], "fmin")  # not "index": <prefix>.sbwt would overwrite the input index.sbwt
    if not os.path.exists(prefix + ".FBV.sdsl"):
        # the index files are inputs of the timed path; any builder that writes the reference's format will do
        ref = reference_binary()
        if ref:
            r = subprocess.run([ref, "build-fmin", "-o", prefix, "-i", ds["sbwt"], "-u", ds["unitigs"]], capture_output=True, text=True)
            if r.returncode != 0:
                raise SystemExit("reference build-fmin failed: " + r.stderr[-500:])
        else:
            import helpers as H

            H.oracle_build(ds["sbwt"], ds["unitigs"], prefix)
    values, last = [], None
    n_runs = args.warmup + args.steps
    budget_s = 150.0 / max(1, n_runs)  # whole run within a few minutes
    threads = cpu_threads()
    sample = int(min(args.reads, max(2000, budget_s * 0.9e6 * threads / (args.read_len - args.k + 1))))
    for i in range(n_runs):
        last = cpu_baseline(prefix, ds, args, all_threads=True, sample_reads=sample)
        if i >= args.warmup:
            values.append(last["value"])
    v = float(np.mean(values))
    kmers = last["detail"]["kmers"]
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": kmers / v * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": workload_name(args), "k": args.k, "t": 1, "type": "rarest", "step": last["sample"]},
            "cpu_baseline": {k_: last[k_] for k_ in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    line["cpu_baseline"]["value"] = v
    line["build"] = reference_build(ds, os.path.join(ds["dir"], "bench_build_ref"))
    print(json.dumps(line))


if __name__ == "__main__":
    main()
