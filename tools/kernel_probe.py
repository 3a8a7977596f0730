#!/usr/bin/env python
"""Quick kernel experiment on the GPU box: device-resident dense / runs search and the two end-to-end calls on one
synthetic workload, CUDA-event times only (no roofline, no CPU baseline).  Used by tools/variant_sweep.sh.

  python tools/kernel_probe.py [--genome G] [--reads R] [--k K] [--steps S] [--no-e2e] [--tag TEXT]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome", type=int, default=5_000_000)
    ap.add_argument("--reads", type=int, default=2_000_000)
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    import torch

    import finito_b200 as fb
    from fin_datasets import make_dataset

    if a.genome >= 100_000_000:
        # large indexes: the device pipeline of tools/big_config.py (bigenome -> fin_spss_build -> fin_sbwt_build -> fin_build)
        import subprocess

        work = "/dev/shm/finwork" if os.path.isdir("/dev/shm") else os.path.join(ROOT, "_work")
        d = os.path.join(work, f"probe_g{a.genome}_k{a.k}")
        os.makedirs(d, exist_ok=True)
        genome, spss, sbwt, prefix, reads_bin = (os.path.join(d, f) for f in ("genome.fna", "spss.fna", "input.sbwt", "fmin", "reads.bin"))
        datagen = os.path.join(ROOT, "tools", "fin_datagen")
        if not os.path.exists(prefix + ".FBV.sdsl"):
            subprocess.run([datagen, "bigenome", genome, str(a.genome), "7", "1", "0.45", "2000"], check=True, capture_output=True)
            fb.spss_build(genome, a.k, spss, flip=True, seed=7)
            fb.sbwt_build(spss, a.k, sbwt, 8)
            fb.build_fmin(sbwt, spss, prefix)
        if not os.path.exists(reads_bin):
            subprocess.run([datagen, "reads", genome, str(a.reads), str(a.read_len), "0.01", "9", reads_bin], check=True, capture_output=True)
        seqs = np.fromfile(reads_bin, dtype=np.uint8)
        offs = np.arange(a.reads + 1, dtype=np.uint64) * np.uint64(a.read_len)
    else:
        ds = make_dataset("bench", genome_len=a.genome, k=a.k, n_reads=a.reads, read_len=a.read_len, seed=20260102,
                          families=max(1, a.genome // 100_000))
        prefix = os.path.join(ds["dir"], "fmin")
        if not os.path.exists(prefix + ".FBV.sdsl"):
            fb.build_fmin(ds["sbwt"], ds["unitigs"], prefix)
        seqs, offs = ds["reads"]
    ix = fb.FinimizerIndex.load(prefix)
    n, nb = a.reads, int(offs[-1])
    nk = n * (a.read_len - a.k + 1)
    d_seqs = torch.from_numpy(seqs).cuda()
    d_offs = torch.from_numpy(offs.astype(np.int64)).cuda()
    d_out = torch.empty((nk, 2), dtype=torch.int32, device="cuda")
    cap = 16 * n
    d_ro = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    d_runs = torch.empty((cap, 4), dtype=torch.int32, device="cuda")
    res = {"tag": a.tag, "genome": a.genome, "reads": n, "k": a.k, "index_mb": ix.info["device_bytes"] / 1e6}
    dense, runs = [], []
    for i in range(a.steps + 2):
        c, _ = ix.search_device(d_seqs.data_ptr(), d_offs.data_ptr(), n, nb, nk, d_out.data_ptr(), 4)
        t = ix.last_timing()
        if i >= 2:
            dense.append(t["search_kernel_ms"])
    for i in range(a.steps + 2):
        n_runs, c2, _ = ix.search_device_runs(d_seqs.data_ptr(), d_offs.data_ptr(), n, nb, d_ro.data_ptr(), d_runs.data_ptr(), cap)
        t = ix.last_timing()
        if i >= 2:
            runs.append(t["search_kernel_ms"])
    assert c == c2
    res.update(dense_kernel_ms=float(np.mean(dense)), runs_kernel_ms=float(np.mean(runs)), dense_gkmers_s=nk / np.mean(dense) / 1e6,
               runs_gkmers_s=nk / np.mean(runs) / 1e6, runs_per_read=n_runs / n, found=c.positive / c.n_queries)
    if not a.no_e2e:
        h_seqs = torch.empty(nb, dtype=torch.uint8).pin_memory()
        h_seqs.numpy()[:] = seqs
        h_out = torch.empty((nk, 2), dtype=torch.int32).pin_memory()
        h_runs = torch.empty((int(n_runs * 1.05) + 1024, 4), dtype=torch.int32).pin_memory()
        h_ro = torch.empty(n + 1, dtype=torch.int64).pin_memory()
        hr = h_runs.numpy().view(np.uint32).reshape(-1).view(fb.RUN_DTYPE)
        ho = h_ro.numpy().view(np.uint64)
        td, tr = [], []
        for i in range(a.steps + 1):
            t0 = time.perf_counter()
            ix.search_batch((h_seqs.numpy(), offs), hit_bits=32, out=h_out.numpy())
            td.append(time.perf_counter() - t0)
        for i in range(a.steps + 1):
            t0 = time.perf_counter()
            ix.search_batch_runs((h_seqs.numpy(), offs), runs=hr, run_offsets=ho)
            tr.append(time.perf_counter() - t0)
        res.update(e2e_dense_ms=float(np.mean(td[1:])) * 1e3, e2e_runs_ms=float(np.mean(tr[1:])) * 1e3)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
