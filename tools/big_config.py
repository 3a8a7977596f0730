#!/usr/bin/env python
"""Large configurations end to end on one B200 (BASELINE.json configs[3]/[4]; SURVEY.md section 8(d) cfg 4/5):

  genome (tools/fin_datagen bigenome) -> fin_spss_build -> fin_sbwt_build -> fin_build -> fin_index_load
  -> reads simulated ON THE DEVICE from the genome (uniform start, strand 1/2, iid substitutions)
  -> fin_search_device / fin_search_device_runs, a fresh batch per step (the "read sweep")
  -> every sampled hit verified against the index's own unitig text on the device.

  python tools/big_config.py --genome 3000000000 --k 31 --reads 10000000 --steps 10 [--work DIR]

Prints one JSON line (also usable as a profiles/ record).  Everything between the genome file and the hits runs through
the C ABI of libfinito_b200.so; torch is only the device RNG / array plumbing of the workload generator and checker."""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def stage(name, log, fn):
    t0 = time.perf_counter()
    r = fn()
    log[name + "_s"] = round(time.perf_counter() - t0, 3)
    print(f"[big_config] {name}: {log[name + '_s']} s", file=sys.stderr, flush=True)
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome", type=int, default=3_000_000_000)
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--reads", type=int, default=10_000_000)
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--error", type=float, default=0.01)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--chromosomes", type=int, default=24)
    ap.add_argument("--repeat-fraction", type=float, default=0.45)
    ap.add_argument("--families", type=int, default=2000)
    ap.add_argument("--seed", type=int, default=20260105)
    ap.add_argument("--work", default=None)
    ap.add_argument("--keep", action="store_true", help="keep the generated files")
    ap.add_argument("--verify", type=int, default=2_000_000, help="hits verified against the unitig text per step")
    a = ap.parse_args()
    import torch

    import finito_b200 as fb

    work = a.work or ("/dev/shm/finwork" if os.path.isdir("/dev/shm") else os.path.join(ROOT, "_work"))
    d = os.path.join(work, f"big_g{a.genome}_k{a.k}_s{a.seed}")
    os.makedirs(d, exist_ok=True)
    genome, spss, sbwt, prefix = (os.path.join(d, f) for f in ("genome.fna", "spss.fna", "input.sbwt", "fmin"))
    log = {"genome_bp": a.genome, "k": a.k, "work": d}
    datagen = os.path.join(ROOT, "tools", "fin_datagen")

    if not os.path.exists(prefix + ".FBV.sdsl"):
        stage("genome_gen", log, lambda: subprocess.run([datagen, "bigenome", genome, str(a.genome), str(a.seed), str(a.chromosomes),
                                                         str(a.repeat_fraction), str(a.families)], check=True))
        log["spss"] = stage("spss_build", log, lambda: fb.spss_build(genome, a.k, spss, flip=True, seed=a.seed))
        log["sbwt"] = stage("sbwt_build", log, lambda: fb.sbwt_build(spss, a.k, sbwt, 8))
        stage("build_fmin", log, lambda: fb.build_fmin(sbwt, spss, prefix))
        log["build_stages"] = fb.build_timing()
        log["build_bp_per_s"] = log["spss"]["bases"] / log["build_fmin_s"]
    ix = stage("index_load", log, lambda: fb.FinimizerIndex.load(prefix))
    info = ix.info
    log["index"] = info
    log["index_files_bytes"] = {s: os.path.getsize(prefix + s) for s in (".O.sdsl", ".FBV.sdsl", ".packed_unitigs.sdsl", ".unitig_endpoints.sdsl",
                                                                          ".Ustart.sdsl", ".LCS.sdsl", ".sbwt")}

    # ---- the genome on the device as 2-bit codes (headers and newlines hold no upper-case ACGT)
    dev = torch.device("cuda")
    lut = torch.full((256,), 255, dtype=torch.uint8)
    for i, ch in enumerate(b"ACGT"):
        lut[ch] = i
    lut = lut.to(dev)
    codes = []
    with open(genome, "rb") as f:
        while True:
            blk = f.read(1 << 28)
            if not blk:
                break
            t = lut[torch.frombuffer(bytearray(blk), dtype=torch.uint8).to(dev).long()]
            codes.append(t[t != 255])
    g = torch.cat(codes)
    del codes
    G = g.numel()
    L, n, k = a.read_len, a.reads, a.k
    per = L - k + 1
    ascii_of = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device=dev)
    gen = torch.Generator(device=dev)
    gen.manual_seed(a.seed)

    def make_reads():
        out = torch.empty((n, L), dtype=torch.uint8, device=dev)
        ar = torch.arange(L, device=dev)
        for a0 in range(0, n, 1 << 20):
            b0 = min(n, a0 + (1 << 20))
            pos = torch.randint(0, G - L, (b0 - a0,), device=dev, generator=gen)
            r = g[pos[:, None] + ar]
            err = torch.rand((b0 - a0, L), device=dev, generator=gen) < a.error
            sub = torch.randint(1, 4, (b0 - a0, L), device=dev, generator=gen, dtype=torch.uint8)
            r = torch.where(err, (r + sub) & 3, r)
            rc = torch.rand((b0 - a0,), device=dev, generator=gen) < 0.5
            r = torch.where(rc[:, None], 3 - r.flip(1), r)
            out[a0:b0] = r
        return out  # codes; ASCII = ascii_of[codes]

    # ---- the index's own unitig text for the checker: 2-bit packed words, exclusive unitig ends
    text_words = torch.from_numpy(np.fromfile(prefix + ".packed_unitigs.sdsl", dtype=np.uint64, offset=8).view(np.int64)).to(dev)
    ends = torch.from_numpy(np.fromfile(prefix + ".unitig_endpoints.sdsl", dtype=np.uint64, offset=9).view(np.int64)).to(dev)
    starts = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), ends[:-1]])

    def verify(read_codes, hits, count):
        found = torch.nonzero(hits[:, 0] != -1).flatten()
        if found.numel() == 0:
            return 0
        pick = found[torch.randint(0, found.numel(), (min(count, found.numel()),), device=dev, generator=gen)]
        r, i = pick // per, pick % per
        u, p = hits[pick, 0].long(), hits[pick, 1].long()
        assert bool((u < ends.numel()).all()) and bool((p >= 0).all())
        t0 = starts[u] + p
        assert bool((t0 + k <= ends[u]).all()), "a hit leaves its unitig"
        ar = torch.arange(k, device=dev)
        tp = t0[:, None] + ar
        tk = (text_words[tp >> 5] >> ((tp & 31) << 1)) & 3
        qk = read_codes[r[:, None], i[:, None] + ar].long()
        fwd = (tk == qk).all(1)
        rcm = (tk == (3 - qk.flip(1))).all(1)
        assert bool((fwd | rcm).all()), "a hit does not point at its k-mer"
        return int(pick.numel())

    offs = (torch.arange(n + 1, dtype=torch.int64, device=dev) * L)
    nk = n * per
    d_out = torch.empty((nk, 2), dtype=torch.int32, device=dev)
    cap = 16 * n
    d_ro = torch.empty(n + 1, dtype=torch.int64, device=dev)
    d_runs = torch.empty((cap, 4), dtype=torch.int32, device=dev)
    dense_ms, runs_ms, found, verified, runs_per_read = [], [], [], 0, []
    for step in range(a.steps + 1):  # step 0 = warm-up
        codes_r = make_reads()
        d_seqs = ascii_of[codes_r.long()].contiguous()
        torch.cuda.synchronize()
        c, _ = ix.search_device(d_seqs.data_ptr(), offs.data_ptr(), n, n * L, nk, d_out.data_ptr(), 4)
        t_dense = ix.last_timing()
        n_runs, c2, _ = ix.search_device_runs(d_seqs.data_ptr(), offs.data_ptr(), n, n * L, d_ro.data_ptr(), d_runs.data_ptr(), cap)
        t_runs = ix.last_timing()
        assert c == c2
        # the runs expand to the dense table (device-side check on this step's answers)
        lens = (d_runs[:n_runs, 1].long() & 0x7FFFFFFF)
        assert int(lens.sum()) == c.positive
        verified += verify(codes_r, d_out, a.verify)
        if step:
            dense_ms.append(t_dense["pipeline_ms"])
            runs_ms.append(t_runs["pipeline_ms"])
            found.append(c.positive / c.n_queries)
            runs_per_read.append(n_runs / n)
        del d_seqs, codes_r
    log.update(reads_per_step=n, steps=a.steps, reads_total=n * a.steps, kmers_per_step=nk,
               dense_ms_per_step=float(np.mean(dense_ms)), runs_ms_per_step=float(np.mean(runs_ms)),
               dense_kmers_per_s=nk / np.mean(dense_ms) * 1e3, runs_kmers_per_s=nk / np.mean(runs_ms) * 1e3,
               found_fraction=float(np.mean(found)), runs_per_read=float(np.mean(runs_per_read)), hits_verified=verified,
               reads="simulated on the device from the genome: uniform start, strand 1/2, iid substitutions at --error, a fresh batch per step")
    print(json.dumps(log))
    ix.close()
    if not a.keep:
        for f in os.listdir(d):
            os.remove(os.path.join(d, f))
        os.rmdir(d)


if __name__ == "__main__":
    main()
