"""Synthetic datasets for tests, smoke() and bench.py, generated with tools/fin_datagen under _work/
(cached by parameters).  Nothing here touches /root/reference."""
from __future__ import annotations

import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATAGEN = os.path.join(ROOT, "tools", "fin_datagen")
WORK = os.environ.get("FIN_WORK", os.path.join(ROOT, "_work"))


def _run(*args):
    if not os.path.exists(DATAGEN):
        subprocess.run(["make", "-C", os.path.join(ROOT, "tools")], check=True, capture_output=True)
    r = subprocess.run([DATAGEN, *[str(a) for a in args]], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"fin_datagen {' '.join(str(a) for a in args)} failed: {r.stderr}")
    return r.stderr


def make_dataset(name, genome_len, k, n_reads, read_len=150, seed=1, error=0.01, families=None, copies=5, rep_len=2000,
                 divergence=0.01):
    """genome -> unitigs (flipped) -> plain-matrix .sbwt, plus simulated reads as (uint8 bases, uint64 offsets)."""
    if families is None:
        families = max(1, genome_len // 100_000)
    rep_len = min(rep_len, max(2 * k, genome_len // 20))
    # every generator parameter is part of the cache tag (a changed parameter must never reuse stale data)
    tag = f"{name}_g{genome_len}_k{k}_r{n_reads}x{read_len}_s{seed}_f{families}"
    if (error, copies, rep_len, divergence) != (0.01, 5, min(2000, max(2 * k, genome_len // 20)), 0.01):
        tag += f"_e{error:g}_c{copies}_l{rep_len}_d{divergence:g}"
    d = os.path.join(WORK, tag)
    os.makedirs(d, exist_ok=True)
    genome, unitigs, sbwt = (os.path.join(d, f) for f in ("genome.fna", "unitigs.fna", "index.sbwt"))
    reads_bin = os.path.join(d, "reads.bin")
    done = os.path.join(d, ".done")
    if not os.path.exists(done):
        _run("genome", genome, genome_len, seed, families, copies, rep_len, divergence)
        _run("unitigs", genome, k, unitigs, seed + 1)
        _run("sbwt", unitigs, k, sbwt, 8)
        _run("reads", genome, n_reads, read_len, error, seed + 2, reads_bin)
        open(done, "w").close()
    seqs = np.fromfile(reads_bin, dtype=np.uint8)
    offs = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(read_len)
    return {"dir": d, "genome": genome, "unitigs": unitigs, "sbwt": sbwt, "reads_bin": reads_bin, "reads": (seqs, offs),
            "k": k, "n_reads": n_reads, "read_len": read_len}
