#!/usr/bin/env python
"""CLI end to end (file in, text out) on the GPU box: finito_b200 search-fmin / build-fmin next to the unmodified reference
binary on the same files, outputs compared byte for byte.  usage: python tools/cli_timing.py [n_reads] [n_ref_reads]"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import finito_b200 as fb  # noqa: E402
from fin_datasets import make_dataset  # noqa: E402

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
n_ref = int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
REF = os.path.join(ROOT, "oracle", "_ref", "finito_ref")
ds = make_dataset("clit", genome_len=5_000_000, k=31, n_reads=n_reads, read_len=150, seed=20260102, families=50)
d = ds["dir"]
seqs, offs = ds["reads"]


def fasta(path, n):
    rows = seqs[: n * 150].reshape(n, 150)
    rec = np.empty((n, 153), dtype=np.uint8)
    rec[:, 0] = ord(">"); rec[:, 1] = ord("\n"); rec[:, 2:152] = rows; rec[:, 152] = ord("\n")
    rec.tofile(path)


def timed(*cmd, env=None):
    t0 = time.perf_counter()
    r = subprocess.run(list(cmd), capture_output=True, text=True, env=env)
    dt = time.perf_counter() - t0
    if r.returncode != 0:
        sys.exit(f"{' '.join(cmd)} failed: {r.stderr[-500:]}")
    for line in r.stderr.splitlines():
        if "pipeline:" in line or "Index loaded" in line or "Running streaming" in line:
            print("      |", line)
    return dt


q_all, q_ref = os.path.join(d, "q_all.fna"), os.path.join(d, "q_ref.fna")
fasta(q_all, n_reads); fasta(q_ref, n_ref)
bp = sum(len(l.strip()) for l in open(ds["unitigs"]) if not l.startswith(">"))
ours, ref = os.path.join(d, "cli_ours"), os.path.join(d, "cli_ref")
t = timed(fb.CLI_PATH, "build-fmin", "-o", ours, "-i", ds["sbwt"], "-u", ds["unitigs"])
print(f"build-fmin  ours      : {t:7.2f} s wall  {bp / t:.3e} unitig bp/s (process start, CUDA context, files in and out)")
t = timed(REF, "build-fmin", "-o", ref, "-i", ds["sbwt"], "-u", ds["unitigs"])
print(f"build-fmin  reference : {t:7.2f} s wall  {bp / t:.3e} unitig bp/s")
same = all(open(ours + s, "rb").read() == open(ref + s, "rb").read() for s in
           (".O.sdsl", ".FBV.sdsl", ".packed_unitigs.sdsl", ".unitig_endpoints.sdsl", ".Ustart.sdsl", ".LCS.sdsl", ".sbwt"))
print("index files byte-identical:", same)
k = 31
for n_dev in sorted({1, fb.device_count()}):
    t = timed(fb.CLI_PATH, "search-fmin", "-i", ours, "-q", q_all, "-o", ours + ".out", env=dict(os.environ, FIN_DEVICES=str(n_dev)))
    print(f"search-fmin ours x{n_dev} GPU: {t:7.2f} s wall  {n_reads * (150 - k + 1) / t:.3e} k-mers/s over {n_reads} reads "
          f"({os.path.getsize(q_all) / 1e9:.2f} GB FASTA in, {os.path.getsize(ours + '.out') / 1e9:.2f} GB text out)")
t = timed(fb.CLI_PATH, "search-fmin", "-i", ours, "-q", q_ref, "-o", ours + ".ref_out")
print(f"search-fmin ours      : {t:7.2f} s wall  {n_ref * (150 - k + 1) / t:.3e} k-mers/s over the {n_ref}-read subset")
t = timed(REF, "search-fmin", "-i", ref, "-q", q_ref, "-o", ref + ".ref_out")
print(f"search-fmin reference : {t:7.2f} s wall  {n_ref * (150 - k + 1) / t:.3e} k-mers/s over the {n_ref}-read subset (1 core)")
print("search output byte-identical on the subset:", open(ours + ".ref_out", "rb").read() == open(ref + ".ref_out", "rb").read())
