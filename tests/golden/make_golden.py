#!/usr/bin/env python
"""Regenerates the golden fixtures under tests/golden/<case>/ by running the UNMODIFIED reference binaries
(oracle/_ref/sbwt_inmem = the reference's in-memory SBWT constructor, oracle/_ref/finito_ref = its CLI) —
only possible in the container that has /root/reference (`make -C oracle ref`).  The fixtures are committed
so that the GPU box, which has neither, can still check against what the reference itself produced.

Per case: unitigs.fna, input.sbwt (with variant header), ref.{O,FBV,packed_unitigs,unitig_endpoints,Ustart,LCS}.sdsl,
ref.sbwt, ref_stats.txt, queries.fna, ref_out.txt (search-fmin text output, both strands merged)."""
import json
import os
import random
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref")
DATAGEN = os.path.join(ROOT, "tools", "fin_datagen")


def run(*cmd, cwd=None):
    r = subprocess.run(list(cmd), cwd=cwd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.exit(f"{' '.join(cmd)} failed:\n{r.stdout}\n{r.stderr}")
    return r


def rc(s):
    return s[::-1].translate(str.maketrans("ACGT", "TGCA"))


def mutate(s, rng, p):
    return "".join(rng.choice([b for b in "ACGT" if b != c]) if rng.random() < p else c for c in s)


def make_case(name, k, unitigs, queries, precalc_k):
    d = os.path.join(HERE, name)
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    with open(os.path.join(d, "unitigs.fna"), "w") as f:
        for i, u in enumerate(unitigs):
            f.write(f">u{i}\n{u}\n")
    with open(os.path.join(d, "queries.fna"), "w") as f:
        for i, q in enumerate(queries):
            f.write(f">q{i}\n{q}\n")
    run(os.path.join(REF, "sbwt_inmem"), "unitigs.fna", "input.sbwt", str(k), str(precalc_k), cwd=d)
    run(os.path.join(REF, "finito_ref"), "build-fmin", "-o", "ref", "-i", "input.sbwt", "-u", "unitigs.fna", cwd=d)
    run(os.path.join(REF, "finito_ref"), "search-fmin", "-i", "ref", "-q", "queries.fna", "-o", "ref_out.txt", cwd=d)
    for junk in ("ref.stats", "refstats.txt"):
        p = os.path.join(d, junk)
        if os.path.exists(p):
            os.remove(p)
    with open(os.path.join(d, "case.json"), "w") as f:
        json.dump({"k": k, "n_unitigs": len(unitigs), "n_queries": len(queries), "precalc_k": precalc_k}, f)


def synthetic_unitigs(tmp, genome_len, k, seed, families):
    os.makedirs(tmp, exist_ok=True)
    run(DATAGEN, "genome", "g.fna", str(genome_len), str(seed), str(families), "3", str(max(2 * k, genome_len // 25)), "0.02",
        cwd=tmp)
    run(DATAGEN, "unitigs", "g.fna", str(k), "u.fna", str(seed + 1), cwd=tmp)
    genome = "".join(l.strip() for l in open(os.path.join(tmp, "g.fna")) if not l.startswith(">"))
    unitigs = [l.strip() for l in open(os.path.join(tmp, "u.fna")) if not l.startswith(">")]
    return genome, unitigs


def main():
    kat = json.load(open(os.path.join(HERE, "kat.json")))
    for case in kat["cases"]:
        qs = [q["q"] for q in case["queries"]] or [case["unitigs"][0]]
        # the reference's own cases plus every unitig and its reverse complement as queries
        qs += case["unitigs"] + [rc(u) for u in case["unitigs"]]
        make_case("kat_" + case["name"], case["k"], case["unitigs"], qs, 0)
    # BASELINE.json configs[0]: the literal ref_implementation/example.fna (three paper-example unitigs), queried
    # against itself and its reverse complements.  k=4 (at k=31 the file holds no k-mer at all).
    example = "/root/reference/ref_implementation/example.fna"
    if os.path.exists(example):
        ex = [l.strip() for l in open(example) if l.strip() and not l.startswith(">")]
        make_case("example_fna_k4", 4, ex, ex + [rc(u) for u in ex], 4)
    if "--only-example" in sys.argv:
        return
    rng = random.Random(20260101)
    tmp = os.path.join(ROOT, "_work", "golden_tmp")
    for name, glen, k, seed, fam, nreads, rlen in [("syn_k31", 12000, 31, 11, 3, 60, 150), ("syn_k63", 6000, 63, 12, 2, 30, 200),
                                                    ("syn_k12", 3000, 12, 13, 2, 40, 60)]:
        genome, unitigs = synthetic_unitigs(os.path.join(tmp, name), glen, k, seed, fam)
        qs = []
        for _ in range(nreads):
            p = rng.randrange(0, len(genome) - rlen)
            r = mutate(genome[p:p + rlen], rng, 0.01)
            qs.append(r if rng.random() < 0.5 else rc(r))
        qs.append(genome[:k - 1])           # shorter than k: empty output line
        qs.append(genome[100:100 + k])      # exactly one k-mer
        qs.append("ACGT" * 20)              # mostly absent
        qs.append(max(unitigs, key=len))    # a whole unitig (long walk)
        qs.append(rc(max(unitigs, key=len)))
        make_case(name, k, unitigs, qs, 2)
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
