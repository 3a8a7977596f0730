#!/usr/bin/env python
"""Adds the golden cases syn_k100 and syn_k200 (k beyond the 64 that tools/fin_datagen builds): the 128- and 256-slot
candidate windows and the byte-compare path for LCS values >= 128.  Strings: random sequences (every k-mer distinct, so
any cut into strings >= k is a disjoint spectrum-preserving set) with planted repeats SHORTER than k-1 — they create long
common suffixes (large LCS values) without creating branches.  Same recipe and binaries as make_golden.py."""
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import make_case, mutate, rc  # noqa: E402


def strings(rng, k, n, rep_len):
    rep = "".join(rng.choice("ACGT") for _ in range(rep_len))
    out = []
    for i in range(n):
        a = "".join(rng.choice("ACGT") for _ in range(rng.randrange(k, 3 * k)))
        b = "".join(rng.choice("ACGT") for _ in range(rng.randrange(k, 2 * k)))
        out.append(a + (rep if i % 2 == 0 else rep[: rep_len // 2]) + b)
    out.append("".join(rng.choice("ACGT") for _ in range(k)))  # a string of exactly one k-mer
    return out


def main():
    rng = random.Random(20260104)
    for name, k, n, rep_len, rlen in (("syn_k100", 100, 8, 80, 260), ("syn_k200", 200, 6, 170, 450)):
        us = strings(rng, k, n, rep_len)
        qs = []
        for _ in range(24):
            u = rng.choice(us[:-1])
            p = rng.randrange(0, max(1, len(u) - rlen))
            r = mutate(u[p:p + rlen], rng, 0.01)
            qs.append(r if rng.random() < 0.5 else rc(r))
        qs += [us[0], rc(us[1]), us[-1], us[-1][:-1], "ACGT" * (k // 2)]
        make_case(name, k, us, qs, 2)
        print(name, "written:", len(us), "strings,", len(qs), "queries")


if __name__ == "__main__":
    main()
