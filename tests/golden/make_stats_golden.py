#!/usr/bin/env python
"""Adds tests/golden/stats_modes.json: the `<t>,<stats>` line the UNMODIFIED reference (oracle/_ref/finito_ref)
appends to <prefix>_stats.txt for `build-fmin --type shortest|verify -t T` (include/build_fmin.hh:95-214,386-395)
on the unitigs of every golden case.  Only runs where /root/reference exists (`make -C oracle ref`)."""
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "finito_ref")
MODES = [("shortest", 1), ("shortest", 2), ("shortest", 4), ("shortest", 16), ("verify", 1), ("verify", 4)]


def main():
    out = {}
    for case in sorted(os.listdir(HERE)):
        d = os.path.join(HERE, case)
        if not os.path.exists(os.path.join(d, "input.sbwt")):
            continue
        n_nodes = os.path.getsize(os.path.join(d, "ref.LCS.sdsl"))  # only used to skip t >= n_nodes (the reference never returns)
        for typ, t in MODES:
            if case.startswith(("kat_", "example_")) and t >= 4:
                continue
            with tempfile.TemporaryDirectory() as tmp:
                r = subprocess.run([REF, "build-fmin", "-o", os.path.join(tmp, "x"), "-i", os.path.join(d, "input.sbwt"), "-u",
                                    os.path.join(d, "unitigs.fna"), "--type", typ, "-t", str(t), "--lcs", os.path.join(d, "ref.LCS.sdsl")],
                                   capture_output=True, text=True, timeout=600)
                if r.returncode != 0:
                    sys.exit(f"{case} {typ} {t}: {r.stderr[-400:]}")
                out.setdefault(case, {})[f"{typ}:{t}"] = open(os.path.join(tmp, "x_stats.txt")).read()
    with open(os.path.join(HERE, "stats_modes.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print(json.dumps(out, indent=1)[:1500])


if __name__ == "__main__":
    main()
