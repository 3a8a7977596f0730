"""Pins the oracle (oracle/oracle.cpp): against the reference's own known-answer vectors (golden/kat.json, from
src/tests.cpp) and against files / output produced by the unmodified reference binaries (golden/<case>/)."""
import os

import numpy as np
import pytest

import helpers as H
from conftest import golden_cases


def load_case_queries(case_dir):
    return H.read_fasta(os.path.join(case_dir, "queries.fna"))


@pytest.mark.parametrize("case", golden_cases())
def test_oracle_build_matches_reference_files(case, tmp_path):
    d = os.path.join(H.GOLDEN, case)
    prefix = str(tmp_path / "orc")
    H.oracle_build(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix)
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(prefix + suf, os.path.join(d, "ref" + suf)), suf


@pytest.mark.parametrize("case", golden_cases())
def test_oracle_search_matches_reference_output(case, tmp_path):
    d = os.path.join(H.GOLDEN, case)
    orc = H.Oracle(os.path.join(d, "ref"))
    out = str(tmp_path / "out.txt")
    orc.search_file("", os.path.join(d, "queries.fna"), out)
    assert H.files_equal(out, os.path.join(d, "ref_out.txt"))
    c = orc.counters()
    assert c["n_deque_underflow"] == 0 and c["n_out_of_domain"] == 0


def colex_rank(unitigs, k, u):
    return sorted(unitigs, key=lambda s: s[:k][::-1]).index(u)


def test_oracle_known_answers(kat):
    lib = H.oracle_lib()
    for case in kat:
        orc = H.Oracle(os.path.join(H.GOLDEN, "kat_" + case["name"], "ref"))
        for which, key in ((0, "lcs"), (1, "fmin"), (2, "offsets"), (3, "text"), (4, "ends"), (5, "ustart")):
            if key in case:
                n = lib.orc_index_member(orc.h, which, None, 0)
                got = np.zeros(n, dtype=np.int64)
                lib.orc_index_member(orc.h, which, got.ctypes.data, n)
                assert got.tolist() == case[key], (case["name"], key)
        for q in case["queries"]:
            s = q["q"].encode()
            out = np.zeros((len(s) - case["k"] + 1, 2), dtype=np.int64)
            nf = lib.orc_search_strand(orc.h, s, len(s), out.ctypes.data)
            if "fwd" in q:
                assert out.tolist() == q["fwd"], case["name"]
                assert nf == sum(1 for r in q["fwd"] if r[0] != -1)
            if "n_found" in q:
                assert nf == q["n_found"]
            if q.get("fwd_unitig_of_query_rank"):
                assert out.tolist() == [[colex_rank(case["unitigs"], case["k"], q["q"]), 0]]
            if "merged" in q:
                seqs, offs = H.batch_of([q["q"]])
                m, _ = orc.search_reads(seqs, offs)
                assert m.tolist() == q["merged"]


@pytest.mark.skipif(not H.have_ref(), reason="reference binaries only exist where /root/reference does")
def test_reference_own_tests_pass(tmp_path):
    import subprocess

    r = subprocess.run([os.path.join(H.REF_DIR, "finito_ref_tests")], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]


@pytest.mark.parametrize("case", golden_cases())
def test_oracle_stats_modes_match_reference(case, tmp_path):
    """--type shortest|verify -t T: the `<t>,<stats>` line written by the unmodified reference binary."""
    import json

    d = os.path.join(H.GOLDEN, case)
    for mode, line in json.load(open(os.path.join(H.GOLDEN, "stats_modes.json")))[case].items():
        typ, t = mode.split(":")
        got = H.oracle_build(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), str(tmp_path / "x"), type=typ, t=int(t))
        assert f"{t},{got}\n" == line, (case, mode)
