"""GPU parity tests of build-fmin through the C ABI: index files byte-identical to the reference's (golden
fixtures) and to the oracle's on seeded synthetic inputs; LCS kernel against the oracle."""
import os

import numpy as np
import pytest

import helpers as H
from conftest import golden_cases
from fin_datasets import make_dataset

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", golden_cases())
def test_golden_index_files(case, gpu_lib, tmp_path):
    fb = gpu_lib
    d = os.path.join(H.GOLDEN, case)
    prefix = str(tmp_path / "gpu")
    assert fb.build_fmin(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix) == ""
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(prefix + suf, os.path.join(d, "ref" + suf)), suf


@pytest.mark.parametrize("case", golden_cases())
def test_golden_lcs(case, gpu_lib):
    fb = gpu_lib
    d = os.path.join(H.GOLDEN, case)
    lcs = fb.lcs_array(os.path.join(d, "input.sbwt"))
    want = H.oracle_lcs(os.path.join(d, "input.sbwt"), len(lcs))
    assert np.array_equal(lcs, want)


@pytest.mark.parametrize("k,genome_len", [(31, 1_000_000), (63, 300_000), (20, 200_000)])
def test_synthetic_build_matches_oracle(k, genome_len, gpu_lib, tmp_path):
    fb = gpu_lib
    ds = make_dataset("gpub", genome_len=genome_len, k=k, n_reads=2000, read_len=150, seed=100 + k)
    gp, op = str(tmp_path / "gpu"), str(tmp_path / "orc")
    fb.build_fmin(ds["sbwt"], ds["unitigs"], gp)
    H.oracle_build(ds["sbwt"], ds["unitigs"], op)
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(gp + suf, op + suf), suf
    # the GPU-built index answers queries like the oracle's
    seqs, offs = ds["reads"]
    want, _ = H.Oracle(op).search_reads(seqs, offs)
    with fb.FinimizerIndex.load(gp) as ix:
        got, _ = ix.search_batch((seqs, offs))
    assert np.array_equal(got, want)


def test_build_with_given_lcs_and_error_paths(gpu_lib, tmp_path):
    fb = gpu_lib
    d = os.path.join(H.GOLDEN, "syn_k31")
    prefix = str(tmp_path / "gpu")
    fb.build_fmin(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix, lcs_path=os.path.join(d, "ref.LCS.sdsl"))
    for suf in H.INDEX_SUFFIXES:
        assert H.files_equal(prefix + suf, os.path.join(d, "ref" + suf)), suf
    with pytest.raises(fb.FinitoError) as e:  # build_fmin.hh:245-247
        fb.build_fmin(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix, t=4)
    assert "t != 1 does not make sense with rarest type" in e.value.message
    with pytest.raises(fb.FinitoError):
        fb.build_fmin(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix, type="nonsense")
    with pytest.raises(fb.FinitoError) as e:  # the index member has no variant header
        fb.build_fmin(os.path.join(d, "ref.sbwt"), os.path.join(d, "unitigs.fna"), prefix)
    assert e.value.status == fb.FIN_ERR_FORMAT


def test_foreign_lcs_file_is_refused(gpu_lib, tmp_path):
    """An LCS that cannot belong to the SBWT (a value >= k, or LCS[0] != 0) would send the word-parallel drop_first_char
    out of bounds (the reference guards it with `first > 0`, common.hh:43): FIN_ERR_FORMAT at build and at load instead."""
    import shutil

    fb = gpu_lib
    d = os.path.join(H.GOLDEN, "syn_k12")  # k = 12: 4-bit LCS entries, value 15 is representable and >= k
    raw = bytearray(open(os.path.join(d, "ref.LCS.sdsl"), "rb").read())
    assert raw[8] == 4
    for name, patch in (("big", (9 + 20, 0xFF)), ("first", (9, raw[9] | 0x01))):
        bad = bytearray(raw)
        bad[patch[0]] = patch[1]
        p = str(tmp_path / f"{name}.LCS.sdsl")
        open(p, "wb").write(bad)
        with pytest.raises(fb.FinitoError) as e:
            fb.build_fmin(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), str(tmp_path / "x"), lcs_path=p)
        assert e.value.status == fb.FIN_ERR_FORMAT
        ix_prefix = str(tmp_path / name)
        for suf in H.INDEX_SUFFIXES:
            shutil.copy(os.path.join(d, "ref" + suf), ix_prefix + suf)
        open(ix_prefix + ".LCS.sdsl", "wb").write(bad)
        with pytest.raises(fb.FinitoError) as e:
            fb.FinimizerIndex.load(ix_prefix)
        assert e.value.status == fb.FIN_ERR_FORMAT


@pytest.mark.parametrize("case", golden_cases())
def test_golden_stats_modes(case, gpu_lib, tmp_path):
    """BASELINE.json configs[2]: --type shortest|verify -t T print statistics only (build_fmin.hh:95-214)."""
    import json

    fb = gpu_lib
    d = os.path.join(H.GOLDEN, case)
    for mode, line in json.load(open(os.path.join(H.GOLDEN, "stats_modes.json")))[case].items():
        typ, t = mode.split(":")
        prefix = str(tmp_path / f"gpu_{typ}_{t}")
        got = fb.build_fmin(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix, type=typ, t=int(t))
        assert f"{t},{got}\n" == line, (case, mode)
        assert open(prefix + "_stats.txt").read() == line
        assert os.path.exists(prefix + ".LCS.sdsl") and not os.path.exists(prefix + ".FBV.sdsl")


@pytest.mark.parametrize("k,genome_len,t", [(31, 1_000_000, 4), (63, 300_000, 2), (20, 200_000, 16)])
def test_synthetic_stats_modes_match_oracle(k, genome_len, t, gpu_lib, tmp_path):
    fb = gpu_lib
    ds = make_dataset("gpub", genome_len=genome_len, k=k, n_reads=2000, read_len=150, seed=100 + k)
    for typ in ("shortest", "verify"):
        want = H.oracle_build(ds["sbwt"], ds["unitigs"], str(tmp_path / f"orc_{typ}"), type=typ, t=t)
        assert fb.build_fmin(ds["sbwt"], ds["unitigs"], str(tmp_path / f"gpu_{typ}"), type=typ, t=t) == want, typ


def test_stats_modes_reject_out_of_domain_t(gpu_lib, tmp_path):
    fb = gpu_lib
    d = os.path.join(H.GOLDEN, "kat_paper_example")
    for t in (0, 128, 18):  # 18 columns: with t >= n_nodes the reference's shrink loop never terminates
        with pytest.raises(fb.FinitoError) as e:
            fb.build_fmin(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), str(tmp_path / "x"), type="shortest", t=t)
        assert e.value.status == fb.FIN_ERR_DOMAIN
