"""GPU parity of fin_sbwt_build (SBWT construction on the device): byte-identical files against
the reference's in-memory constructor (golden input.sbwt files were written by oracle/_ref/sbwt_inmem =
SBWT/include/sbwt/NodeBOSSInMemoryConstructor.hh; the binary itself also runs on the GPU box where it was shipped)."""
import json
import os
import subprocess

import pytest

import helpers as H
from conftest import golden_cases
from fin_datasets import make_dataset

pytestmark = pytest.mark.gpu
SBWT_INMEM = os.path.join(H.REF_DIR, "sbwt_inmem")


@pytest.mark.parametrize("case", golden_cases())
def test_golden_sbwt_files(case, gpu_lib, tmp_path):
    fb = gpu_lib
    d = os.path.join(H.GOLDEN, case)
    meta = json.load(open(os.path.join(d, "case.json")))
    out = str(tmp_path / "gpu.sbwt")
    if meta["k"] > 64:
        with pytest.raises(fb.FinitoError) as e:
            fb.sbwt_build(os.path.join(d, "unitigs.fna"), meta["k"], out, meta["precalc_k"])
        assert e.value.status == fb.FIN_ERR_LIMIT
        return
    t = fb.sbwt_build(os.path.join(d, "unitigs.fna"), meta["k"], out, meta["precalc_k"])
    assert H.files_equal(out, os.path.join(d, "input.sbwt"))
    assert t["n_nodes"] == H.sbwt_header(out)["n_nodes"]


@pytest.mark.parametrize("k,genome_len,precalc", [(31, 1_000_000, 8), (63, 300_000, 8), (17, 200_000, 0), (32, 100_000, 4), (33, 100_000, 4)])
def test_synthetic_sbwt_matches_reference_constructor(k, genome_len, precalc, gpu_lib, tmp_path):
    fb = gpu_lib
    ds = make_dataset("gpusbwt", genome_len=genome_len, k=k, n_reads=100, read_len=150, seed=300 + k)
    out = str(tmp_path / "gpu.sbwt")
    fb.sbwt_build(ds["unitigs"], k, out, precalc)
    if os.path.exists(SBWT_INMEM):
        ref = str(tmp_path / "ref.sbwt")
        r = subprocess.run([SBWT_INMEM, ds["unitigs"], ref, str(k), str(precalc)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert H.files_equal(out, ref)
    if precalc == 8:  # the generator's own writer (pinned to the reference constructor by tests/test_sbwt_writer.py)
        assert H.files_equal(out, ds["sbwt"])


def test_sbwt_of_a_genome_with_repeated_kmers(gpu_lib, tmp_path):
    """Not a disjoint SPSS: the genome itself (repeat families => duplicate k-mers) plus short and lower-case records."""
    fb = gpu_lib
    ds = make_dataset("gpusbwt", genome_len=200_000, k=21, n_reads=100, read_len=150, seed=77)
    genome = H.read_fasta(ds["genome"])[0]
    fa = str(tmp_path / "in.fna")
    with open(fa, "w") as f:
        f.write(">g\n" + genome + "\n>short\nACGTACGT\n>lower\n" + genome[1000:1500].lower() + "\n>again\n" + genome[5000:9000] + "\n")
    out = str(tmp_path / "gpu.sbwt")
    fb.sbwt_build(fa, 21, out, 6)
    if os.path.exists(SBWT_INMEM):
        ref = str(tmp_path / "ref.sbwt")
        r = subprocess.run([SBWT_INMEM, fa, ref, "21", "6"], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert H.files_equal(out, ref)
    with pytest.raises(fb.FinitoError) as e:
        bad = str(tmp_path / "bad.fna")
        open(bad, "w").write(">x\nACGTNNACGTACGTACGTAGCTAGCTAGCATCGATCG\n")
        fb.sbwt_build(bad, 5, out, 2)
    assert e.value.status == fb.FIN_ERR_DOMAIN


def test_sbwt_then_build_fmin_then_search_end_to_end(gpu_lib, tmp_path):
    """unitigs -> fin_sbwt_build -> fin_build -> fin_search: the whole chain on the device against the oracle."""
    import numpy as np

    fb = gpu_lib
    ds = make_dataset("gpusbwt", genome_len=300_000, k=31, n_reads=3000, read_len=150, seed=9)
    sb = str(tmp_path / "input.sbwt")
    fb.sbwt_build(ds["unitigs"], 31, sb, 8)
    gp, op = str(tmp_path / "gpu"), str(tmp_path / "orc")
    fb.build_fmin(sb, ds["unitigs"], gp)
    H.oracle_build(ds["sbwt"], ds["unitigs"], op)
    for suf in H.INDEX_SUFFIXES:
        assert H.files_equal(gp + suf, op + suf), suf
    seqs, offs = ds["reads"]
    want, _ = H.Oracle(op).search_reads(seqs, offs)
    with fb.FinimizerIndex.load(gp) as ix:
        got, _ = ix.search_batch((seqs, offs))
    assert np.array_equal(got, want)


def _canon(s):
    r = s[::-1].translate(str.maketrans("ACGT", "TGCA"))
    return min(s, r)


@pytest.mark.parametrize("k,flip", [(31, True), (21, False), (33, True)])
def test_spss_is_a_disjoint_cover_of_the_canonical_kmers(k, flip, gpu_lib, tmp_path):
    fb = gpu_lib
    ds = make_dataset("gpusbwt", genome_len=120_000, k=21, n_reads=100, read_len=150, seed=78)
    genome = H.read_fasta(ds["genome"])[0]
    comp = str.maketrans("ACGT", "TGCA")
    fa = str(tmp_path / "in.fna")
    with open(fa, "w") as f:  # two records, the second holds a reverse-complemented stretch of the first and a short one
        f.write(">a\n" + genome + "\n>b\n" + genome[2000:9000][::-1].translate(comp) + genome[50_000:50_300] + "\n>short\nACGT\n")
    out = str(tmp_path / "spss.fna")
    info = fb.spss_build(fa, k, out, flip=flip, seed=5)
    pieces = H.read_fasta(out)
    assert info["n_strings"] == len(pieces) and info["bases"] == sum(len(p) for p in pieces)
    assert all(len(p) >= k for p in pieces)
    want = set()
    for s in H.read_fasta(fa):
        want.update(_canon(s[i:i + k]) for i in range(len(s) - k + 1))
    got = [_canon(p[i:i + k]) for p in pieces for i in range(len(p) - k + 1)]
    assert len(got) == len(set(got)) == info["n_kmers"]  # every canonical k-mer once
    assert set(got) == want
    if flip:
        assert any(p in genome for p in pieces if len(p) > 100) and any(p not in genome for p in pieces if len(p) > 100)


def test_spss_sbwt_build_search_chain_matches_oracle(gpu_lib, tmp_path):
    """genome -> fin_spss_build -> fin_sbwt_build -> fin_build -> search, the oracle (and, where shipped, the reference
    binary) on the same files."""
    import numpy as np

    fb = gpu_lib
    ds = make_dataset("gpusbwt", genome_len=400_000, k=31, n_reads=4000, read_len=150, seed=10)
    spss, sb = str(tmp_path / "spss.fna"), str(tmp_path / "input.sbwt")  # not gpu.sbwt: build-fmin writes <prefix>.sbwt
    fb.spss_build(ds["genome"], 31, spss, flip=True, seed=3)
    fb.sbwt_build(spss, 31, sb, 8)
    gp, op = str(tmp_path / "gpu"), str(tmp_path / "orc")
    fb.build_fmin(sb, spss, gp)
    H.oracle_build(sb, spss, op)
    for suf in H.INDEX_SUFFIXES:
        assert H.files_equal(gp + suf, op + suf), suf
    if H.have_ref():
        r = H.ref_run("build-fmin", "-o", str(tmp_path / "ref"), "-i", sb, "-u", spss)
        assert r.returncode == 0, r.stderr
        for suf in H.INDEX_SUFFIXES:
            assert H.files_equal(gp + suf, str(tmp_path / "ref") + suf), suf
    seqs, offs = ds["reads"]
    want, wc = H.Oracle(op).search_reads(seqs, offs)
    with fb.FinimizerIndex.load(gp) as ix:
        got, c = ix.search_batch((seqs, offs))
    assert np.array_equal(got, want) and c.positive == int(wc[2])
    assert 0.6 < c.positive / c.n_queries < 0.8
