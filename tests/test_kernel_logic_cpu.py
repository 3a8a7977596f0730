"""The per-thread bodies of the CUDA kernels (finito_b200/csrc/*.cuh compiled for the host by tests/emu) against
the oracle and the reference's golden files.  This checks the kernel LOGIC on the CPU-only box; the GPU parity
tests proper are test_gpu_*.py.  The emulation is test infrastructure, not a product path."""
import os

import numpy as np
import pytest

import helpers as H
from conftest import golden_cases
from fin_datasets import make_dataset


@pytest.mark.parametrize("case", golden_cases())
@pytest.mark.parametrize("runs", [False, True], ids=["dense", "runs"])
@pytest.mark.parametrize("prefilter_s", [0, 3, 8], ids=["nofilter", "s3", "s8"])
def test_search_bodies_match_reference_output(case, runs, prefilter_s):
    d = os.path.join(H.GOLDEN, case)
    reads = H.read_fasta(os.path.join(d, "queries.fna"))
    seqs, offs = H.batch_of(reads)
    got, counts, _, flags = H.emu_search(os.path.join(d, "ref"), seqs, offs, runs=runs, prefilter_s=prefilter_s)
    assert flags == 0
    want = H.parse_output(os.path.join(d, "ref_out.txt"))
    assert np.array_equal(got, want)
    assert counts[2] == int((want[:, 0] != -1).sum())


@pytest.mark.parametrize("case", golden_cases())
@pytest.mark.parametrize("per_chunk", [0, 1, 5])
def test_build_bodies_match_reference_files(case, per_chunk, tmp_path):
    d = os.path.join(H.GOLDEN, case)
    prefix = str(tmp_path / "emu")
    H.emu_build(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix, per_chunk)
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(prefix + suf, os.path.join(d, "ref" + suf)), suf


@pytest.mark.parametrize("k,genome_len", [(31, 150_000), (63, 60_000), (17, 40_000)])
def test_search_and_build_bodies_match_oracle_on_synthetic(k, genome_len, tmp_path):
    ds = make_dataset("cpu", genome_len=genome_len, k=k, n_reads=400, read_len=150, seed=k)
    orc_prefix = str(tmp_path / "orc")
    H.oracle_build(ds["sbwt"], ds["unitigs"], orc_prefix)
    emu_prefix = str(tmp_path / "emu")
    H.emu_build(ds["sbwt"], ds["unitigs"], emu_prefix, 0)
    for suf in H.INDEX_SUFFIXES:
        assert H.files_equal(emu_prefix + suf, orc_prefix + suf), suf
    seqs, offs = ds["reads"]
    orc = H.Oracle(orc_prefix)
    want, wc = orc.search_reads(seqs, offs)
    oc = orc.counters()
    got, gc, acc, flags = H.emu_search(orc_prefix, seqs, offs)
    assert flags == 0 and oc["n_deque_underflow"] == 0 and oc["n_out_of_domain"] == 0
    assert np.array_equal(got, want) and np.array_equal(gc, wc)
    got_r, gc_r, _, flags_r = H.emu_search(orc_prefix, seqs, offs, runs=True)  # run records -> fin_expand_runs
    assert flags_r == 0 and np.array_equal(got_r, want) and np.array_equal(gc_r, wc)
    # strand prefilter: same answers and counters although about half of the strands are never streamed
    for runs in (False, True):
        streamed = np.zeros(1, dtype=np.uint64)
        got_f, gc_f, acc_f, flags_f = H.emu_search(orc_prefix, seqs, offs, runs=runs, prefilter_s=10, streamed=streamed)
        assert flags_f == 0 and np.array_equal(got_f, want) and np.array_equal(gc_f, wc)
        assert int(streamed[0]) < 0.7 * 2 * (len(offs) - 1) and acc_f["n_rank"] < 0.7 * acc["n_rank"]
    # the access counters behind the algorithmic-bytes figure agree between the oracle and the kernel bodies
    for name in ("n_rank", "n_lcs_reads", "n_ustart_reads", "n_anchor", "n_anchor_branch", "n_walk_words", "n_bases", "n_kmers"):
        assert acc[name] == oc[name], name
    assert 0.5 < wc[2] / wc[3] < 0.95  # most k-mers of 1 %-error reads are found


@pytest.mark.parametrize("runs", [False, True], ids=["dense", "runs"])
def test_search_bodies_long_reads_cross_found_mask_words(runs, tmp_path):
    """Reads of 40..700 bases with errors on either strand: the forward strand's found mask spans up to 11 words, runs
    get cut by the merge rule in the middle of a walk (search_fmin.hh:54-60)."""
    k = 31
    ds = make_dataset("cpu", genome_len=150_000, k=k, n_reads=400, read_len=150, seed=k)
    orc_prefix = str(tmp_path / "orc")
    H.oracle_build(ds["sbwt"], ds["unitigs"], orc_prefix)
    genome = "".join(H.read_fasta(ds["genome"]))
    unitigs = H.read_fasta(ds["unitigs"])
    rng = np.random.default_rng(7)
    comp = str.maketrans("ACGT", "TGCA")
    reads = []
    for i in range(300):
        n = int(rng.integers(40, 700))
        p = int(rng.integers(0, len(genome) - n))
        r = list(genome[p:p + n])
        for j in np.nonzero(rng.random(n) < 0.01)[0]:
            r[j] = "ACGT"[("ACGT".index(r[j]) + 1 + int(rng.integers(0, 3))) % 4]
        r = "".join(r)
        reads.append(r if rng.random() < 0.5 else r[::-1].translate(comp))
    # chimeras: a forward piece of one unitig glued to the reverse complement of another (both strands hit in one read)
    long_u = sorted(unitigs, key=len)[-12:]
    for a, b in zip(long_u[::2], long_u[1::2]):
        reads.append(a[:200] + b[:200][::-1].translate(comp) + a[100:260])
    seqs, offs = H.batch_of(reads)
    want, wc = H.Oracle(orc_prefix).search_reads(seqs, offs)
    got, gc, _, flags = H.emu_search(orc_prefix, seqs, offs, runs=runs)
    assert flags == 0 and np.array_equal(got, want) and np.array_equal(gc, wc)
    got, gc, _, flags = H.emu_search(orc_prefix, seqs, offs, runs=runs, prefilter_s=10)
    assert flags == 0 and np.array_equal(got, want) and np.array_equal(gc, wc)
    assert wc[0] > 0 and wc[1] > 0


def stats_modes():
    import json

    return json.load(open(os.path.join(H.GOLDEN, "stats_modes.json")))


@pytest.mark.parametrize("case", golden_cases())
@pytest.mark.parametrize("per_chunk", [0, 1, 7])
def test_stats_mode_bodies_match_reference(case, per_chunk):
    """--type shortest|verify -t T (build_fmin.hh:95-214): the stats line the reference binary wrote."""
    d = os.path.join(H.GOLDEN, case)
    for mode, line in stats_modes()[case].items():
        typ, t = mode.split(":")
        if typ == "verify" and per_chunk:
            continue
        got = H.emu_stats(os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), typ, int(t), per_chunk)
        assert f"{t},{got}\n" == line, (case, mode)


@pytest.mark.parametrize("k,genome_len,t", [(31, 150_000, 4), (63, 60_000, 2), (17, 40_000, 9)])
def test_stats_mode_bodies_match_oracle_on_synthetic(k, genome_len, t, tmp_path):
    ds = make_dataset("cpu", genome_len=genome_len, k=k, n_reads=400, read_len=150, seed=k)
    for typ in ("shortest", "verify"):
        want = H.oracle_build(ds["sbwt"], ds["unitigs"], str(tmp_path / f"orc_{typ}"), type=typ, t=t)
        assert H.emu_stats(ds["sbwt"], ds["unitigs"], typ, t) == want, typ


@pytest.mark.parametrize("k", [3, 5, 9, 17, 32, 33, 64])
@pytest.mark.parametrize("seed,genome_len,error", [(1, 300, 0.05), (2, 1200, 0.2), (3, 4000, 0.02), (4, 700, 0.5)])
def test_closed_forms_on_tiny_repetitive_inputs(k, seed, genome_len, error, tmp_path):
    """The closed forms of the shrink and contraction loops against the oracle's character-by-character loops where they are
    stressed most: tiny genomes made of repeats (LCS values at both extremes, intervals that widen to the whole range, new_len
    reaching 0) and reads with up to 50 % errors (a contraction at almost every character); answers, the four counters and
    the reference's access tally must agree, dense and as runs."""
    if genome_len < 4 * k:
        genome_len = 4 * k + 50
    ds = make_dataset("tiny", genome_len=genome_len, k=k, n_reads=60, read_len=max(k + 5, 90), seed=100 * k + seed, error=error,
                      families=3, copies=4, rep_len=max(2 * k, genome_len // 6), divergence=0.03)
    prefix = str(tmp_path / "orc")
    H.oracle_build(ds["sbwt"], ds["unitigs"], prefix)
    seqs, offs = ds["reads"]
    orc = H.Oracle(prefix)
    want, wc = orc.search_reads(seqs, offs)
    oc = orc.counters()
    got, gc, acc, flags = H.emu_search(prefix, seqs, offs)
    assert flags == 0 and np.array_equal(got, want) and np.array_equal(gc, wc)
    for name in ("n_rank", "n_lcs_reads", "n_ustart_reads", "n_anchor", "n_anchor_branch", "n_walk_words"):
        assert acc[name] == oc[name], name
    got_r, gc_r, _, flags_r = H.emu_search(prefix, seqs, offs, runs=True, prefilter_s=min(k, 4))
    assert flags_r == 0 and np.array_equal(got_r, want) and np.array_equal(gc_r, wc)
    assert wc[2] > 0 or error >= 0.2  # the reads do hit the index
