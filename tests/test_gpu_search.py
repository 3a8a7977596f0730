"""GPU parity tests of search-fmin, through the C ABI (finito_b200.lib()).  Bit-exact against the reference's
golden outputs and against the oracle on seeded synthetic inputs."""
import os

import numpy as np
import pytest

import helpers as H
from conftest import golden_cases
from fin_datasets import make_dataset

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", golden_cases())
def test_golden_search_output(case, gpu_lib):
    fb = gpu_lib
    d = os.path.join(H.GOLDEN, case)
    reads = H.read_fasta(os.path.join(d, "queries.fna"))
    want = H.parse_output(os.path.join(d, "ref_out.txt"))
    with fb.FinimizerIndex.load(os.path.join(d, "ref")) as ix:
        got, counts = ix.search_batch(reads)
        assert np.array_equal(got, want)
        assert counts.positive == int((want[:, 0] != -1).sum())
        got32, _ = ix.search_batch(reads, hit_bits=32)
        assert np.array_equal(got32.astype(np.int64), want)
        k = ix.info["k"]
        text = fb.format_hits(got, [max(0, len(r) - k + 1) for r in reads])
        assert text == open(os.path.join(d, "ref_out.txt")).read()


def colex_rank(unitigs, k, u):
    return sorted(unitigs, key=lambda s: s[:k][::-1]).index(u)


def test_reference_known_answers(kat, gpu_lib):
    fb = gpu_lib
    for case in kat:
        with fb.FinimizerIndex.load(os.path.join(H.GOLDEN, "kat_" + case["name"], "ref")) as ix:
            for q in case["queries"]:
                if "fwd" in q:
                    assert [list(t) for t in ix.search(q["q"])] == q["fwd"], case["name"]
                if q.get("fwd_unitig_of_query_rank"):
                    assert ix.search(q["q"]) == [(colex_rank(case["unitigs"], case["k"], q["q"]), 0)]
                if "merged" in q:
                    assert [list(t) for t in ix.search_both_strands(q["q"])] == q["merged"]


@pytest.fixture(scope="module")
def synthetic(gpu_lib, tmp_path_factory):
    ds = make_dataset("gpu", genome_len=1_000_000, k=31, n_reads=20_000, read_len=150, seed=5)
    prefix = str(tmp_path_factory.mktemp("idx") / "orc")
    H.oracle_build(ds["sbwt"], ds["unitigs"], prefix)
    return ds, prefix


def test_synthetic_matches_oracle(synthetic, gpu_lib):
    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    orc = H.Oracle(prefix)
    want, wc = orc.search_reads(seqs, offs)
    with fb.FinimizerIndex.load(prefix) as ix:
        got, c = ix.search_batch((seqs, offs))
        assert np.array_equal(got, want)
        assert [c.found_fwd, c.found_rc, c.positive, c.n_queries] == [int(x) for x in wc]
        got32, c32 = ix.search_batch((seqs, offs), hit_bits=32)
        assert np.array_equal(got32.astype(np.int64), want) and c32 == c


def test_ragged_and_edge_inputs(synthetic, gpu_lib):
    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    base = [bytes(seqs[int(offs[i]):int(offs[i + 1])]).decode() for i in range(64)]
    k = ds["k"]
    reads = []
    for i, r in enumerate(base):
        reads.append(r[: (i * 7) % 151])  # lengths 0..150, including 0, < k, == k
    reads += ["", "A", base[0][:k - 1], base[1][:k], base[2].lower(), base[3] * 40, "ACGT" * 100]
    s2, o2 = H.batch_of(reads)
    orc = H.Oracle(prefix)
    want, wc = orc.search_reads(s2, o2)
    with fb.FinimizerIndex.load(prefix) as ix:
        got, c = ix.search_batch(reads)
        assert np.array_equal(got, want)
        assert c.n_queries == int(wc[3]) and c.positive == int(wc[2])
        empty, c0 = ix.search_batch([])
        assert empty.shape == (0, 2) and c0.n_queries == 0
        with pytest.raises(fb.FinitoError) as e:
            ix.search_batch(["ACGTNACGT" * 10])
        assert e.value.status == fb.FIN_ERR_DOMAIN
        # the handle stays usable after an error
        again, _ = ix.search_batch(reads)
        assert np.array_equal(again, want)


def test_batch_split_is_invisible(synthetic, gpu_lib):
    """Sub-batching inside fin_search_batch must not change anything: one call == many small calls."""
    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    n = 3000
    with fb.FinimizerIndex.load(prefix) as ix:
        whole, cw = ix.search_batch((seqs[: int(offs[n])], offs[: n + 1]))
        parts, found = [], 0
        for a in range(0, n, 700):
            b = min(n, a + 700)
            sub = seqs[int(offs[a]):int(offs[b])]
            h, c = ix.search_batch((sub, offs[a:b + 1] - offs[a]))
            parts.append(h)
            found += c.positive
        assert np.array_equal(np.concatenate(parts), whole) and found == cw.positive


def test_long_queries_unitigs_against_themselves(synthetic, gpu_lib):
    """Config #1 plumbing: every unitig queried against its own index is found at (its colex rank, 0..)."""
    fb = gpu_lib
    ds, prefix = synthetic
    unitigs = H.read_fasta(ds["unitigs"])
    k = ds["k"]
    order = sorted(range(len(unitigs)), key=lambda i: unitigs[i][:k][::-1])
    rank = {u: r for r, u in enumerate(order)}
    with fb.FinimizerIndex.load(prefix) as ix:
        hits, c = ix.search_batch(unitigs)
    pos = 0
    for i, u in enumerate(unitigs):
        n = len(u) - k + 1
        assert (hits[pos:pos + n, 0] == rank[i]).all()
        assert (hits[pos:pos + n, 1] == np.arange(n)).all()
        pos += n
    assert c.positive == c.n_queries == pos


def test_access_counters_equal_the_oracles_tally(synthetic, gpu_lib, monkeypatch):
    """Counting variant, strand prefilter off: the kernel tallies exactly the ranks / LCS values / Ustart bits / anchors / walk
    words the reference's loops touch, and says how many of them its closed forms never issue."""
    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    n = 3000
    sub, so = seqs[: int(offs[n])], offs[: n + 1]
    orc = H.Oracle(prefix)
    want, _ = orc.search_reads(sub, so)
    oc = orc.counters()
    monkeypatch.setenv("FIN_NO_PREFILTER", "1")  # read when the index is loaded
    with fb.FinimizerIndex.load(prefix) as ix:
        ix.set_counting(True)
        got, _ = ix.search_batch((sub, so))
        acc = ix.access_counters()
    assert np.array_equal(got, want)
    for name in H.COUNTER_NAMES[:8]:
        assert acc[name] == oc[name], name
    assert 0 < acc["n_lcs_skipped"] < acc["n_lcs_reads"] and 0 < acc["n_rank_skipped"] < acc["n_rank"]


def test_device_resident_entry_point(synthetic, gpu_lib):
    import torch

    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    with fb.FinimizerIndex.load(prefix) as ix:
        want, wc = ix.search_batch((seqs, offs), hit_bits=32)
        d_seqs = torch.from_numpy(seqs).cuda()
        d_offs = torch.from_numpy(offs.astype(np.int64)).cuda()
        d_out = torch.empty((want.shape[0], 2), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        c, ms = ix.search_device(d_seqs.data_ptr(), d_offs.data_ptr(), len(offs) - 1, len(seqs), want.shape[0],
                                 d_out.data_ptr(), 4)
        assert ms > 0 and c == wc
        assert np.array_equal(d_out.cpu().numpy(), want)


def test_bad_batches_are_refused_and_the_handle_survives(synthetic, gpu_lib):
    """Capacity and offset checks happen per sub-batch inside the pipeline: the call reports FIN_ERR_ARG and the handle works on."""
    import ctypes as C

    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    n = 5000
    sub_seqs, sub_offs = np.ascontiguousarray(seqs[: int(offs[n])]), np.ascontiguousarray(offs[: n + 1])
    with fb.FinimizerIndex.load(prefix) as ix:
        want, wc = ix.search_batch((sub_seqs, sub_offs))
        total = want.shape[0]
        out = np.empty((total, 2), dtype=np.int64)
        c = fb._Counts()
        st = fb.lib().fin_search_batch(ix._h, sub_seqs.ctypes.data, sub_offs.ctypes.data, n, out.ctypes.data, total - 1, C.byref(c))
        assert st == fb.FIN_ERR_ARG and b"out_capacity" in fb.lib().fin_last_error()
        bad = sub_offs.copy()
        bad[100], bad[101] = bad[101], bad[100]
        st = fb.lib().fin_search_batch(ix._h, sub_seqs.ctypes.data, bad.ctypes.data, n, out.ctypes.data, total, C.byref(c))
        assert st == fb.FIN_ERR_ARG and b"non-decreasing" in fb.lib().fin_last_error()
        again, ac = ix.search_batch((sub_seqs, sub_offs), out=out)
        assert np.array_equal(again, want) and ac == wc


# ---- run-length hits (fin_search_batch_runs / fin_search_device_runs / fin_expand_runs) ---------------------------
def check_runs_canonical(run_offsets, runs, kmers_per_read):
    """sorted by first k-mer inside each read, disjoint, inside the read, non-empty"""
    ro = run_offsets.astype(np.int64)
    assert ro[0] == 0 and ro[-1] == len(runs) and (np.diff(ro) >= 0).all()
    first = runs["first_kmer"].astype(np.int64)
    length = (runs["len_rc"] & 0x7FFFFFFF).astype(np.int64)
    assert (length > 0).all()
    read_of = np.repeat(np.arange(len(ro) - 1), np.diff(ro))
    assert (first + length <= np.asarray(kmers_per_read, dtype=np.int64)[read_of]).all()
    same = read_of[1:] == read_of[:-1]
    assert (first[1:][same] >= (first + length)[:-1][same]).all()


@pytest.mark.parametrize("case", golden_cases())
def test_golden_search_output_as_runs(case, gpu_lib):
    fb = gpu_lib
    d = os.path.join(H.GOLDEN, case)
    reads = H.read_fasta(os.path.join(d, "queries.fna"))
    want = H.parse_output(os.path.join(d, "ref_out.txt"))
    seqs, offs = H.batch_of(reads)
    with fb.FinimizerIndex.load(os.path.join(d, "ref")) as ix:
        k = ix.info["k"]
        ro, runs, c = ix.search_batch_runs((seqs, offs))
        check_runs_canonical(ro, runs, [max(0, len(r) - k + 1) for r in reads])
        assert np.array_equal(fb.expand_runs(offs, k, ro, runs), want)
        assert np.array_equal(fb.expand_runs(offs, k, ro, runs, hit_bits=32, threads=3).astype(np.int64), want)
        assert c.positive == int((want[:, 0] != -1).sum()) == int((runs["len_rc"] & 0x7FFFFFFF).sum())


def test_runs_match_oracle_and_dense_path(synthetic, gpu_lib):
    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    want, wc = H.Oracle(prefix).search_reads(seqs, offs)
    with fb.FinimizerIndex.load(prefix) as ix:
        k = ix.info["k"]
        ro, runs, c = ix.search_batch_runs((seqs, offs))
        assert [c.found_fwd, c.found_rc, c.positive, c.n_queries] == [int(x) for x in wc]
        check_runs_canonical(ro, runs, np.maximum(np.diff(offs.astype(np.int64)) - k + 1, 0))
        assert np.array_equal(fb.expand_runs(offs, k, ro, runs, threads=4), want)
        assert len(runs) < want.shape[0] // 10  # the point of the format: far fewer records than k-mer positions
        # a too small caller buffer: FIN_ERR_ARG, *n_runs says how much a retry needs, the handle works on
        import ctypes as C

        small = np.empty(len(runs) // 2, dtype=fb.RUN_DTYPE)
        ro2 = np.empty(len(offs), dtype=np.uint64)
        cc, nr = fb._Counts(), C.c_uint64()
        st = fb.lib().fin_search_batch_runs(ix._h, seqs.ctypes.data, offs.ctypes.data, len(offs) - 1, ro2.ctypes.data,
                                            small.ctypes.data, len(small), C.byref(nr), C.byref(cc))
        assert st == fb.FIN_ERR_ARG and nr.value == len(runs) and b"runs_capacity" in fb.lib().fin_last_error()
        ro3, runs3, c3 = ix.search_batch_runs((seqs, offs))
        assert np.array_equal(ro3, ro) and np.array_equal(runs3, runs) and c3 == c


def test_runs_ragged_long_and_chimeric_reads(synthetic, gpu_lib):
    """0..150-base reads, whole unitigs (long walks, many found-mask words), chimeras that hit on both strands."""
    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    k = ds["k"]
    base = [bytes(seqs[int(offs[i]):int(offs[i + 1])]).decode() for i in range(64)]
    unitigs = sorted(H.read_fasta(ds["unitigs"]), key=len)[-10:]
    comp = str.maketrans("ACGT", "TGCA")
    reads = [r[: (i * 7) % 151] for i, r in enumerate(base)] + ["", "A", base[0][:k - 1], base[1][:k], "ACGT" * 100]
    reads += unitigs + [u[::-1].translate(comp) for u in unitigs]
    reads += [a[:300] + b[:300][::-1].translate(comp) + a[150:400] for a, b in zip(unitigs[::2], unitigs[1::2])]
    s2, o2 = H.batch_of(reads)
    want, wc = H.Oracle(prefix).search_reads(s2, o2)
    with fb.FinimizerIndex.load(prefix) as ix:
        ro, runs, c = ix.search_batch_runs((s2, o2))
        assert np.array_equal(fb.expand_runs(o2, k, ro, runs), want)
        assert [c.found_fwd, c.found_rc, c.positive, c.n_queries] == [int(x) for x in wc]
        dense, cd = ix.search_batch((s2, o2))
        assert np.array_equal(dense, want) and cd == c
        ro0, runs0, c0 = ix.search_batch_runs([])
        assert ro0.tolist() == [0] and len(runs0) == 0 and c0.n_queries == 0
        with pytest.raises(fb.FinitoError) as e:
            ix.search_batch_runs(["ACGTNACGT" * 10])
        assert e.value.status == fb.FIN_ERR_DOMAIN


def test_device_resident_runs_entry_point(synthetic, gpu_lib):
    import torch

    fb = gpu_lib
    ds, prefix = synthetic
    seqs, offs = ds["reads"]
    n = len(offs) - 1
    with fb.FinimizerIndex.load(prefix) as ix:
        k = ix.info["k"]
        want, wc = ix.search_batch((seqs, offs), hit_bits=32)
        d_seqs = torch.from_numpy(seqs).cuda()
        d_offs = torch.from_numpy(offs.astype(np.int64)).cuda()
        cap = 16 * n
        d_ro = torch.empty(n + 1, dtype=torch.int64, device="cuda")
        d_runs = torch.empty((cap, 4), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        n_runs, c, ms = ix.search_device_runs(d_seqs.data_ptr(), d_offs.data_ptr(), n, len(seqs), d_ro.data_ptr(), d_runs.data_ptr(), cap)
        assert ms > 0 and c == wc and 0 < n_runs <= cap
        ro = d_ro.cpu().numpy().view(np.uint64)
        runs = d_runs[:n_runs].cpu().numpy().view(np.uint32).reshape(-1).view(fb.RUN_DTYPE)
        assert np.array_equal(fb.expand_runs(offs, k, ro, runs, hit_bits=32), want)
        with pytest.raises(fb.FinitoError) as e:  # capacity too small
            ix.search_device_runs(d_seqs.data_ptr(), d_offs.data_ptr(), n, len(seqs), d_ro.data_ptr(), d_runs.data_ptr(), n_runs - 1)
        assert e.value.status == fb.FIN_ERR_ARG
        # offsets the device entry points cannot trust: not ending at n_bases / decreasing -> FIN_ERR_ARG, no out-of-bounds access
        bad = offs.astype(np.int64).copy()
        bad[-1] += 5
        d_bad = torch.from_numpy(bad).cuda()
        d_out = torch.empty((want.shape[0], 2), dtype=torch.int32, device="cuda")
        with pytest.raises(fb.FinitoError) as e:
            ix.search_device(d_seqs.data_ptr(), d_bad.data_ptr(), n, len(seqs), want.shape[0], d_out.data_ptr(), 4)
        assert e.value.status == fb.FIN_ERR_ARG
        with pytest.raises(fb.FinitoError) as e:  # hit buffer smaller than the batch needs
            ix.search_device(d_seqs.data_ptr(), d_offs.data_ptr(), n, len(seqs), want.shape[0] - 1, d_out.data_ptr(), 4)
        assert e.value.status == fb.FIN_ERR_ARG
        c2, _ = ix.search_device(d_seqs.data_ptr(), d_offs.data_ptr(), n, len(seqs), want.shape[0], d_out.data_ptr(), 4)
        assert c2 == wc and np.array_equal(d_out.cpu().numpy(), want)
