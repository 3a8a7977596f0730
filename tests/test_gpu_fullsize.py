"""BASELINE.json-sized inputs (configs[1]: 5 Mbp genome, k=31, 150 bp reads at 1 % error; configs[3]'s k=63) checked
through properties that do not need the oracle to finish the whole job: every reported (unitig, offset) is verified
against the unitig text itself, unitigs queried against their own index land on (their colex rank, 0..n), the build is
idempotent and independent of how the stream is sharded, and an oracle-sized sample of the same reads is bit-exact."""
import os

import numpy as np
import pytest

import helpers as H
from fin_datasets import make_dataset

pytestmark = pytest.mark.gpu

COMP = bytes.maketrans(b"ACGT", b"TGCA")


def colex_ranked(unitigs, k):
    order = sorted(range(len(unitigs)), key=lambda i: unitigs[i][:k][::-1])
    return [unitigs[i] for i in order]


# configs[1] (5 Mbp, k=31), a small k=63 case, and configs[3]'s shape: 100 Mbp, k=63 (MAX_KMER_LENGTH-sized keys: 128-bit k-mers in
# fin_sbwt_build, the 64-slot candidate window in the search) - that one is built by the device pipeline
PARAMS, IDS = [(31, 5_000_000, 1_000_000), (63, 2_000_000, 300_000), (63, 100_000_000, 300_000)], ["config2_k31", "k63", "config4_k63_100Mbp"]
if os.environ.get("FIN_FULLSIZE_GENOME"):  # one-off runs at other sizes, e.g. 200000000 for an index far beyond the L2
    PARAMS.append((int(os.environ.get("FIN_FULLSIZE_K", "31")), int(os.environ["FIN_FULLSIZE_GENOME"]), 1_000_000))
    IDS.append("env_size")


def device_built_dataset(fb, d, genome_len, k, n_reads):
    """genome (tools/fin_datagen bigenome, 30 % repeats) -> fin_spss_build -> fin_sbwt_build; reads from the same genome"""
    import subprocess

    d = str(d)
    genome, spss, sbwt, reads_bin = (os.path.join(d, f) for f in ("genome.fna", "unitigs.fna", "index.sbwt", "reads.bin"))
    datagen = os.path.join(H.TOOLS, "fin_datagen")
    subprocess.run([datagen, "bigenome", genome, str(genome_len), "20260104", "1", "0.30", str(max(2000, genome_len // 50_000))], check=True,
                   capture_output=True)
    fb.spss_build(genome, k, spss, flip=True, seed=7)
    fb.sbwt_build(spss, k, sbwt, 8)
    subprocess.run([datagen, "reads", genome, str(n_reads), "150", "0.01", "9", reads_bin], check=True, capture_output=True)
    seqs = np.fromfile(reads_bin, dtype=np.uint8)
    offs = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(150)
    return {"dir": d, "genome": genome, "unitigs": spss, "sbwt": sbwt, "reads": (seqs, offs), "k": k, "n_reads": n_reads, "read_len": 150}


@pytest.fixture(scope="module", params=PARAMS, ids=IDS)
def big(request, gpu_lib, tmp_path_factory):
    k, genome_len, n_reads = request.param
    if genome_len >= 50_000_000:
        ds = device_built_dataset(gpu_lib, tmp_path_factory.mktemp("big"), genome_len, k, n_reads)
    else:
        ds = make_dataset("full", genome_len=genome_len, k=k, n_reads=n_reads, read_len=150, seed=20260102 + k,
                          families=max(1, genome_len // 100_000))
    prefix = str(tmp_path_factory.mktemp("full") / "gpu")
    gpu_lib.build_fmin(ds["sbwt"], ds["unitigs"], prefix)
    return ds, prefix


def test_every_hit_points_at_its_kmer(big, gpu_lib):
    fb = gpu_lib
    ds, prefix = big
    k = ds["k"]
    seqs, offs = ds["reads"]
    with fb.FinimizerIndex.load(prefix) as ix:
        hits, c = ix.search_batch((seqs, offs), hit_bits=32)
        info = ix.info
    n_reads, rl = ds["n_reads"], ds["read_len"]
    per = rl - k + 1
    assert hits.shape[0] == n_reads * per == c.n_queries
    found = hits[:, 0] != -1
    assert int(found.sum()) == c.positive
    assert 0.6 < c.positive / c.n_queries < 0.8 if k == 31 else 0.4 < c.positive / c.n_queries < 0.7  # ~0.99^k
    assert (hits[~found] == -1).all()
    assert hits[found, 0].max() < info["n_unitigs"] and hits[found, 1].min() >= 0
    unitigs = [u.encode() for u in colex_ranked(H.read_fasta(ds["unitigs"]), k)]
    rng = np.random.default_rng(1)
    raw = seqs.tobytes()
    for idx in rng.choice(np.flatnonzero(found), size=100_000, replace=False):
        r, i = divmod(int(idx), per)
        q = raw[r * rl + i: r * rl + i + k]
        u, p = int(hits[idx, 0]), int(hits[idx, 1])
        t = unitigs[u][p:p + k]
        assert t == q or t == q.translate(COMP)[::-1], (r, i, u, p)


def test_sample_is_bit_exact_against_the_oracle(big, gpu_lib, tmp_path):
    fb = gpu_lib
    ds, prefix = big
    seqs, offs = ds["reads"]
    n = 5000
    sub = (seqs[: int(offs[n])], offs[: n + 1])
    want, wc = H.Oracle(prefix).search_reads(*sub)
    with fb.FinimizerIndex.load(prefix) as ix:
        got, c = ix.search_batch(sub)
    assert np.array_equal(got, want)
    assert [c.found_fwd, c.found_rc, c.positive, c.n_queries] == [int(x) for x in wc]


def test_unitigs_find_themselves(big, gpu_lib):
    fb = gpu_lib
    ds, prefix = big
    k = ds["k"]
    ranked = colex_ranked(H.read_fasta(ds["unitigs"]), k)
    with fb.FinimizerIndex.load(prefix) as ix:
        hits, c = ix.search_batch(ranked, hit_bits=32)
    lens = np.array([len(u) - k + 1 for u in ranked])
    assert c.positive == c.n_queries == int(lens.sum())
    assert np.array_equal(hits[:, 0], np.repeat(np.arange(len(ranked)), lens))
    starts = np.concatenate([[0], np.cumsum(lens)[:-1]])
    assert np.array_equal(hits[:, 1], np.arange(int(lens.sum())) - np.repeat(starts, lens))


def test_build_is_idempotent_and_shard_independent(big, gpu_lib, tmp_path):
    from test_gpu_distributed import build_on_one_gpu

    fb = gpu_lib
    ds, prefix = big
    again = str(tmp_path / "again")
    fb.build_fmin(ds["sbwt"], ds["unitigs"], again, lcs_path=prefix + ".LCS.sdsl")
    sharded = str(tmp_path / "sharded")
    build_on_one_gpu(fb, ds["sbwt"], ds["unitigs"], sharded, 4)
    for suf in H.INDEX_SUFFIXES:
        assert H.files_equal(again + suf, prefix + suf), suf
        assert H.files_equal(sharded + suf, prefix + suf), suf
