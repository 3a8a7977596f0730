"""The finito_b200 CLI (finito_b200/csrc/cli_main.cpp) as a drop-in for the reference's `build-fmin` / `search-fmin`
(src/main.cpp:13-59): same command lines, same files, same text.  CPU part: usage and error behaviour.  GPU part: run
side by side with the UNMODIFIED reference binary (oracle/_ref/finito_ref, prebuilt, travels to the GPU box) on
BASELINE.json configs[1]-shaped input and compare every output byte."""
import os
import subprocess

import numpy as np
import pytest

import finito_b200 as fb
import helpers as H
from fin_datasets import make_dataset

CLI = fb.CLI_PATH


def run(*args, cwd=None):
    return subprocess.run([CLI, *args], cwd=cwd, capture_output=True, text=True)


def test_usage_and_exit_codes_like_the_reference():
    r = run()
    assert r.returncode == 1 and "Available commands" in r.stderr and "build-fmin" in r.stderr  # src/main.cpp:21-31
    for cmd in ("build-fmin", "search-fmin"):
        r = run(cmd)
        assert r.returncode == 1 and "Usage:" in r.stderr and r.stdout == ""  # search_fmin.hh:148-151, build_fmin.hh:327-330
        assert run(cmd, "-h").returncode == 1
    r = run("nonsense")
    assert r.returncode == 1 and "Invalid command: nonsense" in r.stderr  # src/main.cpp:49-57
    if H.have_ref():
        ref = H.ref_run()
        assert "Available commands" in ref.stderr and ref.returncode == 1


def test_unreadable_inputs_fail_cleanly(tmp_path):
    d = os.path.join(H.GOLDEN, "syn_k31")
    r = run("build-fmin", "-o", str(tmp_path / "x"), "-i", "/nonexistent.sbwt", "-u", os.path.join(d, "unitigs.fna"))
    assert r.returncode == 1 and not os.path.exists(tmp_path / "x.FBV.sdsl")
    r = run("search-fmin", "-i", os.path.join(d, "ref"), "-q", "/nonexistent.fna")
    assert r.returncode == 1 and r.stdout == ""


def write_fasta(path, seqs, offs, n):
    with open(path, "wb") as f:
        raw = seqs.tobytes()
        for i in range(n):
            f.write(b">r%d\n" % i)
            f.write(raw[int(offs[i]):int(offs[i + 1])])
            f.write(b"\n")


@pytest.mark.gpu
@pytest.mark.skipif(not H.have_ref(), reason="needs the prebuilt reference binary oracle/_ref/finito_ref")
@pytest.mark.parametrize("k,genome_len,n_reads", [(31, 2_000_000, 20_000), (63, 500_000, 5_000)])
def test_cli_is_byte_identical_to_the_reference_binary(k, genome_len, n_reads, tmp_path):
    ds = make_dataset("cli", genome_len=genome_len, k=k, n_reads=n_reads, read_len=150, seed=77 + k)
    seqs, offs = ds["reads"]
    q = str(tmp_path / "reads.fna")
    write_fasta(q, seqs, offs, n_reads)
    ours, ref = str(tmp_path / "ours"), str(tmp_path / "ref")
    r = run("build-fmin", "-o", ours, "-i", ds["sbwt"], "-u", ds["unitigs"])
    assert r.returncode == 0, r.stderr[-500:]
    assert r.stdout == "String appended to the file successfully.\n"  # build_fmin.hh:396
    rr = H.ref_run("build-fmin", "-o", ref, "-i", ds["sbwt"], "-u", ds["unitigs"])
    assert rr.returncode == 0, rr.stderr[-500:]
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(ours + suf, ref + suf), suf
    # each binary searches the OTHER one's index: the files are interchangeable
    r = run("search-fmin", "-i", ref, "-q", q, "-o", ours + ".out")
    assert r.returncode == 0, r.stderr[-500:]
    rr = H.ref_run("search-fmin", "-i", ours, "-q", q, "-o", ref + ".out")
    assert rr.returncode == 0, rr.stderr[-500:]
    assert H.files_equal(ours + ".out", ref + ".out")
    # stdout mode
    r = subprocess.run([CLI, "search-fmin", "-i", ours, "-q", q], capture_output=True)
    assert r.returncode == 0 and r.stdout == open(ref + ".out", "rb").read()
    # stats modes print the same line
    for typ, t in (("shortest", 4),):  # verify is O(k^2) per window in the reference: covered by golden/stats_modes.json
        a = run("build-fmin", "-o", ours + typ, "-i", ds["sbwt"], "-u", ds["unitigs"], "--type", typ, "-t", str(t), "--lcs", ours + ".LCS.sdsl")
        b = H.ref_run("build-fmin", "-o", ref + typ, "-i", ds["sbwt"], "-u", ds["unitigs"], "--type", typ, "-t", str(t), "--lcs", ours + ".LCS.sdsl")
        assert a.returncode == 0 and b.returncode == 0, (a.stderr[-300:], b.stderr[-300:])
        assert open(ours + typ + "_stats.txt").read() == open(ref + typ + "_stats.txt").read(), typ


@pytest.mark.gpu
def test_cli_search_sharded_over_all_gpus_prints_the_same(tmp_path):
    """search-fmin cuts every batch into one block of reads per visible GPU (FIN_DEVICES caps it): same bytes out."""
    if fb.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    ds = make_dataset("cli", genome_len=2_000_000, k=31, n_reads=20_000, read_len=150, seed=108)
    seqs, offs = ds["reads"]
    q = str(tmp_path / "reads.fna")
    write_fasta(q, seqs, offs, 20_000)
    prefix = str(tmp_path / "ix")
    assert run("build-fmin", "-o", prefix, "-i", ds["sbwt"], "-u", ds["unitigs"]).returncode == 0
    outs = []
    for n in ("1", str(fb.device_count())):
        o = str(tmp_path / f"out{n}.txt")
        r = subprocess.run([CLI, "search-fmin", "-i", prefix, "-q", q, "-o", o], capture_output=True, text=True,
                           env=dict(os.environ, FIN_DEVICES=n))
        assert r.returncode == 0, r.stderr[-400:]
        if n != "1":
            assert f"on {n} GPUs" in r.stderr
        outs.append(open(o, "rb").read())
    assert outs[0] == outs[1] and len(outs[0]) > 0


@pytest.mark.gpu
def test_cli_input_formats_fastq_gzip_and_file_lists(tmp_path):
    """The reader's other inputs (SBWT/src/SeqIO.cpp:19-52: format from the extension, .gz through zlib) and the `.txt`
    list mode of -q / -o (search_fmin.hh:153-172), against the text the reference wrote for the same queries."""
    import gzip

    d = os.path.join(H.GOLDEN, "syn_k31")
    want = open(os.path.join(d, "ref_out.txt"), "rb").read()
    reads = H.read_fasta(os.path.join(d, "queries.fna"))
    fq = str(tmp_path / "q.fastq")
    with open(fq, "w") as f:
        for i, r in enumerate(reads):
            f.write(f"@r{i}\n{r.lower() if i % 3 == 0 else r}\n+\n{'I' * len(r)}\n")
    fagz, fqgz = str(tmp_path / "q.fna.gz"), str(tmp_path / "q.fq.gz")
    with gzip.open(fagz, "wb") as f:
        f.write(open(os.path.join(d, "queries.fna"), "rb").read())
    with gzip.open(fqgz, "wb") as f:
        f.write(open(fq, "rb").read())
    for q in (fq, fagz, fqgz):
        out = str(tmp_path / "out.txt")
        r = run("search-fmin", "-i", os.path.join(d, "ref"), "-q", q, "-o", out)
        assert r.returncode == 0, (q, r.stderr[-300:])
        assert open(out, "rb").read() == want, q
    # list mode: two query files, two output files
    qlist, olist = str(tmp_path / "queries.txt"), str(tmp_path / "outs.txt")
    open(qlist, "w").write(f"{fq}\n{fagz}\n")
    open(olist, "w").write(f"{tmp_path / 'o1.txt'}\n{tmp_path / 'o2.txt'}\n")
    r = run("search-fmin", "-i", os.path.join(d, "ref"), "-q", qlist, "-o", olist)
    assert r.returncode == 0, r.stderr[-300:]
    assert open(tmp_path / "o1.txt", "rb").read() == want and open(tmp_path / "o2.txt", "rb").read() == want
    open(olist, "w").write(f"{tmp_path / 'o1.txt'}\n")
    r = run("search-fmin", "-i", os.path.join(d, "ref"), "-q", qlist, "-o", olist)
    assert r.returncode == 1 and "does not match" in r.stderr
    # unitigs as gzipped FASTA for build-fmin
    ugz = str(tmp_path / "unitigs.fna.gz")
    with gzip.open(ugz, "wb") as f:
        f.write(open(os.path.join(d, "unitigs.fna"), "rb").read())
    prefix = str(tmp_path / "fromgz")
    assert run("build-fmin", "-o", prefix, "-i", os.path.join(d, "input.sbwt"), "-u", ugz).returncode == 0
    for suf in H.INDEX_SUFFIXES:
        assert H.files_equal(prefix + suf, os.path.join(d, "ref" + suf)), suf


@pytest.mark.gpu
@pytest.mark.skipif(not H.have_ref(), reason="needs the prebuilt reference binary oracle/_ref/finito_ref")
def test_cli_one_million_reads_of_config_2_against_the_reference(tmp_path):
    """BASELINE.json configs[1] at 1 M reads (120 M k-mer positions, 1.3 GB of text): the CLI's output against the
    unmodified reference binary run as one process per host core over consecutive shards of the same reads
    (the reference is single-threaded, search_fmin.hh:43), sha256 of the concatenated text."""
    import hashlib
    from concurrent.futures import ThreadPoolExecutor

    n = 1_000_000
    ds = make_dataset("bench", genome_len=5_000_000, k=31, n_reads=n, read_len=150, seed=20260102, families=50)
    seqs, offs = ds["reads"]
    prefix = str(tmp_path / "ix")
    assert run("build-fmin", "-o", prefix, "-i", ds["sbwt"], "-u", ds["unitigs"]).returncode == 0
    rec = np.empty((n, 150 + 3), dtype=np.uint8)
    rec[:, 0], rec[:, 1], rec[:, -1] = ord(">"), ord("\n"), ord("\n")
    rec[:, 2:152] = seqs.reshape(n, 150)
    q = str(tmp_path / "reads.fna")
    rec.tofile(q)
    out = str(tmp_path / "ours.txt")
    r = run("search-fmin", "-i", prefix, "-q", q, "-o", out)
    assert r.returncode == 0, r.stderr[-500:]
    shards = max(1, len(os.sched_getaffinity(0)))
    per = (n + shards - 1) // shards
    paths = []
    for s in range(shards):
        a, b = s * per, min(n, (s + 1) * per)
        if a < b:
            p = str(tmp_path / f"shard{s}.fna")
            rec[a:b].tofile(p)
            paths.append(p)
    with ThreadPoolExecutor(len(paths)) as ex:
        rcs = list(ex.map(lambda p: H.ref_run("search-fmin", "-i", prefix, "-q", p, "-o", p + ".out").returncode, paths))
    assert rcs == [0] * len(paths)
    want = hashlib.sha256()
    for p in paths:
        with open(p + ".out", "rb") as f:
            for blk in iter(lambda: f.read(1 << 24), b""):
                want.update(blk)
    got = hashlib.sha256()
    with open(out, "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            got.update(blk)
    assert got.hexdigest() == want.hexdigest()


@pytest.mark.gpu
def test_cli_streams_many_batches_and_appends(tmp_path):
    """More reads than one batch of the pipeline (2^19) with ragged lengths, multi-line FASTA records and an output that is
    opened in append mode by the shell (no positional writes there): same text as the library call."""
    ds = make_dataset("cli", genome_len=2_000_000, k=31, n_reads=20_000, read_len=150, seed=108)
    seqs, offs = ds["reads"]
    prefix = str(tmp_path / "ix")
    assert run("build-fmin", "-o", prefix, "-i", ds["sbwt"], "-u", ds["unitigs"]).returncode == 0
    raw = seqs.tobytes()
    rng = np.random.default_rng(3)
    reads = []
    for i in range(600_000):
        j = int(rng.integers(0, 20_000))
        r = raw[int(offs[j]):int(offs[j + 1])]
        reads.append(r[: int(rng.integers(1, 151))] if i % 5 == 0 else r)
    q = str(tmp_path / "many.fna")
    with open(q, "wb") as f:
        for i, r in enumerate(reads):
            f.write(b">x\n" + (r[:70] + b"\n" + r[70:] if i % 3 == 0 and len(r) > 70 else r) + b"\n")
    s2, o2 = H.batch_of(reads)
    with fb.FinimizerIndex.load(prefix) as ix:
        hits, _ = ix.search_batch((s2, o2))
        k = ix.info["k"]
    want = fb.format_hits(hits, [max(0, len(r) - k + 1) for r in reads]).encode()
    out = str(tmp_path / "out.txt")
    r = run("search-fmin", "-i", prefix, "-q", q, "-o", out)
    assert r.returncode == 0, r.stderr[-400:]
    assert open(out, "rb").read() == want
    app = str(tmp_path / "app.txt")
    open(app, "wb").write(b"header line\n")
    rc = subprocess.run(f"{CLI} search-fmin -i {prefix} -q {q} >> {app}", shell=True, capture_output=True).returncode
    assert rc == 0 and open(app, "rb").read() == b"header line\n" + want
