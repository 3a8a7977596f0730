"""The incremental FASTA / FASTQ (.gz) reader behind the CLI and build-fmin (finito_b200/csrc/seqio.hpp) against the
behaviour of the reference's reader (SBWT/include/sbwt/SeqIO.hh:190-292): headers dropped, FASTA sequences may span lines,
FASTQ is four lines per record, bases upper-cased, '\\r' kept, an empty sequence is an error — with refill sizes that put a
chunk boundary inside every kind of token, and batch limits that cut between any two records."""
import gzip
import random

import pytest

import helpers as H


def records(rng, n):
    out = []
    for i in range(n):
        ln = rng.choice([1, 2, 7, 31, 64, 150, 151, 400])
        out.append("".join(rng.choice("ACGTacgt") for _ in range(ln)))
    return out


def write_fasta(path, recs, rng, multiline):
    with open(path, "w") as f:
        for i, r in enumerate(recs):
            f.write(f">read {i} some text\n")
            if multiline and len(r) > 3:
                w = rng.choice([1, 3, 60, 70])
                f.write("\n".join(r[j:j + w] for j in range(0, len(r), w)) + "\n")
            else:
                f.write(r + "\n")


def write_fastq(path, recs, last_newline=True):
    with open(path, "w") as f:
        for i, r in enumerate(recs):
            tail = "\n" if (last_newline or i + 1 < len(recs)) else ""
            f.write(f"@r{i}\n{r}\n+\n{'I' * len(r)}{tail}")


@pytest.mark.parametrize("chunk", [16, 17, 64, 1000, 16 << 20])
def test_reader_is_independent_of_refills_and_batch_limits(chunk, tmp_path):
    rng = random.Random(chunk)
    recs = records(rng, 300)
    want = [r.upper().encode() for r in recs]
    fa, faml, fq, fq2 = (str(tmp_path / n) for n in ("a.fna", "b.fasta", "c.fastq", "d.fq"))
    write_fasta(fa, recs, rng, False)
    write_fasta(faml, recs, rng, True)
    write_fastq(fq, recs)
    write_fastq(fq2, recs, last_newline=False)
    for p in (fa, faml, fq, fq2):
        got, nb = H.emu_read_sequences(p, chunk)
        assert got == want and nb == 1, p
        gz = p.rsplit(".", 1)[0] + "_z." + p.rsplit(".", 1)[1] + ".gz"
        with gzip.open(gz, "wb") as f:
            f.write(open(p, "rb").read())
        got, _ = H.emu_read_sequences(gz, chunk)
        assert got == want, gz
        got, nb = H.emu_read_sequences(p, chunk, max_reads=7)
        assert got == want and nb == (len(want) + 6) // 7
        got, nb = H.emu_read_sequences(p, chunk, max_bases=500)
        assert got == want and nb > 10


def test_reader_errors_like_the_reference(tmp_path):
    p = str(tmp_path / "empty_seq.fna")
    open(p, "w").write(">a\nACGT\n>b\n>c\nAC\n")
    with pytest.raises(RuntimeError, match="empty sequence"):
        H.emu_read_sequences(p, 16)
    p = str(tmp_path / "nohdr.fna")
    open(p, "w").write("ACGT\n")
    with pytest.raises(RuntimeError, match="does not start with"):
        H.emu_read_sequences(p)
    p = str(tmp_path / "x.txtt")
    open(p, "w").write(">a\nA\n")
    with pytest.raises(RuntimeError, match="Unknown file format"):
        H.emu_read_sequences(p)
    p = str(tmp_path / "cr.fna")  # '\r' is not stripped by the reference's reader: it stays in the sequence
    open(p, "wb").write(b">a\r\nACGT\r\n")
    got, _ = H.emu_read_sequences(p, 16)
    assert got == [b"ACGT\r"]
