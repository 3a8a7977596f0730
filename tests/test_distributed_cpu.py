"""world_size-2 (and 3) runs of finito_b200/distributed.py on the CPU with the gloo backend: the host side of the
multi-GPU path (read sharding + counter reduction for search-fmin, the all-to-all tuple exchange and slice gather of
build-fmin).  The per-rank compute is the CPU emulation of the kernel bodies / the oracle — test infrastructure; on a
GPU box the same drivers run over finito_b200.ShardedBuilder / FinimizerIndex with NCCL (tests/test_gpu_distributed.py)."""
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as H
from conftest import golden_cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _init(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    dist.init_process_group("gloo", rank=rank, world_size=world)


def _build_worker(rank, world, port, sbwt, unitigs, prefix, q):
    try:
        _init(rank, world, port)
        from finito_b200.distributed import build_fmin_sharded

        b = H.EmuShardedBuilder(sbwt, unitigs, prefix, rank, world)
        q.put((rank, build_fmin_sharded(b)))
        b.close()
        dist.destroy_process_group()
    except Exception as e:  # surface the failure in the parent
        q.put((rank, repr(e)))
        raise


class _OracleIndex:
    """search_batch of a rank, answered by the oracle (stand-in for FinimizerIndex on a box without GPU)."""

    def __init__(self, prefix):
        self.orc = H.Oracle(prefix)

    def search_batch(self, reads, hit_bits=64):
        from finito_b200 import SearchCounts

        seqs, offs = reads
        hits, c = self.orc.search_reads(np.ascontiguousarray(seqs), np.ascontiguousarray(offs))
        return hits, SearchCounts(*(int(x) for x in c))


def _search_worker(rank, world, port, prefix, reads_npz, q):
    try:
        _init(rank, world, port)
        from finito_b200.distributed import search_sharded

        z = np.load(reads_npz)
        all_hits, counts, mine = search_sharded(_OracleIndex(prefix), z["seqs"], z["offs"])
        q.put((rank, all_hits, counts, mine.shape[0]))
        dist.destroy_process_group()
    except Exception as e:
        q.put((rank, repr(e), None, None))
        raise


def _spawn(fn, world, args):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() * 7 + world) % 2000
    procs = [ctx.Process(target=fn, args=(r, world, port, *args, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return sorted(out, key=lambda x: x[0])


@pytest.mark.parametrize("case,world", [("syn_k31", 2), ("syn_k63", 2), ("syn_k12", 3), ("kat_paper_example", 2), ("kat_walk", 3)])
def test_sharded_build_writes_the_reference_files(case, world, tmp_path):
    d = os.path.join(H.GOLDEN, case)
    prefix = str(tmp_path / "shard")
    res = _spawn(_build_worker, world, (os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix))
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(prefix + suf, os.path.join(d, "ref" + suf)), suf
    stats = [r[1] for r in res]
    assert all(isinstance(s, dict) for s in stats), stats
    assert sum(s["tuples_sent"] for s in stats) == sum(s["tuples_received"] for s in stats)
    assert sum(s["slice_finimizers"] for s in stats) == stats[0]["finimizers"]
    if case.startswith("syn"):
        assert all(s["bases_streamed"] > 0 for s in stats)  # every rank had a share of the stream


def test_shard_ranges_cover_everything_once():
    from finito_b200.distributed import shard_range

    for n in (0, 1, 2, 7, 100, 101):
        for world in (1, 2, 3, 8):
            blocks = [shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_search_equals_single_process(world, tmp_path):
    d = os.path.join(H.GOLDEN, "syn_k31")
    reads = H.read_fasta(os.path.join(d, "queries.fna"))
    seqs, offs = H.batch_of(reads)
    npz = str(tmp_path / "reads.npz")
    np.savez(npz, seqs=seqs, offs=offs)
    res = _spawn(_search_worker, world, (os.path.join(d, "ref"), npz))
    want = H.parse_output(os.path.join(d, "ref_out.txt"))
    assert np.array_equal(res[0][1], want)          # rank 0 holds all hits in input order
    assert all(r[1] is None for r in res[1:])
    _, wc = H.Oracle(os.path.join(d, "ref")).search_reads(seqs, offs)
    for r in res:
        assert r[2] == [int(x) for x in wc]           # every rank holds the summed counters
    assert sum(r[3] for r in res) == want.shape[0]


def _stats_worker(rank, world, port, sbwt, unitigs, typ, t, q):
    try:
        _init(rank, world, port)
        from finito_b200.distributed import stats_sharded

        s = H.EmuShardedStats(sbwt, unitigs, typ, t, rank, world)
        q.put((rank, stats_sharded(s)))
        s.close()
        dist.destroy_process_group()
    except Exception as e:
        q.put((rank, repr(e)))
        raise


@pytest.mark.parametrize("case,world", [("syn_k12", 3), ("kat_paper_example", 2)])
def test_sharded_stats_modes_print_the_reference_line(case, world):
    """build-fmin --type shortest|verify -t T sharded over `world` ranks (one all-reduce(MIN) of the first-window tables, one
    all-gather of the keys): the `<t>,<stats>` line the unmodified reference wrote (tests/golden/stats_modes.json)."""
    import json

    d = os.path.join(H.GOLDEN, case)
    for mode, line in json.load(open(os.path.join(H.GOLDEN, "stats_modes.json")))[case].items():
        typ, t = mode.split(":")
        res = _spawn(_stats_worker, world, (os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), typ, int(t)))
        for r in res:
            assert f"{t},{r[1]}\n" == line, (case, mode, r)
