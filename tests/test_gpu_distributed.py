"""GPU tests of the multi-GPU path through the C ABI's fin_builder_* stages and
finito_b200/distributed.py.  On a one-GPU box the `world` ranks of the sharded build all run on device 0 and the
all-to-all is done by hand with device copies (the CUDA export / merge / slice kernels see real multi-rank data);
with >= 2 GPUs the same build runs as one process per GPU over NCCL."""
import os
import sys

import numpy as np
import pytest

import helpers as H
from conftest import golden_cases
from fin_datasets import make_dataset

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_on_one_gpu(fb, sbwt, unitigs, prefix, world):
    """`world` ShardedBuilders on device 0; the exchange of distributed.build_fmin_sharded done with torch copies."""
    import torch

    bs = [fb.ShardedBuilder(sbwt, unitigs, prefix, r, world, device=0) for r in range(world)]
    sends, counts = [], []
    for b in bs:
        n = b.stream()
        t = torch.empty((max(1, n), 2), dtype=torch.int64, device="cuda")
        c = b.export(t.data_ptr(), n, world)
        assert int(c.sum()) == n
        sends.append(t[:n])
        counts.append([int(x) for x in c])
    bits_all, offs_all, stats = [], [], []
    for g, b in enumerate(bs):
        parts = []
        for r in range(world):
            a = sum(counts[r][:g])
            parts.append(sends[r][a:a + counts[r][g]])
        recv = torch.cat(parts) if parts else torch.empty((0, 2), dtype=torch.int64, device="cuda")
        recv = recv.contiguous() if recv.shape[0] else torch.empty((1, 2), dtype=torch.int64, device="cuda")
        n_recv = sum(counts[r][g] for r in range(world))
        n_slice = b.merge(recv.data_ptr(), n_recv)
        bits = torch.zeros(int(b.info["slice_words"]), dtype=torch.int64, device="cuda")
        offs = torch.zeros(max(1, n_slice), dtype=torch.int32, device="cuda")
        b.slice(bits.data_ptr(), offs.data_ptr())
        torch.cuda.synchronize()
        bits_all.append(bits.cpu().numpy().view(np.uint64))
        offs_all.append(offs[:n_slice].cpu().numpy().view(np.uint32))
        stats.append((b.info["bases_streamed"], n_slice, n_recv))
    bs[0].finish(np.concatenate(bits_all), np.concatenate(offs_all))
    for b in bs:
        b.close()
    return stats


@pytest.mark.parametrize("case", golden_cases())
@pytest.mark.parametrize("world", [1, 2, 5])
def test_staged_build_golden(case, world, gpu_lib, tmp_path):
    fb = gpu_lib
    d = os.path.join(H.GOLDEN, case)
    prefix = str(tmp_path / "gpu")
    build_on_one_gpu(fb, os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), prefix, world)
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(prefix + suf, os.path.join(d, "ref" + suf)), (suf, world)


@pytest.mark.parametrize("k,genome_len,world", [(31, 1_000_000, 8), (63, 300_000, 3)])
def test_staged_build_synthetic_matches_oracle(k, genome_len, world, gpu_lib, tmp_path):
    fb = gpu_lib
    ds = make_dataset("gpub", genome_len=genome_len, k=k, n_reads=2000, read_len=150, seed=100 + k)
    gp, op = str(tmp_path / "gpu"), str(tmp_path / "orc")
    stats = build_on_one_gpu(fb, ds["sbwt"], ds["unitigs"], gp, world)
    H.oracle_build(ds["sbwt"], ds["unitigs"], op)
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(gp + suf, op + suf), suf
    shares = [s[0] for s in stats]
    assert min(shares) > 0 and max(shares) < 1.5 * (sum(shares) / world)  # the stream is balanced over the ranks


def _nccl_worker(rank, world, port, sbwt, unitigs, prefix, reads_npz, index_prefix, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        sys.path.insert(0, ROOT)
        import torch
        import torch.distributed as dist

        import finito_b200 as fb
        from finito_b200.distributed import build_fmin_sharded, search_sharded

        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        b = fb.ShardedBuilder(sbwt, unitigs, prefix, rank, world, device=rank)
        st = build_fmin_sharded(b)
        b.close()
        z = np.load(reads_npz)
        with fb.FinimizerIndex.load(index_prefix, device=rank) as ix:
            all_hits, counts, _ = search_sharded(ix, z["seqs"], z["offs"])
        q.put((rank, st, all_hits, counts))
        dist.destroy_process_group()
    except Exception as e:
        q.put((rank, repr(e), None, None))
        raise


def test_nccl_build_and_search_over_all_gpus(gpu_lib, tmp_path):
    import torch
    import torch.multiprocessing as mp

    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs (the one-GPU box covers the stages with test_staged_build_*)")
    ds = make_dataset("gpu", genome_len=1_000_000, k=31, n_reads=20_000, read_len=150, seed=5)
    gp, op = str(tmp_path / "gpu"), str(tmp_path / "orc")
    H.oracle_build(ds["sbwt"], ds["unitigs"], op)
    seqs, offs = ds["reads"]
    npz = str(tmp_path / "reads.npz")
    np.savez(npz, seqs=seqs, offs=offs)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, ds["sbwt"], ds["unitigs"], gp, npz, op, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=600) for _ in range(world)), key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for suf in H.INDEX_SUFFIXES + ["_stats.txt"]:
        assert H.files_equal(gp + suf, op + suf), suf
    want, wc = H.Oracle(op).search_reads(seqs, offs)
    assert np.array_equal(res[0][2], want)
    assert all(r[3] == [int(x) for x in wc] for r in res)


def stats_on_one_gpu(fb, sbwt, unitigs, typ, t, world):
    """`world` ShardedStats on device 0; the all-reduce(MIN) and the all-gather of distributed.stats_sharded done with torch ops."""
    import torch

    ss = [fb.ShardedStats(sbwt, unitigs, typ, t, r, world, device=0) for r in range(world)]
    tables = []
    for s in ss:
        n = s.first_pass()
        if n:
            tb = torch.empty(n, dtype=torch.int32, device="cuda")
            s.table_out(tb.data_ptr())
            tables.append(tb.to(torch.int64) & 0xFFFFFFFF)
    if tables:
        red = torch.stack(tables).min(dim=0).values.to(torch.int32)
        for s in ss:
            s.table_in(red.data_ptr())
    keys = []
    for s in ss:
        n = s.second_pass()
        kb = torch.empty(max(1, n), dtype=torch.int64, device="cuda")
        s.keys(kb.data_ptr())
        keys.append(kb[:n])
    allk = torch.cat(keys).contiguous()
    torch.cuda.synchronize()
    out = [s.finish(allk.data_ptr(), int(allk.numel())) for s in ss]
    shares = [int(x.numel()) for x in keys]
    for s in ss:
        s.close()
    assert len(set(out)) == 1
    return out[0], shares


@pytest.mark.parametrize("case", golden_cases())
@pytest.mark.parametrize("world", [1, 3])
def test_staged_stats_golden(case, world, gpu_lib):
    import json

    d = os.path.join(H.GOLDEN, case)
    for mode, line in json.load(open(os.path.join(H.GOLDEN, "stats_modes.json")))[case].items():
        typ, t = mode.split(":")
        got, _ = stats_on_one_gpu(gpu_lib, os.path.join(d, "input.sbwt"), os.path.join(d, "unitigs.fna"), typ, int(t), world)
        assert f"{t},{got}\n" == line, (case, mode, world)


@pytest.mark.parametrize("typ,t,world", [("shortest", 4, 8), ("verify", 4, 5), ("shortest", 1, 2)])
def test_staged_stats_synthetic_matches_oracle(typ, t, world, gpu_lib, tmp_path):
    ds = make_dataset("gpub", genome_len=1_000_000, k=31, n_reads=2000, read_len=150, seed=131)
    want = H.oracle_build(ds["sbwt"], ds["unitigs"], str(tmp_path / "orc"), type=typ, t=t)
    got, shares = stats_on_one_gpu(gpu_lib, ds["sbwt"], ds["unitigs"], typ, t, world)
    assert got == want
    assert min(shares) > 0  # every rank produced keys
    assert gpu_lib.build_fmin(ds["sbwt"], ds["unitigs"], str(tmp_path / "one"), type=typ, t=t) == want


def _nccl_stats_worker(rank, world, port, sbwt, unitigs, typ, t, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        sys.path.insert(0, ROOT)
        import torch
        import torch.distributed as dist

        import finito_b200 as fb
        from finito_b200.distributed import stats_sharded

        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        s = fb.ShardedStats(sbwt, unitigs, typ, t, rank, world, device=rank)
        q.put((rank, stats_sharded(s)))
        s.close()
        dist.destroy_process_group()
    except Exception as e:
        q.put((rank, repr(e)))
        raise


def test_nccl_stats_over_all_gpus(gpu_lib, tmp_path):
    import torch
    import torch.multiprocessing as mp

    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs >= 2 GPUs (the one-GPU box covers the stages with test_staged_stats_*)")
    ds = make_dataset("gpub", genome_len=1_000_000, k=31, n_reads=2000, read_len=150, seed=131)
    for typ, t in (("shortest", 4), ("verify", 2)):
        want = H.oracle_build(ds["sbwt"], ds["unitigs"], str(tmp_path / f"orc{typ}"), type=typ, t=t)
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        port = 29500 + (os.getpid() + t) % 2000
        procs = [ctx.Process(target=_nccl_stats_worker, args=(r, world, port, ds["sbwt"], ds["unitigs"], typ, t, q)) for r in range(world)]
        for p in procs:
            p.start()
        res = [q.get(timeout=600) for _ in range(world)]
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        assert all(r[1] == want for r in res), (typ, res)
