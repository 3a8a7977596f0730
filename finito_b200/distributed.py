"""Multi-GPU drivers of the finimizer path: one process per GPU, torch.distributed for the plumbing.

search-fmin  — reads are independent (include/search_fmin.hh:43 has no cross-read state but counters): every rank loads
               the whole index, takes a contiguous block of the reads, and only the four counters are summed.  No
               data-path collective.
build-fmin   — the stream over the permuted unitigs (include/FinimizerIndex.hh:292-299) is split over the ranks; the
               (finimizer column, overwrite-rule key) tuples meet at the owner of the column in ONE all-to-all, the
               owners fold them with the same max as the single-GPU table, and rank 0 gathers the column slices and
               writes the files.  `builder` is a finito_b200.ShardedBuilder (device buffers, NCCL) — or, in the CPU
               tests, an object with the same methods over host memory (gloo).

build-fmin --type shortest|verify — every rank streams a share of the unitigs (shortest) or of the text positions
               (verify); the per-column "first window" tables meet in ONE all-reduce(MIN), the key sets in ONE all-gather,
               and every rank can then sort + unique + sum them (the reference's std::set, include/build_fmin.hh:134-214).

Nothing here computes: the kernels are behind the C ABI.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block of `n_items` for `rank`: sizes differ by at most one, blocks are in rank order."""
    base, extra = divmod(n_items, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_reads(seqs: np.ndarray, offsets: np.ndarray, rank: int, world: int):
    """This rank's block of a read batch as (bases, offsets rebased to 0, index of its first read)."""
    a, b = shard_range(len(offsets) - 1, rank, world)
    lo, hi = int(offsets[a]), int(offsets[b])
    return seqs[lo:hi], (offsets[a:b + 1] - offsets[a]).astype(np.uint64), a


def search_sharded(index, seqs: np.ndarray, offsets: np.ndarray, group=None, hit_bits: int = 64, gather: bool = True):
    """search-fmin over all ranks of `group`: each searches its block of reads on its own GPU against its own copy of
    the index.  Returns (hits of ALL reads in input order on rank 0 — None elsewhere or with gather=False —,
    the counters summed over the ranks, this rank's hits)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    my_seqs, my_offs, _ = shard_reads(seqs, offsets, rank, world)
    hits, c = index.search_batch((my_seqs, my_offs), hit_bits=hit_bits)
    dev = _collective_device(group)
    counts = torch.tensor([c.found_fwd, c.found_rc, c.positive, c.n_queries], dtype=torch.int64, device=dev)
    dist.all_reduce(counts, group=group)
    all_hits = None
    if gather:
        sizes = torch.zeros(world, dtype=torch.int64, device=dev)
        sizes[rank] = hits.shape[0]
        dist.all_reduce(sizes, group=group)
        sizes = [int(x) for x in sizes.cpu()]
        mine = torch.from_numpy(np.ascontiguousarray(hits)).to(dev)
        pad = torch.zeros((max(sizes), 2), dtype=mine.dtype, device=dev)
        pad[: mine.shape[0]] = mine
        parts = [torch.empty_like(pad) for _ in range(world)] if rank == 0 else None
        dist.gather(pad, parts, dst=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        if rank == 0:
            all_hits = np.concatenate([p[:n].cpu().numpy() for p, n in zip(parts, sizes)])
    return all_hits, [int(x) for x in counts.cpu()], hits


def _collective_device(group) -> torch.device:
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")


def build_fmin_sharded(builder, group=None) -> dict:
    """Drives one rank's builder through stream -> all-to-all -> merge -> gather -> (rank 0) finish.
    Returns sizes and timings of this rank's share."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if builder.memory == "cuda" else torch.device("cpu")
    info = builder.info
    n_tuples = builder.stream()

    # ---- the exchange: counts first, then the tuples themselves ({u64 column, u64 key} as 2 x int64)
    send = torch.empty((max(1, n_tuples), 2), dtype=torch.int64, device=dev)
    send_counts = builder.export(send.data_ptr(), n_tuples, world)
    assert int(send_counts.sum()) == n_tuples
    sc = torch.from_numpy(send_counts.astype(np.int64)).to(dev)
    rc = torch.empty_like(sc)
    dist.all_to_all_single(rc, sc, group=group)
    recv_counts = [int(x) for x in rc.cpu()]
    n_recv = sum(recv_counts)
    recv = torch.empty((max(1, n_recv), 2), dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv[:n_recv], send[:n_tuples], output_split_sizes=recv_counts,
                           input_split_sizes=[int(x) for x in send_counts], group=group)
    if dev.type == "cuda":
        torch.cuda.synchronize()
    n_slice = builder.merge(recv.data_ptr(), n_recv)

    # ---- gather the column slices on rank 0: fmin words (fixed size) and offsets (ragged -> padded)
    words = int(info["slice_words"])
    bits = torch.zeros(words, dtype=torch.int64, device=dev)
    offs = torch.zeros(max(1, n_slice), dtype=torch.int32, device=dev)
    builder.slice(bits.data_ptr(), offs.data_ptr())
    ns = torch.zeros(world, dtype=torch.int64, device=dev)
    ns[rank] = n_slice
    dist.all_reduce(ns, group=group)
    ns = [int(x) for x in ns.cpu()]
    pad = torch.zeros(max(1, max(ns)), dtype=torch.int32, device=dev)
    pad[:n_slice] = offs[:n_slice]
    root = dist.get_global_rank(group, 0) if group is not None else 0
    bit_parts = [torch.empty_like(bits) for _ in range(world)] if rank == 0 else None
    off_parts = [torch.empty_like(pad) for _ in range(world)] if rank == 0 else None
    dist.gather(bits, bit_parts, dst=root, group=group)
    dist.gather(pad, off_parts, dst=root, group=group)
    if rank == 0:
        bits_all = np.concatenate([p.cpu().numpy().view(np.uint64) for p in bit_parts])
        offs_all = np.concatenate([p[:n].cpu().numpy().view(np.uint32) for p, n in zip(off_parts, ns)])
        builder.finish(bits_all, offs_all)
    dist.barrier(group=group)
    return {"rank": rank, "world": world, "tuples_sent": n_tuples, "tuples_received": n_recv, "slice_finimizers": n_slice,
            "finimizers": sum(ns), "bases_streamed": int(info["bases_streamed"]), "stream_ms": float(builder.info["stream_ms"])}


def stats_sharded(stats, group=None, write_line: bool = False) -> str:
    """Drives one rank's ShardedStats (or the CPU stand-in of the tests) through first pass -> all-reduce(MIN) of the tables ->
    second pass -> all-gather of the keys -> finish.  Returns the stats string (the same on every rank)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if stats.memory == "cuda" else torch.device("cpu")
    n_table = stats.first_pass()
    if n_table:
        table = torch.empty(n_table, dtype=torch.int32, device=dev)  # u32 entries, 0xFFFFFFFF = no such window: unsigned MIN
        stats.table_out(table.data_ptr())
        as_i64 = table.to(torch.int64) & 0xFFFFFFFF
        dist.all_reduce(as_i64, op=dist.ReduceOp.MIN, group=group)
        table = as_i64.to(torch.int32)  # wraps the values >= 2^31 back to their bit pattern
        stats.table_in(table.data_ptr())
    n_keys = stats.second_pass()
    mine = torch.empty(max(1, n_keys), dtype=torch.int64, device=dev)
    stats.keys(mine.data_ptr())
    sizes = torch.zeros(world, dtype=torch.int64, device=dev)
    sizes[rank] = n_keys
    dist.all_reduce(sizes, group=group)
    sizes = [int(x) for x in sizes.cpu()]
    pad = torch.zeros(max(1, max(sizes)), dtype=torch.int64, device=dev)
    pad[:n_keys] = mine[:n_keys]
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    all_keys = torch.cat([p[:n] for p, n in zip(parts, sizes)]).contiguous()
    if dev.type == "cuda":
        torch.cuda.synchronize()
    return stats.finish(all_keys.data_ptr(), int(all_keys.numel()), write_line=write_line and rank == 0)
