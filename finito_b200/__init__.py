"""finito_b200 — ctypes binding of the C ABI in include/finito_b200.h.

The product is the CUDA library ``finito_b200/lib/libfinito_b200.so`` (built by
``__graft_entry__.build()`` / ``make -C finito_b200/csrc``) and the ``finito_b200`` CLI next to it.
This module only mirrors the reference's object API (``FinimizerIndex::load`` / ``search``,
include/FinimizerIndex.hh:119,209 and the ``build-fmin`` driver, include/build_fmin.hh:302) on top of
those entry points so that tests and ``bench.py`` can call them.  There is no Python or CPU
implementation behind it: if the library is missing, or no CUDA device is visible, calls fail loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libfinito_b200.so")
CLI_PATH = os.path.join(_HERE, "lib", "finito_b200")


class FinitoError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"[fin_status {status}] {message}")
        self.status = status
        self.message = message


FIN_OK, FIN_ERR_ARG, FIN_ERR_IO, FIN_ERR_FORMAT, FIN_ERR_CUDA, FIN_ERR_DOMAIN, FIN_ERR_LIMIT = range(7)


class _Info(C.Structure):
    _fields_ = [(n, C.c_int64) for n in
                ("k", "n_nodes", "n_kmers", "n_unitigs", "text_len", "n_finimizers", "device_bytes", "device")]


class _Counts(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("found_fwd", "found_rc", "positive", "n_queries")]


class _Access(C.Structure):
    _fields_ = [(n, C.c_int64) for n in
                ("n_rank", "n_lcs_reads", "n_ustart_reads", "n_anchor", "n_anchor_branch", "n_walk_words", "n_bases",
                 "n_kmers", "n_filter_lookups", "n_lcs_skipped", "n_rank_skipped")]


class _Timing(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("pipeline_ms", "prepare_ms", "search_kernel_ms")]


class _BuilderInfo(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("n_nodes", "slice_columns", "slice_words", "text_len", "n_unitigs", "n_chunks",
                                         "chunk_begin", "chunk_end", "bases_streamed")] + \
               [("rank", C.c_int32), ("world", C.c_int32), ("stream_ms", C.c_float)]


class _SbwtTiming(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("read_ms", "keys_ms", "sort_ms", "nodes_ms", "write_ms", "total_ms")] + \
               [(n, C.c_int64) for n in ("bases", "n_nodes", "n_kmers")]


class _SpssInfo(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("n_strings", "bases", "n_kmers", "input_bases")]


class _BuildTiming(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("load_sbwt_ms", "lcs_ms", "read_unitigs_ms", "prepare_ms", "stream_merge_ms",
                                         "stream_kernel_ms", "write_ms", "total_ms")] + [("unitig_bases", C.c_int64)]


# every symbol include/finito_b200.h declares: name -> (restype, argtypes)
_u64p = C.POINTER(C.c_uint64)
ABI = {
    "fin_last_error": (C.c_char_p, []),
    "fin_abi_version": (C.c_int, []),
    "fin_device_count": (C.c_int, []),
    "fin_index_load": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]),
    "fin_index_free": (None, [C.c_void_p]),
    "fin_index_info_get": (C.c_int, [C.c_void_p, C.POINTER(_Info)]),
    "fin_count_kmers": (C.c_uint64, [C.c_void_p, C.c_uint64, C.c_int64]),
    "fin_search_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                   C.POINTER(_Counts)]),
    "fin_search_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                     C.POINTER(_Counts)]),
    "fin_search_batch32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                     C.POINTER(_Counts)]),
    "fin_search_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p,
                                    C.c_int, C.POINTER(_Counts), C.POINTER(C.c_float)]),
    "fin_search_device_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p,
                                            C.c_int, C.POINTER(_Counts)]),
    "fin_search_batch_runs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, _u64p,
                                        C.POINTER(_Counts)]),
    "fin_search_device_runs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p,
                                         C.c_uint64, _u64p, C.POINTER(_Counts), C.POINTER(C.c_float)]),
    "fin_expand_runs": (C.c_int, [C.c_void_p, C.c_uint64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int,
                                  C.c_int]),
    "fin_search_access_counters": (C.c_int, [C.c_void_p, C.POINTER(_Access)]),
    "fin_search_set_counting": (C.c_int, [C.c_void_p, C.c_int]),
    "fin_search_last_timing": (C.c_int, [C.c_void_p, C.POINTER(_Timing)]),
    "fin_launch_count": (C.c_uint64, []),
    "fin_probe_random_sectors": (C.c_int, [C.c_int, C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "fin_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "fin_host_free": (None, [C.c_void_p]),
    "fin_sbwt_build": (C.c_int, [C.c_char_p, C.c_int64, C.c_char_p, C.c_int64, C.c_int]),
    "fin_sbwt_build_last_timing": (C.c_int, [C.POINTER(_SbwtTiming)]),
    "fin_spss_build": (C.c_int, [C.c_char_p, C.c_int64, C.c_char_p, C.c_int, C.c_uint64, C.c_int, C.POINTER(_SpssInfo)]),
    "fin_build": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_int, C.c_char_p,
                            C.c_size_t]),
    "fin_build_last_timing": (C.c_int, [C.POINTER(_BuildTiming)]),
    "fin_lcs": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_uint64, _u64p]),
    "fin_stats_open": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_int, C.c_int, C.c_int,
                                 C.POINTER(C.c_void_p)]),
    "fin_stats_free": (None, [C.c_void_p]),
    "fin_stats_first_pass": (C.c_int, [C.c_void_p, _u64p]),
    "fin_stats_table": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "fin_stats_second_pass": (C.c_int, [C.c_void_p, _u64p]),
    "fin_stats_keys": (C.c_int, [C.c_void_p, C.c_void_p]),
    "fin_stats_finish": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_char_p, C.c_size_t]),
    "fin_builder_open": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                   C.POINTER(C.c_void_p)]),
    "fin_builder_free": (None, [C.c_void_p]),
    "fin_builder_info_get": (C.c_int, [C.c_void_p, C.POINTER(_BuilderInfo)]),
    "fin_builder_stream": (C.c_int, [C.c_void_p, _u64p]),
    "fin_builder_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "fin_builder_merge": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, _u64p]),
    "fin_builder_slice": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "fin_builder_finish": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_char_p]),
}

# fin_run (include/finito_b200.h): merged k-mer positions [first_kmer, first_kmer + len) of a read -> one unitig
RUN_DTYPE = np.dtype([("first_kmer", np.uint32), ("len_rc", np.uint32), ("unitig", np.uint32), ("first_offset", np.uint32)])
RUN_RC = 0x80000000
NOALLOC_MAX_READS = 65536  # include/finito_b200.h FIN_NOALLOC_MAX_READS / _BASES
NOALLOC_MAX_BASES = 16777216

_lib = None


def lib() -> C.CDLL:
    """Load libfinito_b200.so (once) and declare every entry point.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FinitoError(FIN_ERR_IO, f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                          "or `make -C finito_b200/csrc` (there is no fallback implementation)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in ABI.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def _check(status: int) -> None:
    if status != FIN_OK:
        raise FinitoError(status, lib().fin_last_error().decode("utf-8", "replace"))


def device_count() -> int:
    return int(lib().fin_device_count())


def launch_count() -> int:
    return int(lib().fin_launch_count())


def _as_batch(reads):
    """list of str/bytes, or (uint8 array, uint64 offsets) -> (uint8 array, uint64 offsets)."""
    if isinstance(reads, tuple):
        seqs, offs = reads
        return np.ascontiguousarray(seqs, dtype=np.uint8), np.ascontiguousarray(offs, dtype=np.uint64)
    bs = [r.encode() if isinstance(r, str) else bytes(r) for r in reads]
    offs = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        offs[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    seqs = np.frombuffer(b"".join(bs), dtype=np.uint8) if bs else np.zeros(0, dtype=np.uint8)
    return seqs, offs


@dataclass
class SearchCounts:
    found_fwd: int
    found_rc: int
    positive: int
    n_queries: int


class FinimizerIndex:
    """FinimizerIndex on one CUDA device (include/FinimizerIndex.hh:105-241)."""

    def __init__(self, handle: int):
        self._h = C.c_void_p(handle)

    @classmethod
    def load(cls, prefix: str, device: int = 0) -> "FinimizerIndex":
        out = C.c_void_p()
        _check(lib().fin_index_load(os.fsencode(prefix), device, C.byref(out)))
        return cls(out.value)

    def close(self) -> None:
        if self._h:
            lib().fin_index_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def info(self) -> dict:
        i = _Info()
        _check(lib().fin_index_info_get(self._h, C.byref(i)))
        return {n: int(getattr(i, n)) for n, _ in _Info._fields_}

    def search_batch(self, reads, hit_bits: int = 64, out: np.ndarray | None = None, forward_only: bool = False):
        """search-fmin for a batch: (hits[n_kmers, 2], SearchCounts); hits are (unitig, offset) or (-1, -1)."""
        seqs, offs = _as_batch(reads)
        n = len(offs) - 1
        k = self.info["k"]
        dtype = np.int64 if hit_bits == 64 else np.int32
        total = None
        if out is None:
            total = int(lib().fin_count_kmers(offs.ctypes.data, n, k))
            out = np.empty((total, 2), dtype=dtype)
        assert out.dtype == dtype and out.flags.c_contiguous  # a caller-supplied buffer is checked by the library (capacity)
        c = _Counts()
        fn = lib().fin_search_batch if hit_bits == 64 else lib().fin_search_batch32
        if forward_only:
            assert hit_bits == 64
            fn = lib().fin_search_forward
        _check(fn(self._h, seqs.ctypes.data, offs.ctypes.data, n, out.ctypes.data, out.size // 2, C.byref(c)))
        if total is None:
            total = int(c.n_queries)  # = sum max(0, len - k + 1), search_fmin.hh:69
        return out[:total], SearchCounts(c.found_fwd, c.found_rc, c.positive, c.n_queries)

    def search_batch_runs(self, reads, runs: np.ndarray | None = None, run_offsets: np.ndarray | None = None):
        """search-fmin for a batch with the answer as run-length records (fin_search_batch_runs):
        (run_offsets[n_reads + 1], runs[n_runs] of RUN_DTYPE, SearchCounts).  A too small `runs` buffer is retried once
        with the size the library reports."""
        seqs, offs = _as_batch(reads)
        n = len(offs) - 1
        if run_offsets is None:
            run_offsets = np.empty(n + 1, dtype=np.uint64)
        if runs is None:
            runs = np.empty(max(1024, 8 * n), dtype=RUN_DTYPE)
        assert runs.dtype == RUN_DTYPE and run_offsets.dtype == np.uint64 and len(run_offsets) >= n + 1
        c, nr = _Counts(), C.c_uint64()
        for attempt in range(2):
            st = lib().fin_search_batch_runs(self._h, seqs.ctypes.data, offs.ctypes.data, n, run_offsets.ctypes.data,
                                             runs.ctypes.data, len(runs), C.byref(nr), C.byref(c))
            if st == FIN_ERR_ARG and nr.value > len(runs) and attempt == 0:
                runs = np.empty(int(nr.value), dtype=RUN_DTYPE)
                continue
            _check(st)
            break
        return run_offsets[: n + 1], runs[: nr.value], SearchCounts(c.found_fwd, c.found_rc, c.positive, c.n_queries)

    def search_device_runs(self, d_seqs: int, d_offsets: int, n_reads: int, n_bases: int, d_run_offsets: int, d_runs: int,
                           runs_capacity: int):
        """Same on device-resident buffers (raw device pointers); returns (n_runs, SearchCounts, kernel milliseconds)."""
        c, ms, nr = _Counts(), C.c_float(), C.c_uint64()
        _check(lib().fin_search_device_runs(self._h, d_seqs, d_offsets, n_reads, n_bases, d_run_offsets, d_runs, runs_capacity,
                                            C.byref(nr), C.byref(c), C.byref(ms)))
        return int(nr.value), SearchCounts(c.found_fwd, c.found_rc, c.positive, c.n_queries), float(ms.value)

    def search(self, query: str):
        """FinimizerIndex::search(query): forward strand only (include/FinimizerIndex.hh:119-185)."""
        hits, _ = self.search_batch([query], forward_only=True)
        return [tuple(int(x) for x in row) for row in hits]

    def search_both_strands(self, query: str):
        """One query, both strands merged — what search-fmin prints for it (include/search_fmin.hh:43-65)."""
        hits, _ = self.search_batch([query])
        return [tuple(int(x) for x in row) for row in hits]

    def search_device(self, d_seqs: int, d_offsets: int, n_reads: int, n_bases: int, total_kmers: int, d_out: int,
                      hit_bytes: int = 4):
        """Same on device-resident buffers (raw device pointers); returns (SearchCounts, kernel milliseconds)."""
        c = _Counts()
        ms = C.c_float()
        _check(lib().fin_search_device(self._h, d_seqs, d_offsets, n_reads, n_bases, total_kmers, d_out, hit_bytes,
                                       C.byref(c), C.byref(ms)))
        return SearchCounts(c.found_fwd, c.found_rc, c.positive, c.n_queries), float(ms.value)

    def search_device_noalloc(self, d_seqs: int, d_offsets: int, n_reads: int, n_bases: int, total_kmers: int, d_out: int,
                              hit_bytes: int = 4):
        """search_device through fin_search_device_noalloc: scratch in the library's fixed static device array, no
        allocation; at most NOALLOC_MAX_READS reads / NOALLOC_MAX_BASES bases per call.  Returns SearchCounts."""
        c = _Counts()
        _check(lib().fin_search_device_noalloc(self._h, d_seqs, d_offsets, n_reads, n_bases, total_kmers, d_out, hit_bytes,
                                               C.byref(c)))
        return SearchCounts(c.found_fwd, c.found_rc, c.positive, c.n_queries)

    def set_counting(self, enabled: bool) -> None:
        _check(lib().fin_search_set_counting(self._h, int(enabled)))

    def last_timing(self) -> dict:
        t = _Timing()
        _check(lib().fin_search_last_timing(self._h, C.byref(t)))
        return {n: float(getattr(t, n)) for n, _ in _Timing._fields_}

    def access_counters(self) -> dict:
        a = _Access()
        _check(lib().fin_search_access_counters(self._h, C.byref(a)))
        return {n: int(getattr(a, n)) for n, _ in _Access._fields_}


def build_fmin(sbwt_path: str, unitig_path: str, out_prefix: str, type: str = "rarest", t: int = 1,
               lcs_path: str | None = None, device: int = 0) -> str:
    """build-fmin (include/build_fmin.hh:302-401); returns the stats string (empty for rarest)."""
    buf = C.create_string_buffer(512)
    _check(lib().fin_build(os.fsencode(sbwt_path), os.fsencode(unitig_path), os.fsencode(out_prefix), type.encode(), t,
                           os.fsencode(lcs_path) if lcs_path else None, device, buf, len(buf)))
    return buf.value.decode()


def sbwt_build(unitig_path: str, k: int, out_path: str, precalc_k: int = 8, device: int = 0) -> dict:
    """`sbwt build --variant plain-matrix -k K` on the device (fin_sbwt_build); returns the stage timings."""
    _check(lib().fin_sbwt_build(os.fsencode(unitig_path), k, os.fsencode(out_path), precalc_k, device))
    t = _SbwtTiming()
    _check(lib().fin_sbwt_build_last_timing(C.byref(t)))
    return {n: (int if n in ("bases", "n_nodes", "n_kmers") else float)(getattr(t, n)) for n, _ in _SbwtTiming._fields_}


def spss_build(in_path: str, k: int, out_path: str, flip: bool = True, seed: int = 1, device: int = 0) -> dict:
    """Disjoint SPSS of the canonical k-mers of a FASTA file (fin_spss_build): the stand-in for ggcat + unitig_flipper."""
    i = _SpssInfo()
    _check(lib().fin_spss_build(os.fsencode(in_path), k, os.fsencode(out_path), int(flip), seed, device, C.byref(i)))
    return {n: int(getattr(i, n)) for n, _ in _SpssInfo._fields_}


class ShardedBuilder:
    """One rank of build-fmin --type rarest sharded over GPUs (fin_builder_* in include/finito_b200.h; the stages of
    FinimizerIndexBuilder, include/FinimizerIndex.hh:273-389).  All buffers are raw DEVICE pointers except the
    per-rank counts and what finish() takes; finito_b200/distributed.py drives it with torch.distributed."""

    memory = "cuda"  # where export/merge/slice buffers live

    def __init__(self, sbwt_path: str, unitig_path: str, out_prefix: str, rank: int, world: int, device: int = 0,
                 lcs_path: str | None = None):
        out = C.c_void_p()
        _check(lib().fin_builder_open(os.fsencode(sbwt_path), os.fsencode(unitig_path),
                                      os.fsencode(lcs_path) if lcs_path else None, os.fsencode(out_prefix), device, rank, world,
                                      C.byref(out)))
        self._h = C.c_void_p(out.value)
        self.out_prefix = out_prefix

    def close(self) -> None:
        if self._h:
            lib().fin_builder_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def info(self) -> dict:
        i = _BuilderInfo()
        _check(lib().fin_builder_info_get(self._h, C.byref(i)))
        return {n: getattr(i, n) for n, _ in _BuilderInfo._fields_}

    def stream(self) -> int:
        n = C.c_uint64()
        _check(lib().fin_builder_stream(self._h, C.byref(n)))
        return int(n.value)

    def export(self, tuples_ptr: int, capacity: int, world: int) -> np.ndarray:
        counts = np.zeros(world, dtype=np.uint64)
        _check(lib().fin_builder_export(self._h, tuples_ptr, capacity, counts.ctypes.data))
        return counts

    def merge(self, tuples_ptr: int, n: int) -> int:
        out = C.c_uint64()
        _check(lib().fin_builder_merge(self._h, tuples_ptr, n, C.byref(out)))
        return int(out.value)

    def slice(self, bits_ptr: int, offsets_ptr: int) -> None:
        _check(lib().fin_builder_slice(self._h, bits_ptr, offsets_ptr))

    def finish(self, bits_all: np.ndarray, offsets_all: np.ndarray) -> None:
        bits_all = np.ascontiguousarray(bits_all, dtype=np.uint64)
        offsets_all = np.ascontiguousarray(offsets_all, dtype=np.uint32)
        _check(lib().fin_builder_finish(self._h, bits_all.ctypes.data, bits_all.size, offsets_all.ctypes.data, offsets_all.size,
                                        os.fsencode(self.out_prefix)))


def expand_runs(read_offsets: np.ndarray, k: int, run_offsets: np.ndarray, runs: np.ndarray, hit_bits: int = 64,
                out: np.ndarray | None = None, threads: int = 1) -> np.ndarray:
    """fin_expand_runs: runs -> the (unitig, offset) table of search_batch, on the host."""
    offs = np.ascontiguousarray(read_offsets, dtype=np.uint64)
    ro = np.ascontiguousarray(run_offsets, dtype=np.uint64)
    runs = np.ascontiguousarray(runs, dtype=RUN_DTYPE)
    n = len(offs) - 1
    dtype = np.int64 if hit_bits == 64 else np.int32
    if out is None:
        out = np.empty((int(lib().fin_count_kmers(offs.ctypes.data, n, k)), 2), dtype=dtype)
    assert out.dtype == dtype and out.flags.c_contiguous
    _check(lib().fin_expand_runs(offs.ctypes.data, n, k, ro.ctypes.data, runs.ctypes.data, out.ctypes.data, out.size // 2,
                                 hit_bits // 8, threads))
    return out


class ShardedStats:
    """One rank of build-fmin --type shortest|verify sharded over GPUs (fin_stats_* in include/finito_b200.h;
    include/build_fmin.hh:95-214).  Buffers are raw DEVICE pointers; finito_b200/distributed.py drives it."""

    memory = "cuda"

    def __init__(self, sbwt_path: str, unitig_path: str, type: str, t: int, rank: int, world: int, device: int = 0,
                 out_prefix: str | None = None, lcs_path: str | None = None):
        out = C.c_void_p()
        _check(lib().fin_stats_open(os.fsencode(sbwt_path), os.fsencode(unitig_path), os.fsencode(lcs_path) if lcs_path else None,
                                    os.fsencode(out_prefix) if out_prefix else None, type.encode(), t, device, rank, world, C.byref(out)))
        self._h = C.c_void_p(out.value)

    def close(self) -> None:
        if self._h:
            lib().fin_stats_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def first_pass(self) -> int:
        n = C.c_uint64()
        _check(lib().fin_stats_first_pass(self._h, C.byref(n)))
        return int(n.value)

    def table_out(self, ptr: int) -> None:
        _check(lib().fin_stats_table(self._h, ptr, 0))

    def table_in(self, ptr: int) -> None:
        _check(lib().fin_stats_table(self._h, ptr, 1))

    def second_pass(self) -> int:
        n = C.c_uint64()
        _check(lib().fin_stats_second_pass(self._h, C.byref(n)))
        return int(n.value)

    def keys(self, ptr: int) -> None:
        _check(lib().fin_stats_keys(self._h, ptr))

    def finish(self, keys_ptr: int, n_keys: int, write_line: bool = False) -> str:
        buf = C.create_string_buffer(512)
        _check(lib().fin_stats_finish(self._h, keys_ptr, n_keys, int(write_line), buf, len(buf)))
        return buf.value.decode()


def probe_random_sectors(device: int = 0, table_bytes: int = 8 << 30, repeats: int = 3, bytes_per_access: int = 32) -> float:
    """GB/s of independent random 32-byte sector reads (fin_probe_random_sectors): the HBM random-access roofline."""
    out = C.c_double()
    _check(lib().fin_probe_random_sectors(device, table_bytes, repeats, bytes_per_access, C.byref(out)))
    return float(out.value)


def build_timing() -> dict:
    """Stages of the last build_fmin() call on this thread (fin_build_last_timing)."""
    t = _BuildTiming()
    _check(lib().fin_build_last_timing(C.byref(t)))
    return {n: (int if n == "unitig_bases" else float)(getattr(t, n)) for n, _ in _BuildTiming._fields_}


def lcs_array(sbwt_path: str, with_variant_header: bool = True, device: int = 0) -> np.ndarray:
    """LCS bytes of an SBWT file (include/lcs_basic_parallel_algorithm.hpp:52-120)."""
    n = C.c_uint64()
    cap = max(1, os.path.getsize(sbwt_path) * 2)  # 4 bit rows of n bits each: n < 2 * file bytes
    out = np.empty(cap, dtype=np.uint8)
    _check(lib().fin_lcs(os.fsencode(sbwt_path), int(with_variant_header), device, out.ctypes.data, cap, C.byref(n)))
    return out[: n.value].copy()


def format_hits(hits: np.ndarray, kmers_per_read) -> str:
    """Text of search-fmin's output file for already computed hits (include/search_fmin.hh:54-65)."""
    lines, pos = [], 0
    for n in kmers_per_read:
        rows = hits[pos:pos + n]
        lines.append(" ".join(f"({int(u)},{int(p)})" for u, p in rows))
        pos += n
    return "\n".join(lines) + ("\n" if lines else "")
