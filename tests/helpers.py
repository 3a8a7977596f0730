"""Test-side bindings: the oracle (oracle/liboracle.so), the reference binaries (oracle/_ref/, when built)
and the CPU emulation of the kernel bodies (tests/emu/libfin_emu.so).  None of this is product code."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
EMU_DIR = os.path.join(ROOT, "tests", "emu")
GOLDEN = os.path.join(ROOT, "tests", "golden")
TOOLS = os.path.join(ROOT, "tools")

INDEX_SUFFIXES = [".O.sdsl", ".FBV.sdsl", ".packed_unitigs.sdsl", ".unitig_endpoints.sdsl", ".Ustart.sdsl", ".LCS.sdsl",
                  ".sbwt"]


def _newer(target, *sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def oracle_lib() -> C.CDLL:
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    if not _newer(so, os.path.join(ORACLE_DIR, "oracle.cpp")):
        subprocess.run(["make", "-C", ORACLE_DIR, "oracle"], check=True, capture_output=True)
    lib = C.CDLL(so)
    lib.orc_last_error.restype = C.c_char_p
    lib.orc_index_load.restype = C.c_void_p
    lib.orc_index_load.argtypes = [C.c_char_p]
    lib.orc_index_free.argtypes = [C.c_void_p]
    lib.orc_index_info.argtypes = [C.c_void_p, C.c_void_p]
    lib.orc_search_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    lib.orc_search_strand.restype = C.c_int64
    lib.orc_search_strand.argtypes = [C.c_void_p, C.c_char_p, C.c_int64, C.c_void_p]
    lib.orc_search_file.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p]
    lib.orc_build_fmin.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_char_p,
                                   C.c_int64]
    lib.orc_lcs.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_int64]
    lib.orc_counters_get.argtypes = [C.c_void_p]
    lib.orc_index_member.restype = C.c_int64
    lib.orc_index_member.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64]
    lib.orc_member_width.argtypes = [C.c_void_p, C.c_int]
    return lib


COUNTER_NAMES = ["n_rank", "n_lcs_reads", "n_ustart_reads", "n_anchor", "n_anchor_branch", "n_walk_words", "n_bases",
                 "n_kmers", "n_deque_underflow", "n_out_of_domain", "n_drop_calls", "reserved"]


class Oracle:
    """CPU restatement of FinimizerIndex (oracle/oracle.cpp)."""

    def __init__(self, prefix: str):
        self.lib = oracle_lib()
        self.h = self.lib.orc_index_load(os.fsencode(prefix))
        if not self.h:
            raise RuntimeError("oracle: " + self.lib.orc_last_error().decode())
        info = np.zeros(6, dtype=np.int64)
        self.lib.orc_index_info(self.h, info.ctypes.data)
        self.k, self.n_nodes, self.n_kmers, self.n_unitigs, self.text_len, self.n_fmin = (int(x) for x in info)

    def close(self):
        if self.h:
            self.lib.orc_index_free(self.h)
            self.h = None

    def search_reads(self, seqs: np.ndarray, offs: np.ndarray):
        n = len(offs) - 1
        lens = np.diff(offs.astype(np.int64))
        total = int(np.maximum(lens - self.k + 1, 0).sum())
        out = np.empty((total, 2), dtype=np.int64)
        counts = np.zeros(4, dtype=np.int64)
        o64 = np.ascontiguousarray(offs, dtype=np.int64)
        s = np.ascontiguousarray(seqs, dtype=np.uint8)
        self.lib.orc_counters_reset()
        rc = self.lib.orc_search_reads(self.h, s.ctypes.data, o64.ctypes.data, n, out.ctypes.data, counts.ctypes.data)
        if rc:
            raise RuntimeError("oracle: " + self.lib.orc_last_error().decode())
        return out, counts

    def counters(self) -> dict:
        c = np.zeros(12, dtype=np.int64)
        self.lib.orc_counters_get(c.ctypes.data)
        return dict(zip(COUNTER_NAMES, (int(x) for x in c)))

    def search_file(self, index_prefix: str, query: str, out_path: str):
        counts = np.zeros(4, dtype=np.int64)
        rc = self.lib.orc_search_file(self.h, os.fsencode(index_prefix), os.fsencode(query), os.fsencode(out_path),
                                      counts.ctypes.data)
        if rc:
            raise RuntimeError("oracle: " + self.lib.orc_last_error().decode())
        return counts


def oracle_build(sbwt, unitigs, prefix, type="rarest", t=1, lcs=None) -> str:
    lib = oracle_lib()
    buf = C.create_string_buffer(512)
    rc = lib.orc_build_fmin(os.fsencode(sbwt), os.fsencode(unitigs), os.fsencode(prefix), type.encode(), t,
                            os.fsencode(lcs) if lcs else None, buf, 512)
    if rc:
        raise RuntimeError("oracle: " + lib.orc_last_error().decode())
    return buf.value.decode()


def oracle_lcs(sbwt, n_nodes, with_variant=True) -> np.ndarray:
    lib = oracle_lib()
    out = np.zeros(n_nodes, dtype=np.uint8)
    if lib.orc_lcs(os.fsencode(sbwt), int(with_variant), out.ctypes.data, n_nodes):
        raise RuntimeError("oracle: " + lib.orc_last_error().decode())
    return out


def emu_lib() -> C.CDLL:
    # FIN_EMU_ASAN=1 (tests/test_emu_sanitized.py): the kernel bodies under AddressSanitizer + UBSan on the host
    asan = os.environ.get("FIN_EMU_ASAN") == "1"
    so = os.path.join(EMU_DIR, "libfin_emu_asan.so" if asan else "libfin_emu.so")
    csrc = os.path.join(ROOT, "finito_b200", "csrc")
    srcs = [os.path.join(EMU_DIR, "emu.cpp")] + [os.path.join(csrc, f) for f in os.listdir(csrc)
                                                 if f.endswith((".cuh", ".hpp"))]
    if not _newer(so, *srcs):
        flags = ["-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-fno-omit-frame-pointer"] if asan else ["-O2"]
        subprocess.run(["/usr/bin/g++", "-std=c++17", *flags, "-fPIC", "-shared", "-Wno-unknown-pragmas", "-o", so,
                        os.path.join(EMU_DIR, "emu.cpp"), "-lz", "-pthread"], check=True, capture_output=True)
    lib = C.CDLL(so)
    lib.emu_last_error.restype = C.c_char_p
    lib.emu_search.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p,
                               C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    lib.emu_lcs.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_uint64]
    lib.emu_read_sequences.restype = C.c_int64
    lib.emu_read_sequences.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                       C.c_void_p]
    lib.emu_build.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint64]
    lib.emu_builder_open.restype = C.c_void_p
    lib.emu_builder_open.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int]
    lib.emu_builder_free.argtypes = [C.c_void_p]
    lib.emu_builder_info.argtypes = [C.c_void_p, C.c_void_p]
    lib.emu_builder_stream.argtypes = [C.c_void_p, C.c_void_p]
    lib.emu_builder_export.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.emu_builder_merge.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.emu_builder_slice.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.emu_builder_finish.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
    lib.emu_stats.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_uint64, C.c_char_p, C.c_uint64]
    lib.emu_stats_open.restype = C.c_void_p
    lib.emu_stats_open.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.c_int, C.c_int]
    lib.emu_stats_free.argtypes = [C.c_void_p]
    lib.emu_stats_first_pass.argtypes = [C.c_void_p, C.c_void_p]
    lib.emu_stats_table.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    lib.emu_stats_second_pass.argtypes = [C.c_void_p, C.c_void_p]
    lib.emu_stats_keys.argtypes = [C.c_void_p, C.c_void_p]
    lib.emu_stats_finish.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_char_p, C.c_uint64]
    return lib


def emu_search(prefix, seqs, offs, n_strands=2, runs=False, prefilter_s=0, streamed=None):
    """prefilter_s > 0: with the strand prefilter on s-mers (DeviceIndex builds it at load); `streamed`: a one-element
    uint64 array that receives the number of strands that were streamed"""
    lib = emu_lib()
    offs = np.ascontiguousarray(offs, dtype=np.uint64)
    seqs = np.ascontiguousarray(seqs, dtype=np.uint8)
    n = len(offs) - 1
    k = sbwt_header(prefix + ".sbwt", with_variant=False)["k"]
    lens = np.diff(offs.astype(np.int64))
    total = int(np.maximum(lens - k + 1, 0).sum())
    out = np.empty((total, 2), dtype=np.int64)
    counts = np.zeros(4, dtype=np.int64)
    access = np.zeros(8, dtype=np.int64)
    flags = np.zeros(1, dtype=np.uint32)
    if lib.emu_search(os.fsencode(prefix), seqs.ctypes.data, offs.ctypes.data, n, out.ctypes.data, counts.ctypes.data,
                      access.ctypes.data, flags.ctypes.data, n_strands, int(runs), None, prefilter_s,
                      streamed.ctypes.data if streamed is not None else None):
        raise RuntimeError("emu: " + lib.emu_last_error().decode())
    return out, counts, dict(zip(COUNTER_NAMES[:8], (int(x) for x in access))), int(flags[0])


def emu_read_sequences(path, chunk_bytes=16 << 20, max_reads=1 << 62, max_bases=1 << 62):
    """The product's incremental sequence reader (finito_b200/csrc/seqio.hpp) -> (list of byte strings, number of batches)."""
    lib = emu_lib()
    cap = max(1024, 40 * os.path.getsize(path))  # gzip of a repetitive file can expand a lot
    seqs = np.zeros(cap, dtype=np.uint8)
    offs = np.zeros(cap // 2 + 2, dtype=np.uint64)
    nb = np.zeros(1, dtype=np.uint64)
    n = lib.emu_read_sequences(os.fsencode(path), chunk_bytes, max_reads, max_bases, seqs.ctypes.data, cap, offs.ctypes.data, len(offs),
                               nb.ctypes.data)
    if n < 0:
        raise RuntimeError("emu: " + lib.emu_last_error().decode())
    raw = seqs.tobytes()
    return [raw[int(offs[i]):int(offs[i + 1])] for i in range(n)], int(nb[0])


def emu_build(sbwt, unitigs, prefix, per_chunk=0):
    lib = emu_lib()
    if lib.emu_build(os.fsencode(sbwt), os.fsencode(unitigs), os.fsencode(prefix), per_chunk):
        raise RuntimeError("emu: " + lib.emu_last_error().decode())


class EmuShardedBuilder:
    """Same methods as finito_b200.ShardedBuilder over HOST memory, the stream run by the kernel bodies on the CPU
    (tests/emu).  Test infrastructure: lets gloo tests drive finito_b200/distributed.py without a GPU."""

    memory = "cpu"

    def __init__(self, sbwt, unitigs, out_prefix, rank, world):
        self.lib = emu_lib()
        self.h = self.lib.emu_builder_open(os.fsencode(sbwt), os.fsencode(unitigs), os.fsencode(out_prefix), rank, world)
        if not self.h:
            raise RuntimeError("emu: " + self.lib.emu_last_error().decode())
        self.out_prefix = out_prefix

    def _ok(self, rc):
        if rc:
            raise RuntimeError("emu: " + self.lib.emu_last_error().decode())

    @property
    def info(self):
        v = np.zeros(9, dtype=np.int64)
        self.lib.emu_builder_info(self.h, v.ctypes.data)
        names = ("n_nodes", "slice_columns", "slice_words", "text_len", "n_unitigs", "n_chunks", "chunk_begin", "chunk_end",
                 "bases_streamed")
        d = dict(zip(names, (int(x) for x in v)))
        d["stream_ms"] = 0.0
        return d

    def stream(self):
        n = np.zeros(1, dtype=np.uint64)
        self._ok(self.lib.emu_builder_stream(self.h, n.ctypes.data))
        return int(n[0])

    def export(self, ptr, capacity, world):
        counts = np.zeros(world, dtype=np.uint64)
        self._ok(self.lib.emu_builder_export(self.h, ptr, capacity, counts.ctypes.data))
        return counts

    def merge(self, ptr, n):
        out = np.zeros(1, dtype=np.uint64)
        self._ok(self.lib.emu_builder_merge(self.h, ptr, n, out.ctypes.data))
        return int(out[0])

    def slice(self, bits_ptr, offs_ptr):
        self.lib.emu_builder_slice(self.h, bits_ptr, offs_ptr)

    def finish(self, bits_all, offs_all):
        bits_all = np.ascontiguousarray(bits_all, dtype=np.uint64)
        offs_all = np.ascontiguousarray(offs_all, dtype=np.uint32)
        self._ok(self.lib.emu_builder_finish(self.h, bits_all.ctypes.data, bits_all.size, offs_all.ctypes.data, offs_all.size))

    def close(self):
        if self.h:
            self.lib.emu_builder_free(self.h)
            self.h = None


class EmuShardedStats:
    """Same methods as finito_b200.ShardedStats over HOST memory, the streams run by the kernel bodies on the CPU
    (tests/emu).  Test infrastructure: lets gloo tests drive finito_b200.distributed.stats_sharded without a GPU."""

    memory = "cpu"

    def __init__(self, sbwt, unitigs, type, t, rank, world):
        self.lib = emu_lib()
        self.h = self.lib.emu_stats_open(os.fsencode(sbwt), os.fsencode(unitigs), type.encode(), t, rank, world)
        if not self.h:
            raise RuntimeError("emu: " + self.lib.emu_last_error().decode())

    def _ok(self, rc):
        if rc:
            raise RuntimeError("emu: " + self.lib.emu_last_error().decode())

    def first_pass(self):
        n = np.zeros(1, dtype=np.uint64)
        self._ok(self.lib.emu_stats_first_pass(self.h, n.ctypes.data))
        return int(n[0])

    def table_out(self, ptr):
        self._ok(self.lib.emu_stats_table(self.h, ptr, 0))

    def table_in(self, ptr):
        self._ok(self.lib.emu_stats_table(self.h, ptr, 1))

    def second_pass(self):
        n = np.zeros(1, dtype=np.uint64)
        self._ok(self.lib.emu_stats_second_pass(self.h, n.ctypes.data))
        return int(n[0])

    def keys(self, ptr):
        self._ok(self.lib.emu_stats_keys(self.h, ptr))

    def finish(self, keys_ptr, n_keys, write_line=False):
        buf = C.create_string_buffer(512)
        self._ok(self.lib.emu_stats_finish(self.h, keys_ptr, n_keys, buf, 512))
        return buf.value.decode()

    def close(self):
        if self.h:
            self.lib.emu_stats_free(self.h)
            self.h = None


def emu_stats(sbwt, unitigs, type, t, per_chunk=0) -> str:
    lib = emu_lib()
    buf = C.create_string_buffer(512)
    if lib.emu_stats(os.fsencode(sbwt), os.fsencode(unitigs), type.encode(), t, per_chunk, buf, 512):
        raise RuntimeError("emu: " + lib.emu_last_error().decode())
    return buf.value.decode()


def sbwt_header(path, with_variant=True) -> dict:
    """k, n_nodes, n_kmers, precalc_k from the tail of an .sbwt file (SBWT.hh:395-424)."""
    with open(path, "rb") as f:
        f.seek(-32, os.SEEK_END)
        t = np.frombuffer(f.read(32), dtype=np.int64)
    return {"precalc_k": int(t[0]), "n_nodes": int(t[1]), "n_kmers": int(t[2]), "k": int(t[3])}


def have_ref() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "finito_ref"))


def ref_run(*args, cwd=None):
    return subprocess.run([os.path.join(REF_DIR, "finito_ref"), *args], cwd=cwd, capture_output=True, text=True)


def batch_of(reads):
    bs = [r.encode() if isinstance(r, str) else bytes(r) for r in reads]
    offs = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        offs[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    return np.frombuffer(b"".join(bs), dtype=np.uint8), offs


def read_fasta(path):
    seqs, cur = [], []
    with open(path) as f:
        for line in f:
            if line.startswith(">"):
                if cur:
                    seqs.append("".join(cur))
                cur = []
            else:
                cur.append(line.strip().upper())
    if cur:
        seqs.append("".join(cur))
    return seqs


def parse_output(path):
    """search-fmin text output -> int64 array [n, 2]"""
    rows = []
    with open(path) as f:
        for line in f:
            for tok in line.split():
                u, p = tok.strip("()").split(",")
                rows.append((int(u), int(p)))
    return np.array(rows, dtype=np.int64).reshape(-1, 2)


def files_equal(a, b) -> bool:
    with open(a, "rb") as fa, open(b, "rb") as fb:
        return fa.read() == fb.read()
