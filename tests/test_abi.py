"""The C-ABI library loads, exports every symbol include/finito_b200.h declares, and has no CPU fallback."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import finito_b200 as fb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "finito_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fin_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert header_functions() == sorted(fb.ABI)


def test_library_exports_every_symbol():
    lib = fb.lib()
    for name in header_functions():
        assert hasattr(lib, name), name
    assert lib.fin_abi_version() == 3


def test_count_kmers_is_host_arithmetic():
    offs = np.array([0, 10, 10, 40, 45, 200], dtype=np.uint64)
    assert fb.lib().fin_count_kmers(offs.ctypes.data, 5, 31) == 0 + 0 + 0 + 0 + 125
    assert fb.lib().fin_count_kmers(offs.ctypes.data, 5, 4) == 7 + 0 + 27 + 2 + 152


def test_expand_runs_is_host_arithmetic():
    """fin_expand_runs needs no device: two reads, a forward run, a reverse-complement run and the gaps between them."""
    offs = np.array([0, 10, 17], dtype=np.uint64)  # k = 4: 7 and 4 k-mer positions
    runs = np.array([(1, 3, 5, 10), (5, 2 | fb.RUN_RC, 9, 7), (0, 4 | fb.RUN_RC, 2, 3)], dtype=fb.RUN_DTYPE)
    ro = np.array([0, 2, 3], dtype=np.uint64)
    want = [[-1, -1], [5, 10], [5, 11], [5, 12], [-1, -1], [9, 7], [9, 6], [2, 3], [2, 2], [2, 1], [2, 0]]
    for bits in (64, 32):
        for threads in (1, 4):
            assert fb.expand_runs(offs, 4, ro, runs, hit_bits=bits, threads=threads).tolist() == want
    bad = runs.copy()
    bad[1]["first_kmer"] = 2  # overlaps the first run
    with pytest.raises(fb.FinitoError) as e:
        fb.expand_runs(offs, 4, ro, bad)
    assert e.value.status == fb.FIN_ERR_ARG
    bad = runs.copy()
    bad[2]["first_offset"] = 2  # a reverse run that would go below offset 0
    with pytest.raises(fb.FinitoError):
        fb.expand_runs(offs, 4, ro, bad)


def test_no_cpu_fallback_without_device():
    if fb.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    prefix = os.path.join(ROOT, "tests", "golden", "kat_paper_example", "ref")
    with pytest.raises(fb.FinitoError) as e:
        fb.FinimizerIndex.load(prefix)
    assert e.value.status == fb.FIN_ERR_CUDA
    with pytest.raises(fb.FinitoError) as e:
        fb.build_fmin(os.path.join(os.path.dirname(prefix), "input.sbwt"), os.path.join(os.path.dirname(prefix), "unitigs.fna"),
                      "/tmp/fin_should_not_exist")
    assert e.value.status == fb.FIN_ERR_CUDA
    assert not os.path.exists("/tmp/fin_should_not_exist.FBV.sdsl")


def test_errors_are_reported_not_thrown():
    out = C.c_void_p()
    st = fb.lib().fin_index_load(b"/nonexistent/prefix", 0, C.byref(out))
    assert st == fb.FIN_ERR_IO and b"cannot open" in fb.lib().fin_last_error()
    assert fb.lib().fin_index_load(None, 0, C.byref(out)) == fb.FIN_ERR_ARG


def test_product_does_not_reference_the_oracle():
    bad = []
    for d, _, files in os.walk(os.path.join(ROOT, "finito_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".cpp", ".h")) or f == "Makefile":
                text = open(os.path.join(d, f), errors="replace").read()
                if re.search(r"liboracle|oracle\/|orc_|libfin_emu|dlopen", text):
                    bad.append(f)
    assert not bad, bad
