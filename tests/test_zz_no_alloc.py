"""fin_search_device_noalloc (include/finito_b200.h): the dense device-resident search with its scratch in the library's
fixed static __device__ array - no allocation inside the call.  Checked through the C ABI against fin_search_device
(the existing entry point) bit for bit (integer hits: tolerance 0), on calls in a row that share the array, and
against the oracle on a golden case.  Collected last on purpose (the file name sorts behind every other test file)."""
import os

import numpy as np
import pytest

import helpers as H
from fin_datasets import make_dataset

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def workload(gpu_lib, tmp_path_factory):
    ds = make_dataset("gpu", genome_len=1_000_000, k=31, n_reads=20_000, read_len=150, seed=5)
    prefix = str(tmp_path_factory.mktemp("idx_noalloc") / "fmin")
    gpu_lib.build_fmin(ds["sbwt"], ds["unitigs"], prefix)
    return ds, prefix


def _device_batch(torch, seqs, offs):
    d_seqs = torch.from_numpy(np.ascontiguousarray(seqs)).cuda()
    d_offs = torch.from_numpy(np.ascontiguousarray(offs).astype(np.int64)).cuda()
    return d_seqs, d_offs


def _slice_reads(seqs, offs, r0, r1):
    o = offs[r0:r1 + 1].astype(np.uint64)
    return seqs[int(o[0]):int(o[-1])].copy(), o - o[0]


def test_two_calls_in_a_row_equal_the_allocating_entry_point(workload, gpu_lib):
    import torch

    fb = gpu_lib
    ds, prefix = workload
    seqs, offs = ds["reads"]
    k = 31
    # three different batches through the ONE shared array, large -> small -> ragged: stale contents of a call would show
    ragged_seqs = np.concatenate([seqs[:40], seqs[150:150 + k - 1], seqs[300:300 + k], seqs[450:1050]])
    ragged_offs = np.array([0, 40, 40, 40 + k - 1, 40 + 2 * k - 1, 40 + 2 * k - 1 + 600], dtype=np.uint64)
    batches = [_slice_reads(seqs, offs, 0, 20_000), _slice_reads(seqs, offs, 7_000, 7_900), (ragged_seqs, ragged_offs),
               _slice_reads(seqs, offs, 0, 20_000)]
    with fb.FinimizerIndex.load(prefix) as ix:
        for hit_bytes, dt in ((4, torch.int32), (8, torch.int64)):
            for bs, bo in batches:
                n, nb = len(bo) - 1, len(bs)
                nk = int(np.maximum(np.diff(bo.astype(np.int64)) - k + 1, 0).sum())
                d_seqs, d_offs = _device_batch(torch, bs, bo)
                d_want = torch.full((max(nk, 1), 2), -7, dtype=dt, device="cuda")
                d_got = torch.full((max(nk, 1), 2), -9, dtype=dt, device="cuda")
                torch.cuda.synchronize()
                wc, _ = ix.search_device(d_seqs.data_ptr(), d_offs.data_ptr(), n, nb, nk, d_want.data_ptr(), hit_bytes)
                c = ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), n, nb, nk, d_got.data_ptr(), hit_bytes)
                assert c == wc
                assert torch.equal(d_got[:nk], d_want[:nk])  # integer hits: bit-exact
        # the calls allocate nothing: the driver's free-memory figure does not move across warm calls
        bs, bo = batches[0]
        n, nb = len(bo) - 1, len(bs)
        nk = n * (150 - k + 1)
        d_seqs, d_offs = _device_batch(torch, bs, bo)
        d_got = torch.empty((nk, 2), dtype=torch.int32, device="cuda")
        ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), n, nb, nk, d_got.data_ptr(), 4)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        for _ in range(3):
            ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), n, nb, nk, d_got.data_ptr(), 4)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0


def test_limits_and_bad_batches_are_refused(workload, gpu_lib):
    import torch

    fb = gpu_lib
    ds, prefix = workload
    seqs, offs = ds["reads"]
    bs, bo = _slice_reads(seqs, offs, 0, 100)
    n, nb = 100, len(bs)
    nk = n * (150 - 31 + 1)
    d_seqs, d_offs = _device_batch(torch, bs, bo)
    d_out = torch.empty((nk, 2), dtype=torch.int32, device="cuda")
    with fb.FinimizerIndex.load(prefix) as ix:
        with pytest.raises(fb.FinitoError) as e:  # beyond the static scratch: refused before anything is launched
            ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), fb.NOALLOC_MAX_READS + 1, nb, nk, d_out.data_ptr(), 4)
        assert e.value.status == fb.FIN_ERR_LIMIT and "at most" in str(e.value)
        with pytest.raises(fb.FinitoError) as e:
            ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), n, fb.NOALLOC_MAX_BASES + 1, nk, d_out.data_ptr(), 4)
        assert e.value.status == fb.FIN_ERR_LIMIT
        with pytest.raises(fb.FinitoError):  # d_out too small, as fin_search_device
            ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), n, nb, nk - 1, d_out.data_ptr(), 4)
        d_bad = d_offs.clone()
        d_bad[3] = d_bad[5]
        with pytest.raises(fb.FinitoError):  # offsets that are not non-decreasing
            ix.search_device_noalloc(d_seqs.data_ptr(), d_bad.data_ptr(), n, nb, nk, d_out.data_ptr(), 4)
        # an empty batch, and the handle and the array still work afterwards
        c0 = ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), 0, 0, 0, d_out.data_ptr(), 4)
        assert c0.n_queries == 0
        want, wc = ix.search_batch((bs, bo), hit_bits=32)
        c = ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), n, nb, nk, d_out.data_ptr(), 4)
        assert c == wc and np.array_equal(d_out.cpu().numpy(), want)


def test_golden_case_against_the_oracle(gpu_lib):
    import torch

    fb = gpu_lib
    case = os.path.join(H.GOLDEN, "example_fna_k4")
    prefix = os.path.join(case, "ref")
    if not os.path.exists(prefix + ".FBV.sdsl"):
        pytest.skip("golden case layout changed")
    orc = H.Oracle(prefix)
    reads = [b"ACGTACGTTTGACCA", b"TTTT", b"ACG", b"GATTACAGATTACA"]
    seqs = np.frombuffer(b"".join(reads), dtype=np.uint8).copy()
    offs = np.cumsum([0] + [len(r) for r in reads]).astype(np.uint64)
    want, wc = orc.search_reads(seqs, offs)
    d_seqs, d_offs = _device_batch(torch, seqs, offs)
    d_out = torch.empty((max(1, want.shape[0]), 2), dtype=torch.int64, device="cuda")
    with fb.FinimizerIndex.load(prefix) as ix:
        for _ in range(2):
            c = ix.search_device_noalloc(d_seqs.data_ptr(), d_offs.data_ptr(), len(reads), len(seqs), want.shape[0], d_out.data_ptr(), 8)
            assert np.array_equal(d_out.cpu().numpy()[:want.shape[0]], want)
            assert [c.found_fwd, c.found_rc, c.positive, c.n_queries] == [int(x) for x in wc]
