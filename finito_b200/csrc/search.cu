// search-fmin on the device: index upload (device layout of device_index.cuh) and the batch pipeline
// host buffers -> H2D -> pack -> search -> D2H that stands in for run_fmin_queries_streaming
// (include/search_fmin.hh:33-84) minus the text formatting, which stays on the host (cli_main.cpp).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include <cub/device/device_scan.cuh>

#include "engine.hpp"
#include "search_kernels.cuh"

namespace fin {

thread_local uint64_t g_launches = 0;

// -------------------------------------------------------------------------------------------------------
// per-stream staging of the host pipeline
// -------------------------------------------------------------------------------------------------------
struct SearchSlot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    DevBuf<uint8_t> ascii, scan_tmp;
    DevBuf<uint64_t> packed, offsets, out_offsets;
    DevBuf<uint8_t> hits;
    DevBuf<uint8_t> totals;
    SearchTotals* h_totals = nullptr;  // pinned
    std::vector<uint64_t> h_offsets;   // rebased read offsets of the sub-batch
    uint64_t* h_offsets_pinned = nullptr;
    size_t h_offsets_cap = 0;
    bool busy = false;
    ~SearchSlot() {
        if (stream) cudaStreamDestroy(stream);
        if (done) cudaEventDestroy(done);
        if (h_totals) cudaFreeHost(h_totals);
        if (h_offsets_pinned) cudaFreeHost(h_offsets_pinned);
    }
};

template <class T> static DevBuf<T> upload(const T* src, size_t n) {
    DevBuf<T> d(n);
    if (n) FIN_CUDA(cudaMemcpy(d.p, src, n * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}

DeviceIndex::DeviceIndex(const IndexHost& h, int device) : device_(device) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) throw cuda_error("no CUDA device is visible");
    if (device < 0 || device >= n_dev) throw arg_error("CUDA device ordinal out of range");
    FIN_CUDA(cudaSetDevice(device));

    const HostLayout L = build_host_layout(h, false);
    lines_ = index_lines_device(h);
    ustart_ = upload(L.ustart.data(), L.ustart.size());
    fmin_ = upload(L.fmin.data(), L.fmin.size());
    tstart_ = upload(L.tstart.data(), L.tstart.size());
    text_ = upload(L.text.data(), L.text.size());
    offsets_ = upload(L.offsets.data(), L.offsets.size());
    ends_ = upload(L.ends.data(), L.ends.size());

    view_.lines = reinterpret_cast<const uint4*>(lines_.p);
    view_.ustart = reinterpret_cast<const ulonglong2*>(ustart_.p);
    view_.fmin = reinterpret_cast<const ulonglong2*>(fmin_.p);
    view_.tstart = reinterpret_cast<const ulonglong2*>(tstart_.p);
    view_.offsets = offsets_.p;
    view_.ends = ends_.p;
    view_.text = text_.p;
    view_.n_nodes = (idx_t)L.n_nodes;
    view_.text_len = (idx_t)L.text_len;
    view_.n_unitigs = (uint32_t)L.n_unitigs;
    view_.n_fmin = (uint32_t)L.n_fmin;
    view_.k = (int32_t)L.k;
    fits32_ = L.n_unitigs < 0x7FFFFFFFull && L.max_unitig_len < 0x7FFFFFFFull;

    info_.k = L.k;
    info_.n_nodes = (int64_t)L.n_nodes;
    info_.n_kmers = L.n_kmers;
    info_.n_unitigs = (int64_t)L.n_unitigs;
    info_.text_len = (int64_t)L.text_len;
    info_.n_finimizers = (int64_t)L.n_fmin;
    info_.device = device;
    info_.device_bytes = (int64_t)(lines_.bytes() + ustart_.bytes() + fmin_.bytes() + tstart_.bytes() +
                                   offsets_.bytes() + ends_.bytes() + text_.bytes());
    const char* e = getenv("FIN_COUNTERS");
    count_accesses = e && *e && *e != '0';
}

DeviceIndex::~DeviceIndex() {
    cudaSetDevice(device_);
    slots_.clear();
}

// -------------------------------------------------------------------------------------------------------
// kernel launches
// -------------------------------------------------------------------------------------------------------
static int window_capacity(int k) { return k <= 32 ? 32 : k <= 64 ? 64 : k <= 128 ? 128 : 256; }

// persistent grid: as many blocks as are resident at once (148 SMs x blocks per SM), never more than the reads need
template <int KCAP, class HitT, bool COUNT>
static void launch_search_k(const IndexView& v, const uint64_t* packed, const uint64_t* offsets, uint64_t n_reads,
                            const uint64_t* out_offsets, HitPair<HitT>* o, SearchTotals* totals, int n_strands, cudaStream_t st) {
    const int threads = FIN_SEARCH_THREADS;
#ifdef FIN_DQ_SHARED
    const size_t dq_smem = KCAP == 32 ? (size_t)threads * KCAP * sizeof(uint64_t) : 0;  // the candidate rings
#else
    const size_t dq_smem = 0;
#endif
    static thread_local int resident[64] = {0};  // per device
    int dev = 0;
    FIN_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw arg_error("CUDA device ordinal out of range");
    if (!resident[dev]) {
        int per_sm = 0, sms = 0;
        if (dq_smem == 0)  // no shared memory: all of the unified L1/shared array as L1 (the index lines live there)
            FIN_CUDA(cudaFuncSetAttribute(search_reads_kernel<KCAP, HitT, COUNT>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                          (int)cudaSharedmemCarveoutMaxL1));
        if (dq_smem > 48 * 1024)
            FIN_CUDA(cudaFuncSetAttribute(search_reads_kernel<KCAP, HitT, COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dq_smem));
        FIN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, search_reads_kernel<KCAP, HitT, COUNT>, threads, dq_smem));
        FIN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        resident[dev] = std::max(1, per_sm) * std::max(1, sms);
    }
    const uint64_t need = (n_reads + threads - 1) / threads;
    const unsigned blocks = (unsigned)std::min<uint64_t>(need, (uint64_t)resident[dev]);
    search_reads_kernel<KCAP, HitT, COUNT><<<blocks, threads, dq_smem, st>>>(v, packed, offsets, n_reads, out_offsets, o, totals, n_strands);
    g_launches++;
    FIN_CUDA(cudaGetLastError());
}

template <class HitT, bool COUNT>
static void launch_search_t(const IndexView& v, const uint64_t* packed, const uint64_t* offsets, uint64_t n_reads,
                            const uint64_t* out_offsets, void* out, SearchTotals* totals, int n_strands, cudaStream_t st) {
    HitPair<HitT>* o = reinterpret_cast<HitPair<HitT>*>(out);
    switch (window_capacity(v.k)) {
        case 32: launch_search_k<32, HitT, COUNT>(v, packed, offsets, n_reads, out_offsets, o, totals, n_strands, st); break;
        case 64: launch_search_k<64, HitT, COUNT>(v, packed, offsets, n_reads, out_offsets, o, totals, n_strands, st); break;
        case 128: launch_search_k<128, HitT, COUNT>(v, packed, offsets, n_reads, out_offsets, o, totals, n_strands, st); break;
        default: launch_search_k<256, HitT, COUNT>(v, packed, offsets, n_reads, out_offsets, o, totals, n_strands, st); break;
    }
}

// pack + k-mer counts + exclusive scan + search on one stream; every pointer is device memory
static void launch_pipeline(const IndexView& v, const uint8_t* d_ascii, uint64_t n_bases, const uint64_t* d_offsets,
                            uint64_t n_reads, uint64_t* d_packed, uint64_t* d_out_offsets, void* d_scan_tmp,
                            size_t scan_tmp_bytes, void* d_out, int hit_bytes, SearchTotals* d_totals, bool count,
                            int n_strands, cudaStream_t st, cudaEvent_t before_search = nullptr) {
    if (n_reads == 0) return;
    const uint64_t n_words = (n_bases + 31) / 32 + 1;
    {
        const unsigned threads = 256;
        pack_reads_kernel<<<(unsigned)((n_words + threads - 1) / threads), threads, 0, st>>>(d_ascii, n_bases, d_packed, n_words,
                                                                                             d_totals);
        g_launches++;
        kmer_counts_kernel<<<(unsigned)((n_reads + threads - 1) / threads), threads, 0, st>>>(d_offsets, n_reads, v.k,
                                                                                              d_out_offsets);
        g_launches++;
        FIN_CUDA(cudaGetLastError());
    }
    FIN_CUDA(cub::DeviceScan::ExclusiveSum(d_scan_tmp, scan_tmp_bytes, d_out_offsets, d_out_offsets, (int64_t)n_reads, st));
    g_launches += 2;
    if (before_search) FIN_CUDA(cudaEventRecord(before_search, st));
    if (hit_bytes == 4) {
        if (count) launch_search_t<int32_t, true>(v, d_packed, d_offsets, n_reads, d_out_offsets, d_out, d_totals, n_strands, st);
        else launch_search_t<int32_t, false>(v, d_packed, d_offsets, n_reads, d_out_offsets, d_out, d_totals, n_strands, st);
    } else {
        if (count) launch_search_t<int64_t, true>(v, d_packed, d_offsets, n_reads, d_out_offsets, d_out, d_totals, n_strands, st);
        else launch_search_t<int64_t, false>(v, d_packed, d_offsets, n_reads, d_out_offsets, d_out, d_totals, n_strands, st);
    }
}

static size_t scan_tmp_bytes_for(uint64_t n) {
    size_t bytes = 0;
    uint64_t* dummy = nullptr;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, dummy, dummy, (int64_t)n);
    return bytes + 256;
}

static void check_flags(unsigned flags) {
    if (flags & kErrBadBase)
        throw domain_error("Error: unknown character in the query (only A, C, G, T are inside the parity domain, common.hh:108-111)");
    if (flags & kErrReadTooLong) throw limit_error("a query sequence is longer than 16777215 bases");
    if (flags & kErrWindowUnderflow)
        throw domain_error("candidate window underflow: the index is not a finimizer index of a disjoint spectrum-preserving string set");
    if (flags & kErrLookupRange) throw domain_error("finimizer lookup out of range: index files are inconsistent");
}

static void accumulate(const SearchTotals& t, fin_search_counts& c, fin_access_counters& a) {
    c.found_fwd += (int64_t)t.found_fwd;
    c.found_rc += (int64_t)t.found_rc;
    c.positive += (int64_t)t.positive;
    c.n_queries += (int64_t)t.n_queries;
    a.n_rank += (int64_t)t.n_rank;
    a.n_lcs_reads += (int64_t)t.n_lcs_reads;
    a.n_ustart_reads += (int64_t)t.n_ustart_reads;
    a.n_anchor += (int64_t)t.n_anchor;
    a.n_anchor_branch += (int64_t)t.n_anchor_branch;
    a.n_walk_words += (int64_t)t.n_walk_words;
    a.n_bases += (int64_t)t.n_bases;
    a.n_kmers += (int64_t)t.n_kmers;
}

// -------------------------------------------------------------------------------------------------------
// device-resident batch
// -------------------------------------------------------------------------------------------------------
void DeviceIndex::search_device(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases,
                                uint64_t total_kmers, void* d_out, int hit_bytes, fin_search_counts* counts, float* kernel_ms) {
    if (hit_bytes != 4 && hit_bytes != 8) throw arg_error("hit_bytes must be 4 or 8");
    if (hit_bytes == 4 && !fits32_) throw limit_error("this index needs 64-bit hits");
    (void)total_kmers;
    FIN_CUDA(cudaSetDevice(device_));
    dev_packed_.ensure((n_bases + 31) / 32 + 2);
    dev_out_offsets_.ensure(n_reads + 1);
    const size_t tmp = scan_tmp_bytes_for(n_reads);
    dev_scan_tmp_.ensure(tmp);
    dev_totals_.ensure(sizeof(SearchTotals));
    FIN_CUDA(cudaMemsetAsync(dev_totals_.p, 0, sizeof(SearchTotals), 0));
    cudaEvent_t e0, e1, em;
    FIN_CUDA(cudaEventCreate(&e0));
    FIN_CUDA(cudaEventCreate(&e1));
    FIN_CUDA(cudaEventCreate(&em));
    FIN_CUDA(cudaEventRecord(e0, 0));
    if (n_reads == 0) FIN_CUDA(cudaEventRecord(em, 0));
    launch_pipeline(view_, (const uint8_t*)d_seqs, n_bases, d_read_offsets, n_reads, dev_packed_.p, dev_out_offsets_.p,
                    dev_scan_tmp_.p, tmp, d_out, hit_bytes, (SearchTotals*)dev_totals_.p, count_accesses, 2, 0, em);
    FIN_CUDA(cudaEventRecord(e1, 0));
    SearchTotals t;
    FIN_CUDA(cudaMemcpy(&t, dev_totals_.p, sizeof t, cudaMemcpyDeviceToHost));
    FIN_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    FIN_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    last_timing.pipeline_ms = ms;
    FIN_CUDA(cudaEventElapsedTime(&last_timing.prepare_ms, e0, em));
    FIN_CUDA(cudaEventElapsedTime(&last_timing.search_kernel_ms, em, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaEventDestroy(em);
    if (kernel_ms) *kernel_ms = ms;
    check_flags(t.error_flags);
    fin_search_counts c{};
    last_access = fin_access_counters{};
    accumulate(t, c, last_access);
    if (counts) *counts = c;
}

// -------------------------------------------------------------------------------------------------------
// host batch: sub-batches round-robin over three streams, so that the H2D copy of sub-batch i+1 and the
// D2H copy of sub-batch i-1 overlap the kernels of sub-batch i.
// -------------------------------------------------------------------------------------------------------
static constexpr uint64_t kSubBases = 48ull << 20;  // bases per sub-batch
static constexpr uint64_t kSubReads = 1ull << 19;   // reads per sub-batch
static constexpr int kSlots = 3;

template <class HitT>
void DeviceIndex::search_host(const char* seqs, const uint64_t* read_offsets, uint64_t n_reads, HitT* out_pairs,
                              uint64_t out_capacity, fin_search_counts* counts, int n_strands) {
    if (sizeof(HitT) == 4 && !fits32_) throw limit_error("this index needs 64-bit hits (use fin_search_batch)");
    FIN_CUDA(cudaSetDevice(device_));
    const int k = view_.k;
    // the offsets are validated and the k-mer positions counted sub-batch by sub-batch while the GPU already works on
    // the previous ones: no pass over the whole batch stands between the call and the first copy

    if (slots_.empty()) {
        for (int i = 0; i < kSlots; i++) {
            std::unique_ptr<SearchSlot> s(new SearchSlot());
            FIN_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
            FIN_CUDA(cudaEventCreateWithFlags(&s->done, cudaEventDisableTiming));
            FIN_CUDA(cudaHostAlloc((void**)&s->h_totals, sizeof(SearchTotals), cudaHostAllocDefault));
            s->totals.alloc(sizeof(SearchTotals));
            slots_.push_back(std::move(s));
        }
    }
    fin_search_counts c{};
    last_access = fin_access_counters{};
    unsigned flags = 0;

    auto retire = [&](SearchSlot& s) {
        if (!s.busy) return;
        FIN_CUDA(cudaEventSynchronize(s.done));
        flags |= s.h_totals->error_flags;
        accumulate(*s.h_totals, c, last_access);
        s.busy = false;
    };

    uint64_t r0 = 0, out_pos = 0;
    int turn = 0;
    while (r0 < n_reads) {
        // sub-batch [r0, r1): at most kSubReads reads and kSubBases bases (always at least one read)
        uint64_t r1 = std::min(n_reads, r0 + kSubReads);
        const uint64_t base0 = read_offsets[r0];
        if (read_offsets[r1] - base0 > kSubBases) {
            const uint64_t* lo = read_offsets + r0 + 1;
            const uint64_t* hi = read_offsets + r1 + 1;
            const uint64_t* p = std::upper_bound(lo, hi, base0 + kSubBases);  // first offset beyond the cap
            r1 = std::max<uint64_t>(r0 + 1, (uint64_t)(p - read_offsets) - 1);
        }
        const uint64_t nr = r1 - r0, nb = read_offsets[r1] - base0;
        uint64_t nkm = 0;
        bool ordered = true;
        for (uint64_t i = r0; i < r1; i++) {
            ordered &= read_offsets[i + 1] >= read_offsets[i];
            const uint64_t len = read_offsets[i + 1] - read_offsets[i];
            nkm += len >= (uint64_t)k ? len - (uint64_t)k + 1 : 0;
        }
        if (!ordered || out_pos + nkm > out_capacity) {
            for (auto& sl : slots_) {  // let the sub-batches in flight finish before reporting
                if (sl->busy) cudaEventSynchronize(sl->done);
                sl->busy = false;
            }
            if (!ordered) throw arg_error("read_offsets must be non-decreasing");
            throw arg_error("out_capacity is smaller than the number of k-mer positions of the batch");
        }

        SearchSlot& s = *slots_[turn % kSlots];
        turn++;
        retire(s);
        s.ascii.ensure(nb + 64);
        s.packed.ensure((nb + 31) / 32 + 2);
        s.offsets.ensure(nr + 1);
        s.out_offsets.ensure(nr + 1);
        const size_t tmp = scan_tmp_bytes_for(nr);
        s.scan_tmp.ensure(tmp);
        s.hits.ensure(std::max<uint64_t>(nkm, 1) * 2 * sizeof(HitT));
        if (s.h_offsets_cap < nr + 1) {
            if (s.h_offsets_pinned) cudaFreeHost(s.h_offsets_pinned);
            s.h_offsets_pinned = nullptr;
            s.h_offsets_cap = std::max<uint64_t>(nr + 1, kSubReads + 1);
            FIN_CUDA(cudaHostAlloc((void**)&s.h_offsets_pinned, s.h_offsets_cap * 8, cudaHostAllocDefault));
        }
        for (uint64_t i = 0; i <= nr; i++) s.h_offsets_pinned[i] = read_offsets[r0 + i] - base0;

        FIN_CUDA(cudaMemsetAsync(s.totals.p, 0, sizeof(SearchTotals), s.stream));
        FIN_CUDA(cudaMemcpyAsync(s.ascii.p, seqs + base0, nb, cudaMemcpyHostToDevice, s.stream));
        FIN_CUDA(cudaMemcpyAsync(s.offsets.p, s.h_offsets_pinned, (nr + 1) * 8, cudaMemcpyHostToDevice, s.stream));
        launch_pipeline(view_, s.ascii.p, nb, s.offsets.p, nr, s.packed.p, s.out_offsets.p, s.scan_tmp.p, tmp, s.hits.p,
                        (int)sizeof(HitT), (SearchTotals*)s.totals.p, count_accesses, n_strands, s.stream);
        if (nkm)
            FIN_CUDA(cudaMemcpyAsync(out_pairs + 2 * out_pos, s.hits.p, nkm * 2 * sizeof(HitT), cudaMemcpyDeviceToHost, s.stream));
        FIN_CUDA(cudaMemcpyAsync(s.h_totals, s.totals.p, sizeof(SearchTotals), cudaMemcpyDeviceToHost, s.stream));
        FIN_CUDA(cudaEventRecord(s.done, s.stream));
        s.busy = true;
        out_pos += nkm;
        r0 = r1;
    }
    for (auto& s : slots_) retire(*s);
    check_flags(flags);
    if (counts) *counts = c;
}

template void DeviceIndex::search_host<int32_t>(const char*, const uint64_t*, uint64_t, int32_t*, uint64_t, fin_search_counts*, int);
template void DeviceIndex::search_host<int64_t>(const char*, const uint64_t*, uint64_t, int64_t*, uint64_t, fin_search_counts*, int);

}  // namespace fin
