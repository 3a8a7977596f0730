// search-fmin on the device: index upload (device layout of device_index.cuh) and the batch pipeline
// host buffers -> H2D -> pack -> search -> D2H that stands in for run_fmin_queries_streaming
// (include/search_fmin.hh:33-84) minus the text formatting, which stays on the host (cli_main.cpp).
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <thread>

#include "engine.hpp"
#include "runs_host.hpp"
#include "search_kernels.cuh"

namespace fin {

thread_local uint64_t g_launches = 0;

// -------------------------------------------------------------------------------------------------------
// per-stream staging of the host pipeline
// -------------------------------------------------------------------------------------------------------
struct SearchSlot {
    cudaStream_t stream = nullptr;
    cudaStream_t prep_stream = nullptr;             // copy-in + packing + prefilter of the runs pipeline (high priority)
    cudaEvent_t prepped = nullptr;
    cudaEvent_t counted = nullptr, done = nullptr;  // kernels + totals copy finished / everything finished
    DevBuf<uint8_t> ascii, scan_tmp, sort_tmp, strand_flags;
    DevBuf<uint64_t> packed, offsets, out_offsets;
    DevBuf<uint32_t> fmask;
    DevBuf<uint8_t> hits;
    DevBuf<uint64_t> keys_a, keys_b, run_offsets;  // runs mode: appended / sorted keys, CSR offsets
    DevBuf<RunRec> recs_a, recs_b;
    DevBuf<uint8_t> totals;
    SearchTotals* h_totals = nullptr;  // pinned
    uint64_t* h_offsets_pinned = nullptr;
    size_t h_offsets_cap = 0;
    bool busy = false;
    // the sub-batch in flight (runs mode needs it again when the record count is known)
    uint64_t r0 = 0, nr = 0, nb = 0, nkm = 0, base0 = 0;
    ~SearchSlot() {
        if (stream) cudaStreamDestroy(stream);
        if (prep_stream) cudaStreamDestroy(prep_stream);
        if (prepped) cudaEventDestroy(prepped);
        if (counted) cudaEventDestroy(counted);
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
    NvtxRange range("fin:index_load:device_layout");
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
    view_.smer = nullptr;
    view_.smer_s = 0;
    {   // strand prefilter table: s ~ log4(text length) + 1 (a few percent of the 4^s codes occur), at most 16 and at most k
        const char* off = getenv("FIN_NO_PREFILTER");
        if (!(off && *off && *off != '0')) {
            int s = 1;
            while (s < 16 && (1ull << (2 * s)) < L.text_len) s++;
            s = std::min<int>(std::min(16, s + 1), (int)L.k);
            if (const char* e = getenv("FIN_PREFILTER_S")) if (atoi(e) >= 4 && atoi(e) <= 16) s = std::min<int>(atoi(e), (int)L.k);
            const uint64_t words = std::max<uint64_t>(1, (1ull << (2 * s)) / 32);
            smer_.alloc(words);
            FIN_CUDA(cudaMemset(smer_.p, 0, words * 4));
            if (L.text_len >= (uint64_t)s) {
                const unsigned threads = 256;
                smer_table_kernel<<<(unsigned)((L.text_len + threads - 1) / threads), threads>>>(text_.p, L.text_len, s, smer_.p);
                g_launches++;
                FIN_CUDA(cudaGetLastError());
                FIN_CUDA(cudaDeviceSynchronize());
            }
            view_.smer = smer_.p;
            view_.smer_s = s;
        }
    }
    fits32_ = L.n_unitigs < 0x7FFFFFFFull && L.max_unitig_len < 0x7FFFFFFFull;

    info_.k = L.k;
    info_.n_nodes = (int64_t)L.n_nodes;
    info_.n_kmers = L.n_kmers;
    info_.n_unitigs = (int64_t)L.n_unitigs;
    info_.text_len = (int64_t)L.text_len;
    info_.n_finimizers = (int64_t)L.n_fmin;
    info_.device = device;
    info_.device_bytes = (int64_t)(lines_.bytes() + ustart_.bytes() + fmin_.bytes() + tstart_.bytes() +
                                   offsets_.bytes() + ends_.bytes() + text_.bytes() + smer_.bytes());
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
template <int KCAP, class HitT, bool COUNT, bool RUNS, int G>
static void launch_search_g(const IndexView& v, const uint64_t* packed, const uint64_t* offsets, uint64_t n_reads,
                            const SearchOut<HitT>& out, SearchTotals* totals, int n_strands, cudaStream_t st) {
    const int threads = FIN_SEARCH_THREADS;
    static thread_local int resident[64] = {0};  // per device
    int dev = 0;
    FIN_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw arg_error("CUDA device ordinal out of range");
    auto kernel = search_reads_kernel<KCAP, HitT, COUNT, RUNS, G>;
    if (!resident[dev]) {
        int per_sm = 0, sms = 0;
        // no shared memory: all of the unified L1/shared array as L1 (the index lines live there)
        FIN_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxL1));
        FIN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0));
        FIN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        resident[dev] = std::max(1, per_sm) * std::max(1, sms);
    }
    const uint64_t need = (n_reads * (uint64_t)G + threads - 1) / threads;
    const unsigned blocks = (unsigned)std::min<uint64_t>(need, (uint64_t)resident[dev]);
    kernel<<<blocks, threads, 0, st>>>(v, packed, offsets, n_reads, out, totals, n_strands);
    g_launches++;
    FIN_CUDA(cudaGetLastError());
}

// FIN_COOP_GROUP=2|4: the warp-cooperative variant (G lanes per chain), dense int32 hits without counters, k <= 64 only
static int coop_group() {
    static const int g = [] {
        const char* e = getenv("FIN_COOP_GROUP");
        const int v = e ? atoi(e) : 1;
        return v == 2 || v == 4 ? v : 1;
    }();
    return g;
}

template <int KCAP, class HitT, bool COUNT, bool RUNS>
static void launch_search_k(const IndexView& v, const uint64_t* packed, const uint64_t* offsets, uint64_t n_reads,
                            const SearchOut<HitT>& out, SearchTotals* totals, int n_strands, cudaStream_t st) {
    if constexpr (KCAP <= 64 && sizeof(HitT) == 4 && !COUNT && !RUNS) {
        if (coop_group() == 2) return launch_search_g<KCAP, HitT, COUNT, RUNS, 2>(v, packed, offsets, n_reads, out, totals, n_strands, st);
        if (coop_group() == 4) return launch_search_g<KCAP, HitT, COUNT, RUNS, 4>(v, packed, offsets, n_reads, out, totals, n_strands, st);
    }
    launch_search_g<KCAP, HitT, COUNT, RUNS, 1>(v, packed, offsets, n_reads, out, totals, n_strands, st);
}

template <class HitT, bool COUNT, bool RUNS>
static void launch_search_t(const IndexView& v, const uint64_t* packed, const uint64_t* offsets, uint64_t n_reads,
                            const SearchOut<HitT>& out, SearchTotals* totals, int n_strands, cudaStream_t st) {
    switch (window_capacity(v.k)) {
        case 32: launch_search_k<32, HitT, COUNT, RUNS>(v, packed, offsets, n_reads, out, totals, n_strands, st); break;
        case 64: launch_search_k<64, HitT, COUNT, RUNS>(v, packed, offsets, n_reads, out, totals, n_strands, st); break;
        case 128: launch_search_k<128, HitT, COUNT, RUNS>(v, packed, offsets, n_reads, out, totals, n_strands, st); break;
        default: launch_search_k<256, HitT, COUNT, RUNS>(v, packed, offsets, n_reads, out, totals, n_strands, st); break;
    }
}

// words of the forward strand's found mask for a batch (SearchOut::fmask)
static uint64_t fmask_words(uint64_t n_reads, uint64_t total_kmers) { return (total_kmers >> 5) + n_reads + 2; }

// what the pipeline writes: dense hits (hit_bytes 4 / 8, d_hits) or runs (hit_bytes 0, sink)
struct PipelineOut {
    int hit_bytes = 4;
    void* d_hits = nullptr;
    uint32_t* d_fmask = nullptr;
    uint8_t* d_strand_flags = nullptr;  // n_reads bytes when the index has a strand prefilter
    bool filter_done = false;           // the strand prefilter of this batch was already launched (launch_filter)
    RunSink sink{nullptr, nullptr, nullptr, 0};
};

// pack + k-mer counts + exclusive scan on one stream; every pointer is device memory
static void launch_prepare(const IndexView& v, const uint8_t* d_ascii, uint64_t n_bases, const uint64_t* d_offsets, uint64_t n_reads,
                           uint64_t* d_packed, uint64_t* d_out_offsets, void* d_scan_tmp, size_t scan_tmp_bytes,
                           SearchTotals* d_totals, cudaStream_t st) {
    const uint64_t n_words = (n_bases + 31) / 32 + 1;
    const unsigned threads = 256;
    pack_reads_kernel<<<(unsigned)((n_words + threads - 1) / threads), threads, 0, st>>>(d_ascii, n_bases, d_packed, n_words, d_totals);
    g_launches++;
    kmer_counts_kernel<<<(unsigned)((n_reads + 1 + threads - 1) / threads), threads, 0, st>>>(d_offsets, n_reads, n_bases, v.k,
                                                                                                d_out_offsets, d_totals);
    g_launches++;
    FIN_CUDA(cudaGetLastError());
    FIN_CUDA(cub::DeviceScan::ExclusiveSum(d_scan_tmp, scan_tmp_bytes, d_out_offsets, d_out_offsets, (int64_t)(n_reads + 1), st));
    g_launches += 2;
}

// the strand prefilter of a batch (runs mode: hit_bytes 0, nothing to fill; dense: the reads no strand of which is streamed
// get their (-1,-1) here)
static void launch_filter(const IndexView& v, const uint64_t* d_packed, const uint64_t* d_offsets, uint64_t n_reads,
                          const uint64_t* d_out_offsets, const PipelineOut& po, SearchTotals* d_totals, bool count, int n_strands,
                          cudaStream_t st) {
    if (n_reads == 0 || !v.smer || !po.d_strand_flags) return;
    const unsigned ft = 128, fb = (unsigned)((n_reads + ft - 1) / ft);
    if (po.hit_bytes == 8)
        strand_filter_kernel<int64_t><<<fb, ft, 0, st>>>(v, d_packed, d_offsets, n_reads, n_strands, po.d_strand_flags, d_out_offsets,
                                                         reinterpret_cast<HitPair<int64_t>*>(po.d_hits), count ? d_totals : nullptr);
    else
        strand_filter_kernel<int32_t><<<fb, ft, 0, st>>>(v, d_packed, d_offsets, n_reads, n_strands, po.d_strand_flags, d_out_offsets,
                                                         po.hit_bytes == 4 ? reinterpret_cast<HitPair<int32_t>*>(po.d_hits) : nullptr,
                                                         count ? d_totals : nullptr);
    g_launches++;
    FIN_CUDA(cudaGetLastError());
}

static void launch_search(const IndexView& v, const uint64_t* d_packed, const uint64_t* d_offsets, uint64_t n_reads,
                          const uint64_t* d_out_offsets, const PipelineOut& po, SearchTotals* d_totals, bool count, int n_strands,
                          cudaStream_t st) {
    if (n_reads == 0) return;
    const uint8_t* sf = v.smer ? po.d_strand_flags : nullptr;
    if (!po.filter_done) launch_filter(v, d_packed, d_offsets, n_reads, d_out_offsets, po, d_totals, count, n_strands, st);
    if (po.hit_bytes == 8) {
        SearchOut<int64_t> out{reinterpret_cast<HitPair<int64_t>*>(po.d_hits), d_out_offsets, po.d_fmask, po.sink, sf};
        if (count) launch_search_t<int64_t, true, false>(v, d_packed, d_offsets, n_reads, out, d_totals, n_strands, st);
        else launch_search_t<int64_t, false, false>(v, d_packed, d_offsets, n_reads, out, d_totals, n_strands, st);
        return;
    }
    SearchOut<int32_t> out{reinterpret_cast<HitPair<int32_t>*>(po.d_hits), d_out_offsets, po.d_fmask, po.sink, sf};
    if (po.hit_bytes == 4) {
        if (count) launch_search_t<int32_t, true, false>(v, d_packed, d_offsets, n_reads, out, d_totals, n_strands, st);
        else launch_search_t<int32_t, false, false>(v, d_packed, d_offsets, n_reads, out, d_totals, n_strands, st);
    } else {
        if (count) launch_search_t<int32_t, true, true>(v, d_packed, d_offsets, n_reads, out, d_totals, n_strands, st);
        else launch_search_t<int32_t, false, true>(v, d_packed, d_offsets, n_reads, out, d_totals, n_strands, st);
    }
}

// runs mode, after the record count is known: sort by (read, first k-mer), CSR offsets (+ base)
static size_t sort_tmp_bytes_for(uint64_t n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const RunRec*)nullptr, (RunRec*)nullptr,
                                    (int64_t)n);
    return bytes + 256;
}

static void launch_run_sort(const uint64_t* keys_in, uint64_t* keys_out, const RunRec* recs_in, RunRec* recs_out, uint64_t n_runs,
                            uint64_t n_reads, uint64_t base, uint64_t* d_run_offsets, void* d_tmp, size_t tmp_bytes, cudaStream_t st) {
    if (n_runs) {
        const int end_bit = std::min(64, kRunKeyPosBits + std::max(1, bit_length(n_reads)));
        FIN_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, keys_in, keys_out, recs_in, recs_out, (int64_t)n_runs, 0, end_bit, st));
        g_launches += (end_bit + 7) / 8 + 1;
    }
    const unsigned threads = 256;
    run_bounds_kernel<<<(unsigned)((n_runs + 1 + threads - 1) / threads), threads, 0, st>>>(keys_out, n_runs, n_reads, base, d_run_offsets);
    g_launches++;
    FIN_CUDA(cudaGetLastError());
}

static size_t scan_tmp_bytes_for(uint64_t n) {
    size_t bytes = 0;
    uint64_t* dummy = nullptr;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, dummy, dummy, (int64_t)n);
    return bytes + 256;
}

static void check_flags(unsigned flags) {
    if (flags & kErrBadOffsets) throw arg_error("read_offsets must be non-decreasing and end at n_bases");
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
    a.n_filter_lookups += (int64_t)t.n_filter;
    a.n_lcs_skipped += (int64_t)t.n_skip_lcs;
    a.n_rank_skipped += (int64_t)t.n_skip_rank;
}

// -------------------------------------------------------------------------------------------------------
// device-resident batch
// -------------------------------------------------------------------------------------------------------
struct EventTimer {  // three events on the default stream: start, before the search kernel, end
    cudaEvent_t e0 = nullptr, em = nullptr, e1 = nullptr;
    EventTimer() {
        FIN_CUDA(cudaEventCreate(&e0));
        FIN_CUDA(cudaEventCreate(&em));
        FIN_CUDA(cudaEventCreate(&e1));
    }
    ~EventTimer() {
        cudaEventDestroy(e0);
        cudaEventDestroy(em);
        cudaEventDestroy(e1);
    }
};

// prepare + the host-side check of what the device found in the caller's offsets; returns the k-mer positions
uint64_t DeviceIndex::prepare_device_batch(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases) {
    dev_packed_.ensure((n_bases + 31) / 32 + 2);
    dev_out_offsets_.ensure(n_reads + 2);
    const size_t tmp = scan_tmp_bytes_for(n_reads + 1);
    dev_scan_tmp_.ensure(tmp);
    dev_totals_.ensure(sizeof(SearchTotals));
    FIN_CUDA(cudaMemsetAsync(dev_totals_.p, 0, sizeof(SearchTotals), 0));
    launch_prepare(view_, (const uint8_t*)d_seqs, n_bases, d_read_offsets, n_reads, dev_packed_.p, dev_out_offsets_.p, dev_scan_tmp_.p,
                   tmp, (SearchTotals*)dev_totals_.p, 0);
    uint64_t total = 0;
    unsigned flags = 0;
    FIN_CUDA(cudaMemcpyAsync(&total, dev_out_offsets_.p + n_reads, 8, cudaMemcpyDeviceToHost, 0));
    FIN_CUDA(cudaMemcpyAsync(&flags, &((SearchTotals*)dev_totals_.p)->error_flags, 4, cudaMemcpyDeviceToHost, 0));
    FIN_CUDA(cudaStreamSynchronize(0));
    check_flags(flags & kErrBadOffsets);
    if ((total >> 5) + n_reads >= 0xFFFFFFF0ull) throw limit_error("batch too large for one call: split it");
    dev_fmask_.ensure(fmask_words(n_reads, total));
    dev_strand_flags_.ensure(n_reads + 1);
    return total;
}

void DeviceIndex::finish_device_batch(EventTimer& t, fin_search_counts* counts, float* kernel_ms, SearchTotals& tot) {
    FIN_CUDA(cudaEventRecord(t.e1, 0));
    FIN_CUDA(cudaMemcpy(&tot, dev_totals_.p, sizeof tot, cudaMemcpyDeviceToHost));
    FIN_CUDA(cudaEventSynchronize(t.e1));
    float ms = 0;
    FIN_CUDA(cudaEventElapsedTime(&ms, t.e0, t.e1));
    last_timing.pipeline_ms = ms;
    FIN_CUDA(cudaEventElapsedTime(&last_timing.prepare_ms, t.e0, t.em));
    FIN_CUDA(cudaEventElapsedTime(&last_timing.search_kernel_ms, t.em, t.e1));
    if (kernel_ms) *kernel_ms = ms;
    check_flags(tot.error_flags);
    fin_search_counts c{};
    last_access = fin_access_counters{};
    accumulate(tot, c, last_access);
    if (counts) *counts = c;
}

void DeviceIndex::search_device(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases,
                                uint64_t total_kmers, void* d_out, int hit_bytes, fin_search_counts* counts, float* kernel_ms) {
    if (hit_bytes != 4 && hit_bytes != 8) throw arg_error("hit_bytes must be 4 or 8");
    if (hit_bytes == 4 && !fits32_) throw limit_error("this index needs 64-bit hits");
    FIN_CUDA(cudaSetDevice(device_));
    EventTimer t;
    FIN_CUDA(cudaEventRecord(t.e0, 0));
    const uint64_t total = prepare_device_batch(d_seqs, d_read_offsets, n_reads, n_bases);
    if (total > total_kmers) throw arg_error("total_kmers (the capacity of d_out) is smaller than the number of k-mer positions of the batch");
    FIN_CUDA(cudaEventRecord(t.em, 0));
    PipelineOut po;
    po.hit_bytes = hit_bytes;
    po.d_hits = d_out;
    po.d_fmask = dev_fmask_.p;
    po.d_strand_flags = dev_strand_flags_.p;
    launch_search(view_, dev_packed_.p, d_read_offsets, n_reads, dev_out_offsets_.p, po, (SearchTotals*)dev_totals_.p, count_accesses, 2, 0);
    SearchTotals tot;
    finish_device_batch(t, counts, kernel_ms, tot);
}

uint64_t DeviceIndex::search_device_runs(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases,
                                         uint64_t* d_run_offsets, fin_run* d_runs, uint64_t runs_capacity, fin_search_counts* counts,
                                         float* kernel_ms) {
    static_assert(sizeof(fin_run) == sizeof(RunRec), "fin_run and RunRec are the same record");
    FIN_CUDA(cudaSetDevice(device_));
    EventTimer t;
    FIN_CUDA(cudaEventRecord(t.e0, 0));
    prepare_device_batch(d_seqs, d_read_offsets, n_reads, n_bases);
    dev_keys_a_.ensure(runs_capacity + 1);
    dev_keys_b_.ensure(runs_capacity + 1);
    dev_recs_a_.ensure((runs_capacity + 1) * sizeof(RunRec));
    FIN_CUDA(cudaEventRecord(t.em, 0));
    PipelineOut po;
    po.hit_bytes = 0;
    po.d_fmask = dev_fmask_.p;
    po.d_strand_flags = dev_strand_flags_.p;
    SearchTotals* dt = (SearchTotals*)dev_totals_.p;
    po.sink = RunSink{dev_keys_a_.p, (RunRec*)dev_recs_a_.p, &dt->n_runs, runs_capacity};
    launch_search(view_, dev_packed_.p, d_read_offsets, n_reads, dev_out_offsets_.p, po, dt, count_accesses, 2, 0);
    uint64_t n_runs = 0;
    FIN_CUDA(cudaMemcpy(&n_runs, &dt->n_runs, 8, cudaMemcpyDeviceToHost));
    if (n_runs <= runs_capacity) {
        const size_t tmp = sort_tmp_bytes_for(n_runs);
        dev_sort_tmp_.ensure(tmp);
        launch_run_sort(dev_keys_a_.p, dev_keys_b_.p, (const RunRec*)dev_recs_a_.p, (RunRec*)d_runs, n_runs, n_reads, 0, d_run_offsets,
                        dev_sort_tmp_.p, tmp, 0);
    }
    SearchTotals tot;
    finish_device_batch(t, counts, kernel_ms, tot);
    return n_runs;
}

// -------------------------------------------------------------------------------------------------------
// device-resident batch without any allocation (include/finito_b200.h fin_search_device_noalloc)
// -------------------------------------------------------------------------------------------------------
// Everything search_device keeps in DevBufs (2-bit reads, output offsets, found mask, strand flags, totals) is carved out
// of ONE file-scope device array of fixed size; the cub scan (whose temporary storage has no compile-time bound) is
// replaced by a single-block scan.  Every GPU that loads the library holds the array (8 MiB) until the process exits.
// All calls share it: it keeps the last call's contents (the totals are cleared here, the other parts are fully written
// before they are read), and two calls that overlap on one GPU race.
constexpr uint64_t kNoAllocMaxReads = FIN_NOALLOC_MAX_READS;
constexpr uint64_t kNoAllocMaxBases = FIN_NOALLOC_MAX_BASES;
constexpr size_t noalloc_round(size_t b) { return (b + 255) & ~(size_t)255; }
constexpr size_t kNaTotals = 0;
constexpr size_t kNaOutOffsets = kNaTotals + noalloc_round(sizeof(SearchTotals));
constexpr size_t kNaStrandFlags = kNaOutOffsets + noalloc_round((kNoAllocMaxReads + 2) * 8);
constexpr size_t kNaPacked = kNaStrandFlags + noalloc_round(kNoAllocMaxReads + 1);
constexpr size_t kNaFmask = kNaPacked + noalloc_round((kNoAllocMaxBases / 32 + 2) * 8);
constexpr size_t kNaEnd = kNaFmask + noalloc_round(((kNoAllocMaxBases >> 5) + kNoAllocMaxReads + 2) * 4);
constexpr size_t kNoAllocScratchBytes = (size_t)8 << 20;
static_assert(kNaEnd <= kNoAllocScratchBytes, "the scratch parts must fit the static array");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// exclusive prefix sum of a[0..n) in place, one block: every thread owns a contiguous slice
static __global__ void __launch_bounds__(1024) scan_offsets_block_kernel(uint64_t* __restrict__ a, uint32_t n) {
    __shared__ uint64_t part[1024];
    const uint32_t per = (n + blockDim.x - 1) / blockDim.x;
    const uint32_t lo = min(n, threadIdx.x * per), hi = min(n, lo + per);
    uint64_t sum = 0;
    for (uint32_t i = lo; i < hi; i++) sum += a[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (uint32_t d = 1; d < blockDim.x; d <<= 1) {  // Hillis-Steele over the 1024 slice sums
        const uint64_t add = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
        __syncthreads();
        part[threadIdx.x] += add;
        __syncthreads();
    }
    uint64_t run = part[threadIdx.x] - sum;
    for (uint32_t i = lo; i < hi; i++) {
        const uint64_t x = a[i];
        a[i] = run;
        run += x;
    }
}

// device address of the array on `device` (current): the symbol's name is only a host-side stand-in, never a pointer
// a kernel may be given.  Resolved once per device.
static unsigned char* noalloc_scratch_on(int device) {
    static std::mutex mu;
    static unsigned char* ptr[64] = {nullptr};
    if (device < 0 || device >= 64) throw arg_error("CUDA device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[device]) {
        void* p = nullptr;
        FIN_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p || ((uintptr_t)p & 255)) throw cuda_error("cudaGetSymbolAddress returned an unusable address for the static scratch");
        ptr[device] = (unsigned char*)p;
    }
    return ptr[device];
}

void DeviceIndex::search_device_noalloc(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases,
                                        uint64_t total_kmers, void* d_out, int hit_bytes, fin_search_counts* counts) {
    if (hit_bytes != 4 && hit_bytes != 8) throw arg_error("hit_bytes must be 4 or 8");
    if (hit_bytes == 4 && !fits32_) throw limit_error("this index needs 64-bit hits");
    if (n_reads > kNoAllocMaxReads || n_bases > kNoAllocMaxBases)
        throw limit_error("fin_search_device_noalloc: at most " + std::to_string(kNoAllocMaxReads) + " reads and " +
                          std::to_string(kNoAllocMaxBases) + " bases per call (the static scratch is fixed); got " +
                          std::to_string(n_reads) + " reads, " + std::to_string(n_bases) + " bases");
    FIN_CUDA(cudaSetDevice(device_));
    unsigned char* base = noalloc_scratch_on(device_);
    SearchTotals* d_totals = (SearchTotals*)(base + kNaTotals);
    uint64_t* d_out_offsets = (uint64_t*)(base + kNaOutOffsets);
    uint64_t* d_packed = (uint64_t*)(base + kNaPacked);
    FIN_CUDA(cudaMemsetAsync(d_totals, 0, sizeof(SearchTotals), 0));
    const uint64_t n_words = (n_bases + 31) / 32 + 1;
    const unsigned threads = 256;
    pack_reads_kernel<<<(unsigned)((n_words + threads - 1) / threads), threads, 0, 0>>>((const uint8_t*)d_seqs, n_bases, d_packed, n_words,
                                                                                         d_totals);
    kmer_counts_kernel<<<(unsigned)((n_reads + 1 + threads - 1) / threads), threads, 0, 0>>>(d_read_offsets, n_reads, n_bases, view_.k,
                                                                                              d_out_offsets, d_totals);
    scan_offsets_block_kernel<<<1, 1024, 0, 0>>>(d_out_offsets, (uint32_t)(n_reads + 1));
    g_launches += 3;
    FIN_CUDA(cudaGetLastError());
    uint64_t total = 0;
    unsigned flags = 0;
    FIN_CUDA(cudaMemcpyAsync(&total, d_out_offsets + n_reads, 8, cudaMemcpyDeviceToHost, 0));
    FIN_CUDA(cudaMemcpyAsync(&flags, &d_totals->error_flags, 4, cudaMemcpyDeviceToHost, 0));
    FIN_CUDA(cudaStreamSynchronize(0));
    check_flags(flags & kErrBadOffsets);
    if (total > total_kmers) throw arg_error("total_kmers (the capacity of d_out) is smaller than the number of k-mer positions of the batch");
    PipelineOut po;
    po.hit_bytes = hit_bytes;
    po.d_hits = d_out;
    po.d_fmask = (uint32_t*)(base + kNaFmask);
    po.d_strand_flags = base + kNaStrandFlags;
    launch_search(view_, d_packed, d_read_offsets, n_reads, d_out_offsets, po, d_totals, count_accesses, 2, 0);
    SearchTotals tot;
    FIN_CUDA(cudaMemcpy(&tot, d_totals, sizeof tot, cudaMemcpyDeviceToHost));
    check_flags(tot.error_flags);
    fin_search_counts c{};
    last_access = fin_access_counters{};
    accumulate(tot, c, last_access);
    if (counts) *counts = c;
}

// -------------------------------------------------------------------------------------------------------
// host batch: sub-batches round-robin over three streams, so that the H2D copy of sub-batch i+1 and the
// D2H copy of sub-batch i-1 overlap the kernels of sub-batch i.
// -------------------------------------------------------------------------------------------------------
// Dense hits: 8 bytes per k-mer leave over PCIe, so small sub-batches (the copies of one hide behind the kernels of
// the next).  Runs: the copies are small and what counts is the tail of every persistent kernel (a lane owns a read
// for ~2.5 ms at 150 bp; every extra sub-batch costs 3-4 ms, profiles/r2_sweep_runs_subbatch.log), so few sub-batches that
// double from 512 k to 4 M reads.  FIN_SUB_FIRST / FIN_SUB_CAP (runs), FIN_SUB_READS / FIN_SUB_MBASES (both) override.
struct SubBatchLimits {
    uint64_t reads, bases;
};
static SubBatchLimits sub_batch_limits(bool runs, int turn) {
    // runs: sub-batches double, so that the first copy-in (which nothing hides) stays short
    uint64_t first = 1ull << 19, cap = 1ull << 22;
    if (runs) {
        if (const char* e = getenv("FIN_SUB_FIRST")) if (atoll(e) > 0) first = (uint64_t)atoll(e);
        if (const char* e = getenv("FIN_SUB_CAP")) if (atoll(e) > 0) cap = (uint64_t)atoll(e);
    }
    SubBatchLimits l = runs ? SubBatchLimits{std::min<uint64_t>(cap, first << std::min(turn, 8)), cap * 192}
                            : SubBatchLimits{1ull << 19, 48ull << 20};
    if (const char* e = getenv("FIN_SUB_READS")) if (atoll(e) > 0) l.reads = (uint64_t)atoll(e);
    if (const char* e = getenv("FIN_SUB_MBASES")) if (atoll(e) > 0) l.bases = (uint64_t)atoll(e) << 20;
    return l;
}
static constexpr int kSlots = 4;

void DeviceIndex::ensure_slots() {
    if (slots_.empty()) {
        for (int i = 0; i < kSlots; i++) {
            std::unique_ptr<SearchSlot> s(new SearchSlot());
            FIN_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
            int lo_prio = 0, hi_prio = 0;
            FIN_CUDA(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
            FIN_CUDA(cudaStreamCreateWithPriority(&s->prep_stream, cudaStreamNonBlocking, hi_prio));
            FIN_CUDA(cudaEventCreateWithFlags(&s->prepped, cudaEventDisableTiming));
            FIN_CUDA(cudaEventCreateWithFlags(&s->counted, cudaEventDisableTiming));
            FIN_CUDA(cudaEventCreateWithFlags(&s->done, cudaEventDisableTiming));
            FIN_CUDA(cudaHostAlloc((void**)&s->h_totals, sizeof(SearchTotals), cudaHostAllocDefault));
            s->totals.alloc(sizeof(SearchTotals));
            slots_.push_back(std::move(s));
        }
    }
    // a previous call that ended with an exception may have left sub-batches in flight: let them finish and forget them,
    // so that their copies no longer write into the caller's buffers and their totals never reach this call's counters
    for (auto& s : slots_) {
        if (s->busy) {
            cudaStreamSynchronize(s->prep_stream);
            cudaStreamSynchronize(s->stream);
        }
        s->busy = false;
    }
}

// drains the pipeline when a call leaves with an exception (the caller's buffers must not be written after the return)
struct SlotDrain {
    std::vector<std::unique_ptr<SearchSlot>>& slots;
    bool armed = true;
    ~SlotDrain() {
        if (!armed) return;
        for (auto& s : slots) {
            if (s->busy) {
                cudaStreamSynchronize(s->prep_stream);
                cudaStreamSynchronize(s->stream);
            }
            s->busy = false;
        }
    }
};

// sub-batch [r0, r1): at most kSubReads reads and kSubBases bases (always at least one read); validates the offsets
// of the sub-batch and counts its k-mer positions
static uint64_t cut_sub_batch(const uint64_t* read_offsets, uint64_t n_reads, uint64_t r0, int k, const SubBatchLimits lim, uint64_t& nkm,
                              uint64_t& max_len) {
    const uint64_t kSubBases = lim.bases;
    uint64_t r1 = std::min(n_reads, r0 + lim.reads);
    const uint64_t base0 = read_offsets[r0];
    if (read_offsets[r1] >= base0 && read_offsets[r1] - base0 > kSubBases) {
        const uint64_t* lo = read_offsets + r0 + 1;
        const uint64_t* hi = read_offsets + r1 + 1;
        const uint64_t* p = std::upper_bound(lo, hi, base0 + kSubBases);  // first offset beyond the cap
        r1 = std::max<uint64_t>(r0 + 1, (uint64_t)(p - read_offsets) - 1);
    }
    nkm = 0;
    bool ordered = true;
    for (uint64_t i = r0; i < r1; i++) {
        ordered &= read_offsets[i + 1] >= read_offsets[i];
        const uint64_t len = read_offsets[i + 1] - read_offsets[i];
        max_len = std::max(max_len, len);
        nkm += len >= (uint64_t)k ? len - (uint64_t)k + 1 : 0;
    }
    if (!ordered) throw arg_error("read_offsets must be non-decreasing");
    return r1;
}

// H2D of a sub-batch and its preparation kernels on the slot's stream
static void stage_sub_batch(const IndexView& v, SearchSlot& s, const char* seqs, const uint64_t* read_offsets, uint64_t r0, uint64_t r1,
                            uint64_t nkm, cudaStream_t st) {
    NvtxRange range("fin:search:stage_sub_batch");
    const uint64_t base0 = read_offsets[r0];
    const uint64_t nr = r1 - r0, nb = read_offsets[r1] - base0;
    s.r0 = r0; s.nr = nr; s.nb = nb; s.nkm = nkm; s.base0 = base0;
    s.ascii.ensure(nb + 64);
    s.packed.ensure((nb + 31) / 32 + 2);
    s.offsets.ensure(nr + 1);
    s.out_offsets.ensure(nr + 2);
    s.fmask.ensure(fmask_words(nr, nkm));
    s.strand_flags.ensure(nr + 1);
    const size_t tmp = scan_tmp_bytes_for(nr + 1);
    s.scan_tmp.ensure(tmp);
    if (s.h_offsets_cap < nr + 1) {
        if (s.h_offsets_pinned) cudaFreeHost(s.h_offsets_pinned);
        s.h_offsets_pinned = nullptr;
        s.h_offsets_cap = nr + 1 + nr / 4;
        FIN_CUDA(cudaHostAlloc((void**)&s.h_offsets_pinned, s.h_offsets_cap * 8, cudaHostAllocDefault));
    }
    for (uint64_t i = 0; i <= nr; i++) s.h_offsets_pinned[i] = read_offsets[r0 + i] - base0;
    FIN_CUDA(cudaMemsetAsync(s.totals.p, 0, sizeof(SearchTotals), st));
    FIN_CUDA(cudaMemcpyAsync(s.ascii.p, seqs + base0, nb, cudaMemcpyHostToDevice, st));
    FIN_CUDA(cudaMemcpyAsync(s.offsets.p, s.h_offsets_pinned, (nr + 1) * 8, cudaMemcpyHostToDevice, st));
    launch_prepare(v, s.ascii.p, nb, s.offsets.p, nr, s.packed.p, s.out_offsets.p, s.scan_tmp.p, tmp, (SearchTotals*)s.totals.p, st);
}

template <class HitT>
void DeviceIndex::search_host(const char* seqs, const uint64_t* read_offsets, uint64_t n_reads, HitT* out_pairs,
                              uint64_t out_capacity, fin_search_counts* counts, int n_strands) {
    if (sizeof(HitT) == 4 && !fits32_) throw limit_error("this index needs 64-bit hits (use fin_search_batch)");
    FIN_CUDA(cudaSetDevice(device_));
    const int k = view_.k;
    // the offsets are validated and the k-mer positions counted sub-batch by sub-batch while the GPU already works on
    // the previous ones: no pass over the whole batch stands between the call and the first copy
    ensure_slots();
    SlotDrain drain{slots_};
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

    uint64_t r0 = 0, out_pos = 0, max_len = 0;
    int turn = 0;
    while (r0 < n_reads) {
        const SubBatchLimits lim = sub_batch_limits(false, turn);
        uint64_t nkm = 0;
        const uint64_t r1 = cut_sub_batch(read_offsets, n_reads, r0, k, lim, nkm, max_len);
        if (out_pos + nkm > out_capacity) throw arg_error("out_capacity is smaller than the number of k-mer positions of the batch");
        SearchSlot& s = *slots_[turn % kSlots];
        turn++;
        retire(s);
        s.hits.ensure(std::max<uint64_t>(nkm, 1) * 2 * sizeof(HitT));
        stage_sub_batch(view_, s, seqs, read_offsets, r0, r1, nkm, s.stream);
        PipelineOut po;
        po.hit_bytes = (int)sizeof(HitT);
        po.d_hits = s.hits.p;
        po.d_fmask = s.fmask.p;
        po.d_strand_flags = s.strand_flags.p;
        launch_search(view_, s.packed.p, s.offsets.p, s.nr, s.out_offsets.p, po, (SearchTotals*)s.totals.p, count_accesses, n_strands, s.stream);
        if (nkm)
            FIN_CUDA(cudaMemcpyAsync(out_pairs + 2 * out_pos, s.hits.p, nkm * 2 * sizeof(HitT), cudaMemcpyDeviceToHost, s.stream));
        FIN_CUDA(cudaMemcpyAsync(s.h_totals, s.totals.p, sizeof(SearchTotals), cudaMemcpyDeviceToHost, s.stream));
        FIN_CUDA(cudaEventRecord(s.done, s.stream));
        s.busy = true;
        out_pos += nkm;
        r0 = r1;
    }
    for (auto& s : slots_) retire(*s);
    drain.armed = false;
    check_flags(flags);
    if (counts) *counts = c;
}

template void DeviceIndex::search_host<int32_t>(const char*, const uint64_t*, uint64_t, int32_t*, uint64_t, fin_search_counts*, int);
template void DeviceIndex::search_host<int64_t>(const char*, const uint64_t*, uint64_t, int64_t*, uint64_t, fin_search_counts*, int);

// -------------------------------------------------------------------------------------------------------
// host batch, answers as runs.  Per sub-batch t:
//   stage(t)  on the slot's high-priority prep stream: copy in, pack, offsets, strand prefilter        -> event prepped
//   search(t) on the slot's stream, after prepped(t): the persistent search kernel, totals to the host -> event counted
//   finish(t) once the record count is on the host: sort by (read, first k-mer), CSR offsets, copy out -> event done
// A persistent search kernel owns every SM until its last blocks retire, so whatever else has to run before the next
// search kernel runs in that tail.  stage(t+1) is therefore issued BEFORE search(t): its copy-in overlaps search(t-1) and
// its small kernels take the first SM slots that search(t-1) frees (stream priority), while search(t) - whose own staging
// finished a whole sub-batch earlier - fills the rest of the tail.  finish(t-1) is issued after search(t), so the GPU
// always has the next search queued while the host waits for a count.
// -------------------------------------------------------------------------------------------------------
uint64_t DeviceIndex::search_host_runs(const char* seqs, const uint64_t* read_offsets, uint64_t n_reads, uint64_t* run_offsets,
                                       fin_run* runs, uint64_t runs_capacity, fin_search_counts* counts) {
    FIN_CUDA(cudaSetDevice(device_));
    const int k = view_.k;
    ensure_slots();
    SlotDrain drain{slots_};
    fin_search_counts c{};
    last_access = fin_access_counters{};
    unsigned flags = 0;
    uint64_t run_pos = 0;

    auto ensure_runs = [&](SearchSlot& s, uint64_t n) {
        s.keys_a.ensure(n);
        s.keys_b.ensure(n);
        s.recs_a.ensure(n);
        s.recs_b.ensure(n);
    };
    auto pipeline_out = [&](SearchSlot& s) {
        PipelineOut po;
        po.hit_bytes = 0;
        po.d_fmask = s.fmask.p;
        po.d_strand_flags = s.strand_flags.p;
        po.filter_done = true;
        po.sink = RunSink{s.keys_a.p, s.recs_a.p, &((SearchTotals*)s.totals.p)->n_runs, std::min<uint64_t>(s.keys_a.n, s.recs_a.n)};
        return po;
    };
    auto stage = [&](SearchSlot& s, uint64_t r0, uint64_t r1, uint64_t nkm) {
        // The slot's previous sub-batch was searched and its finish() issued kSlots - 1 turns ago: the staging buffers are
        // free (the host saw `counted`), and what finish() still reads (the appended records) is only written again by
        // the next search on the slot's own stream, i.e. behind it.  No wait on `done` here: the sort of finish() sits
        // behind the running persistent kernel and would stall the host for a whole sub-batch.
        ensure_runs(s, nkm / 32 + 4096);  // ~4 runs per 150 bp read; more than that is searched again with the exact room
        stage_sub_batch(view_, s, seqs, read_offsets, r0, r1, nkm, s.prep_stream);
        launch_filter(view_, s.packed.p, s.offsets.p, s.nr, s.out_offsets.p, pipeline_out(s), (SearchTotals*)s.totals.p, count_accesses, 2,
                      s.prep_stream);
        FIN_CUDA(cudaEventRecord(s.prepped, s.prep_stream));
        s.busy = true;
    };
    auto search = [&](SearchSlot& s) {
        FIN_CUDA(cudaStreamWaitEvent(s.stream, s.prepped, 0));
        launch_search(view_, s.packed.p, s.offsets.p, s.nr, s.out_offsets.p, pipeline_out(s), (SearchTotals*)s.totals.p, count_accesses, 2,
                      s.stream);
        FIN_CUDA(cudaMemcpyAsync(s.h_totals, s.totals.p, sizeof(SearchTotals), cudaMemcpyDeviceToHost, s.stream));
        FIN_CUDA(cudaEventRecord(s.counted, s.stream));
    };
    static const bool trace = getenv("FIN_TRACE") != nullptr;
    const auto t_enter = std::chrono::steady_clock::now();
    auto since = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_enter).count(); };
    auto finish = [&](SearchSlot& s) {
        NvtxRange range("fin:search:runs_sort_and_copy_out");
        const double w0 = since();
        FIN_CUDA(cudaEventSynchronize(s.counted));
        if (trace) fprintf(stderr, "[fin trace] finish r0=%llu nr=%llu: waited %.2f ms for the count (at %.2f ms)\n", (unsigned long long)s.r0,
                           (unsigned long long)s.nr, since() - w0, since());
        uint64_t n = s.h_totals->n_runs;
        if (n > std::min<uint64_t>(s.keys_a.n, s.recs_a.n)) {  // more runs than the slot had room for: search again with enough
            const unsigned first_flags = s.h_totals->error_flags;  // what packing and the offset check found
            FIN_CUDA(cudaStreamSynchronize(s.stream));
            ensure_runs(s, n + n / 8 + 1024);
            SearchTotals* dt = (SearchTotals*)s.totals.p;
            FIN_CUDA(cudaMemsetAsync(&dt->found_fwd, 0, sizeof(SearchTotals), s.stream));
            launch_search(view_, s.packed.p, s.offsets.p, s.nr, s.out_offsets.p, pipeline_out(s), dt, false, 2, s.stream);
            FIN_CUDA(cudaMemcpyAsync(s.h_totals, s.totals.p, sizeof(SearchTotals), cudaMemcpyDeviceToHost, s.stream));
            FIN_CUDA(cudaStreamSynchronize(s.stream));
            flags |= first_flags;
            n = s.h_totals->n_runs;
        }
        flags |= s.h_totals->error_flags;
        accumulate(*s.h_totals, c, last_access);
        const size_t tmp = sort_tmp_bytes_for(n);
        s.sort_tmp.ensure(tmp);
        s.run_offsets.ensure(s.nr + 2);
        launch_run_sort(s.keys_a.p, s.keys_b.p, s.recs_a.p, s.recs_b.p, n, s.nr, run_pos, s.run_offsets.p, s.sort_tmp.p, tmp, s.stream);
        FIN_CUDA(cudaMemcpyAsync(run_offsets + s.r0, s.run_offsets.p, (s.nr + 1) * 8, cudaMemcpyDeviceToHost, s.stream));
        if (n && run_pos + n <= runs_capacity)
            FIN_CUDA(cudaMemcpyAsync(runs + run_pos, s.recs_b.p, n * sizeof(RunRec), cudaMemcpyDeviceToHost, s.stream));
        FIN_CUDA(cudaEventRecord(s.done, s.stream));
        run_pos += n;
    };

    if (n_reads == 0) run_offsets[0] = 0;
    // the sub-batch plan of the call (offsets validated, k-mer positions counted) is made as it goes: cut(t+1) while the GPU works
    uint64_t next_r0 = 0, max_len = 0;
    int planned = 0;
    auto cut_next = [&](uint64_t& r0, uint64_t& r1, uint64_t& nkm) {
        if (next_r0 >= n_reads) return false;
        r0 = next_r0;
        r1 = cut_sub_batch(read_offsets, n_reads, r0, k, sub_batch_limits(true, planned), nkm, max_len);
        next_r0 = r1;
        planned++;
        return true;
    };
    uint64_t r0 = 0, r1 = 0, nkm = 0;
    int t = 0;
    SearchSlot* prev = nullptr;
    bool have = cut_next(r0, r1, nkm);
    if (have) stage(*slots_[0], r0, r1, nkm);
    while (have) {
        SearchSlot& cur = *slots_[t % kSlots];
        uint64_t a0 = 0, a1 = 0, akm = 0;
        const bool more = cut_next(a0, a1, akm);
        const double s0 = since();
        if (more) stage(*slots_[(t + 1) % kSlots], a0, a1, akm);  // before search(t): see the note above
        const double s1 = since();
        search(cur);
        if (trace) fprintf(stderr, "[fin trace] t=%d: stage(t+1) %.2f ms on the host, search(t) enqueued at %.2f ms\n", t, s1 - s0, since());
        if (prev) finish(*prev);
        prev = &cur;
        have = more;
        t++;
    }
    if (prev) finish(*prev);
    for (auto& s : slots_) {
        if (s->busy) FIN_CUDA(cudaEventSynchronize(s->done));
        s->busy = false;
    }
    drain.armed = false;
    if (trace) fprintf(stderr, "[fin trace] done at %.2f ms\n", since());
    check_flags(flags);
    if (counts) *counts = c;
    return run_pos;
}

}  // namespace fin
