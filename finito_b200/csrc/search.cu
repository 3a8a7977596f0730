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
// rank sectors
// -------------------------------------------------------------------------------------------------------
std::vector<uint64_t> make_rank_sectors(const uint64_t* const* rows, int n_rows, uint64_t n_bits, uint64_t* totals) {
    const uint64_t n_blocks = n_bits / kRankBlockBits + 1;
    const uint64_t n_words = (n_bits + 63) >> 6;
    std::vector<uint64_t> s(n_blocks * 4 * (uint64_t)n_rows, 0);
    for (int r = 0; r < n_rows; r++) {
        uint64_t cum = 0;
        const uint64_t* w = rows[r];
        const uint64_t tail_bits = n_bits & 63;
        for (uint64_t b = 0; b < n_blocks; b++) {
            uint64_t* sec = &s[(b * (uint64_t)n_rows + (uint64_t)r) * 4];
            sec[0] = cum;
            for (int j = 0; j < 3; j++) {
                const uint64_t wi = b * 3 + (uint64_t)j;
                uint64_t x = wi < n_words ? w[wi] : 0;
                if (tail_bits && wi == n_words - 1) x &= (1ULL << tail_bits) - 1;  // bits past the end never count
                sec[1 + j] = x;
                cum += (uint64_t)__builtin_popcountll(x);
            }
        }
        if (totals) totals[r] = cum;
    }
    return s;
}

// -------------------------------------------------------------------------------------------------------
// index files (FinimizerIndex.hh:187-241)
// -------------------------------------------------------------------------------------------------------
IndexHost load_index_files(const std::string& prefix) {
    IndexHost ix;
    ix.lcs = load_packed_file(prefix + ".LCS.sdsl", 0);
    ix.fmin = load_packed_file(prefix + ".FBV.sdsl", 1);
    ix.offsets = load_packed_file(prefix + ".O.sdsl", 0);
    ix.text = load_packed_file(prefix + ".packed_unitigs.sdsl", 2);
    ix.ends = load_packed_file(prefix + ".unitig_endpoints.sdsl", 0);
    ix.ustart = load_packed_file(prefix + ".Ustart.sdsl", 1);
    ix.sbwt = load_sbwt_file(prefix + ".sbwt", false);
    const uint64_t N = (uint64_t)ix.sbwt.n_nodes;
    if (ix.lcs.size() != N || ix.fmin.size() != N || ix.ustart.size() != N)
        throw format_error("index files under " + prefix + " disagree on the number of SBWT columns");
    return ix;
}

void save_index_files(const IndexHost& ix, const std::string& prefix) {
    save_packed_file(prefix + ".O.sdsl", ix.offsets, true);
    save_packed_file(prefix + ".FBV.sdsl", ix.fmin, false);
    save_packed_file(prefix + ".packed_unitigs.sdsl", ix.text, false);
    save_packed_file(prefix + ".unitig_endpoints.sdsl", ix.ends, true);
    save_packed_file(prefix + ".Ustart.sdsl", ix.ustart, false);
    save_packed_file(prefix + ".LCS.sdsl", ix.lcs, true);
    const std::string path = prefix + ".sbwt";
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw io_error("cannot write " + path);
    const size_t n = ix.sbwt.body.size();
    const bool ok = fwrite(ix.sbwt.body.data(), 1, n, f) == n;
    if (fclose(f) != 0 || !ok) throw io_error("short write on " + path);
}

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

    const uint64_t N = (uint64_t)h.sbwt.n_nodes, T = h.text.size(), U = h.ends.size(), F = h.offsets.size();
    const uint64_t lim = 0xFFFFFFFFull - 256;
    if (N >= lim || T >= lim) throw limit_error("index too large for the 32-bit column/text positions of this build");
    if (N < 2) throw domain_error("SBWT with fewer than 2 columns");
    if (U == 0 || (U && h.ends.get(U - 1) != T)) throw format_error("unitig end points do not match the packed text");

    {  // subset matrix
        const uint64_t* rows[4] = {h.sbwt.bits[0].words.data(), h.sbwt.bits[1].words.data(), h.sbwt.bits[2].words.data(),
                                   h.sbwt.bits[3].words.data()};
        std::vector<uint64_t> s = make_rank_sectors(rows, 4, N, nullptr);
        matrix_ = upload(s.data(), s.size());
    }
    {
        const uint64_t* r[1] = {h.ustart.words.data()};
        std::vector<uint64_t> s = make_rank_sectors(r, 1, N, nullptr);
        ustart_ = upload(s.data(), s.size());
    }
    {
        const uint64_t* r[1] = {h.fmin.words.data()};
        uint64_t total = 0;
        std::vector<uint64_t> s = make_rank_sectors(r, 1, N, &total);
        fmin_ = upload(s.data(), s.size());
        (void)total;
    }
    std::vector<uint32_t> ends(U);
    {
        std::vector<uint64_t> ts((T >> 6) + 2, 0);
        uint64_t prev = 0;
        for (uint64_t u = 0; u < U; u++) {
            const uint64_t e = h.ends.get(u);
            if (e <= prev || e > T) throw format_error("unitig end points are not strictly increasing");
            ts[prev >> 6] |= 1ULL << (prev & 63);
            ends[u] = (uint32_t)e;
            prev = e;
        }
        const uint64_t* r[1] = {ts.data()};
        std::vector<uint64_t> s = make_rank_sectors(r, 1, T, nullptr);
        tstart_ = upload(s.data(), s.size());
    }
    ends_ = upload(ends.data(), ends.size());
    {
        std::vector<uint8_t> l(N + 64, 0);
        for (uint64_t i = 0; i < N; i++) l[i] = (uint8_t)h.lcs.get(i);
        lcs_ = upload(l.data(), l.size());
    }
    {
        std::vector<uint32_t> o(F + 1, 0);
        for (uint64_t i = 0; i < F; i++) {
            const uint64_t v = h.offsets.get(i);
            if (v >= lim) throw limit_error("finimizer offset beyond 32 bits");
            o[i] = (uint32_t)v;
        }
        offsets_ = upload(o.data(), o.size());
    }
    {
        const uint64_t nw = (T + 31) / 32 + 2;
        std::vector<uint64_t> t(nw, 0);
        memcpy(t.data(), h.text.words.data(), std::min<uint64_t>(nw, h.text.words.size()) * 8);
        text_ = upload(t.data(), t.size());
    }

    view_.matrix = reinterpret_cast<const ulonglong2*>(matrix_.p);
    view_.ustart = reinterpret_cast<const ulonglong2*>(ustart_.p);
    view_.fmin = reinterpret_cast<const ulonglong2*>(fmin_.p);
    view_.tstart = reinterpret_cast<const ulonglong2*>(tstart_.p);
    view_.lcs = lcs_.p;
    view_.offsets = offsets_.p;
    view_.ends = ends_.p;
    view_.text = text_.p;
    for (int c = 0; c < 4; c++) view_.C[c] = (idx_t)h.sbwt.C[c];
    view_.n_nodes = (idx_t)N;
    view_.text_len = (idx_t)T;
    view_.n_unitigs = (uint32_t)U;
    view_.n_fmin = (uint32_t)F;
    view_.k = (int32_t)h.sbwt.k;

    uint64_t max_len = 0, prev = 0;
    for (uint64_t u = 0; u < U; u++) {
        max_len = std::max<uint64_t>(max_len, ends[u] - prev);
        prev = ends[u];
    }
    fits32_ = U < 0x7FFFFFFFull && max_len < 0x7FFFFFFFull;

    info_.k = h.sbwt.k;
    info_.n_nodes = (int64_t)N;
    info_.n_kmers = h.sbwt.n_kmers;
    info_.n_unitigs = (int64_t)U;
    info_.text_len = (int64_t)T;
    info_.n_finimizers = (int64_t)F;
    info_.device = device;
    info_.device_bytes = (int64_t)(matrix_.bytes() + ustart_.bytes() + fmin_.bytes() + tstart_.bytes() + lcs_.bytes() +
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

template <class HitT, bool COUNT>
static void launch_search_t(const IndexView& v, const uint64_t* packed, const uint64_t* offsets, uint64_t n_reads,
                            const uint64_t* out_offsets, void* out, SearchTotals* totals, cudaStream_t st) {
    const unsigned threads = 128;
    const unsigned blocks = (unsigned)((n_reads + threads - 1) / threads);
    HitPair<HitT>* o = reinterpret_cast<HitPair<HitT>*>(out);
    switch (window_capacity(v.k)) {
        case 32: search_reads_kernel<32, HitT, COUNT><<<blocks, threads, 0, st>>>(v, packed, offsets, n_reads, out_offsets, o, totals); break;
        case 64: search_reads_kernel<64, HitT, COUNT><<<blocks, threads, 0, st>>>(v, packed, offsets, n_reads, out_offsets, o, totals); break;
        case 128: search_reads_kernel<128, HitT, COUNT><<<blocks, threads, 0, st>>>(v, packed, offsets, n_reads, out_offsets, o, totals); break;
        default: search_reads_kernel<256, HitT, COUNT><<<blocks, threads, 0, st>>>(v, packed, offsets, n_reads, out_offsets, o, totals); break;
    }
    g_launches++;
    FIN_CUDA(cudaGetLastError());
}

// pack + k-mer counts + exclusive scan + search on one stream; every pointer is device memory
static void launch_pipeline(const IndexView& v, const uint8_t* d_ascii, uint64_t n_bases, const uint64_t* d_offsets,
                            uint64_t n_reads, uint64_t* d_packed, uint64_t* d_out_offsets, void* d_scan_tmp,
                            size_t scan_tmp_bytes, void* d_out, int hit_bytes, SearchTotals* d_totals, bool count,
                            cudaStream_t st) {
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
    if (hit_bytes == 4) {
        if (count) launch_search_t<int32_t, true>(v, d_packed, d_offsets, n_reads, d_out_offsets, d_out, d_totals, st);
        else launch_search_t<int32_t, false>(v, d_packed, d_offsets, n_reads, d_out_offsets, d_out, d_totals, st);
    } else {
        if (count) launch_search_t<int64_t, true>(v, d_packed, d_offsets, n_reads, d_out_offsets, d_out, d_totals, st);
        else launch_search_t<int64_t, false>(v, d_packed, d_offsets, n_reads, d_out_offsets, d_out, d_totals, st);
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
    cudaEvent_t e0, e1;
    FIN_CUDA(cudaEventCreate(&e0));
    FIN_CUDA(cudaEventCreate(&e1));
    FIN_CUDA(cudaEventRecord(e0, 0));
    launch_pipeline(view_, (const uint8_t*)d_seqs, n_bases, d_read_offsets, n_reads, dev_packed_.p, dev_out_offsets_.p,
                    dev_scan_tmp_.p, tmp, d_out, hit_bytes, (SearchTotals*)dev_totals_.p, count_accesses, 0);
    FIN_CUDA(cudaEventRecord(e1, 0));
    SearchTotals t;
    FIN_CUDA(cudaMemcpy(&t, dev_totals_.p, sizeof t, cudaMemcpyDeviceToHost));
    FIN_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    FIN_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
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
                              uint64_t out_capacity, fin_search_counts* counts) {
    if (sizeof(HitT) == 4 && !fits32_) throw limit_error("this index needs 64-bit hits (use fin_search_batch)");
    FIN_CUDA(cudaSetDevice(device_));
    const int k = view_.k;
    const uint64_t total = fin_count_kmers(read_offsets, n_reads, k);
    if (total > out_capacity) throw arg_error("out_capacity is smaller than the number of k-mer positions of the batch");
    for (uint64_t i = 0; i < n_reads; i++)
        if (read_offsets[i + 1] < read_offsets[i]) throw arg_error("read_offsets must be non-decreasing");

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
        for (uint64_t i = r0; i < r1; i++) {
            const uint64_t len = read_offsets[i + 1] - read_offsets[i];
            if (len >= (uint64_t)k) nkm += len - (uint64_t)k + 1;
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
                        (int)sizeof(HitT), (SearchTotals*)s.totals.p, count_accesses, s.stream);
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

template void DeviceIndex::search_host<int32_t>(const char*, const uint64_t*, uint64_t, int32_t*, uint64_t, fin_search_counts*);
template void DeviceIndex::search_host<int64_t>(const char*, const uint64_t*, uint64_t, int64_t*, uint64_t, fin_search_counts*);

}  // namespace fin
