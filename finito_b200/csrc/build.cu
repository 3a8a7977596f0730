// build-fmin on the device.
//   lcs_round_kernel     : one round of the label-propagation LCS algorithm — update + propagate fused
//                          (include/lcs_basic_parallel_algorithm.hpp:10-27, rounds :80-109)
//   first_kmer_kernel    : SBWT column of every unitig's first k-mer = colex sort key and Ustart mark
//                          (include/PackedStrings.hh:104-137, SBWT.hh:320-347)
//   build_stream_kernel  : FinimizerIndexBuilder::add_sequence (include/FinimizerIndex.hh:321-389) over chunks
//                          of unitigs; the order-dependent overwrite rule (:370-378) becomes one atomicMax
//   finalize kernels     : fmin bit vector + finimizer offsets in column order (FinimizerIndex.hh:301-306)
// Host code around them reproduces FinimizerIndexBuilder's constructor (:273-319), the build_fmin driver
// (include/build_fmin.hh:302-401) and FinimizerIndex::serialize (:187-207).
#include <algorithm>
#include <cstring>
#include <numeric>

#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "engine.hpp"
#include "search_kernels.cuh"
#include "seqio.hpp"

namespace fin {

// =====================================================================================================
// B1  LCS
// =====================================================================================================

// ones per 64-column word of each row -> (after an exclusive scan) destination of the word's first edge
__global__ void row_popcount_kernel(const uint64_t* __restrict__ rows, uint64_t n_words, uint32_t* __restrict__ counts) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;  // over 4 * n_words
    if (i < 4 * n_words) counts[i] = (uint32_t)__popcll(rows[i]);
}

__global__ void lcs_init_kernel(uint8_t* __restrict__ last, uint8_t* __restrict__ lcs, uint64_t n, uint32_t c0, uint32_t c1,
                                uint32_t c2, uint32_t c3, uint8_t k) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // first column from the C array (lcs_basic_parallel_algorithm.hpp:29-50): 0 = '$', 1..4 = A,C,G,T
    last[i] = i < c0 ? 0 : i < c1 ? 1 : i < c2 ? 2 : i < c3 ? 3 : 4;
    lcs[i] = k;  // "not set yet" (:76-78)
}

// One thread owns 64 consecutive columns (one word of each row).
//   update   : lcs[i] = round where still unset and last[i] != last[i-1]            (:21-27)
//   propagate: next[C[c] + rank_c(i)] = last[i] for every set bit (i, c)            (:10-19)
// word_dst[c * n_words + w] = rank_c(64 w), so the edges of a word land on consecutive bytes.
__global__ void __launch_bounds__(128) lcs_round_kernel(const uint64_t* __restrict__ rows, const uint32_t* __restrict__ word_dst,
                                                        const uint8_t* __restrict__ last, uint8_t* __restrict__ next,
                                                        uint8_t* __restrict__ lcs, uint64_t n, uint64_t n_words, uint32_t c0,
                                                        uint32_t c1, uint32_t c2, uint32_t c3, uint8_t k, uint8_t round) {
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint64_t i0 = w * 64;
    if (w == 0) next[0] = 0;  // the root has no incoming edge (:86)
    // buffers are padded to a multiple of 64 bytes, so whole-word vector loads are in bounds
    uint32_t lv[16], cv[16];
    {
        const uint4* lp = reinterpret_cast<const uint4*>(last + i0);
        const uint4* cp = reinterpret_cast<const uint4*>(lcs + i0);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 a = lp[q], b = cp[q];
            lv[4 * q] = a.x; lv[4 * q + 1] = a.y; lv[4 * q + 2] = a.z; lv[4 * q + 3] = a.w;
            cv[4 * q] = b.x; cv[4 * q + 1] = b.y; cv[4 * q + 2] = b.z; cv[4 * q + 3] = b.w;
        }
    }
    uint32_t prev = i0 ? last[i0 - 1] : 0xFFu;  // column 0 always differs from its (absent) predecessor
    bool changed = false;
#pragma unroll
    for (int j = 0; j < 16; j++) {
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const uint32_t ch = (lv[j] >> (8 * b)) & 0xFFu;
            const uint32_t cur = (cv[j] >> (8 * b)) & 0xFFu;
            if (cur == k && ch != prev && i0 + 4 * j + b < n) {
                cv[j] = (cv[j] & ~(0xFFu << (8 * b))) | ((uint32_t)round << (8 * b));
                changed = true;
            }
            prev = ch;
        }
    }
    if (changed) {
        uint4* cp = reinterpret_cast<uint4*>(lcs + i0);
#pragma unroll
        for (int q = 0; q < 4; q++) cp[q] = make_uint4(cv[4 * q], cv[4 * q + 1], cv[4 * q + 2], cv[4 * q + 3]);
    }
    const uint32_t base[4] = {c0, c1, c2, c3};
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint64_t bits = rows[(uint64_t)c * n_words + w];
        if (!bits) continue;
        uint8_t* dst = next + base[c] + word_dst[(uint64_t)c * n_words + w];
#pragma unroll
        for (int j = 0; j < 16; j++) {
#pragma unroll
            for (int b = 0; b < 4; b++) {
                if ((bits >> (4 * j + b)) & 1ULL) *dst++ = (uint8_t)((lv[j] >> (8 * b)) & 0xFFu);
            }
        }
    }
}

struct SbwtRowsDevice {  // the four rows as plain words, row-major: rows[c * n_words + w]
    DevBuf<uint64_t> rows;
    uint64_t n_words = 0;
};

static SbwtRowsDevice upload_rows(const SbwtHost& s) {
    SbwtRowsDevice d;
    const uint64_t N = (uint64_t)s.n_nodes;
    d.n_words = (N + 63) / 64;
    std::vector<uint64_t> h(4 * d.n_words, 0);
    const uint64_t tail = N & 63;
    for (int c = 0; c < 4; c++) {
        memcpy(&h[(uint64_t)c * d.n_words], s.bits[c].words.data(), d.n_words * 8);
        if (tail) h[(uint64_t)c * d.n_words + d.n_words - 1] &= (1ULL << tail) - 1;
    }
    d.rows.alloc(h.size());
    FIN_CUDA(cudaMemcpy(d.rows.p, h.data(), h.size() * 8, cudaMemcpyHostToDevice));
    return d;
}

static void check_sbwt_limits(const SbwtHost& s) {
    if ((uint64_t)s.n_nodes >= 0xFFFFFFFFull - 256) throw limit_error("SBWT too large for the 32-bit columns of this build");
    if (s.n_nodes < 2) throw domain_error("SBWT with fewer than 2 columns");
}

// LCS bytes on the device (padded to a multiple of 64 + 64)
static DevBuf<uint8_t> lcs_on_device(const SbwtHost& s, const SbwtRowsDevice& rows) {
    const uint64_t N = (uint64_t)s.n_nodes, nw = rows.n_words;
    const uint64_t padded = nw * 64 + 64;
    DevBuf<uint8_t> last(padded), next(padded), lcs(padded);
    DevBuf<uint32_t> word_dst(4 * nw);
    const unsigned T = 256;
    row_popcount_kernel<<<(unsigned)((4 * nw + T - 1) / T), T>>>(rows.rows.p, nw, word_dst.p);
    g_launches++;
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, word_dst.p, word_dst.p, (int64_t)nw);
    DevBuf<uint8_t> tmp(tmp_bytes + 256);
    for (int c = 0; c < 4; c++) {
        size_t tb = tmp_bytes;
        FIN_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, word_dst.p + (uint64_t)c * nw, word_dst.p + (uint64_t)c * nw, (int64_t)nw));
        g_launches++;
    }
    const uint32_t c0 = (uint32_t)s.C[0], c1 = (uint32_t)s.C[1], c2 = (uint32_t)s.C[2], c3 = (uint32_t)s.C[3];
    FIN_CUDA(cudaMemset(last.p, 0, padded));
    FIN_CUDA(cudaMemset(next.p, 0, padded));
    FIN_CUDA(cudaMemset(lcs.p, 0, padded));
    lcs_init_kernel<<<(unsigned)((N + T - 1) / T), T>>>(last.p, lcs.p, N, c0, c1, c2, c3, (uint8_t)s.k);
    g_launches++;
    uint8_t* a = last.p;
    uint8_t* b = next.p;
    for (int64_t round = 0; round < s.k; round++) {
        lcs_round_kernel<<<(unsigned)((nw + 127) / 128), 128>>>(rows.rows.p, word_dst.p, a, b, lcs.p, N, nw, c0, c1, c2, c3,
                                                                (uint8_t)s.k, (uint8_t)round);
        g_launches++;
        std::swap(a, b);
    }
    FIN_CUDA(cudaGetLastError());
    FIN_CUDA(cudaDeviceSynchronize());
    return lcs;
}

static void select_device(int device) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) throw cuda_error("no CUDA device is visible");
    if (device < 0 || device >= n_dev) throw arg_error("CUDA device ordinal out of range");
    FIN_CUDA(cudaSetDevice(device));
}

std::vector<uint8_t> lcs_device(const SbwtHost& s, int device) {
    select_device(device);
    check_sbwt_limits(s);
    SbwtRowsDevice rows = upload_rows(s);
    DevBuf<uint8_t> d = lcs_on_device(s, rows);
    std::vector<uint8_t> h((size_t)s.n_nodes);
    FIN_CUDA(cudaMemcpy(h.data(), d.p, h.size(), cudaMemcpyDeviceToHost));
    return h;
}

// =====================================================================================================
// B2  first k-mer of every unitig -> SBWT column
// =====================================================================================================
struct BuildView {  // the part of IndexView the build kernels need
    const ulonglong2* __restrict__ matrix;
    const uint8_t* __restrict__ lcs;
    idx_t C[4];
    idx_t n_nodes;
    int32_t k;
};

__device__ __forceinline__ IndexView as_index_view(const BuildView& b) {
    IndexView v{};
    v.matrix = b.matrix;
    v.lcs = b.lcs;
    for (int c = 0; c < 4; c++) v.C[c] = b.C[c];
    v.n_nodes = b.n_nodes;
    v.k = b.k;
    return v;
}

__device__ __forceinline__ int text_base(const uint64_t* __restrict__ text, uint64_t pos) {
    return (int)((__ldg(text + (pos >> 5)) >> (2 * (pos & 31))) & 3);
}

// columns[u] = column of text[starts[u] .. starts[u]+k), 0xFFFFFFFF when the k-mer is not a single SBWT column
__global__ void first_kmer_kernel(const BuildView bv, const uint64_t* __restrict__ text, const uint64_t* __restrict__ starts,
                                  uint32_t n_unitigs, uint32_t* __restrict__ columns) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_unitigs) return;
    const IndexView ix = as_index_view(bv);
    const uint64_t s = starts[u];
    Ival I{0, ix.n_nodes - 1};
    for (int i = 0; i < ix.k && !is_empty(I); i++) I = extend(ix, I, text_base(text, s + (uint64_t)i));
    columns[u] = (!is_empty(I) && I.l == I.r) ? I.l : 0xFFFFFFFFu;
}

// =====================================================================================================
// B3  add_sequence over chunks of unitigs
// =====================================================================================================
struct BuildChunk {
    uint32_t text_begin;   // global text position of the first streamed base (inside the unitig)
    uint32_t first_pos;    // position of that base in its unitig
    uint32_t n_stream;     // bases to stream
    uint32_t first_emit;   // unitig-local END position of the first k-mer this chunk reports
    uint32_t unitig_start; // global text position of the unitig
};

// table[colex]: 0 = unset; (end_local << 32) | (0xFFFFFFFF - global_off) for end_local > 0; global_off + 1 otherwise.
// atomicMax over it is the closed form of the overwrite rule (FinimizerIndex.hh:370-378): the winner is the
// first event (= smallest global offset) among those with the largest unitig-local end, and with all ends
// equal to 0 — where the reference's "0 means unset" makes every event overwrite — the last one.
__device__ __forceinline__ unsigned long long fmin_key(uint32_t end_local, uint32_t global_off) {
    return end_local ? (((unsigned long long)end_local << 32) | (unsigned long long)(0xFFFFFFFFu - global_off))
                     : (unsigned long long)global_off + 1ULL;
}

template <int KCAP>
__global__ void __launch_bounds__(128) build_stream_kernel(const BuildView bv, const uint64_t* __restrict__ text,
                                                           const BuildChunk* __restrict__ chunks, uint32_t n_chunks,
                                                           unsigned long long* __restrict__ table,
                                                           unsigned int* __restrict__ error_flags) {
    const uint32_t cid = blockIdx.x * blockDim.x + threadIdx.x;
    if (cid >= n_chunks) return;
    const IndexView ix = as_index_view(bv);
    const BuildChunk ch = chunks[cid];
    const int k = ix.k;
    const idx_t N = ix.n_nodes;
    uint64_t dq[KCAP];
    int dq_head = 0, dq_n = 0;
    unsigned dummy = 0, err = 0;
    Ival I{0, N - 1};
    // positions are unitig-local; `start` is the left end of the current finimizer window
    int64_t start = ch.first_pos;
    unsigned long long last_key = 0;  // consecutive windows mostly share their finimizer: one atomic per change
    idx_t last_colex = 0;
    bool have_last = false;
    uint64_t word = 0;
    for (uint32_t s = 0; s < ch.n_stream; s++) {
        const uint64_t tp = (uint64_t)ch.text_begin + s;
        if (s == 0 || (tp & 31) == 0) word = __ldg(text + (tp >> 5));
        const int c = (int)((word >> (2 * (tp & 31))) & 3);
        const int64_t end = (int64_t)ch.first_pos + s;
        I = extend(ix, I, c);  // :346, unitigs are in the SBWT so this cannot become empty
        if (is_empty(I)) {
            err |= kErrLookupRange;
            break;
        }
        const int64_t kmer = end - (k - 1);  // start of the window that ends here
        if (I.l == I.r) {  // :349-366
            uint64_t cur;
            do {
                // candidate positions are kept relative to the chunk so that they fit the 24-bit field
                cur = cand_pack((int)(end - start + 1), I.l, (int)s);
                start++;
                I = drop_first_char<false>(ix, (int)(end - start + 1), I, dummy);
            } while (I.l == I.r);
            while (dq_n > 0 && (int64_t)ch.first_pos + cand_start(dq[dq_head]) < kmer) {
                dq_head = (dq_head + 1) & (KCAP - 1);
                dq_n--;
            }
            while (dq_n > 0 && dq[(dq_head + dq_n - 1) & (KCAP - 1)] > cur) dq_n--;
            if (dq_n == KCAP) {
                err |= kErrWindowUnderflow;
                dq_n--;
            }
            dq[(dq_head + dq_n) & (KCAP - 1)] = cur;
            dq_n++;
        }
        if (end >= (int64_t)ch.first_emit) {
            while (dq_n > 0 && (int64_t)ch.first_pos + cand_start(dq[dq_head]) < kmer) {
                dq_head = (dq_head + 1) & (KCAP - 1);
                dq_n--;
            }
            if (dq_n == 0) {  // the reference would record its sentinel here (:328,:384): not a DSPSS of this SBWT
                err |= kErrWindowUnderflow;
            } else {
                const uint64_t w = dq[dq_head];
                const uint32_t end_local = ch.first_pos + (uint32_t)cand_end(w);
                const unsigned long long key = fmin_key(end_local, ch.unitig_start + end_local);
                const idx_t colex = cand_colex(w);
                if (!have_last || colex != last_colex || key != last_key) {
                    atomicMax(table + colex, key);
                    last_key = key;
                    last_colex = colex;
                    have_last = true;
                }
            }
        }
    }
    if (err) atomicOr(error_flags, err);
}

// =====================================================================================================
// finalize: fmin bits, offsets in column order
// =====================================================================================================
__global__ void fmin_bits_kernel(const unsigned long long* __restrict__ table, uint64_t n, uint32_t* __restrict__ bits32,
                                 uint32_t* __restrict__ offs_or_sentinel) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long key = i < n ? table[i] : 0ULL;
    const unsigned mask = __ballot_sync(0xFFFFFFFFu, key != 0ULL);
    if ((threadIdx.x & 31) == 0 && i < ((n + 31) / 32) * 32) bits32[i >> 5] = mask;
    if (i < n) {
        uint32_t off = 0xFFFFFFFFu;
        if (key) off = (key >> 32) ? 0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFu) : (uint32_t)(key - 1);
        offs_or_sentinel[i] = off;
    }
}

struct NotSentinel {
    __host__ __device__ bool operator()(const uint32_t& v) const { return v != 0xFFFFFFFFu; }
};

// =====================================================================================================
// host driver
// =====================================================================================================
static int window_capacity(int k) { return k <= 32 ? 32 : k <= 64 ? 64 : k <= 128 ? 128 : 256; }

static std::string float_to_string(float x) {  // std::to_string(float)
    char buf[64];
    snprintf(buf, sizeof buf, "%f", (double)x);
    return buf;
}

struct RarestResult {
    IndexHost index;
};

static void build_rarest(SbwtHost&& sbwt, PackedVec&& lcs_file, const SeqBatch& unitigs, const std::string& out_prefix) {
    const int k = (int)sbwt.k;
    const uint64_t N = (uint64_t)sbwt.n_nodes, U = unitigs.size();
    if (U == 0) throw domain_error("no unitigs in the input");
    if (lcs_file.size() != N) throw format_error("LCS length differs from the number of SBWT columns");
    const uint64_t T = unitigs.seqs.size();
    if (T >= 0xFFFFFFFFull - 256) throw limit_error("unitig text too large for the 32-bit positions of this build");

    // ---- device copy of the SBWT rank sectors and the LCS bytes
    const uint64_t* rows[4] = {sbwt.bits[0].words.data(), sbwt.bits[1].words.data(), sbwt.bits[2].words.data(),
                               sbwt.bits[3].words.data()};
    std::vector<uint64_t> sectors = make_rank_sectors(rows, 4, N, nullptr);
    DevBuf<uint64_t> d_matrix(sectors.size());
    FIN_CUDA(cudaMemcpy(d_matrix.p, sectors.data(), sectors.size() * 8, cudaMemcpyHostToDevice));
    std::vector<uint64_t>().swap(sectors);
    DevBuf<uint8_t> d_lcs(N + 64);
    {
        std::vector<uint8_t> l(N + 64, 0);
        for (uint64_t i = 0; i < N; i++) l[i] = (uint8_t)lcs_file.get(i);
        FIN_CUDA(cudaMemcpy(d_lcs.p, l.data(), l.size(), cudaMemcpyHostToDevice));
    }
    BuildView bv{};
    bv.matrix = reinterpret_cast<const ulonglong2*>(d_matrix.p);
    bv.lcs = d_lcs.p;
    for (int c = 0; c < 4; c++) bv.C[c] = (idx_t)sbwt.C[c];
    bv.n_nodes = (idx_t)N;
    bv.k = k;

    // ---- pack the unitigs in file order (PackedStrings.hh:53-57) and find the column of each first k-mer
    auto pack = [&](const std::vector<uint64_t>& order, PackedVec& text, std::vector<uint64_t>& starts) {
        text.assign(T, 2);
        starts.resize(U);
        uint64_t pos = 0;
        for (uint64_t j = 0; j < U; j++) {
            const uint64_t u = order[j];
            const uint64_t b = unitigs.offsets[u], e = unitigs.offsets[u + 1];
            starts[j] = pos;
            for (uint64_t i = b; i < e; i++) {
                int code;
                switch (unitigs.seqs[i]) {
                    case 'A': code = 0; break;
                    case 'C': code = 1; break;
                    case 'G': code = 2; break;
                    case 'T': code = 3; break;
                    default: throw domain_error("Invalid character");  // PackedStrings.hh:57
                }
                text.words[pos >> 5] |= (uint64_t)code << (2 * (pos & 31));
                pos++;
            }
        }
    };
    for (uint64_t u = 0; u < U; u++)
        if (unitigs.offsets[u + 1] - unitigs.offsets[u] < (uint64_t)k)
            throw domain_error("a unitig is shorter than k (the reference reads out of bounds there, PackedStrings.hh:116)");

    std::vector<uint64_t> order(U), starts;
    std::iota(order.begin(), order.end(), 0);
    PackedVec text;
    pack(order, text, starts);
    DevBuf<uint64_t> d_text(text.words.size() + 1);
    FIN_CUDA(cudaMemset(d_text.p, 0, d_text.bytes()));
    FIN_CUDA(cudaMemcpy(d_text.p, text.words.data(), text.words.size() * 8, cudaMemcpyHostToDevice));
    DevBuf<uint64_t> d_starts(U);
    FIN_CUDA(cudaMemcpy(d_starts.p, starts.data(), U * 8, cudaMemcpyHostToDevice));
    DevBuf<uint32_t> d_cols(U);
    first_kmer_kernel<<<(unsigned)((U + 127) / 128), 128>>>(bv, d_text.p, d_starts.p, (uint32_t)U, d_cols.p);
    g_launches++;
    FIN_CUDA(cudaGetLastError());
    std::vector<uint32_t> cols(U);
    FIN_CUDA(cudaMemcpy(cols.data(), d_cols.p, U * 4, cudaMemcpyDeviceToHost));

    // ---- permute_unitigs: colex order of the first k-mers == order of their SBWT columns (PackedStrings.hh:121)
    PackedVec ustart;
    ustart.assign(N, 1);
    for (uint64_t u = 0; u < U; u++) {
        if (cols[u] == 0xFFFFFFFFu) throw domain_error("the first k-mer of a unitig is not in the SBWT (PackedStrings.hh:129-131)");
        ustart.words[cols[u] >> 6] |= 1ULL << (cols[u] & 63);
    }
    std::sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) { return cols[a] != cols[b] ? cols[a] < cols[b] : a < b; });
    pack(order, text, starts);
    FIN_CUDA(cudaMemcpy(d_text.p, text.words.data(), text.words.size() * 8, cudaMemcpyHostToDevice));
    PackedVec ends;
    ends.assign(U, 64);  // width stays 64 (PackedStrings.hh:44 passes the bit length as the default VALUE)
    for (uint64_t j = 0; j < U; j++) ends.words[j] = starts[j] + (unitigs.offsets[order[j] + 1] - unitigs.offsets[order[j]]);

    // ---- chunks of k-mer windows, each with 2k bases of warm-up context (DESIGN.md "chunked streams")
    uint64_t total_windows = 0;
    for (uint64_t j = 0; j < U; j++) total_windows += ends.words[j] - starts[j] - (uint64_t)k + 1;
    uint64_t per_chunk = std::min<uint64_t>(1024, std::max<uint64_t>(64, total_windows / (148ull * 1024)));
    const uint64_t warm = 2 * (uint64_t)k;
    std::vector<BuildChunk> chunks;
    for (uint64_t j = 0; j < U; j++) {
        const uint64_t len = ends.words[j] - starts[j], nwin = len - (uint64_t)k + 1;
        for (uint64_t w0 = 0; w0 < nwin; w0 += per_chunk) {
            const uint64_t w1 = std::min(nwin, w0 + per_chunk);
            const uint64_t first = w0 > warm ? w0 - warm : 0;  // first streamed unitig position
            BuildChunk c;
            c.first_pos = (uint32_t)first;
            c.text_begin = (uint32_t)(starts[j] + first);
            c.n_stream = (uint32_t)(w1 + (uint64_t)k - 1 - first);
            c.first_emit = (uint32_t)(w0 + (uint64_t)k - 1);
            c.unitig_start = (uint32_t)starts[j];
            if (c.n_stream >= (1u << 24)) throw limit_error("internal: build chunk too long");
            chunks.push_back(c);
        }
    }
    DevBuf<BuildChunk> d_chunks(chunks.size());
    FIN_CUDA(cudaMemcpy(d_chunks.p, chunks.data(), chunks.size() * sizeof(BuildChunk), cudaMemcpyHostToDevice));
    DevBuf<unsigned long long> d_table(N);
    FIN_CUDA(cudaMemset(d_table.p, 0, d_table.bytes()));
    DevBuf<unsigned int> d_err(1);
    FIN_CUDA(cudaMemset(d_err.p, 0, 4));
    {
        const unsigned blocks = (unsigned)((chunks.size() + 127) / 128);
        const uint32_t nc = (uint32_t)chunks.size();
        switch (window_capacity(k)) {
            case 32: build_stream_kernel<32><<<blocks, 128>>>(bv, d_text.p, d_chunks.p, nc, d_table.p, d_err.p); break;
            case 64: build_stream_kernel<64><<<blocks, 128>>>(bv, d_text.p, d_chunks.p, nc, d_table.p, d_err.p); break;
            case 128: build_stream_kernel<128><<<blocks, 128>>>(bv, d_text.p, d_chunks.p, nc, d_table.p, d_err.p); break;
            default: build_stream_kernel<256><<<blocks, 128>>>(bv, d_text.p, d_chunks.p, nc, d_table.p, d_err.p); break;
        }
        g_launches++;
        FIN_CUDA(cudaGetLastError());
    }
    unsigned err = 0;
    FIN_CUDA(cudaMemcpy(&err, d_err.p, 4, cudaMemcpyDeviceToHost));
    if (err)
        throw domain_error("the unitigs are not a disjoint spectrum-preserving string set of this SBWT "
                           "(a unitig k-mer is missing from the SBWT or a window has no unique substring)");

    // ---- fmin bit vector and offsets in column order (FinimizerIndex.hh:301-306)
    const uint64_t n_words32 = (N + 31) / 32;
    DevBuf<uint32_t> d_bits(n_words32 + 2), d_offs(N), d_sel(N);
    DevBuf<uint64_t> d_nsel(1);
    FIN_CUDA(cudaMemset(d_bits.p, 0, d_bits.bytes()));
    {
        const uint64_t threads_total = n_words32 * 32;
        fmin_bits_kernel<<<(unsigned)((threads_total + 255) / 256), 256>>>(d_table.p, N, d_bits.p, d_offs.p);
        g_launches++;
        FIN_CUDA(cudaGetLastError());
    }
    {
        size_t tmp_bytes = 0;
        cub::DeviceSelect::If(nullptr, tmp_bytes, d_offs.p, d_sel.p, d_nsel.p, (int64_t)N, NotSentinel());
        DevBuf<uint8_t> tmp(tmp_bytes + 256);
        FIN_CUDA(cub::DeviceSelect::If(tmp.p, tmp_bytes, d_offs.p, d_sel.p, d_nsel.p, (int64_t)N, NotSentinel()));
        g_launches += 2;
    }
    uint64_t n_fmin = 0;
    FIN_CUDA(cudaMemcpy(&n_fmin, d_nsel.p, 8, cudaMemcpyDeviceToHost));
    std::vector<uint32_t> offs(n_fmin);
    if (n_fmin) FIN_CUDA(cudaMemcpy(offs.data(), d_sel.p, n_fmin * 4, cudaMemcpyDeviceToHost));
    PackedVec fmin;
    fmin.assign(N, 1);
    {
        std::vector<uint32_t> b32(n_words32 + 2, 0);
        FIN_CUDA(cudaMemcpy(b32.data(), d_bits.p, n_words32 * 4, cudaMemcpyDeviceToHost));
        memcpy(fmin.words.data(), b32.data(), std::min<size_t>(n_words32 * 4, fmin.words.size() * 8));
    }
    uint64_t max_off = 0;
    for (uint32_t v : offs) max_off = std::max<uint64_t>(max_off, v);
    PackedVec offsets;
    offsets.assign(n_fmin, (uint8_t)bit_length(max_off));  // :301 width from the largest offset
    if (offsets.width == 0) {  // all offsets 0: sdsl stores width 0 as ... not reachable with k >= 1 unitigs of length >= k > 1
        offsets.width = 1;
        offsets.n_bits = n_fmin;
        offsets.words.assign(offsets.disk_words() + 1, 0);
    }
    for (uint64_t i = 0; i < n_fmin; i++) offsets.put(i, offs[i]);

    IndexHost ix;
    ix.sbwt = std::move(sbwt);
    ix.lcs = std::move(lcs_file);
    ix.fmin = std::move(fmin);
    ix.offsets = std::move(offsets);
    ix.text = std::move(text);
    ix.ends = std::move(ends);
    ix.ustart = std::move(ustart);
    save_index_files(ix, out_prefix);
}

std::string build_fmin_device(const std::string& sbwt_path, const std::string& unitig_path, const std::string& out_prefix,
                              const std::string& type, int64_t t_arg, const std::string& lcs_path, int device) {
    const signed char t = (signed char)t_arg;  // build_fmin.hh:333 narrows to char
    std::vector<std::string> files;
    if (has_suffix(unitig_path, ".txt")) files = read_lines(unitig_path);  // :338
    else files.push_back(unitig_path);
    for (const std::string& f : files) {
        FILE* fp = fopen(f.c_str(), "rb");
        if (!fp) throw io_error("Error: cannot read file " + f);
        fclose(fp);
    }
    SbwtHost sbwt = load_sbwt_file(sbwt_path, true);
    if (type != "rarest" && type != "shortest" && type != "verify") throw arg_error("Error: unknown type: " + type);
    select_device(device);
    check_sbwt_limits(sbwt);

    std::string lcs_file = lcs_path;
    if (lcs_file.empty()) {  // :373-380: compute, save, then load it back like the reference does
        lcs_file = out_prefix + ".LCS.sdsl";
        SbwtRowsDevice rows = upload_rows(sbwt);
        DevBuf<uint8_t> d = lcs_on_device(sbwt, rows);
        std::vector<uint8_t> h((size_t)sbwt.n_nodes);
        FIN_CUDA(cudaMemcpy(h.data(), d.p, h.size(), cudaMemcpyDeviceToHost));
        PackedVec v;
        v.assign((uint64_t)sbwt.n_nodes, (uint8_t)bit_length((uint64_t)(sbwt.k - 1)));  // lcs_basic_parallel_algorithm.hpp:115
        if (v.width == 0) throw domain_error("k = 1 is outside the parity domain");
        for (uint64_t i = 0; i < h.size(); i++) v.put(i, h[i]);
        save_packed_file(lcs_file, v, true);
    }
    PackedVec lcs = load_packed_file(lcs_file, 0);

    std::string result;
    if (files.empty()) throw arg_error("no unitig file given");
    if (files.size() > 1)
        throw domain_error("more than one unitig file: the reference moves its SBWT into the first builder (build_fmin.hh:281-300)");
    SeqBatch unitigs;
    append_sequences(files[0], unitigs);
    if (type == "rarest") {
        if (t != 1) throw domain_error("t != 1 does not make sense with rarest type");  // :245-247
        build_rarest(std::move(sbwt), std::move(lcs), unitigs, out_prefix);
        result = "";
    } else {
        throw arg_error("--type " + type + ": the stats-only modes are not built yet in this library");
    }
    const std::string stats_path = out_prefix + "_stats.txt";  // :386-395
    FILE* sf = fopen(stats_path.c_str(), "ab");
    if (!sf) throw io_error("Error: Unable to open file.");
    const std::string line = std::to_string((int)t) + "," + result + "\n";
    fwrite(line.data(), 1, line.size(), sf);
    fclose(sf);
    (void)float_to_string;
    return result;
}

}  // namespace fin
