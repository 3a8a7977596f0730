// Kernels of the build-fmin path.  The per-thread bodies are FIN_D functions so that tests/emu can run the
// same source on the host; build.cu holds the host driver.
//   lcs_round_warp_kernel: one round of the label-propagation LCS algorithm — update + propagate fused, one warp per
//                          2048 columns, edges compacted in shared memory and written out as 16-byte stores
//                          (include/lcs_basic_parallel_algorithm.hpp:10-27, rounds :80-109); lcs_round_word is the
//                          same round for one 64-column word, the form tests/emu runs on the host
//   first_kmer_kernel    : SBWT column of every unitig's first k-mer = colex sort key and Ustart mark
//                          (include/PackedStrings.hh:104-137, SBWT.hh:320-347)
//   build_stream_kernel  : FinimizerIndexBuilder::add_sequence (include/FinimizerIndex.hh:321-389) over chunks
//                          of unitigs; the order-dependent overwrite rule (:370-378) becomes one atomicMax
#pragma once
#include "search_kernels.cuh"
#include "build_host.hpp"  // BuildChunk

namespace fin {

#ifdef FIN_EMU
inline void fin_atomic_max(unsigned long long* p, unsigned long long v) { if (*p < v) *p = v; }
inline void fin_atomic_or(unsigned int* p, unsigned int v) { *p |= v; }
#else
FIN_D void fin_atomic_max(unsigned long long* p, unsigned long long v) { atomicMax(p, v); }
FIN_D void fin_atomic_or(unsigned int* p, unsigned int v) { atomicOr(p, v); }
#endif

// =====================================================================================================
// B1  LCS
// =====================================================================================================
struct LcsArgs {
    const uint64_t* __restrict__ rows;      // rows[c * n_words + w]: word w of row c, bits past n cleared
    const uint32_t* __restrict__ word_dst;  // word_dst[c * n_words + w] = rank_c(64 w)
    uint64_t n, n_words;
    uint32_t C[4];
    uint8_t k;
};

// first column from the C array (lcs_basic_parallel_algorithm.hpp:29-50): 0 = '$', 1..4 = A,C,G,T
FIN_D uint8_t lcs_first_column(const LcsArgs& a, uint64_t i) {
    return i < a.C[0] ? 0 : i < a.C[1] ? 1 : i < a.C[2] ? 2 : i < a.C[3] ? 3 : 4;
}

// One thread owns 64 consecutive columns (one word of each row).
//   update   : lcs[i] = round where still unset (== k) and last[i] != last[i-1]     (:21-27)
//   propagate: next[C[c] + rank_c(i)] = last[i] for every set bit (i, c)            (:10-19)
// The edges of a word land on consecutive bytes of `next`.  All three byte arrays are padded to
// 64 * n_words + 64 bytes, so the whole-word vector accesses stay in bounds.
FIN_D void lcs_round_word(const LcsArgs& a, const uint8_t* __restrict__ last, uint8_t* __restrict__ next,
                          uint8_t* __restrict__ lcs, const uint64_t w, const uint8_t round) {
    const uint64_t i0 = w * 64;
    if (w == 0) next[0] = 0;  // the root has no incoming edge (:86)
    uint32_t lv[16], cv[16];
    {
        const uint4* lp = reinterpret_cast<const uint4*>(last + i0);
        const uint4* cp = reinterpret_cast<const uint4*>(lcs + i0);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 x = lp[q], y = cp[q];
            lv[4 * q] = x.x; lv[4 * q + 1] = x.y; lv[4 * q + 2] = x.z; lv[4 * q + 3] = x.w;
            cv[4 * q] = y.x; cv[4 * q + 1] = y.y; cv[4 * q + 2] = y.z; cv[4 * q + 3] = y.w;
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
            if (cur == a.k && ch != prev && i0 + 4 * j + b < a.n) {
                cv[j] = (cv[j] & ~(0xFFu << (8 * b))) | ((uint32_t)round << (8 * b));
                changed = true;
            }
            prev = ch;
        }
    }
    if (changed) {
        uint4* cp = reinterpret_cast<uint4*>(lcs + i0);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint4 y;
            y.x = cv[4 * q]; y.y = cv[4 * q + 1]; y.z = cv[4 * q + 2]; y.w = cv[4 * q + 3];
            cp[q] = y;
        }
    }
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint64_t bits = a.rows[(uint64_t)c * a.n_words + w];
        if (!bits) continue;
        uint8_t* dst = next + a.C[c] + a.word_dst[(uint64_t)c * a.n_words + w];
#pragma unroll
        for (int j = 0; j < 16; j++) {
#pragma unroll
            for (int b = 0; b < 4; b++) {
                if ((bits >> (4 * j + b)) & 1ULL) *dst++ = (uint8_t)((lv[j] >> (8 * b)) & 0xFFu);
            }
        }
    }
}

// =====================================================================================================
// B2  first k-mer of every unitig -> SBWT column
// =====================================================================================================
FIN_D int text_base(const uint64_t* __restrict__ text, uint64_t pos) {
    return (int)((__ldg(text + (pos >> 5)) >> (2 * (pos & 31))) & 3);
}

// column of text[s .. s+k), 0xFFFFFFFF when the k-mer is not a single SBWT column (SBWT.hh:320-347)
FIN_D uint32_t first_kmer_column(const IndexView& ix, const uint64_t* __restrict__ text, const uint64_t s) {
    Ival I{0, ix.n_nodes - 1};
    for (int i = 0; i < ix.k && !is_empty(I); i++) I = extend(ix, I, text_base(text, s + (uint64_t)i));
    return (!is_empty(I) && I.l == I.r) ? I.l : 0xFFFFFFFFu;
}

// =====================================================================================================
// B3  add_sequence over chunks of unitigs
// =====================================================================================================

// table[colex]: 0 = unset; (end_local << 32) | (0xFFFFFFFF - global_off) for end_local > 0; global_off + 1 otherwise.
// A max over it is the closed form of the overwrite rule (FinimizerIndex.hh:370-378): the winner is the
// first event (= smallest global offset) among those with the largest unitig-local end, and with all ends
// equal to 0 — where the reference's "0 means unset" makes every event overwrite — the last one.
FIN_D unsigned long long fmin_key(uint32_t end_local, uint32_t global_off) {
    return end_local ? (((unsigned long long)end_local << 32) | (unsigned long long)(0xFFFFFFFFu - global_off))
                     : (unsigned long long)global_off + 1ULL;
}
FIN_D uint32_t fmin_key_offset(unsigned long long key) {  // key != 0
    return (key >> 32) ? 0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFu) : (uint32_t)(key - 1);
}

template <int KCAP>
FIN_D void build_stream_chunk(const IndexView& ix, const uint64_t* __restrict__ text, const BuildChunk ch,
                              unsigned long long* __restrict__ table, unsigned int* __restrict__ error_flags) {
    const int k = ix.k;
    const idx_t N = ix.n_nodes;
    uint64_t dq[KCAP];
    int dq_head = 0, dq_n = 0;
    unsigned dummy = 0, err = 0;
    Ival I{0, N - 1};
    int64_t start = ch.first_pos;     // unitig-local left end of the current finimizer window
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
        if (I.l == I.r) {                    // :349-366
            uint64_t cur;
            do {
                // candidate ends are kept relative to the chunk so that they fit the 24-bit field
                cur = cand_pack((int)(end - start + 1), I.l, (int)s);
                start++;
                I = drop_first_char<(KCAP > 128), false>(ix, (int)(end - start + 1), I, dummy);
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
                    fin_atomic_max(table + colex, key);
                    last_key = key;
                    last_colex = colex;
                    have_last = true;
                }
            }
        }
    }
    if (err) fin_atomic_or(error_flags, err);
}

// =====================================================================================================
// B5  build_shortest_streaming_search (include/build_fmin.hh:134-200) over chunks — statistics only
// =====================================================================================================
// Candidates are triggered by freq <= t instead of freq == 1 and ordered by (len, freq, colex, end) (:161).  Packed
// so that the unsigned order is that tuple order: len:8 | freq:8 | colex:32 | end:16 (end relative to the chunk,
// < 2^16 by the chunk plan; freq <= t <= 127 because the reference narrows t to char, :333).  The reference's
// clear()/pop_back()/pop_front() dance (:168-176, :192-195) is a monotone window minimum whenever every window
// holds a candidate, which is the case for unitigs of the SBWT (DESIGN.md "stats modes").
FIN_D uint64_t scand_pack(int len, uint32_t freq, idx_t colex, uint32_t end_rel) {
    return ((uint64_t)(uint32_t)len << 56) | ((uint64_t)freq << 48) | ((uint64_t)colex << 16) | (uint64_t)end_rel;
}
FIN_D int scand_len(uint64_t c) { return (int)(c >> 56); }
FIN_D idx_t scand_colex(uint64_t c) { return (idx_t)(c >> 16); }
FIN_D int scand_end(uint64_t c) { return (int)(c & 0xFFFFu); }
FIN_D int scand_start(uint64_t c) { return scand_end(c) - scand_len(c) + 1; }
// the (len, freq, colex) tuple of the statistics set (:183), as one sortable key
FIN_D uint64_t stat_key(uint64_t c) { return c >> 16; }
constexpr uint64_t kNoStatKey = ~0ULL;

// The insertion rule (:182-187) depends on the order of the windows: (len,freq,colex) is inserted while
// fmin_found[colex] is unset, and fmin_found[colex] is set by the first window whose finimizer ends at or after
// position k-1 of its unitig.  Closed form: with order(window) = global text position of the window in FILE order,
// a window inserts its tuple iff order <= first_flag[colex] = min order over the flagged windows of that colex.
//   PASS 0: first_flag[colex] = atomicMin(order) over flagged windows; runs per chunk counted
//   PASS 1: one key per run of equal window minima: the tuple if it is inserted, kNoStatKey otherwise
// Returns the number of runs of the chunk.
#ifdef FIN_EMU
inline void fin_atomic_min(uint32_t* p, uint32_t v) { if (*p > v) *p = v; }
#else
FIN_D void fin_atomic_min(uint32_t* p, uint32_t v) { atomicMin(p, v); }
#endif

template <int KCAP, int PASS>
FIN_D uint32_t shortest_stream_chunk(const IndexView& ix, const uint64_t* __restrict__ text, const BuildChunk ch, const uint32_t t,
                                     uint32_t* __restrict__ first_flag, uint64_t* __restrict__ out_keys,
                                     unsigned int* __restrict__ error_flags) {
    const int k = ix.k;
    const idx_t N = ix.n_nodes;
    uint64_t dq[KCAP];
    int dq_head = 0, dq_n = 0;
    unsigned dummy = 0, err = 0;
    Ival I{0, N - 1};
    int64_t start = ch.first_pos;
    uint64_t last_w = kNoStatKey, word = 0;
    uint32_t n_runs = 0;
    for (uint32_t s = 0; s < ch.n_stream; s++) {
        const uint64_t tp = (uint64_t)ch.text_begin + s;
        if (s == 0 || (tp & 31) == 0) word = __ldg(text + (tp >> 5));
        const int c = (int)((word >> (2 * (tp & 31))) & 3);
        const int64_t end = (int64_t)ch.first_pos + s;
        I = extend(ix, I, c);  // :150
        if (is_empty(I)) {
            err |= kErrLookupRange;
            break;
        }
        const int64_t kmer = end - (k - 1);
        if (I.r - I.l < t) {  // freq <= t (:154)
            uint64_t cur;
            do {              // :155-166 shrink to the shortest suffix that is still rare enough
                cur = scand_pack((int)(end - start + 1), I.r - I.l + 1, I.l, s);
                start++;
                I = drop_first_char<(KCAP > 128), false>(ix, (int)(end - start + 1), I, dummy);
            } while (I.r - I.l < t);
            while (dq_n > 0 && (int64_t)ch.first_pos + scand_start(dq[dq_head]) < kmer) {
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
        if (end >= (int64_t)ch.first_emit) {  // :181
            while (dq_n > 0 && (int64_t)ch.first_pos + scand_start(dq[dq_head]) < kmer) {
                dq_head = (dq_head + 1) & (KCAP - 1);
                dq_n--;
            }
            if (dq_n == 0) {  // the reference would index fmin_found with its sentinel n_nodes here
                err |= kErrWindowUnderflow;
            } else {
                const uint64_t w = dq[dq_head];
                if (w != last_w) {
                    last_w = w;
                    const uint32_t order = ch.unitig_start + (uint32_t)kmer;
                    const idx_t colex = scand_colex(w);
                    if (PASS == 0) {
                        const bool flagged = ch.first_pos + (uint32_t)scand_end(w) >= (uint32_t)(k - 1);  // :184
                        if (flagged) fin_atomic_min(first_flag + colex, order);
                    } else {
                        out_keys[n_runs] = order <= first_flag[colex] ? stat_key(w) : kNoStatKey;
                    }
                    n_runs++;
                }
            }
        }
    }
    if (err) fin_atomic_or(error_flags, err);
    return n_runs;
}

// =====================================================================================================
// B6  verify_shortest_streaming_search (include/build_fmin.hh:95-132) — statistics only
// =====================================================================================================
// The reference tries all k(k+1)/2 substrings of every window.  For a fixed start the tuple (len, freq, colex, end)
// grows with the end, so only the SHORTEST prefix with freq <= t of every start can win, and that prefix does not
// depend on the window: step 1 computes it once per text position, step 2 takes, per window, the minimum over its k
// starts of those that end inside the window.  The set only keeps (len, freq, colex), so ties on `end` are moot.
// rec = stat key of the shortest prefix of text[p ..) with freq <= t, at most `limit` long; kNoStatKey if none.
FIN_D uint64_t shortest_rare_prefix(const IndexView& ix, const uint64_t* __restrict__ text, const uint64_t p, const int limit,
                                    const uint32_t t, unsigned& err) {
    Ival I{0, ix.n_nodes - 1};
    for (int j = 0; j < limit; j++) {
        I = extend(ix, I, text_base(text, p + (uint64_t)j));
        if (is_empty(I)) {  // the reference would continue with {-1,-1} and freq 1: a unitig that is not in the SBWT
            err |= kErrLookupRange;
            return kNoStatKey;
        }
        if (I.r - I.l < t) return stat_key(scand_pack(j + 1, I.r - I.l + 1, I.l, 0));
    }
    return kNoStatKey;
}

// window [p, p+k): min over its starts s of rec[s] with s + len(rec[s]) <= p + k
FIN_D uint64_t verify_window_min(const uint64_t* __restrict__ rec, const uint64_t p, const int k) {
    uint64_t best = kNoStatKey;
    for (int j = 0; j < k; j++) {
        const uint64_t r = rec[p + (uint64_t)j];
        if (r != kNoStatKey && j + (int)(r >> 40) <= k && r < best) best = r;
    }
    return best;
}

#ifdef __CUDACC__
// ones per 64-column word of each row -> (after an exclusive scan) destination of the word's first edge
static __global__ void row_popcount_kernel(const uint64_t* __restrict__ rows, uint64_t n_words, uint32_t* __restrict__ counts) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;  // over 4 * n_words
    if (i < 4 * n_words) counts[i] = (uint32_t)__popcll(rows[i]);
}

static __global__ void lcs_init_kernel(const LcsArgs a, uint8_t* __restrict__ last, uint8_t* __restrict__ lcs) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    last[i] = lcs_first_column(a, i);
    lcs[i] = a.k;  // "not set yet" (:76-78)
}

// The same round, one WARP per 32 consecutive words (2048 columns).  lcs_round_word's scatter issues up to 256 byte
// stores per thread, every one of them a separate L2 transaction of a few useful bytes (5 % of the HBM rate).  The
// edges of 32 consecutive words with label c land on ONE contiguous stretch of `next` (C[c] + rank_c is monotone), so
// the lanes compact their bytes into shared memory at the offset their word's rank gives them and the warp then copies
// the stretch out with aligned 16-byte stores (byte stores only at its two ragged ends).
constexpr int kLcsWarpsPerBlock = 4;
constexpr int kLcsStageBytes = 2048 + 32;  // 32 words x 64 edges, + the 16-byte misalignment of the destination

static __global__ void __launch_bounds__(32 * kLcsWarpsPerBlock) lcs_round_warp_kernel(const LcsArgs a, const uint8_t* __restrict__ last,
                                                                                     uint8_t* __restrict__ next, uint8_t* __restrict__ lcs,
                                                                                     const uint8_t round) {
    __shared__ __align__(16) uint8_t stage[kLcsWarpsPerBlock][4][kLcsStageBytes];
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint64_t w0 = ((uint64_t)blockIdx.x * kLcsWarpsPerBlock + warp) * 32u;
    if (w0 >= a.n_words) return;  // the whole warp
    const uint64_t w = w0 + lane;
    const bool live = w < a.n_words;
    const uint64_t i0 = w * 64;
    if (w == 0) next[0] = 0;  // the root has no incoming edge (:86)
    uint32_t lv[16];
#pragma unroll
    for (int j = 0; j < 16; j++) lv[j] = 0;
    if (live) {  // update (:21-27), as in lcs_round_word
        uint32_t cv[16];
        const uint4* lp = reinterpret_cast<const uint4*>(last + i0);
        const uint4* cp = reinterpret_cast<const uint4*>(lcs + i0);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 x = lp[q], y = cp[q];
            lv[4 * q] = x.x; lv[4 * q + 1] = x.y; lv[4 * q + 2] = x.z; lv[4 * q + 3] = x.w;
            cv[4 * q] = y.x; cv[4 * q + 1] = y.y; cv[4 * q + 2] = y.z; cv[4 * q + 3] = y.w;
        }
        uint32_t prev = i0 ? last[i0 - 1] : 0xFFu;
        bool changed = false;
#pragma unroll
        for (int j = 0; j < 16; j++) {
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const uint32_t ch = (lv[j] >> (8 * b)) & 0xFFu;
                const uint32_t cur = (cv[j] >> (8 * b)) & 0xFFu;
                if (cur == a.k && ch != prev && i0 + 4 * j + b < a.n) {
                    cv[j] = (cv[j] & ~(0xFFu << (8 * b))) | ((uint32_t)round << (8 * b));
                    changed = true;
                }
                prev = ch;
            }
        }
        if (changed) {
            uint4* cq = reinterpret_cast<uint4*>(lcs + i0);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                uint4 y;
                y.x = cv[4 * q]; y.y = cv[4 * q + 1]; y.z = cv[4 * q + 2]; y.w = cv[4 * q + 3];
                cq[q] = y;
            }
        }
    }
    // propagate (:10-19)
    uint64_t base[4];
    unsigned total[4];
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint64_t bits = live ? a.rows[(uint64_t)c * a.n_words + w] : 0;
        const uint64_t my_dst = (uint64_t)a.C[c] + (live ? a.word_dst[(uint64_t)c * a.n_words + w] : 0u);
        const uint64_t b0 = __shfl_sync(0xFFFFFFFFu, my_dst, 0);
        base[c] = b0;
        const unsigned head = (unsigned)(b0 & 15u);
        const unsigned off = live ? (unsigned)(my_dst - b0) : 0u;
        total[c] = __reduce_max_sync(0xFFFFFFFFu, live ? off + (unsigned)__popcll(bits) : 0u);
        uint8_t* sp = stage[warp][c] + head + off;
#pragma unroll
        for (int j = 0; j < 16; j++) {
#pragma unroll
            for (int b = 0; b < 4; b++) {
                if ((bits >> (4 * j + b)) & 1ULL) *sp++ = (uint8_t)((lv[j] >> (8 * b)) & 0xFFu);
            }
        }
    }
    __syncwarp();
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const unsigned head = (unsigned)(base[c] & 15u), end = head + total[c];  // bytes [head, end) of the stage are valid
        uint8_t* g = next + (base[c] - head);                                     // 16-byte aligned
        const uint8_t* sp = stage[warp][c];
        for (unsigned q = lane; q * 16u < end; q += 32u) {
            const unsigned lo = q * 16u, hi = lo + 16u;
            if (lo >= head && hi <= end) {
                *reinterpret_cast<uint4*>(g + lo) = *reinterpret_cast<const uint4*>(sp + lo);
            } else {
                for (unsigned x = lo < head ? head : lo; x < (hi < end ? hi : end); x++) g[x] = sp[x];
            }
        }
    }
}

static __global__ void first_kmer_kernel(const IndexView ix, const uint64_t* __restrict__ text,
                                         const uint64_t* __restrict__ starts, uint32_t n_unitigs, uint32_t* __restrict__ columns) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u < n_unitigs) columns[u] = first_kmer_column(ix, text, starts[u]);
}

template <int KCAP>
__global__ void __launch_bounds__(128) build_stream_kernel(const IndexView ix, const uint64_t* __restrict__ text,
                                                           const BuildChunk* __restrict__ chunks, uint32_t n_chunks,
                                                           unsigned long long* __restrict__ table,
                                                           unsigned int* __restrict__ error_flags) {
    const uint32_t cid = blockIdx.x * blockDim.x + threadIdx.x;
    if (cid < n_chunks) build_stream_chunk<KCAP>(ix, text, chunks[cid], table, error_flags);
}

// fmin bit vector (one ballot per 32 columns) and the offset of every column (0xFFFFFFFF where unset)
static __global__ void fmin_bits_kernel(const unsigned long long* __restrict__ table, uint64_t n, uint32_t* __restrict__ bits32,
                                        uint32_t* __restrict__ offs_or_sentinel) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long key = i < n ? table[i] : 0ULL;
    const unsigned mask = __ballot_sync(0xFFFFFFFFu, key != 0ULL);
    if ((threadIdx.x & 31) == 0 && i < ((n + 31) / 32) * 32) bits32[i >> 5] = mask;
    if (i < n) offs_or_sentinel[i] = key ? fmin_key_offset(key) : 0xFFFFFFFFu;
}

// ---- preparation on the device: bit packing, the 128-byte lines, the permuted 2-bit text -------------------------
// sdsl int_vector layout (int_vector.hpp:624-640): element i occupies bits [i*w, (i+1)*w) LSB-first.  One thread
// assembles one 64-bit word of the packed vector from the elements that overlap it.
template <class T>
__global__ void pack_bits_kernel(const T* __restrict__ in, uint64_t n, uint32_t width, uint64_t* __restrict__ words, uint64_t n_words) {
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint64_t bit0 = w * 64;
    uint64_t i = bit0 / width;
    uint64_t out = 0;
    for (; i < n && i * width < bit0 + 64; i++) {
        const uint64_t v = (uint64_t)in[i], b = i * width;
        out |= b >= bit0 ? v << (b - bit0) : v >> (bit0 - b);
    }
    words[w] = out;
}

// the inverse, for widths <= 8 (an LCS file given with --lcs): one thread per element
// also checks what the word-parallel drop_first_char relies on (search_kernels.cuh): LCS[0] = 0 and every value < k
// (the reference guards the scan with `first > 0`, common.hh:43; a foreign or corrupt file must not send it out of bounds)
static __global__ void unpack_bits_to_bytes_kernel(const uint64_t* __restrict__ words, uint64_t n, uint32_t width, uint8_t* __restrict__ out,
                                                   uint32_t k, unsigned int* __restrict__ bad) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t bit = i * width, wi = bit >> 6, off = bit & 63;
    uint64_t v = words[wi] >> off;
    if (off + width > 64) v |= words[wi + 1] << (64 - off);
    v &= (1ULL << width) - 1ULL;
    out[i] = (uint8_t)v;
    if (v >= k || (i == 0 && v != 0)) atomicOr(bad, 1u);
}

// clears the bits past n in the last word of each of the 4 rows (they never count, SBWT.hh:395-424 stores whole words)
static __global__ void mask_row_tails_kernel(uint64_t* __restrict__ rows, uint64_t n_words, uint64_t n) {
    const uint32_t c = threadIdx.x;
    if (c < 4 && (n & 63)) rows[(uint64_t)c * n_words + n_words - 1] &= (1ULL << (n & 63)) - 1ULL;
}

// the lines of device_index.cuh: one thread per line.  ustart / fmin = the words of those bit vectors, or null
// (build-fmin streams before they exist): zeros then.
static __global__ void make_lines_kernel(const LcsArgs a, const uint8_t* __restrict__ lcs_bytes, const uint64_t* __restrict__ ustart,
                                         const uint64_t* __restrict__ fmin, uint64_t n_lines, uint4* __restrict__ lines) {
    const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_lines) return;
    uint4* line = lines + b * 8;
    const uint64_t us = ustart && b < a.n_words ? ustart[b] : 0, fm = fmin && b < a.n_words ? fmin[b] : 0;
    const uint32_t extra[4] = {(uint32_t)us, (uint32_t)(us >> 32), (uint32_t)fm, (uint32_t)(fm >> 32)};
#pragma unroll
    for (int c = 0; c < 4; c++) {
        uint64_t bits = 0;
        uint32_t cum;
        if (b < a.n_words) {
            bits = a.rows[(uint64_t)c * a.n_words + b];
            cum = a.C[c] + a.word_dst[(uint64_t)c * a.n_words + b];
        } else {  // past the end: rank of the whole row
            cum = a.C[c] + a.word_dst[(uint64_t)c * a.n_words + a.n_words - 1] +
                  (uint32_t)__popcll(a.rows[(uint64_t)c * a.n_words + a.n_words - 1]);
        }
        line[c] = make_uint4((uint32_t)bits, (uint32_t)(bits >> 32), cum, extra[c]);
    }
    // lcs_bytes is padded with zeros to 64 * n_words + 64 bytes
    const uint4* src = reinterpret_cast<const uint4*>(lcs_bytes + b * 64);
#pragma unroll
    for (int q = 0; q < 4; q++) line[4 + q] = b <= a.n_words ? src[q] : make_uint4(0, 0, 0, 0);
}

FIN_D int ascii_code(uint32_t ch, bool& bad) {  // PackedStrings.hh:53-57: A,C,G,T only (the reader upper-cased already)
    bad |= !(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T');
    return (int)(((ch >> 1) ^ (ch >> 2)) & 3u);
}

// SBWT column of the first k-mer of every unitig straight from the ASCII text (PackedStrings.hh:116,129-131)
static __global__ void first_kmer_ascii_kernel(const IndexView ix, const char* __restrict__ ascii, const uint64_t* __restrict__ offsets,
                                               uint32_t n_unitigs, uint32_t* __restrict__ columns, unsigned int* __restrict__ error_flags) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_unitigs) return;
    const char* s = ascii + offsets[u];
    Ival I{0, ix.n_nodes - 1};
    bool bad = false;
    for (int i = 0; i < ix.k && !is_empty(I); i++) I = extend(ix, I, ascii_code((uint8_t)s[i], bad));
    columns[u] = (!bad && !is_empty(I) && I.l == I.r) ? I.l : 0xFFFFFFFFu;
    if (bad) atomicOr(error_flags, kErrBadBase);
}

// concatenation of the unitigs in permuted order, 2 bits per base (PackedStrings.hh:36-64): one thread per output word.
// starts[j] = text position of the j-th unitig of the new order, src[j] = offset of its characters in `ascii`.
static __global__ void pack_text_kernel(const char* __restrict__ ascii, const uint64_t* __restrict__ starts, const uint64_t* __restrict__ src,
                                        uint32_t n_unitigs, uint64_t text_len, uint64_t* __restrict__ words, uint64_t n_words,
                                        unsigned int* __restrict__ error_flags) {
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint64_t p0 = w * 32;
    uint64_t out = 0;
    if (p0 < text_len) {
        uint32_t lo = 0, hi = n_unitigs;  // last unitig whose start is <= p0
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (starts[mid] <= p0) lo = mid;
            else hi = mid;
        }
        uint32_t j = lo;
        uint64_t next_start = j + 1 < n_unitigs ? starts[j + 1] : text_len;
        const char* s = ascii + src[j] - starts[j];  // s[p] is the character of text position p
        bool bad = false;
        const uint64_t p1 = min(p0 + 32, text_len);
        for (uint64_t p = p0; p < p1; p++) {
            if (p == next_start) {
                j++;
                next_start = j + 1 < n_unitigs ? starts[j + 1] : text_len;
                s = ascii + src[j] - starts[j];
            }
            out |= (uint64_t)ascii_code((uint8_t)s[p], bad) << (2 * (p - p0));
        }
        if (bad) atomicOr(error_flags, kErrBadBase);
    }
    words[w] = out;
}

// ---- statistics modes (B5, B6) ------------------------------------------------------------------------------
template <int KCAP, int PASS>
__global__ void __launch_bounds__(128) shortest_stream_kernel(const IndexView ix, const uint64_t* __restrict__ text,
                                                              const BuildChunk* __restrict__ chunks, uint32_t n_chunks, uint32_t t,
                                                              uint32_t* __restrict__ first_flag, uint32_t* __restrict__ runs,
                                                              uint64_t* __restrict__ out_keys, unsigned int* __restrict__ error_flags) {
    const uint32_t cid = blockIdx.x * blockDim.x + threadIdx.x;
    if (cid >= n_chunks) return;
    if (PASS == 0) runs[cid] = shortest_stream_chunk<KCAP, 0>(ix, text, chunks[cid], t, first_flag, nullptr, error_flags);
    else shortest_stream_chunk<KCAP, 1>(ix, text, chunks[cid], t, first_flag, out_keys + runs[cid], error_flags);  // runs = offsets now
}

// index of the unitig that holds text position p: first u with ends[u] > p
FIN_D uint32_t unitig_of(const uint32_t* __restrict__ ends, uint32_t n_unitigs, uint32_t p) {
    uint32_t lo = 0, hi = n_unitigs;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(ends + mid) > p) hi = mid;
        else lo = mid + 1;
    }
    return lo;
}

// text positions [p_begin, p_end): rec[p - p_begin]
static __global__ void rare_prefix_kernel(const IndexView ix, const uint64_t* __restrict__ text, const uint32_t* __restrict__ ends,
                                          uint32_t n_unitigs, uint32_t p_begin, uint32_t p_end, uint32_t t, uint64_t* __restrict__ rec,
                                          unsigned int* __restrict__ error_flags) {
    const uint64_t p = (uint64_t)p_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= p_end) return;
    const uint32_t e = __ldg(ends + unitig_of(ends, n_unitigs, (uint32_t)p));
    unsigned err = 0;
    rec[p - p_begin] = shortest_rare_prefix(ix, text, p, (int)min((uint32_t)ix.k, e - (uint32_t)p), t, err);
    if (err) atomicOr(error_flags, err);
}

// windows that start in [p_begin, p_end): keys[p - p_begin]; rec is indexed from p_begin and reaches k positions past p_end
static __global__ void verify_window_kernel(const uint64_t* __restrict__ rec, const uint32_t* __restrict__ ends, uint32_t n_unitigs,
                                            uint32_t p_begin, uint32_t p_end, int k, uint64_t* __restrict__ keys,
                                            unsigned int* __restrict__ error_flags) {
    const uint64_t p = (uint64_t)p_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= p_end) return;
    const uint32_t e = __ldg(ends + unitig_of(ends, n_unitigs, (uint32_t)p));
    uint64_t key = kNoStatKey;
    if (p + (uint64_t)k <= e) {
        key = verify_window_min(rec, p - p_begin, k);
        if (key == kNoStatKey) atomicOr(error_flags, kErrWindowUnderflow);  // the reference inserts its sentinel (k+1, N, N)
    }
    keys[p - p_begin] = key;
}

// sums[0] += valid keys, sums[1] += their freq, sums[2] += their len   (print_finimizer_stats, common.hh:188-197)
static __global__ void stat_sum_kernel(const uint64_t* __restrict__ keys, uint64_t n, unsigned long long* __restrict__ sums) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t key = i < n ? keys[i] : kNoStatKey;
    const bool valid = key != kNoStatKey;
    unsigned cnt = valid ? 1u : 0u, fr = valid ? (unsigned)((key >> 32) & 0xFFu) : 0u, ln = valid ? (unsigned)(key >> 40) : 0u;
    cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
    fr = __reduce_add_sync(0xFFFFFFFFu, fr);
    ln = __reduce_add_sync(0xFFFFFFFFu, ln);
    if ((threadIdx.x & 31) == 0 && cnt) {
        atomicAdd(sums, (unsigned long long)cnt);
        atomicAdd(sums + 1, (unsigned long long)fr);
        atomicAdd(sums + 2, (unsigned long long)ln);
    }
}

struct NotSentinel {
    __host__ __device__ bool operator()(const uint32_t& v) const { return v != 0xFFFFFFFFu; }
};
#endif  // __CUDACC__

}  // namespace fin
