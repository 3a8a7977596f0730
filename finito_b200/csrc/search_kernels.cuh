// Kernels of the search-fmin path.
//   pack_reads_kernel   : ASCII -> 2 bits per base + validation      (SBWT/include/sbwt/SeqIO.hh:232 upper-casing,
//                                                                     common.hh:50-58 char_idx, :106-111 validation)
//   kmer_counts_kernel  : k-mer positions per read                   (search_fmin.hh:69)
//   search_reads_kernel : persistent warps; every lane runs one chain = one read, forward strand then reverse
//                         complement, as a flattened state machine (struct Chain below) whose iteration is the
//                         same instruction sequence for every lane whatever its phase:
//                         rarest_fmin_streaming_search (common.hh:78-186) fused with the resolve loop of
//                         FinimizerIndex::search (FinimizerIndex.hh:148-183), walk_in_unitigs (:47-102),
//                         the dictionary lookups (common.hh:61-72), global_offset_to_local_offset
//                         (PackedStrings.hh:91-100) and the strand merge (search_fmin.hh:54-60).
#pragma once
#include "device_index.cuh"

namespace fin {

struct SearchTotals {  // device-side accumulators of one launch
    unsigned long long found_fwd, found_rc, positive, n_queries;
    unsigned long long n_rank, n_lcs_reads, n_ustart_reads, n_anchor, n_anchor_branch, n_walk_words, n_bases, n_kmers;
    unsigned int error_flags;  // bit 0: non-ACGT base; bit 1: candidate window underflow; bit 2: lookup out of range
    unsigned int pad;
    unsigned long long next_read;  // work counter of the persistent search kernel
    unsigned long long n_runs;     // runs mode: records appended (RunSink::count)
    unsigned long long n_filter;   // s-mer table lookups of the strand prefilter (counting mode only)
    // of n_lcs_reads / n_rank (the reference's tally): accesses of drops and retried extensions that the closed forms of
    // the shrink and contraction loops never issue (counting mode only)
    unsigned long long n_skip_lcs, n_skip_rank;
};

enum : unsigned { kErrBadBase = 1u, kErrWindowUnderflow = 2u, kErrLookupRange = 4u, kErrReadTooLong = 8u, kErrBadOffsets = 16u };

struct ReadCounters {  // per-thread tallies, reduced per warp at the end of the kernel
    unsigned found[2] = {0, 0}, positive = 0, n_q = 0, err = 0;
    unsigned c_rank = 0, c_lcs = 0, c_ustart = 0, c_anchor = 0, c_branch = 0, c_walk = 0, c_bases = 0, c_kmers = 0;
    unsigned c_skip_lcs = 0, c_skip_rank = 0;  // part of c_lcs / c_rank that the closed forms skip
};

#ifdef __CUDACC__
// ---- ASCII -> 2-bit.  One thread packs 32 consecutive bases (two 128-bit loads) into one u64. ----------
// code = ((c>>1) ^ (c>>2)) & 3 maps A,C,G,T (either case) to 0,1,2,3.
static __global__ void pack_reads_kernel(const uint8_t* __restrict__ ascii, uint64_t n_bases, uint64_t* __restrict__ packed,
                                  uint64_t n_words, SearchTotals* __restrict__ totals) {
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint64_t base0 = w * 32;
    uint32_t chunk[8];
    if (base0 + 32 <= n_bases && ((reinterpret_cast<uintptr_t>(ascii) & 15) == 0)) {
        const uint4 v0 = __ldg(reinterpret_cast<const uint4*>(ascii + base0));
        const uint4 v1 = __ldg(reinterpret_cast<const uint4*>(ascii + base0) + 1);
        chunk[0] = v0.x; chunk[1] = v0.y; chunk[2] = v0.z; chunk[3] = v0.w;
        chunk[4] = v1.x; chunk[5] = v1.y; chunk[6] = v1.z; chunk[7] = v1.w;
    } else {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            uint32_t x = 0;
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const uint64_t i = base0 + j * 4 + b;
                const uint32_t ch = i < n_bases ? ascii[i] : (uint32_t)'A';
                x |= ch << (8 * b);
            }
            chunk[j] = x;
        }
    }
    uint64_t out = 0;
    bool bad = false;
#pragma unroll
    for (int j = 0; j < 8; j++) {
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const uint32_t ch = ((chunk[j] >> (8 * b)) & 0xFFu) & ~32u;
            bad |= !(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T');
            const uint64_t code = ((ch >> 1) ^ (ch >> 2)) & 3u;
            out |= code << (2 * (j * 4 + b));
        }
    }
    packed[w] = out;
    if (bad) atomicOr(&totals->error_flags, kErrBadBase);
}

// k-mer positions per read (counts[n_reads] = 0 so that the exclusive scan ends with the total); also checks that the
// offsets are non-decreasing and end at n_bases (the device entry points cannot trust their caller's arrays).
static __global__ void kmer_counts_kernel(const uint64_t* __restrict__ read_offsets, uint64_t n_reads, uint64_t n_bases, int k,
                                   uint64_t* __restrict__ counts, SearchTotals* __restrict__ totals) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r > n_reads) return;
    if (r == n_reads) {
        counts[r] = 0;
        if (read_offsets[r] != n_bases) atomicOr(&totals->error_flags, kErrBadOffsets);
        return;
    }
    const uint64_t a = read_offsets[r], b = read_offsets[r + 1];
    if (b < a || b > n_bases) {
        atomicOr(&totals->error_flags, kErrBadOffsets);
        counts[r] = 0;
        return;
    }
    const uint64_t len = b - a;
    counts[r] = len >= (uint64_t)k ? len - (uint64_t)k + 1 : 0;
}

// the s-mers of the text: one thread per text position (windows that span two unitigs only add bits: harmless)
static __global__ void smer_table_kernel(const uint64_t* __restrict__ text, const uint64_t text_len, const int s, uint32_t* __restrict__ bits) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p + (uint64_t)s > text_len) return;
    const uint64_t w0 = text[p >> 5], w1 = text[(p >> 5) + 1];  // the text array is padded by one word
    const int sh = 2 * (int)(p & 31);
    const uint64_t v = sh ? (w0 >> sh) | (w1 << (64 - sh)) : w0;
    const uint32_t code = (uint32_t)v & (s >= 16 ? 0xFFFFFFFFu : ((1u << (2 * s)) - 1u));
    atomicOr(bits + (code >> 5), 1u << (code & 31));
}
#endif  // __CUDACC__

// ---- strand prefilter ---------------------------------------------------------------------------------------------
// Every k-mer of the index is a substring of the unitig text, so all of its k - s + 1 s-mers are set in IndexView::smer.
// A strand of a read that nowhere has k - s + 1 consecutive present s-mers therefore holds no k-mer of the index:
// the reference would answer (-1,-1) for every position of that strand (FinimizerIndex.hh:150-152) after streaming
// through all of it; here the strand is not streamed at all.  No false negatives by construction, false positives only
// cost time.  In the benchmark workload a read matches the index on one strand, so about half of the streaming work
// (the more expensive half: a non-matching strand contracts at every character) disappears.
// s-mer code = sum base[i] << 2i, the bit order of the packed text and of the packed reads.
enum : unsigned { kStrandFwd = 1u, kStrandRc = 2u };

FIN_D uint32_t smer_mask(const int s) { return s >= 16 ? 0xFFFFFFFFu : ((1u << (2 * s)) - 1u); }
FIN_D bool smer_present(const uint32_t* __restrict__ bits, const uint32_t code) { return (__ldg(bits + (code >> 5)) >> (code & 31)) & 1u; }

// s-mer of the read that starts at base q (absolute base index b = b0 + q), first base in the low bits
FIN_D uint32_t read_smer(const uint64_t* __restrict__ packed, const uint64_t b, const uint32_t mask) {
    const uint64_t w0 = __ldg(packed + (b >> 5)), w1 = __ldg(packed + (b >> 5) + 1);  // the packed array is padded by a word
    const int sh = 2 * (int)(b & 31);
    return (uint32_t)(sh ? (w0 >> sh) | (w1 << (64 - sh)) : w0) & mask;
}
FIN_D uint32_t smer_revcomp(uint32_t x, const int s) {  // reverse the 2-bit groups, complement, right-align
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0F0F0F0Fu) | ((x & 0x0F0F0F0Fu) << 4);
    x = ((x >> 8) & 0x00FF00FFu) | ((x & 0x00FF00FFu) << 8);
    x = (x >> 16) | (x << 16);
    return (~x) >> (32 - 2 * s);
}

// Is there a stretch of `need` consecutive s-mer starts in [0, n_starts) all of which are present?  Checked from the far
// end of each candidate stretch backwards, so that one absent s-mer rules out `need` stretches at once (Boyer-Moore
// style): a strand that does not match costs ~n_starts / need lookups instead of n_starts.
template <bool RC>
FIN_D bool strand_has_candidate(const IndexView& ix, const uint64_t* __restrict__ packed, const uint64_t b0, const int n_starts, const int need,
                                unsigned& lookups) {
    const int s = ix.smer_s;
    const uint32_t mask = smer_mask(s);
    int a = 0;             // the stretch under test starts here
    int known = -1;        // starts in [a, known] are known to be present
    while (a + need <= n_starts) {
        int j = a + need - 1;
        const int stop = known >= a ? known : a - 1;  // everything in (stop, a + need - 1] still has to be looked up
        while (j > stop) {
            uint32_t code = read_smer(packed, b0 + (uint64_t)j, mask);
            if (RC) code = smer_revcomp(code, s);
            lookups++;
            if (!smer_present(ix.smer, code)) break;
            j--;
        }
        if (j <= stop) return true;
        known = a + need - 1;  // (j, a + need - 1] are present, j is absent
        a = j + 1;
    }
    return false;
}

// which strands of the read at bases [b0, b0 + len) of `packed` can hold a k-mer of the index
FIN_D unsigned strand_candidates(const IndexView& ix, const uint64_t* __restrict__ packed, const uint64_t b0, const uint64_t len,
                                 unsigned& lookups) {
    const int s = ix.smer_s, k = ix.k;
    if (ix.smer == nullptr) return kStrandFwd | kStrandRc;
    if (len < (uint64_t)k) return 0;
    const int n_starts = (int)len - s + 1, need = k - s + 1;
    return (strand_has_candidate<false>(ix, packed, b0, n_starts, need, lookups) ? kStrandFwd : 0u) |
           (strand_has_candidate<true>(ix, packed, b0, n_starts, need, lookups) ? kStrandRc : 0u);
}

// ---- drop_first_char (common.hh:38-48), word-parallel ------------------------------------------------------
// The reference widens l downwards while LCS[l] >= new_len and r upwards while LCS[r+1] >= new_len, one element
// per iteration.  Here one step inspects the aligned 8-byte LCS word that holds l (resp. r+1) with byte-parallel
// compares and jumps to the terminating element; a side that runs off its word stays pending for another step.
// LCS[0] = 0 and the LCS pad past n_nodes is 0, so both scans stop inside the array without bound checks
// (new_len >= 1 here; new_len <= 0 is the "whole range" case handled by the callers).
constexpr uint64_t kByteHigh = 0x8080808080808080ULL;

// bit 7 of every byte whose value is < n (1 <= n <= 255).  WIDE = false needs all bytes and n below 129 (k <= 128).
template <bool WIDE> FIN_D uint64_t bytes_below(const uint64_t x, const int n) {
    const uint64_t n8 = (uint64_t)(uint32_t)n * 0x0101010101010101ULL;
    if (!WIDE) {
        return ~((x | kByteHigh) - n8) & kByteHigh;
    } else {
        const uint64_t t = ((x & ~kByteHigh) | kByteHigh) - (n8 & ~kByteHigh);  // bit 7: low 7 bits of x >= those of n
        const uint64_t xh = x & kByteHigh, nh = n8 & kByteHigh;
        const uint64_t ge = (xh & ~nh) | (~(xh ^ nh) & t);
        return ~ge & kByteHigh;
    }
}

enum : unsigned { kDropLeft = 1u, kDropRight = 2u };

// u64 index of the aligned 8-byte LCS word of column p inside the lines array (< n_nodes / 4: fits 32 bits)
FIN_D uint32_t lcs_word_index(const idx_t p) { return ((p >> 6) << 4) | 8u | ((p >> 3) & 7u); }

// 8 LCS bytes around column p, aligned to FOUR bytes inside the LCS half of p's line (two 4-byte loads): a left scan then
// always has 5..8 columns at and below p in its window, a right scan 5..8 at and above p (fewer only in the first / last
// four columns of a line), where the aligned 8-byte word leaves 1..8.  Returns the first column of the window (mod 64).
FIN_D uint32_t lcs_win_left_start(const idx_t p) {
    const uint32_t lo = p & 63u;
    return lo >= 4u ? ((lo - 4u) & ~3u) : 0u;
}
FIN_D uint32_t lcs_win_right_start(const idx_t p) {
    const uint32_t lo = (p & 63u) & ~3u;
    return lo < 56u ? lo : 56u;
}
FIN_D uint64_t lcs_win_load(const uint32_t* __restrict__ l32, const idx_t p, const uint32_t start) {
    const uint32_t idx = ((p >> 6) << 5) | 16u | (start >> 2);  // 32 u32 per line, the LCS bytes are its second half
    return (uint64_t)__ldg(l32 + idx) | ((uint64_t)__ldg(l32 + idx + 1u) << 32);
}
// word steps on a window: j = byte of the window that holds LCS[l] (left) / LCS[r + 1] (right)
template <bool WIDE> FIN_D idx_t drop_left_window(const uint64_t w, const idx_t l, const uint32_t j, const int new_len, bool& found) {
    const uint64_t stop = bytes_below<WIDE>(w, new_len) << (8u * (7u - j));
    found = stop != 0;
    return found ? l - (idx_t)(__clzll((long long)stop) >> 3) : l - j - 1u;
}
template <bool WIDE> FIN_D idx_t drop_right_window(const uint64_t w, const idx_t r, const uint32_t j, const int new_len, bool& found) {
    const uint64_t stop = bytes_below<WIDE>(w, new_len) >> (8u * j);
    found = stop != 0;
    return found ? r + (idx_t)((__ffsll((long long)stop) - 1) >> 3) : r + (8u - j);
}

// Left end: w = the LCS word of column l.  Returns the new l; found = the scan stopped inside this word.
template <bool WIDE> FIN_D idx_t drop_left_word(const uint64_t w, const idx_t l, const int new_len, bool& found) {
    const uint32_t j = l & 7u;
    const uint64_t stop = bytes_below<WIDE>(w, new_len) << (8u * (7u - j));  // bytes above l shifted out
    found = stop != 0;  // always true in the word of column 0: LCS[0] = 0
    return found ? l - (idx_t)(__clzll((long long)stop) >> 3) : l - j - 1u;
}
// Right end: w = the LCS word of column r + 1.  Returns the new r.
template <bool WIDE> FIN_D idx_t drop_right_word(const uint64_t w, const idx_t r, const int new_len, bool& found) {
    const idx_t rp = r + 1;
    const uint32_t j = rp & 7u;
    const uint64_t stop = bytes_below<WIDE>(w, new_len) >> (8u * j);  // bytes below r+1 shifted out
    found = stop != 0;
    return found ? r + (idx_t)((__ffsll((long long)stop) - 1) >> 3) : (rp | 7u);
}

// Left end: w = the LCS word of column I.l.  Clears kDropLeft in pend unless the scan runs off the word.
template <bool WIDE, bool COUNT>
FIN_D void drop_left(const uint64_t w, Ival& I, const int new_len, unsigned& pend, unsigned& n_lcs) {
    bool found;
    const idx_t l = I.l, nl = drop_left_word<WIDE>(w, l, new_len, found);
    if (COUNT) n_lcs += (l - nl) + (found && nl > 0 ? 1u : 0u);
    if (found) pend &= ~kDropLeft;
    I.l = nl;
}

// Right end: w = the LCS word of column I.r + 1.
template <bool WIDE, bool COUNT>
FIN_D void drop_right(const uint64_t w, Ival& I, const int new_len, unsigned& pend, const idx_t n_nodes, unsigned& n_lcs) {
    bool found;
    const idx_t nr = drop_right_word<WIDE>(w, I.r, new_len, found);
    if (COUNT) n_lcs += (nr - I.r) + (found && nr < n_nodes - 1 ? 1u : 0u);
    if (found) pend &= ~kDropRight;
    I.r = nr;
}

// One word step on the pending sides of I; returns the sides that are still pending.
template <bool WIDE, bool COUNT>
FIN_D unsigned drop_step(const IndexView& ix, Ival& I, const int new_len, unsigned pend, unsigned& n_lcs) {
    const uint64_t* __restrict__ lw = reinterpret_cast<const uint64_t*>(ix.lines);
    if (pend & kDropLeft) drop_left<WIDE, COUNT>(__ldg(lw + lcs_word_index(I.l)), I, new_len, pend, n_lcs);
    if (pend & kDropRight) drop_right<WIDE, COUNT>(__ldg(lw + lcs_word_index(I.r + 1)), I, new_len, pend, ix.n_nodes, n_lcs);
    return pend;
}

// the whole of drop_first_char, for the callers that are not state machines (build-fmin kernels)
template <bool WIDE, bool COUNT>
FIN_D Ival drop_first_char(const IndexView& ix, const int new_len, Ival I, unsigned& n_lcs) {
    if (new_len <= 0) {
        I.l = 0;
        I.r = ix.n_nodes - 1;
        return I;
    }
    unsigned pend = kDropLeft | kDropRight;
    while (pend) pend = drop_step<WIDE, COUNT>(ix, I, new_len, pend, n_lcs);
    return I;
}


template <class HitT> struct alignas(2 * sizeof(HitT)) HitPair { HitT unitig, pos; };

// Candidate finimizers are packed into one u64 so that the unsigned order equals the reference's tuple
// order (freq, len, colex, end) with freq == 1 throughout (common.hh:84,147): len:8 | colex:32 | end:24.
FIN_D uint64_t cand_pack(int len, idx_t colex, int end) {
    return ((uint64_t)(uint32_t)len << 56) | ((uint64_t)colex << 24) | (uint64_t)(uint32_t)end;
}
FIN_D int cand_len(uint64_t c) { return (int)(c >> 56); }
FIN_D idx_t cand_colex(uint64_t c) { return (idx_t)(c >> 24); }
FIN_D int cand_end(uint64_t c) { return (int)(c & 0xFFFFFFu); }
FIN_D int cand_start(uint64_t c) { return cand_end(c) - cand_len(c) + 1; }

constexpr int kMaxReadLen = (1 << 24) - 1;

// ---- run-length hits ------------------------------------------------------------------------------------------
// 98 % of the found k-mers of a read are produced by walk_in_unitigs (FinimizerIndex.hh:47-102): consecutive k-mer
// positions that map to consecutive offsets of one unitig.  A run record says exactly that for a maximal stretch the
// search produced in one go: merged k-mer positions [i0, i0 + len) of the read map to `unitig`, position i0 + d at
// offset pos0 + d (forward strand) or pos0 - d (reverse-complement strand: the merged position runs against the
// strand's own direction, search_fmin.hh:54-60).  Runs of one read never overlap (the reverse strand only reports
// positions the forward strand missed), every position outside a run is (-1,-1).  Same layout as fin_run in
// include/finito_b200.h.
struct RunRec {
    uint32_t i0, len_rc /* len | rc << 31 */, unitig, pos0;
};
constexpr uint32_t kRunRc = 0x80000000u;
constexpr int kRunKeyPosBits = 24;  // sort key of a record: read << 24 | i0  (i0 <= kMaxReadLen)

// where the search kernel appends its records: struct of arrays (cub sorts keys + values), one counter
struct RunSink {
    uint64_t* keys;
    RunRec* recs;
    unsigned long long* count;  // records produced (keeps counting past the capacity: the caller retries)
    uint64_t capacity;
};

// does any lane of the (converged) warp want to append?
FIN_D bool run_any(const bool want) {
#ifdef FIN_EMU
    return want;
#else
    return __any_sync(0xFFFFFFFFu, want);
#endif
}

// Appends one record per lane that has one.  Called by every lane of the warp at the same point (converged).
FIN_D void run_append(const RunSink& sink, const bool want, const uint64_t key, const RunRec rec) {
#ifdef FIN_EMU
    if (want) {
        const uint64_t t = (*sink.count)++;
        if (t < sink.capacity) {
            sink.keys[t] = key;
            sink.recs[t] = rec;
        }
    }
#else
    const unsigned m = __ballot_sync(0xFFFFFFFFu, want);
    if (m == 0) return;
    const unsigned lane = threadIdx.x & 31u;
    const int leader = __ffs((int)m) - 1;
    unsigned long long base = 0;
    if ((int)lane == leader) base = atomicAdd(sink.count, (unsigned long long)__popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (want) {
        const uint64_t t = base + (uint64_t)__popc(m & ((1u << lane) - 1u));
        if (t < sink.capacity) {
            sink.keys[t] = key;
            *reinterpret_cast<uint4*>(sink.recs + t) = make_uint4(rec.i0, rec.len_rc, rec.unitig, rec.pos0);
        }
    }
#endif
}

// ---- one chain = one read, both strands, as a flattened state machine ------------------------------------------
// The reference's loop nest (per character: contract-and-retry loops for the finimizer interval and the k-mer
// interval, the shrink loop, the widening loops inside drop_first_char) has data-dependent trip counts: bursts of
// ~20 contractions after every read error, one or two per character on the strand that does not match.  With one
// read per lane and the loops as written, a warp waits for its slowest lane at every character (measured: 6.8 of
// 32 lanes active per instruction).  Here every iteration of ONE loop does, for every lane, the same four things
//   A  extend        : ranks for I and I_kmer with the current character        (common.hh:114, :133)
//   B  decide        : Ustart test, set up the shrink drop and the window drop  (:145-154, :167, :170, :181)
//   F  drop          : one word step of every pending drop_first_char           (common.hh:38-48)
//   D/E shrink check, candidate window, emission, resolve, next character       (:155-164, :173-182, FinimizerIndex.hh:148-183)
// and a lane that needs another contraction or another drop word simply spends its next iteration on it while its
// neighbours move on to their next character: lanes drift apart in position instead of waiting for each other.
// In the matching region a character costs one iteration.
// Every lane of the warp calls step() in every iteration, so the phases can be fenced with warp barriers: they keep
// the compiler from threading a phase into the branches of the one before it (which would execute it once per
// incoming path, each time for a few lanes).
#if defined(FIN_EMU)
#define FIN_PHASE_FENCE() do {} while (0)
#else
#define FIN_PHASE_FENCE() __syncwarp()
#endif

#ifndef FIN_LCS_WIN4
#define FIN_LCS_WIN4 1  // 4-byte aligned LCS windows for the word steps of the search (profiles/r2_sweep_closed_form_shrink.log)
#endif
#ifndef FIN_DQ_BK2
#define FIN_DQ_BK2 1  // cache the entry below the back of the candidate window in a register (measured: profiles/r2_sweep_closed_form_shrink.log)
#endif
enum : int { ST_EXT = 0, ST_EXTK = 1, ST_POST = 2, ST_SHRINK = 3, ST_EMIT = 4, ST_IDLE = 5 };

#ifndef FIN_SEARCH_THREADS
#define FIN_SEARCH_THREADS 128
#endif

// Where a chain's answers go.  Dense: one HitPair per k-mer position (out + out_offsets[read]).  Runs: RunSink.
// Both: `fmask`, one bit per k-mer position = "the forward strand found it", so that the reverse strand can apply
// the merge rule (search_fmin.hh:54-60) without reading the forward strand's hits back.  The current 32-bit word
// lives in a register; words are only spilled to `fmask` for reads with more than 32 k-mer positions.
template <class HitT> struct SearchOut {
    HitPair<HitT>* hits;        // dense mode (RUNS == false)
    const uint64_t* out_offsets;
    uint32_t* fmask;            // (total_kmers >> 5) + n_reads + 1 words; read r starts at (out_offsets[r] >> 5) + r
    RunSink sink;               // runs mode
    const uint8_t* strand_flags;  // per read kStrandFwd | kStrandRc from strand_filter_kernel; nullptr = stream every strand
};

// G > 1 = the warp-cooperative variant (BASELINE.json north_star; SURVEY.md section 7 hard part 1): G adjacent lanes own ONE
// chain.  The control flow and the scalar state are replicated on the G lanes (it costs no issue slots: they execute the
// same instructions anyway), the memory accesses of a step are dealt out - lane q of the group fetches the q-th rank row /
// LCS word and the group exchanges the results with __shfl_sync - and only the group's first lane writes answers and
// tallies counters.  Measured against G = 1 in profiles/r2_sweep_warp_cooperative.log (it loses: chains in flight per SM,
// not loads per chain, are what hides the latency of a dependent walk); kept selectable with FIN_COOP_GROUP=2|4 so that
// the comparison can be repeated, dense int32 hits only.
template <int KCAP, class HitT, bool COUNT, bool RUNS, int G = 1> struct Chain {
    static constexpr bool WIDE = KCAP > 128;
    // the read
    uint64_t b0 = 0;
    HitPair<HitT>* o = nullptr;  // dense mode: the read's slots
    uint32_t rid = 0;            // runs mode: the read's index in the batch
    uint32_t fmw = 0;            // first word of the read in SearchOut::fmask
    uint32_t fm = 0;             // current word of the forward strand's found mask
    int len = 0, nk = 0, strand = 0;
    unsigned sflags = 0;         // strands of this read that are streamed (strand prefilter)
    // stream state (common.hh:80-103)
    int st = ST_IDLE, end = 0, c = 0;
    uint32_t rword = 0;  // the 16 bases of the read around `end`
    int start = 0, kmer_start = 0;
    Ival I{0, 0}, Ik{0, 0};
    unsigned pendI = 0, pendK = 0;  // sides of the pending drops
    int lenI = 0, lenK = 0;         // their new_len
    bool emit = false;              // a complete k-mer window ends at `end`
    bool fresh = false;             // the character at `end` has not been fetched yet
    uint64_t cur = 0;               // the candidate being shrunk
    int dq_head = 0, dq_n = 0;
    int bestU_pos = -1;
    idx_t bestU_col = 0;
    // walk state (walk_in_unitigs): the last answered k-mer ends at text position walk_g
    bool walking = false;
    idx_t walk_g = 0, walk_ustart = 0, walk_uend = 0;
    uint32_t walk_uid = 0;
    uint64_t tword = 0;
    int walk_j = 0;
    int run_len = 0;          // runs mode: length of the open run (it ends at the previous k-mer position)
    uint64_t fr = 0, bk = 0, bk2 = 0;  // front, back and the entry below the back of the candidate window, cached in registers

    FIN_D void begin_strand(const IndexView& ix) {
        start = kmer_start = 0;
        I.l = 0; I.r = ix.n_nodes - 1;
        Ik = I;
        dq_head = dq_n = 0;
        bestU_pos = -1;
        walking = false;
        pendI = pendK = 0;
        end = 0;
        fresh = true;
        run_len = 0;
        st = ST_EXT;
    }

    // 2-bit code of the character at `end` of the current strand (reverse complement: read backwards, complemented)
    FIN_D void fetch_base(const uint64_t* __restrict__ packed) {
        const uint64_t b = b0 + (uint64_t)(strand ? len - 1 - end : end);
        // the 16-base word of the read stays in a register: one load per 16 characters, and the character (which the
        // rank row addresses wait for) is there without a load in between
        if (end == 0 || ((uint32_t)b & 15u) == (strand ? 15u : 0u)) rword = __ldg(reinterpret_cast<const uint32_t*>(packed) + (b >> 4));
        c = (int)((rword >> (2 * (b & 15))) & 3u) ^ (strand ? 3 : 0);
    }

    // takes a new read; returns false (and stays idle) when the read has no k-mer
    FIN_D bool begin_read(const IndexView& ix, const uint64_t base, const uint64_t len64, const uint64_t read,
                          const SearchOut<HitT>& out, ReadCounters& rc, const int n_strands) {
        st = ST_IDLE;
        if (len64 > (uint64_t)kMaxReadLen) {
            rc.err |= kErrReadTooLong;
            return false;
        }
        if (len64 < (uint64_t)ix.k) return false;
        b0 = base;
        len = (int)len64;
        nk = len - ix.k + 1;
        rc.n_q += (unsigned)nk;
        sflags = out.strand_flags ? (unsigned)__ldg(out.strand_flags + read) : (kStrandFwd | kStrandRc);
        if (n_strands < 2) sflags &= kStrandFwd;
        if (sflags == 0) return false;  // no strand can hold a k-mer of the index: every answer is (-1,-1)
        const uint64_t oo = __ldg(out.out_offsets + read);
        if (!RUNS) o = out.hits + oo;
        rid = (uint32_t)read;
        fmw = (uint32_t)((oo >> 5) + read);
        fm = 0;
        strand = (sflags & kStrandFwd) ? 0 : 1;
        begin_strand(ix);
        return true;
    }

    FIN_D void set_drop_I(const int new_len) { lenI = new_len; pendI = kDropLeft | kDropRight; }
    FIN_D void set_drop_K(const int new_len) { lenK = new_len; pendK = kDropLeft | kDropRight; }

    // dq: the KCAP slots of the candidate window, a separate array so that the scalar state stays in registers
    FIN_D void step(const IndexView& ix, const uint64_t* __restrict__ packed, ReadCounters& rc, const int n_strands, uint64_t* dq,
                    const SearchOut<HitT>& out);
};

#if defined(__CUDACC__) && !defined(FIN_EMU)
// cooperative loads of a group of G lanes (G a power of two <= 32): the owner loads, everybody receives
template <int G> FIN_D unsigned coop_sub() { return threadIdx.x & (unsigned)(G - 1); }
template <int G> FIN_D unsigned coop_mask() { return (G >= 32 ? 0xFFFFFFFFu : ((1u << G) - 1u)) << ((threadIdx.x & 31u) & ~(unsigned)(G - 1)); }
template <int G> FIN_D uint32_t coop_u32(const bool cond, const uint32_t v_owner, const int owner) {
    return cond ? __shfl_sync(coop_mask<G>(), v_owner, owner, G) : 0u;
}
template <int G> FIN_D uint64_t coop_u64(const bool cond, const uint64_t v_owner, const int owner) {
    return cond ? __shfl_sync(coop_mask<G>(), v_owner, owner, G) : 0ull;
}
#endif

template <int KCAP, class HitT, bool COUNT, bool RUNS, int G>
FIN_D void Chain<KCAP, HitT, COUNT, RUNS, G>::step(const IndexView& ix, const uint64_t* __restrict__ packed, ReadCounters& rc,
                                                const int n_strands, uint64_t* dq, const SearchOut<HitT>& out) {
    const int k = ix.k;
    const idx_t N = ix.n_nodes;
#if defined(__CUDACC__) && !defined(FIN_EMU)
    const bool leader = G == 1 || coop_sub<G>() == 0;
#else
    const bool leader = true;
#endif

    // ---- A: extend both intervals with the current character (common.hh:114-143).  ST_EXTK = the finimizer
    //         interval is already extended, only the k-mer interval is being contracted and retried.
    bool contrI = false, contrK = false;  // this iteration starts a contraction drop of I / of the k-mer interval
    if ((pendI | pendK) == 0 && st <= ST_EXTK) {
        if (fresh) {
            fetch_base(packed);
            fresh = false;
        }
        const bool both = st == ST_EXT;
        // all four ranks at once (the k-mer interval speculatively: unused when it follows I)
        Ival JI, JK;
#if defined(__CUDACC__) && !defined(FIN_EMU)
        if (G > 1) {  // one rank row per lane of the group, results exchanged
            const unsigned sub = coop_sub<G>();
            const idx_t pos[4] = {I.l, I.r + 1, Ik.l, Ik.r + 1};
            idx_t r[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int owner = q % G;
                const idx_t v = (int)sub == owner ? step_c(ix, pos[q], c) : 0u;
                r[q] = coop_u32<G>(true, v, owner);
            }
            JI.l = r[0]; JI.r = r[1] - 1;
            JK.l = r[2]; JK.r = r[3] - 1;
        } else
#endif
        {
            JI = extend(ix, I, c);
            JK = extend(ix, Ik, c);
        }
        if (COUNT && both) rc.c_rank += 2;
        if (both && is_empty(JI)) {         // :116-126 contract the finimizer interval from the left and retry
            kmer_start = ++start;
            if (start > end) {
                I.l = 0;
                I.r = N - 1;
                Ik = I;
                emit = false;
                st = ST_EMIT;
            } else {
                set_drop_I(end - start);
                contrI = true;
            }
        } else {
            if (both) I = JI;
            if (both && start == kmer_start) {  // :141-143
                Ik = I;
                st = ST_POST;
            } else {                            // :132-140
                if (COUNT) rc.c_rank += 2;
                if (is_empty(JK)) {
                    kmer_start++;
                    set_drop_K(end - kmer_start);
                    contrK = true;
                    st = ST_EXTK;
                } else {
                    Ik = JK;
                    st = ST_POST;
                }
            }
        }
    }

    FIN_PHASE_FENCE();
    // ---- B: both intervals stand for this character.  Decisions first (pure arithmetic), then every load of this
    //         phase and of the drops in one batch, so that the chain waits for memory once, not five times.
    const bool post = st == ST_POST;
    bool shrink0 = false;  // the shrink loop of this character starts in this iteration
    const bool single = post && Ik.l == Ik.r;  // :167 rightmost unitig start among the singleton k-mer intervals
    const idx_t ucol = Ik.l;
    if (post) {
        emit = end - kmer_start + 1 == k;  // :170
        if (emit) set_drop_K(k - 1);       // :181, independent of the shrink below
        if (I.l == I.r) {                  // :145-154 shrink to the shortest unique suffix (closed form in phase F)
            shrink0 = true;
            set_drop_I(1);                 // the real new_len comes out of the two LCS bytes loaded below
            st = ST_SHRINK;
        } else {
            st = ST_EMIT;
        }
    }
    // ---- F: one word step of every pending drop_first_char (common.hh:38-48)
    if (pendI && lenI <= 0) {  // :39-41 nothing left of the string: the whole range
        I.l = 0;
        I.r = N - 1;
        pendI = 0;
    }
    if (pendK && lenK <= 0) {
        Ik.l = 0;
        Ik.r = N - 1;
        pendK = 0;
    }
    {
        const uint64_t* __restrict__ lw = reinterpret_cast<const uint64_t*>(ix.lines);
        const uint32_t* __restrict__ l32 = reinterpret_cast<const uint32_t*>(ix.lines);
        const idx_t il = I.l, ir = I.r + 1, kl = Ik.l, kr = Ik.r + 1;
        uint32_t uw = 0;
        uint64_t w_il = 0, w_ir = 0, w_kl = 0, w_kr = 0;
        uint32_t j_il = il & 7u, j_ir = ir & 7u, j_kl = kl & 7u, j_kr = kr & 7u;  // byte of LCS[il] .. LCS[kr] in its word / window
#if defined(__CUDACC__) && !defined(FIN_EMU)
        if (G > 1) {  // the five loads of the phase dealt out over the lanes of the group
            const int sub = (int)coop_sub<G>();
            const bool c_il = pendI & kDropLeft, c_ir = pendI & kDropRight, c_kl = pendK & kDropLeft, c_kr = pendK & kDropRight;
            uint32_t m_uw = 0;
            uint64_t m_il = 0, m_ir = 0, m_kl = 0, m_kr = 0;
            if (single && sub == 0) m_uw = __ldg(l32 + (((ucol >> 6) << 5) | ((ucol >> 3) & 4u) | 3u));
            if (c_il && sub == 0 % G) m_il = __ldg(lw + lcs_word_index(il));
            if (c_ir && sub == 1 % G) m_ir = __ldg(lw + lcs_word_index(ir));
            if (c_kl && sub == 2 % G) m_kl = __ldg(lw + lcs_word_index(kl));
            if (c_kr && sub == 3 % G) m_kr = __ldg(lw + lcs_word_index(kr));
            uw = coop_u32<G>(single, m_uw, 0);
            w_il = coop_u64<G>(c_il, m_il, 0 % G);
            w_ir = coop_u64<G>(c_ir, m_ir, 1 % G);
            w_kl = coop_u64<G>(c_kl, m_kl, 2 % G);
            w_kr = coop_u64<G>(c_kr, m_kr, 3 % G);
        } else
#endif
        {
            if (single) uw = __ldg(l32 + (((ucol >> 6) << 5) | ((ucol >> 3) & 4u) | 3u));
#if FIN_LCS_WIN4
            const uint32_t a_il = lcs_win_left_start(il), a_ir = lcs_win_right_start(ir);
            j_il = (il & 63u) - a_il; j_ir = (ir & 63u) - a_ir;
            if (pendI & kDropLeft) w_il = lcs_win_load(l32, il, a_il);
            if (pendI & kDropRight) w_ir = lcs_win_load(l32, ir, a_ir);
            const uint32_t a_kl = lcs_win_left_start(kl), a_kr = lcs_win_right_start(kr);
            j_kl = (kl & 63u) - a_kl; j_kr = (kr & 63u) - a_kr;
            if (pendK & kDropLeft) w_kl = lcs_win_load(l32, kl, a_kl);
            if (pendK & kDropRight) w_kr = lcs_win_load(l32, kr, a_kr);
#else
            if (pendI & kDropLeft) w_il = __ldg(lw + lcs_word_index(il));
            if (pendI & kDropRight) w_ir = __ldg(lw + lcs_word_index(ir));
            if (pendK & kDropLeft) w_kl = __ldg(lw + lcs_word_index(kl));
            if (pendK & kDropRight) w_kr = __ldg(lw + lcs_word_index(kr));
#endif
        }
        if (shrink0) {
            // The reference's shrink loop (:146-154) drops one character at a time while the interval stays the single
            // column l: new_len = L0-1, L0-2, ... and a drop leaves {l} alone exactly when LCS[l] < new_len and
            // LCS[l+1] < new_len (LCS[0] = 0 and the pad behind the last column = 0 stand for the two bound checks).
            // So the loop ends at new_len = min(L0 - 1, max(LCS[l], LCS[l+1])), its last candidate is one longer, and
            // only that last drop moves the interval: one word step instead of one iteration per dropped character.
            const int lcs_l = (int)((w_il >> (8u * j_il)) & 0xFFu), lcs_r = (int)((w_ir >> (8u * j_ir)) & 0xFFu);
            const int L0 = end - start + 1;
            const int nl = min(L0 - 1, max(lcs_l, lcs_r));
            if (COUNT) {  // the drops that left {l} alone
                const unsigned n = (unsigned)(L0 - 1 - nl) * ((il > 0 ? 1u : 0u) + (I.r < N - 1 ? 1u : 0u));
                rc.c_lcs += n;
                rc.c_skip_lcs += n;
            }
            cur = cand_pack(nl + 1, I.l, end);
            start = end - nl + 1;
            lenI = nl;
            if (nl <= 0) {  // :39-41
                I.l = 0;
                I.r = N - 1;
                pendI = 0;
            }
        }
        // The contraction loops (:116-126, :136-140) drop one character and retry the extension.  A drop that leaves the
        // interval [l, r] as it is (LCS[l] and LCS[r+1] both below new_len) leads to the same empty extension, so the
        // reference just goes on to the next shorter length: jump straight to the first length that moves the interval,
        // new_len = min(new_len, max(LCS[l], LCS[r+1])) (typically ~18 of the ~19 contractions of a k-mer interval
        // after a read error).
        if (contrI && pendI) {
            const int lcs_l = (int)((w_il >> (8u * j_il)) & 0xFFu), lcs_r = (int)((w_ir >> (8u * j_ir)) & 0xFFu);
            const int nl = min(lenI, max(lcs_l, lcs_r));
            const int skipped = lenI - nl;
            if (COUNT) {
                const unsigned n = (unsigned)skipped * ((il > 0 ? 1u : 0u) + (I.r < N - 1 ? 1u : 0u));
                rc.c_lcs += n;
                rc.c_skip_lcs += n;
                rc.c_rank += 2u * (unsigned)skipped;
                rc.c_skip_rank += 2u * (unsigned)skipped;
            }
            start += skipped;
            kmer_start = start;
            lenI = nl;
            if (nl <= 0) {
                I.l = 0;
                I.r = N - 1;
                pendI = 0;
            }
        }
        if (contrK && pendK) {
            const int lcs_l = (int)((w_kl >> (8u * j_kl)) & 0xFFu), lcs_r = (int)((w_kr >> (8u * j_kr)) & 0xFFu);
            // never past the finimizer interval's own length (in the parity domain the k-mer interval has moved by then)
            const int nl = max(min(lenK, max(lcs_l, lcs_r)), min(lenK, end - start));
            const int skipped = lenK - nl;
            if (COUNT) {
                const unsigned n = (unsigned)skipped * ((kl > 0 ? 1u : 0u) + (Ik.r < N - 1 ? 1u : 0u));
                rc.c_lcs += n;
                rc.c_skip_lcs += n;
                rc.c_rank += 2u * (unsigned)skipped;
                rc.c_skip_rank += 2u * (unsigned)skipped;
            }
            kmer_start += skipped;
            lenK = nl;
            if (nl <= 0) {
                Ik.l = 0;
                Ik.r = N - 1;
                pendK = 0;
            }
        }
        if (single) {
            if (COUNT) rc.c_ustart++;
            if ((uw >> (ucol & 31)) & 1u) {
                bestU_pos = end;
                bestU_col = ucol;
            }
        }
        // all four word steps computed by every lane, committed by select: no divergent regions in this phase
        bool f_il, f_ir, f_kl, f_kr;
        const idx_t n_il = drop_left_window<WIDE>(w_il, I.l, j_il, lenI, f_il), n_ir = drop_right_window<WIDE>(w_ir, I.r, j_ir, lenI, f_ir);
        const idx_t n_kl = drop_left_window<WIDE>(w_kl, Ik.l, j_kl, lenK, f_kl), n_kr = drop_right_window<WIDE>(w_kr, Ik.r, j_kr, lenK, f_kr);
        const bool d_il = pendI & kDropLeft, d_ir = pendI & kDropRight, d_kl = pendK & kDropLeft, d_kr = pendK & kDropRight;
        if (COUNT) {
            if (d_il) rc.c_lcs += (I.l - n_il) + (f_il && n_il > 0 ? 1u : 0u);
            if (d_ir) rc.c_lcs += (n_ir - I.r) + (f_ir && n_ir < N - 1 ? 1u : 0u);
            if (d_kl) rc.c_lcs += (Ik.l - n_kl) + (f_kl && n_kl > 0 ? 1u : 0u);
            if (d_kr) rc.c_lcs += (n_kr - Ik.r) + (f_kr && n_kr < N - 1 ? 1u : 0u);
        }
        I.l = d_il ? n_il : I.l;
        I.r = d_ir ? n_ir : I.r;
        Ik.l = d_kl ? n_kl : Ik.l;
        Ik.r = d_kr ? n_kr : Ik.r;
        pendI &= ~((d_il && f_il ? kDropLeft : 0u) | (d_ir && f_ir ? kDropRight : 0u));
        pendK &= ~((d_kl && f_kl ? kDropLeft : 0u) | (d_kr && f_kr ? kDropRight : 0u));
    }
    FIN_PHASE_FENCE();

    // ---- D: shrink loop condition (:154) and the candidate window (:155-164)
    if (st == ST_SHRINK && pendI == 0) {
        if (I.l == I.r) {
            cur = cand_pack(end - start + 1, I.l, end);
            start++;
            set_drop_I(end - start + 1);
        } else {
            // monotone window: drop the candidates that are larger than the new one, then append it.  Front and
            // back live in registers, so only a pop touches the slots.
            // (the entry below the back is cached too: the usual single pop then ends without waiting for a slot load)
            while (dq_n > 0 && bk > cur) {
                dq_n--;
#if FIN_DQ_BK2
                bk = bk2;
                if (dq_n > 1) bk2 = dq[(dq_head + dq_n - 2) & (KCAP - 1)];
#else
                if (dq_n > 0) bk = dq[(dq_head + dq_n - 1) & (KCAP - 1)];
#endif
            }
            // an empty ring restarts at slot 0: the live entries then stay in the first few slots instead of rotating
            // through all KCAP of them, which keeps the thread-local footprint (and its share of the L1) small
            if (dq_n == 0) dq_head = 0;
            if (dq_n == KCAP) {  // only entries that left the window can fill the ring: expire them now
                while (dq_n > 0 && cand_start(fr) < kmer_start) {
                    dq_head = (dq_head + 1) & (KCAP - 1);
                    dq_n--;
                    if (dq_n > 0) fr = dq[dq_head];
                }
                if (dq_n == KCAP) {  // cannot happen for live entries; keeps memory safe out of domain
                    rc.err |= kErrWindowUnderflow;
                    dq_n--;
                    bk = bk2 = 0;
                }
            }
            dq[(dq_head + dq_n) & (KCAP - 1)] = cur;
            dq_n++;
            bk2 = bk;
            bk = cur;
            if (dq_n == 1) fr = cur;
            st = ST_EMIT;
        }
    }

    FIN_PHASE_FENCE();
    // ---- E: emission (:170-182), resolve (FinimizerIndex.hh:148-183), next character
    bool rec_want = false, strand_done = false;  // runs mode: a run was closed by this position / the strand ends here
    RunRec rec{0, 0, 0, 0};
    if (st == ST_EMIT && (pendI | pendK) == 0) {
        if (COUNT) rc.c_bases++;
        bool present = false;
        uint64_t w = 0;
        // candidates that begin before the window are gone for good (kmer_start never decreases); the reference pops
        // them at its next emission (:173-176), popping them at every character is unobservable
        while (dq_n > 0 && cand_start(fr) < kmer_start) {
            dq_head = (dq_head + 1) & (KCAP - 1);
            dq_n--;
            if (dq_n > 0) fr = dq[dq_head];
        }
        if (dq_n == 0) dq_head = 0;
        if (emit) {
            present = true;
            if (dq_n > 0) {
                w = fr;
            } else {  // the reference would read a stale slot here (BoundedDeque.hh:28-34)
                rc.err |= kErrWindowUnderflow;
                present = false;
            }
            kmer_start++;
        }
        if (end >= k - 1) {
            if (COUNT) rc.c_kmers++;
            // the previous position's answer (meaningful while a run is open)
            const uint32_t p_uid = walk_uid;
            const idx_t p_pos = walk_g - (idx_t)(k - 1) - walk_ustart;
            bool hit = false, walked = false;
            if (walking) {  // walk_in_unitigs: next query base against the next text base
                const idx_t tp = walk_g + 1;
                bool match = false;
                if (tp < walk_uend) {
                    if ((tp & 31) == 0) tword = __ldg(ix.text + (tp >> 5));
                    if (COUNT && (walk_j & 31) == 0) rc.c_walk++;  // one 32-base compare of the reference (:72-100)
                    match = (int)((tword >> (2 * (tp & 31))) & 3) == c;
                }
                if (match) {
                    walk_g = tp;
                    walk_j++;
                    hit = walked = true;
                } else {
                    walking = false;
                }
            }
            if (!hit && present) {
                if (COUNT) rc.c_anchor++;
                idx_t g;
                bool ok = true;
                if (bestU_pos >= cand_end(w)) {  // lookup_from_branch_dictionary (common.hh:61-67)
                    if (COUNT) rc.c_branch++;
                    const uint32_t u = rank1(ix.ustart, bestU_col);
                    const idx_t us = u > 0 ? __ldg(ix.ends + (u - 1 < ix.n_unitigs ? u - 1 : ix.n_unitigs - 1)) : 0;
                    g = us + (idx_t)(k - 1) + (idx_t)(end - bestU_pos);
                } else {  // lookup_from_finimizer_dictionary (common.hh:69-72)
                    const uint32_t id = rank1(ix.fmin, cand_colex(w));
                    if (id >= ix.n_fmin) {
                        ok = false;
                        g = 0;
                    } else {
                        g = __ldg(ix.offsets + id) + (idx_t)(end - cand_end(w));
                    }
                }
                const idx_t gs = g - (idx_t)(k - 1);
                if (!ok || g < (idx_t)(k - 1) || gs >= ix.text_len) {
                    rc.err |= kErrLookupRange;
                } else {  // global_offset_to_local_offset (PackedStrings.hh:91-100)
                    walk_uid = rank1(ix.tstart, gs + 1) - 1;
                    walk_ustart = walk_uid > 0 ? __ldg(ix.ends + walk_uid - 1) : 0;
                    walk_uend = __ldg(ix.ends + walk_uid);
                    walk_g = g;
                    hit = true;
                    walking = g < walk_uend;
                    walk_j = 0;
                    if (walking && g + 1 < walk_uend) tword = __ldg(ix.text + ((g + 1) >> 5));
                }
            }
            // merge (search_fmin.hh:54-60): the forward answer wins; the reverse strand reports position
            // i = nk - 1 - j only where the forward strand's found mask has no bit
            const int j = end - (k - 1);
            const int i = strand ? nk - 1 - j : j;
            bool take = hit;
            if (strand == 0) {
                fm |= (uint32_t)hit << (j & 31);
                if ((j & 31) == 31 && j != nk - 1) {  // the last word stays in the register for the reverse strand
                    if (leader) out.fmask[fmw + (uint32_t)(j >> 5)] = fm;
                    fm = 0;
                }
            } else {
                take = hit && !((fm >> (i & 31)) & 1u);
                if ((i & 31) == 0 && i > 0 && (sflags & kStrandFwd)) fm = out.fmask[fmw + (uint32_t)(i >> 5) - 1u];  // G > 1: written by the leader long ago (a strand earlier)
            }
            rc.found[0] += hit && strand == 0 ? 1u : 0u;
            rc.found[1] += hit && strand != 0 ? 1u : 0u;
            rc.positive += take ? 1u : 0u;
            const idx_t pos = walk_g - (idx_t)(k - 1) - walk_ustart;
            if (!RUNS) {
                HitPair<HitT> h;
                h.unitig = take ? (HitT)walk_uid : (HitT)-1;
                h.pos = take ? (HitT)pos : (HitT)-1;
                // the first streamed strand writes every position, the reverse strand after a forward one only its own hits
                if ((strand == 0 || take || !(sflags & kStrandFwd)) && leader) o[i] = h;
            } else {
                const bool cont = take && walked && run_len > 0;
                if (run_len > 0 && !cont) {  // the open run ended at the previous position j - 1
                    rec_want = true;
                    rec.len_rc = (uint32_t)run_len | (strand ? kRunRc : 0u);
                    rec.unitig = p_uid;
                    rec.i0 = strand ? (uint32_t)(i + 1) : (uint32_t)(j - run_len);
                    rec.pos0 = strand ? (uint32_t)p_pos : (uint32_t)p_pos - (uint32_t)(run_len - 1);
                }
                run_len = take ? (cont ? run_len + 1 : 1) : 0;
            }
        }
        end++;
        if (end == len) {
            strand_done = true;
        } else {
            fresh = true;
            st = ST_EXT;
        }
    }
    if (RUNS) {
        FIN_PHASE_FENCE();
        // the run that is still open when the strand ends (its last position is the strand's last k-mer)
        const bool tail = strand_done && run_len > 0;
        if (run_any(rec_want || tail)) {  // rare: ~2 records per read
            run_append(out.sink, rec_want, ((uint64_t)rid << kRunKeyPosBits) | rec.i0, rec);
            if (tail) {
                const uint32_t pos = walk_g - (idx_t)(k - 1) - walk_ustart;
                rec.len_rc = (uint32_t)run_len | (strand ? kRunRc : 0u);
                rec.unitig = walk_uid;
                rec.i0 = strand ? 0u : (uint32_t)(nk - run_len);
                rec.pos0 = strand ? pos : pos - (uint32_t)(run_len - 1);
            }
            run_append(out.sink, tail, ((uint64_t)rid << kRunKeyPosBits) | rec.i0, rec);
        }
    }
    if (strand_done) {
        strand++;
        if (strand < n_strands && (sflags & kStrandRc)) begin_strand(ix);
        else st = ST_IDLE;
    }
}

#ifdef __CUDACC__
// Persistent warps: a lane that finishes its read takes the next one from a global counter (one atomic per
// group of lanes that run dry in the same iteration), so no lane idles while reads remain.
#ifndef FIN_SEARCH_MIN_BLOCKS
// 7 resident blocks = 72 registers and no spills.  8 blocks (64 registers) were the optimum until the closed forms and the LCS
// windows added state: with 56-68 bytes of spills they now lose 4-9 % (profiles/r2c_sweep_resident_blocks.log); 9-12 blocks lose
// 4-30 % (profiles/r2_sweep_occupancy.log), 6 blocks gain nothing over 7.
#define FIN_SEARCH_MIN_BLOCKS 7
#endif
template <int KCAP, class HitT, bool COUNT, bool RUNS, int G = 1>
__global__ void __launch_bounds__(FIN_SEARCH_THREADS, FIN_SEARCH_MIN_BLOCKS) search_reads_kernel(const IndexView ix, const uint64_t* __restrict__ packed,
                                                           const uint64_t* __restrict__ read_offsets, const uint64_t n_reads,
                                                           const SearchOut<HitT> out, SearchTotals* __restrict__ totals,
                                                           const int n_strands) {
    const unsigned full = 0xFFFFFFFFu;
    const unsigned lane = threadIdx.x & 31u;
    const bool leader = (lane & (unsigned)(G - 1)) == 0;  // first lane of its group of G (G = 1: every lane)
    Chain<KCAP, HitT, COUNT, RUNS, G> ch;
    uint64_t dq[KCAP];
    ReadCounters rc;
    bool dry = false;  // the counter ran past the last read
    while (true) {
        const bool want = ch.st == ST_IDLE && !dry;
        const unsigned m = __ballot_sync(full, want && leader);  // one read per group
        if (m) {
            const int first = __ffs((int)m) - 1;
            unsigned long long base = 0;
            if ((int)lane == first) base = atomicAdd(&totals->next_read, (unsigned long long)__popc(m));
            base = __shfl_sync(full, base, first);
            const unsigned my_leader = lane & ~(unsigned)(G - 1);
            const uint64_t rid = base + (uint64_t)__popc(m & ((1u << my_leader) - 1u));  // the same value on the G lanes of a group
            if (want) {
                if (rid < n_reads) {
                    const uint64_t b0 = __ldg(read_offsets + rid);
                    const uint64_t len64 = __ldg(read_offsets + rid + 1) - b0;
                    ch.begin_read(ix, b0, len64, rid, out, rc, n_strands);
                } else {
                    dry = true;
                }
            }
        }
        if (__all_sync(full, ch.st == ST_IDLE)) {
            if (__all_sync(full, dry)) break;
            continue;
        }
        ch.step(ix, packed, rc, n_strands, dq, out);
    }
    if (!leader) {  // the replicas tallied the same things as their leader
        const unsigned err = rc.err;
        rc = ReadCounters();
        rc.err = err;
    }

    // ---- per-warp reduction of the counters, one atomic per warp and counter
    unsigned long long v_ff = rc.found[0], v_fr = rc.found[1], v_pos = rc.positive, v_q = rc.n_q;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        v_ff += __shfl_down_sync(full, v_ff, s);
        v_fr += __shfl_down_sync(full, v_fr, s);
        v_pos += __shfl_down_sync(full, v_pos, s);
        v_q += __shfl_down_sync(full, v_q, s);
    }
    const unsigned any_err = __reduce_or_sync(full, rc.err);
    if (lane == 0) {
        if (v_ff) atomicAdd(&totals->found_fwd, v_ff);
        if (v_fr) atomicAdd(&totals->found_rc, v_fr);
        if (v_pos) atomicAdd(&totals->positive, v_pos);
        if (v_q) atomicAdd(&totals->n_queries, v_q);
        if (any_err) atomicOr(&totals->error_flags, any_err);
    }
    if (COUNT) {
        unsigned long long cs[8] = {rc.c_rank, rc.c_lcs, rc.c_ustart, rc.c_anchor, rc.c_branch, rc.c_walk, rc.c_bases, rc.c_kmers};
        unsigned long long* dst = &totals->n_rank;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            unsigned long long v = cs[i];
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) v += __shfl_down_sync(full, v, s);
            if (lane == 0 && v) atomicAdd(dst + i, v);
        }
        unsigned long long sk[2] = {rc.c_skip_lcs, rc.c_skip_rank};
#pragma unroll
        for (int i = 0; i < 2; i++) {
            unsigned long long v = sk[i];
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) v += __shfl_down_sync(full, v, s);
            if (lane == 0 && v) atomicAdd(&totals->n_skip_lcs + i, v);
        }
    }
}

// ---- strand prefilter: one thread per read; dense mode also answers the reads no strand of which is streamed ------
template <class HitT>
static __global__ void strand_filter_kernel(const IndexView ix, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ read_offsets,
                                            const uint64_t n_reads, const int n_strands, uint8_t* __restrict__ flags,
                                            const uint64_t* __restrict__ out_offsets, HitPair<HitT>* __restrict__ hits,
                                            SearchTotals* __restrict__ count_into) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reads) return;
    const uint64_t b0 = read_offsets[r], len = read_offsets[r + 1] - b0;
    unsigned lookups = 0;
    unsigned f = strand_candidates(ix, packed, b0, len, lookups);
    if (count_into && lookups) atomicAdd(&count_into->n_filter, (unsigned long long)lookups);
    if (n_strands < 2) f &= kStrandFwd;
    flags[r] = (uint8_t)f;
    if (hits != nullptr && f == 0) {
        HitPair<HitT> none;
        none.unitig = none.pos = (HitT)-1;
        for (uint64_t i = out_offsets[r]; i < out_offsets[r + 1]; i++) hits[i] = none;
    }
}

// ---- after the sort by (read, i0): CSR offsets of the runs per read --------------------------------------------
// run_offsets[r] = base + index of the first record of a read >= r; run_offsets[n_reads] = base + n_runs.
static __global__ void run_bounds_kernel(const uint64_t* __restrict__ keys, const uint64_t n_runs, const uint64_t n_reads,
                                         const uint64_t base, uint64_t* __restrict__ run_offsets) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_runs) return;
    const uint64_t lo = t == 0 ? 0 : (keys[t - 1] >> kRunKeyPosBits) + 1;       // first read not yet covered
    const uint64_t hi = t == n_runs ? n_reads : (keys[t] >> kRunKeyPosBits);   // this record's read (or the end)
    for (uint64_t r = lo; r <= hi; r++) run_offsets[r] = base + t;
}

// ---- runs -> dense table on the device (fin_search_device callers that want the table, tests) -------------------
template <class HitT>
static __global__ void expand_runs_kernel(const uint64_t* __restrict__ keys, const RunRec* __restrict__ recs, const uint64_t n_runs,
                                          const uint64_t* __restrict__ out_offsets, HitPair<HitT>* __restrict__ out) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_runs) return;
    const RunRec r = recs[t];
    HitPair<HitT>* o = out + out_offsets[keys[t] >> kRunKeyPosBits] + r.i0;
    const uint32_t len = r.len_rc & ~kRunRc;
    const bool rcs = r.len_rc & kRunRc;
    for (uint32_t d = 0; d < len; d++) {
        HitPair<HitT> h;
        h.unitig = (HitT)r.unitig;
        h.pos = (HitT)(rcs ? r.pos0 - d : r.pos0 + d);
        o[d] = h;
    }
}
#endif  // __CUDACC__

}  // namespace fin
