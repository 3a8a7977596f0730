// Device-resident layout of the FinimizerIndex and the device functions every kernel shares.
//
// Layout (all little-endian, all arrays padded so that position n itself is addressable):
//   lines   — one 128-byte line per 64 SBWT columns; everything the streaming search touches for a column
//             interval lives in the line(s) of that interval, so one step of a chain costs one new line:
//               bytes   0..63  four 16-byte rows, one per A,C,G,T:
//                              {u64 bits of the row for these 64 columns,
//                               u32 C[c] + ones of the row before the line,
//                               u32 extra: row A/C = Ustart bits 0..31 / 32..63, row G/T = fmin bits 0..31 / 32..63}
//               bytes 64..127  LCS of the 64 columns, one byte each (k <= 255)
//             rank_c(p) and with it SBWT::update_sbwt_interval is ONE 16-byte load and a popcount
//             (replaces 4 x sdsl::bit_vector + 4 x rank_support_v5 + the C array,
//              SBWT/include/sbwt/SubsetMatrixRank.hh:18-37, sdsl/rank_support_v5.hpp:116-134, SBWT.hh:355-370;
//              the LCS int_vector, FinimizerIndex.hh:109; the Ustart bit test of common.hh:167)
//   rank sectors — one 32-byte sector {u64 ones_before_block, u64 bits[3]} per 192 bit positions, for the
//             three rank queries of the (rare) anchor lookups:
//       ustart, fmin : over SBWT columns            (FinimizerIndex.hh:111-115; rank rebuilt at load :219,:235)
//       tstart       : over TEXT positions, bit set at the first base of every unitig; its rank replaces
//                      std::upper_bound over unitigs.ends (PackedStrings.hh:91-100)
//   offsets — u32 per finimizer, in column order       (global_offsets, FinimizerIndex.hh:113)
//   ends    — u32 exclusive end of every unitig        (unitigs.ends, PackedStrings.hh:29)
//   text    — 2 bits per base, 32 bases per u64, same bit order as sdsl::int_vector<2> (PackedStrings.hh:28)
//   smer    — 4^s bits: which s-mers (s ~ log4(text length) + 1, <= 16) occur in the text.  Not part of the reference:
//             a k-mer of the index has all its s-mers in the text, so a read strand without k - s + 1 consecutive
//             present s-mers cannot hold a k-mer of the index and its streaming search is skipped (all (-1,-1))
#pragma once
#include <cstdint>
#ifndef FIN_EMU
#include <cuda_runtime.h>
#define FIN_HD __host__ __device__ __forceinline__
#define FIN_D __device__ __forceinline__
#else
// tests/emu compiles the device functions for the host (debugging aid of the test-suite only, never shipped)
#define FIN_HD inline
#define FIN_D inline
#endif

namespace fin {

typedef uint32_t idx_t;  // SBWT column / text position.  This build supports n_nodes, text_len < 2^32 - 256.

constexpr int kRankBlockBits = 192;
constexpr int kLineColumns = 64;   // columns per 128-byte line
constexpr int kLineU4 = 8;         // uint4 per line
constexpr int kLineU64 = 16;       // u64 per line
constexpr int kLineLcsU64 = 8;     // first u64 of the LCS half of a line

struct IndexView {
    const uint4* __restrict__ lines;        // 8 x uint4 per line
    const ulonglong2* __restrict__ ustart;  // 2 x ulonglong2 per sector
    const ulonglong2* __restrict__ fmin;
    const ulonglong2* __restrict__ tstart;
    const uint32_t* __restrict__ offsets;
    const uint32_t* __restrict__ ends;
    const uint64_t* __restrict__ text;
    idx_t n_nodes;
    idx_t text_len;
    uint32_t n_unitigs;
    uint32_t n_fmin;
    int32_t k;
    // strand prefilter (search_kernels.cuh): bit c of `smer` is set iff the s-mer with code c occurs in `text`;
    // nullptr = no prefilter
    const uint32_t* __restrict__ smer;
    int32_t smer_s;
};

#if defined(__CUDACC__) || defined(FIN_EMU)

struct Ival {  // inclusive SBWT interval; empty <=> l > r
    idx_t l, r;
};

FIN_D uint64_t low_mask(int n) {  // n in [0,64]
    return n >= 64 ? ~0ULL : ((1ULL << n) - 1ULL);
}

// ones in [block start, block start + q) of a sector, q in [0,192)
FIN_D uint32_t sector_count(const ulonglong2 a, const ulonglong2 b, int q) {
    const int q0 = min(q, 64), q1 = min(max(q - 64, 0), 64), q2 = max(q - 128, 0);
    return (uint32_t)a.x + __popcll(a.y & low_mask(q0)) + __popcll(b.x & low_mask(q1)) + __popcll(b.y & low_mask(q2));
}

// rank over a single-row sector array (Ustart / fmin / tstart): ones in [0, p)
FIN_D uint32_t rank1(const ulonglong2* __restrict__ sectors, idx_t p) {
    const idx_t blk = p / kRankBlockBits;
    const int q = (int)(p - blk * kRankBlockBits);
    const ulonglong2 a = __ldg(sectors + 2 * (size_t)blk), b = __ldg(sectors + 2 * (size_t)blk + 1);
    return sector_count(a, b, q);
}

// C[c] + rank_c(p): the column the edge labelled c out of the first column >= p leads to
FIN_D idx_t step_c(const IndexView& ix, const idx_t p, const int c) {
    const uint4 row = __ldg(ix.lines + (((p >> 6) << 3) | (uint32_t)c));  // 32-bit index: n_nodes / 8 < 2^29
    const uint64_t bits = (uint64_t)row.x | ((uint64_t)row.y << 32);
    return row.z + (idx_t)__popcll(bits & ((1ULL << (p & 63)) - 1ULL));
}

// SBWT::update_sbwt_interval for one character (SBWT/include/sbwt/SBWT.hh:355-370):
// I' = [C[c] + rank_c(l), C[c] + rank_c(r+1) - 1].  Two 16-byte loads, the same 32-byte sector whenever l and
// r+1 share a line, which is the common case once the interval is narrow.
FIN_D Ival extend(const IndexView& ix, const Ival I, const int c) {
    Ival J;
    J.l = step_c(ix, I.l, c);
    J.r = step_c(ix, I.r + 1, c) - 1;  // equal ranks give r == l - 1: empty
    return J;
}

// C[c] >= 1 (SBWT.hh:279), so an empty result has l = r + 1 without wrap-around
FIN_D bool is_empty(const Ival I) { return I.l > I.r; }

// Ustart[p] (common.hh:167) from the line of p
FIN_D bool ustart_bit(const IndexView& ix, const idx_t p) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(ix.lines);
    return (__ldg(w + (((p >> 6) << 5) | ((p >> 3) & 4u) | 3u)) >> (p & 31)) & 1u;
}

// LCS[p] from the line of p (build kernels; the search uses the word-parallel drop_step below)
FIN_D int lcs_at(const IndexView& ix, const idx_t p) {
    const uint8_t* b = reinterpret_cast<const uint8_t*>(ix.lines);
    return (int)__ldg(b + (((size_t)(p >> 6) << 7) | 64u | (p & 63u)));
}

#endif  // __CUDACC__ || FIN_EMU

}  // namespace fin
