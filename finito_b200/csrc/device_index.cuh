// Device-resident layout of the FinimizerIndex and the device functions every kernel shares.
//
// Layout (all little-endian, all arrays padded so that position n itself is addressable):
//   rank sectors — the unit is one 32-byte sector {u64 ones_before_block, u64 bits[3]} covering a block of
//                  192 bit positions.  One rank = one sector + at most three masked popcounts.
//       subset matrix : sector (4*block + c) for c in A,C,G,T, so the four rows of a block share one
//                       128-byte line (replaces 4 x sdsl::bit_vector + 4 x rank_support_v5,
//                       SBWT/include/sbwt/SubsetMatrixRank.hh:18-37, sdsl/rank_support_v5.hpp:116-134)
//       Ustart, fmin  : sector (block)                 (FinimizerIndex.hh:111-115; rank rebuilt at load :219,:235)
//       tstart        : sector (block) over TEXT positions, bit set at the first base of every unitig;
//                       rank replaces std::upper_bound over unitigs.ends (PackedStrings.hh:91-100)
//   lcs     — one byte per SBWT column (k <= 255)      (sdsl::int_vector<> LCS, FinimizerIndex.hh:109)
//   offsets — u32 per finimizer, in column order       (global_offsets, FinimizerIndex.hh:113)
//   ends    — u32 exclusive end of every unitig        (unitigs.ends, PackedStrings.hh:29)
//   text    — 2 bits per base, 32 bases per u64, same bit order as sdsl::int_vector<2> (PackedStrings.hh:28)
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

struct IndexView {
    const ulonglong2* __restrict__ matrix;  // 2 x ulonglong2 per sector, 4 sectors per block
    const ulonglong2* __restrict__ ustart;  // 2 x ulonglong2 per sector
    const ulonglong2* __restrict__ fmin;
    const ulonglong2* __restrict__ tstart;
    const uint8_t* __restrict__ lcs;
    const uint32_t* __restrict__ offsets;
    const uint32_t* __restrict__ ends;
    const uint64_t* __restrict__ text;
    idx_t C[4];
    idx_t n_nodes;
    idx_t text_len;
    uint32_t n_unitigs;
    uint32_t n_fmin;
    int32_t k;
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

FIN_D bool bit1(const ulonglong2* __restrict__ sectors, idx_t p) {
    const idx_t blk = p / kRankBlockBits;
    const int q = (int)(p - blk * kRankBlockBits);
    const uint64_t* w = reinterpret_cast<const uint64_t*>(sectors + 2 * (size_t)blk);
    return (__ldg(w + 1 + (q >> 6)) >> (q & 63)) & 1ULL;
}

// SBWT::update_sbwt_interval for one character (SBWT/include/sbwt/SBWT.hh:355-370):
// I' = [C[c] + rank_c(l), C[c] + rank_c(r+1) - 1].  Both ranks come from one sector when l and r+1
// share a 192-column block, which is the common case once the interval is narrow.
FIN_D Ival extend(const IndexView& ix, const Ival I, const int c) {
    const idx_t lp = I.l, rp = I.r + 1;
    const idx_t bl = lp / kRankBlockBits, br = rp / kRankBlockBits;
    const ulonglong2* sl = ix.matrix + 2 * ((size_t)bl * 4 + c);
    const ulonglong2 a = __ldg(sl), b = __ldg(sl + 1);
    const uint32_t rl = sector_count(a, b, (int)(lp - bl * kRankBlockBits));
    uint32_t rr;
    if (br == bl) {
        rr = sector_count(a, b, (int)(rp - br * kRankBlockBits));
    } else {
        const ulonglong2* sr = ix.matrix + 2 * ((size_t)br * 4 + c);
        const ulonglong2 a2 = __ldg(sr), b2 = __ldg(sr + 1);
        rr = sector_count(a2, b2, (int)(rp - br * kRankBlockBits));
    }
    Ival J;
    J.l = ix.C[c] + rl;
    J.r = ix.C[c] + rr - 1;  // rr == rl gives r == l - 1: empty
    return J;
}

// C[c] >= 1 (SBWT.hh:279), so an empty result has l = r + 1 without wrap-around
FIN_D bool is_empty(const Ival I) { return I.l > I.r; }

#endif  // __CUDACC__ || FIN_EMU

}  // namespace fin
