// Kernels of the search-fmin path.
//   pack_reads_kernel   : ASCII -> 2 bits per base + validation      (SBWT/include/sbwt/SeqIO.hh:232 upper-casing,
//                                                                     common.hh:50-58 char_idx, :106-111 validation)
//   kmer_counts_kernel  : k-mer positions per read                   (search_fmin.hh:69)
//   search_reads_kernel : one thread streams one read, forward strand then reverse complement:
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
};

enum : unsigned { kErrBadBase = 1u, kErrWindowUnderflow = 2u, kErrLookupRange = 4u, kErrReadTooLong = 8u };

struct ReadCounters {  // per-thread tallies, reduced per warp at the end of the kernel
    unsigned found[2] = {0, 0}, positive = 0, n_q = 0, err = 0;
    unsigned c_rank = 0, c_lcs = 0, c_ustart = 0, c_anchor = 0, c_branch = 0, c_walk = 0, c_bases = 0, c_kmers = 0;
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

static __global__ void kmer_counts_kernel(const uint64_t* __restrict__ read_offsets, uint64_t n_reads, int k,
                                   uint64_t* __restrict__ counts) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reads) return;
    const uint64_t len = read_offsets[r + 1] - read_offsets[r];
    counts[r] = len >= (uint64_t)k ? len - (uint64_t)k + 1 : 0;
}

#endif  // __CUDACC__

// ---- drop_first_char (common.hh:38-48) ---------------------------------------------------------------------
template <bool COUNT>
FIN_D Ival drop_first_char(const IndexView& ix, const int new_len, Ival I, unsigned& n_lcs) {
    if (new_len <= 0) {
        I.l = 0;
        I.r = ix.n_nodes - 1;
        return I;
    }
    while (I.l > 0) {
        if (COUNT) n_lcs++;
        if ((int)__ldg(ix.lcs + I.l) >= new_len) I.l--;
        else break;
    }
    const idx_t last = ix.n_nodes - 1;
    while (I.r < last) {
        if (COUNT) n_lcs++;
        if ((int)__ldg(ix.lcs + I.r + 1) >= new_len) I.r++;
        else break;
    }
    return I;
}

template <class HitT> struct HitPair { HitT unitig, pos; };

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

// One read, both strands.  KCAP: power of two >= k, capacity of the candidate window (see DESIGN.md
// "candidate window": entries are expired eagerly, which is unobservable because candidate starts are
// non-decreasing along the deque, so at most k live entries exist).
template <int KCAP, class HitT, bool COUNT>
FIN_D void search_one_read(const IndexView& ix, const uint64_t* __restrict__ packed, const uint64_t b0, const uint64_t len64,
                           HitPair<HitT>* __restrict__ o, ReadCounters& rc, const int n_strands = 2) {
    unsigned (&found)[2] = rc.found;
    unsigned &positive = rc.positive, &n_q = rc.n_q, &err = rc.err;
    unsigned &c_rank = rc.c_rank, &c_lcs = rc.c_lcs, &c_ustart = rc.c_ustart, &c_anchor = rc.c_anchor, &c_branch = rc.c_branch,
             &c_walk = rc.c_walk, &c_bases = rc.c_bases, &c_kmers = rc.c_kmers;
    {
        const int k = ix.k;
        const idx_t N = ix.n_nodes;
        if (len64 > (uint64_t)kMaxReadLen) {
            err |= kErrReadTooLong;
        } else if (len64 >= (uint64_t)k) {
            const int len = (int)len64;
            const int nk = len - k + 1;
            n_q = (unsigned)nk;
            uint64_t dq[KCAP];

            for (int strand = 0; strand < n_strands; strand++) {
                int start = 0, kmer_start = 0;
                Ival I{0, N - 1}, Ik{0, N - 1};
                int dq_head = 0, dq_n = 0;
                int bestU_pos = -1;
                idx_t bestU_col = 0;
                // walk state (walk_in_unitigs): last answered k-mer ends at text position walk_g
                bool walking = false;
                idx_t walk_g = 0, walk_ustart = 0, walk_uend = 0;
                uint32_t walk_uid = 0;
                uint64_t tword = 0, qword = 0;
                int walk_j = 0;  // bases compared since the anchor (access counters only)

                for (int end = 0; end < len; end++) {
                    // ---- next base of this strand
                    int c;
                    if (strand == 0) {
                        const uint64_t b = b0 + (uint64_t)end;
                        if (end == 0 || (b & 31) == 0) qword = __ldg(packed + (b >> 5));
                        c = (int)((qword >> (2 * (b & 31))) & 3);
                    } else {
                        const uint64_t b = b0 + (uint64_t)(len - 1 - end);
                        if (end == 0 || (b & 31) == 31) qword = __ldg(packed + (b >> 5));
                        c = 3 - (int)((qword >> (2 * (b & 31))) & 3);
                    }
                    if (COUNT) c_bases++;

                    // ---- (1) finimizer interval: extend, contracting from the left while absent (common.hh:114-126)
                    Ival In = extend(ix, I, c);
                    if (COUNT) c_rank += 2;
                    while (is_empty(In)) {
                        kmer_start = ++start;
                        if (start > end) {
                            In.l = 0;
                            In.r = N - 1;
                            Ik = In;
                            break;
                        }
                        I = drop_first_char<COUNT>(ix, end - start, I, c_lcs);
                        In = extend(ix, I, c);
                        if (COUNT) c_rank += 2;
                        Ik = In;
                    }
                    I = In;
                    // ---- (2) k-mer interval (common.hh:132-143)
                    if (start != kmer_start) {
                        Ival J = extend(ix, Ik, c);
                        if (COUNT) c_rank += 2;
                        while (is_empty(J)) {
                            kmer_start++;
                            Ik = drop_first_char<COUNT>(ix, end - kmer_start, Ik, c_lcs);
                            J = extend(ix, Ik, c);
                            if (COUNT) c_rank += 2;
                        }
                        Ik = J;
                    } else {
                        Ik = I;
                    }
                    // ---- (3) unique: shrink to the shortest unique suffix and record it (common.hh:145-164)
                    if (I.l == I.r) {
                        uint64_t cur;
                        do {
                            cur = cand_pack(end - start + 1, I.l, end);
                            start++;
                            I = drop_first_char<COUNT>(ix, end - start + 1, I, c_lcs);
                        } while (I.l == I.r);
                        while (dq_n > 0 && cand_start(dq[dq_head]) < kmer_start) {  // eager expiry
                            dq_head = (dq_head + 1) & (KCAP - 1);
                            dq_n--;
                        }
                        while (dq_n > 0 && dq[(dq_head + dq_n - 1) & (KCAP - 1)] > cur) dq_n--;
                        if (dq_n == KCAP) {  // cannot happen for live entries; keeps memory safe out of domain
                            err |= kErrWindowUnderflow;
                            dq_n--;
                        }
                        dq[(dq_head + dq_n) & (KCAP - 1)] = cur;
                        dq_n++;
                    }
                    // ---- (4) rightmost unitig start among the singleton k-mer intervals (common.hh:167)
                    if (Ik.l == Ik.r) {
                        if (COUNT) c_ustart++;
                        if (bit1(ix.ustart, Ik.l)) {
                            bestU_pos = end;
                            bestU_col = Ik.l;
                        }
                    }
                    // ---- (5) a complete k-mer ends here (common.hh:170-182)
                    bool present = false;
                    uint64_t w = 0;
                    idx_t kmer_col = 0;
                    if (end - kmer_start + 1 == k) {
                        while (dq_n > 0 && cand_start(dq[dq_head]) < kmer_start) {
                            dq_head = (dq_head + 1) & (KCAP - 1);
                            dq_n--;
                        }
                        present = true;
                        kmer_col = Ik.l;
                        if (dq_n > 0) {
                            w = dq[dq_head];
                        } else {  // the reference would read a stale slot here (BoundedDeque.hh:28-34)
                            err |= kErrWindowUnderflow;
                            present = false;
                        }
                        kmer_start++;
                        Ik = drop_first_char<COUNT>(ix, k - 1, Ik, c_lcs);
                    }
                    (void)kmer_col;

                    // ---- resolve (FinimizerIndex.hh:148-183), in step with the stream
                    if (end >= k - 1) {
                        if (COUNT) c_kmers++;
                        HitT hit_u = (HitT)-1, hit_p = (HitT)-1;
                        bool hit = false;
                        if (walking) {  // walk_in_unitigs: next query base against the next text base
                            const idx_t tp = walk_g + 1;
                            bool match = false;
                            if (tp < walk_uend) {
                                if ((tp & 31) == 0) tword = __ldg(ix.text + (tp >> 5));
                                if (COUNT && (walk_j & 31) == 0) c_walk++;  // one 32-base compare of the reference (:72-100)
                                match = (int)((tword >> (2 * (tp & 31))) & 3) == c;
                            }
                            if (match) {
                                walk_g = tp;
                                walk_j++;
                                hit = true;
                            } else {
                                walking = false;
                            }
                        }
                        if (!hit && present) {
                            if (COUNT) c_anchor++;
                            idx_t g;
                            bool ok = true;
                            if (bestU_pos >= cand_end(w)) {  // lookup_from_branch_dictionary (common.hh:61-67)
                                if (COUNT) c_branch++;
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
                                err |= kErrLookupRange;
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
                        if (hit) {
                            hit_u = (HitT)walk_uid;
                            hit_p = (HitT)(walk_g - (idx_t)(k - 1) - walk_ustart);
                            found[strand]++;
                        }
                        const int j = end - (k - 1);
                        if (strand == 0) {
                            o[j].unitig = hit_u;
                            o[j].pos = hit_p;
                            if (hit) positive++;
                        } else if (hit) {  // merge: the forward answer wins (search_fmin.hh:54-60)
                            const int i = nk - 1 - j;
                            if (o[i].unitig == (HitT)-1) {
                                o[i].unitig = hit_u;
                                o[i].pos = hit_p;
                                positive++;
                            }
                        }
                    }
                }
            }
        }
    }
}

#ifdef __CUDACC__
template <int KCAP, class HitT, bool COUNT>
__global__ void __launch_bounds__(128) search_reads_kernel(const IndexView ix, const uint64_t* __restrict__ packed,
                                                           const uint64_t* __restrict__ read_offsets, const uint64_t n_reads,
                                                           const uint64_t* __restrict__ out_offsets,
                                                           HitPair<HitT>* __restrict__ out, SearchTotals* __restrict__ totals,
                                                           const int n_strands) {
    const uint64_t rid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    ReadCounters rc;
    if (rid < n_reads) {
        const uint64_t b0 = __ldg(read_offsets + rid);
        const uint64_t len64 = __ldg(read_offsets + rid + 1) - b0;
        search_one_read<KCAP, HitT, COUNT>(ix, packed, b0, len64, out + __ldg(out_offsets + rid), rc, n_strands);
    }
    const unsigned err = rc.err;
    const unsigned c_rank = rc.c_rank, c_lcs = rc.c_lcs, c_ustart = rc.c_ustart, c_anchor = rc.c_anchor, c_branch = rc.c_branch,
                   c_walk = rc.c_walk, c_bases = rc.c_bases, c_kmers = rc.c_kmers;
    const unsigned* found = rc.found;
    const unsigned positive = rc.positive, n_q = rc.n_q;

    // ---- per-warp reduction of the counters, one atomic per warp and counter
    const unsigned full = 0xFFFFFFFFu;
    unsigned long long v_ff = found[0], v_fr = found[1], v_pos = positive, v_q = n_q;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        v_ff += __shfl_down_sync(full, v_ff, s);
        v_fr += __shfl_down_sync(full, v_fr, s);
        v_pos += __shfl_down_sync(full, v_pos, s);
        v_q += __shfl_down_sync(full, v_q, s);
    }
    const unsigned any_err = __reduce_or_sync(full, err);
    if ((threadIdx.x & 31) == 0) {
        if (v_ff) atomicAdd(&totals->found_fwd, v_ff);
        if (v_fr) atomicAdd(&totals->found_rc, v_fr);
        if (v_pos) atomicAdd(&totals->positive, v_pos);
        if (v_q) atomicAdd(&totals->n_queries, v_q);
        if (any_err) atomicOr(&totals->error_flags, any_err);
    }
    if (COUNT) {
        unsigned long long cs[8] = {c_rank, c_lcs, c_ustart, c_anchor, c_branch, c_walk, c_bases, c_kmers};
        unsigned long long* dst = &totals->n_rank;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            unsigned long long v = cs[i];
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) v += __shfl_down_sync(full, v, s);
            if ((threadIdx.x & 31) == 0 && v) atomicAdd(dst + i, v);
        }
    }
}
#endif  // __CUDACC__

}  // namespace fin
