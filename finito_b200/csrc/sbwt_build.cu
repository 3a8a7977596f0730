// Plain-matrix SBWT construction on the device: what `sbwt build` (SBWT/src/CLI/sbwt_build.cpp:30 ->
// SBWT/include/sbwt/kmc_construct.hh:201-238, KMC + external-memory sort) and the reference's in-memory constructor
// (SBWT/include/sbwt/NodeBOSSInMemoryConstructor.hh:98-188) compute on the CPU, re-designed as sort + join passes:
//
//   1. every k-mer of the input as an integer key whose order is the colex order (last character most significant),
//      straight out of the 2-bit packed text (one funnel shift per k-mer);
//   2. cub radix sort + unique                                   -> the k-mer nodes in SBWT order
//   3. the same keys rotated by one character (prefix of k-1 characters in the high bits, last character in the low
//      two bits), sorted: the k-mers grouped by their (k-1)-prefix.  A (k-1)-suffix group S of the node list has the
//      outgoing label c iff the rotated list holds (S, c): one binary search per group, sorted queries
//      (NodeBOSSInMemoryConstructor.hh:120-165: edges live on the first node of a suffix group);
//   4. k-mers whose prefix is nobody's suffix have no incoming edge: all their proper prefixes become dummy nodes
//      ($-padded), sorted by (key, length), duplicates folded with their labels OR-ed (:98-118);
//   5. the two sorted lists are merged by rank (binary search of each element in the other list), and every node
//      sets its bits in the four rows and in suffix_group_starts;
//   6. host: C array, rank_support_v5 blobs (sdsl/rank_support_v5.hpp:65-109), k-mer prefix precalc table
//      (SBWT.hh:529-555) and the file (SBWT.hh:395-424) with the "plain-matrix" variant string sbwt_build.cpp writes.
//
// k <= 32 uses 64-bit keys, k <= 64 128-bit keys.  Bit-exact against the reference constructor (tests/test_gpu_sbwt.py).
#include <algorithm>
#include <chrono>
#include <cstring>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>

#include "engine.hpp"
#include "seqio.hpp"

namespace fin {

typedef unsigned __int128 u128;

template <class Key> __host__ __device__ __forceinline__ Key key_low_mask(int bits) {
    return bits >= (int)(8 * sizeof(Key)) ? ~(Key)0 : (((Key)1 << bits) - (Key)1);
}

// 2k bits of the packed text starting at base p (first base in the low bits)
template <class Key> __device__ __forceinline__ Key text_window(const uint64_t* __restrict__ text, uint64_t p, int k);
template <> __device__ __forceinline__ uint64_t text_window<uint64_t>(const uint64_t* __restrict__ text, uint64_t p, int k) {
    const uint64_t w0 = text[p >> 5], w1 = text[(p >> 5) + 1];
    const int sh = 2 * (int)(p & 31);
    const uint64_t v = sh ? (w0 >> sh) | (w1 << (64 - sh)) : w0;
    return v & key_low_mask<uint64_t>(2 * k);
}
template <> __device__ __forceinline__ u128 text_window<u128>(const uint64_t* __restrict__ text, uint64_t p, int k) {
    const uint64_t w0 = text[p >> 5], w1 = text[(p >> 5) + 1], w2 = text[(p >> 5) + 2];
    const int sh = 2 * (int)(p & 31);
    const uint64_t lo = sh ? (w0 >> sh) | (w1 << (64 - sh)) : w0;
    const uint64_t hi = sh ? (w1 >> sh) | (w2 << (64 - sh)) : w1;
    return (((u128)hi << 64) | (u128)lo) & key_low_mask<u128>(2 * k);
}

// ASCII -> 2 bits per base, 32 bases per word; flags a non-ACGT character (outside the parity domain)
static __global__ void pack_text_2bit_kernel(const uint8_t* __restrict__ ascii, uint64_t n, uint64_t* __restrict__ packed, uint64_t n_words,
                                             unsigned* __restrict__ bad) {
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    uint64_t out = 0;
    bool b = false;
    for (int i = 0; i < 32; i++) {
        const uint64_t q = w * 32 + i;
        const uint32_t ch = q < n ? (uint32_t)ascii[q] & ~32u : (uint32_t)'A';
        b |= !(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T');
        out |= (uint64_t)(((ch >> 1) ^ (ch >> 2)) & 3u) << (2 * i);
    }
    packed[w] = out;
    if (b) atomicOr(bad, 1u);
}

// k-mer j of the input: unitig u = the last one with kstart[u] <= j, text position start[u] + (j - kstart[u])
template <class Key>
static __global__ void kmer_keys_kernel(const uint64_t* __restrict__ text, const uint64_t* __restrict__ starts, const uint64_t* __restrict__ kstart,
                                        uint64_t n_seqs, uint64_t n_pos, int k, Key* __restrict__ keys) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_pos) return;
    uint64_t lo = 0, hi = n_seqs;  // invariant: kstart[lo] <= j < kstart[hi]  (kstart[n_seqs] = n_pos)
    while (hi - lo > 1) {
        const uint64_t mid = (lo + hi) >> 1;
        if (kstart[mid] <= j) lo = mid;
        else hi = mid;
    }
    keys[j] = text_window<Key>(text, starts[lo] + (j - kstart[lo]), k);
}

// last character to the low two bits, the (k-1)-prefix above it
template <class Key> static __global__ void rotate_keys_kernel(const Key* __restrict__ keys, uint64_t n, int k, Key* __restrict__ rot) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Key x = keys[i];
    rot[i] = ((x & key_low_mask<Key>(2 * (k - 1))) << 2) | (x >> (2 * (k - 1)));
}

template <class Key> __device__ __forceinline__ uint64_t lower_bound_dev(const Key* __restrict__ a, uint64_t n, const Key x) {
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (a[mid] < x) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}
template <class Key> __device__ __forceinline__ uint64_t upper_bound_dev(const Key* __restrict__ a, uint64_t n, const Key x) {
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if (a[mid] <= x) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

// first node of every (k-1)-suffix group: its outgoing labels = the last characters of the k-mers whose (k-1)-prefix
// is that suffix; those k-mers have an incoming edge
template <class Key>
static __global__ void group_labels_kernel(const Key* __restrict__ keys, const Key* __restrict__ rot, uint64_t n, uint8_t* __restrict__ kflags,
                                           uint8_t* __restrict__ has_pred) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Key S = keys[i] >> 2;
    uint8_t f = 0;
    if (i == 0 || (keys[i - 1] >> 2) != S) {
        uint64_t t = lower_bound_dev<Key>(rot, n, S << 2);
        for (; t < n && (rot[t] >> 2) == S; t++) {
            f |= (uint8_t)(1u << (unsigned)(rot[t] & 3));
            has_pred[t] = 1;
        }
        f |= 16;  // suffix group start
    }
    kflags[i] = f;
}

template <class Key> static __global__ void no_pred_flags_kernel(const uint8_t* __restrict__ has_pred, uint64_t n, uint8_t* __restrict__ sel) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) sel[i] = has_pred[i] ? 0 : 1;
}

// dummy node (m, L): the first L characters of the m-th k-mer without an incoming edge, right-aligned ($-padded on the
// left, the padding reads as 0 and the length breaks ties: $ < A); its label is the k-mer's next character.
// Entry 0 is the root (empty string, no label of its own).
template <class Key>
static __global__ void dummy_nodes_kernel(const Key* __restrict__ nopred_rot, uint64_t m_count, int k, Key* __restrict__ dkey,
                                          uint8_t* __restrict__ dlen, uint8_t* __restrict__ dflag) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m_count * (uint64_t)k + 1) return;
    if (t == 0) {
        dkey[0] = 0;
        dlen[0] = 0;
        dflag[0] = 0;
        return;
    }
    const uint64_t m = (t - 1) / (uint64_t)k;
    const int L = (int)((t - 1) % (uint64_t)k);
    const Key r = nopred_rot[m];
    const Key y = (r >> 2) | ((r & 3) << (2 * (k - 1)));  // undo the rotation
    dkey[t] = L ? (y & key_low_mask<Key>(2 * L)) << (2 * (k - L)) : (Key)0;
    dlen[t] = (uint8_t)L;
    dflag[t] = (uint8_t)(1u << (unsigned)((y >> (2 * L)) & 3));
}

template <class T> static __global__ void gather_kernel(const T* __restrict__ src, const uint32_t* __restrict__ idx, uint64_t n, T* __restrict__ dst) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}
static __global__ void iota_kernel(uint32_t* __restrict__ a, uint64_t n) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = (uint32_t)i;
}

template <class Key>
static __global__ void dummy_heads_kernel(const Key* __restrict__ key, const uint8_t* __restrict__ len, uint64_t n, uint32_t* __restrict__ head) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) head[i] = (i == 0 || key[i] != key[i - 1] || len[i] != len[i - 1]) ? 1u : 0u;
}
// slot = inclusive scan of head - 1: the distinct dummies with their labels OR-ed
template <class Key>
static __global__ void dummy_fold_kernel(const Key* __restrict__ key, const uint8_t* __restrict__ len, const uint8_t* __restrict__ flag,
                                         const uint32_t* __restrict__ slot1, uint64_t n, Key* __restrict__ ukey, uint8_t* __restrict__ ulen,
                                         unsigned* __restrict__ uflag) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t s = slot1[i] - 1;
    if (i == 0 || slot1[i] != slot1[i - 1]) {
        ukey[s] = key[i];
        ulen[s] = len[i];
    }
    atomicOr(uflag + s, (unsigned)flag[i]);
}

// merge by rank and set the bits.  rows: 4 x row_words u64 (as u32 for the atomics), then suffix_group_starts
static __device__ __forceinline__ void set_node_bits(unsigned* __restrict__ rows32, uint64_t row_words, uint64_t pos, unsigned labels, bool group_start) {
    const uint64_t w = pos >> 5;
    const unsigned bit = 1u << (pos & 31);
#pragma unroll
    for (int c = 0; c < 4; c++)
        if (labels & (1u << c)) atomicOr(rows32 + (uint64_t)c * row_words * 2 + w, bit);
    if (group_start) atomicOr(rows32 + 4ull * row_words * 2 + w, bit);
}
template <class Key>
static __global__ void place_kmers_kernel(const Key* __restrict__ keys, const uint8_t* __restrict__ kflags, uint64_t n, const Key* __restrict__ ukey,
                                          uint64_t n_dummies, unsigned* __restrict__ rows32, uint64_t row_words) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t pos = i + upper_bound_dev<Key>(ukey, n_dummies, keys[i]);  // dummies with an equal key are shorter: they sort first
    set_node_bits(rows32, row_words, pos, kflags[i] & 15u, kflags[i] & 16u);
}
template <class Key>
static __global__ void place_dummies_kernel(const Key* __restrict__ ukey, const unsigned* __restrict__ uflag, uint64_t n_dummies,
                                            const Key* __restrict__ keys, uint64_t n, unsigned* __restrict__ rows32, uint64_t row_words) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_dummies) return;
    const uint64_t pos = j + lower_bound_dev<Key>(keys, n, ukey[j]);
    set_node_bits(rows32, row_words, pos, uflag[j] & 15u, true);  // every distinct dummy is a suffix group of its own
}

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
static thread_local fin_sbwt_build_timing g_sbwt_timing{};
fin_sbwt_build_timing last_sbwt_build_timing() { return g_sbwt_timing; }

static unsigned grid_for(uint64_t n, unsigned threads = 256) { return (unsigned)std::max<uint64_t>(1, (n + threads - 1) / threads); }

template <class KeyT, class ValT>
static void sort_pairs(DevBuf<uint8_t>& tmp, const KeyT* kin, KeyT* kout, const ValT* vin, ValT* vout, uint64_t n, int end_bit) {
    size_t bytes = 0;
    FIN_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, bytes, kin, kout, vin, vout, (int64_t)n, 0, end_bit));
    tmp.ensure(bytes + 256);
    FIN_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, kin, kout, vin, vout, (int64_t)n, 0, end_bit));
    g_launches += (end_bit + 7) / 8 + 1;
}
template <class KeyT> static void sort_keys(DevBuf<uint8_t>& tmp, const KeyT* kin, KeyT* kout, uint64_t n, int end_bit) {
    size_t bytes = 0;
    FIN_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, bytes, kin, kout, (int64_t)n, 0, end_bit));
    tmp.ensure(bytes + 256);
    FIN_CUDA(cub::DeviceRadixSort::SortKeys(tmp.p, bytes, kin, kout, (int64_t)n, 0, end_bit));
    g_launches += (end_bit + 7) / 8 + 1;
}

// sdsl::rank_support_v5<1> over a bit vector (sdsl/rank_support_v5.hpp:65-109): per 2048-bit superblock one absolute
// count and five 11-bit relative counts in 12-bit slots
static std::vector<uint64_t> rank_v5_blob(const uint64_t* words, uint64_t n_bits) {
    const uint64_t capacity = ((n_bits + 63) >> 6) << 6;
    std::vector<uint64_t> bb;
    if (n_bits == 0) {
        bb.assign(2, 0);
        return bb;
    }
    bb.assign(((capacity >> 11) + 1) << 1, 0);
    const uint64_t n_words = capacity >> 6;
    uint64_t j = 0, sum = (uint64_t)__builtin_popcountll(words[0]), second = 0, in_block = 1;
    for (uint64_t i = 1; i < n_words; ++i, ++in_block) {
        if (in_block == 32) {
            j += 2;
            bb[j - 1] = second;
            bb[j] = bb[j - 2] + sum;
            second = sum = in_block = 0;
        } else if (in_block % 6 == 0) {
            second |= sum << (60 - 12 * (in_block / 6));
        }
        sum += (uint64_t)__builtin_popcountll(words[i]);
    }
    if (in_block % 6 == 0) second |= sum << (60 - 12 * (in_block / 6));
    if (in_block == 32) {
        j += 2;
        bb[j - 1] = second;
        bb[j] = bb[j - 2] + sum;
        bb[j + 1] = 0;
    } else {
        bb[j + 1] = second;
    }
    return bb;
}

struct SbwtRows {  // the result of the device part, on the host
    std::vector<uint64_t> words;  // 5 x row_words: A, C, G, T, suffix_group_starts
    uint64_t row_words = 0, n_nodes = 0, n_kmers = 0;
};

// reverse complement of a k-mer key (first base in the low bits): reverse the 2-bit groups, complement
__device__ __forceinline__ uint64_t rev_groups64(uint64_t x) {
    x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
    x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
    x = ((x >> 8) & 0x00FF00FF00FF00FFULL) | ((x & 0x00FF00FF00FF00FFULL) << 8);
    x = ((x >> 16) & 0x0000FFFF0000FFFFULL) | ((x & 0x0000FFFF0000FFFFULL) << 16);
    return (x >> 32) | (x << 32);
}
__device__ __forceinline__ uint64_t revcomp_key(uint64_t x, int k) { return (~rev_groups64(x)) >> (64 - 2 * k); }
__device__ __forceinline__ u128 revcomp_key(u128 x, int k) {
    const u128 r = ((u128)rev_groups64((uint64_t)x) << 64) | (u128)rev_groups64((uint64_t)(x >> 64));
    return (~r) >> (128 - 2 * k);
}
template <class Key> static __global__ void canonical_keys_kernel(Key* __restrict__ keys, uint64_t n, int k) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Key x = keys[i], r = revcomp_key(x, k);
    keys[i] = r < x ? r : x;
}

// the k-mers of a batch of sequences as keys on the device, in input order
template <class Key> struct KmerKeys {
    DevBuf<Key> keys;
    std::vector<uint64_t> kstart;  // index of the first k-mer of every sequence; kstart[n_seqs] = n_pos
    uint64_t n_pos = 0;
};
template <class Key> static KmerKeys<Key> make_kmer_keys(const SeqBatch& seqs, int k, bool canonical) {
    KmerKeys<Key> K;
    const uint64_t n_seqs = seqs.size(), text_len = seqs.seqs.size();
    // k-mer positions per sequence (sequences shorter than k hold none, NodeBOSSInMemoryConstructor.hh:104)
    K.kstart.assign(n_seqs + 1, 0);
    for (uint64_t u = 0; u < n_seqs; u++) {
        const uint64_t len = seqs.offsets[u + 1] - seqs.offsets[u];
        K.kstart[u + 1] = K.kstart[u] + (len >= (uint64_t)k ? len - (uint64_t)k + 1 : 0);
    }
    const uint64_t n_pos = K.n_pos = K.kstart[n_seqs];
    if (n_pos == 0) throw domain_error("the input holds no k-mer");
    DevBuf<uint64_t> d_text((text_len + 31) / 32 + 3);
    {
        DevBuf<uint8_t> d_ascii(text_len + 1);
        DevBuf<unsigned> d_bad(1);
        FIN_CUDA(cudaMemcpy(d_ascii.p, seqs.seqs.data(), text_len, cudaMemcpyHostToDevice));
        FIN_CUDA(cudaMemset(d_bad.p, 0, 4));
        FIN_CUDA(cudaMemset(d_text.p, 0, d_text.bytes()));
        const uint64_t nw = (text_len + 31) / 32;
        pack_text_2bit_kernel<<<grid_for(nw), 256>>>(d_ascii.p, text_len, d_text.p, nw, d_bad.p);
        g_launches++;
        unsigned bad = 0;
        FIN_CUDA(cudaMemcpy(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost));
        if (bad) throw domain_error("a sequence holds a character other than A, C, G, T (outside the parity domain)");
    }
    K.keys.alloc(n_pos);
    DevBuf<uint64_t> d_starts(n_seqs + 1), d_kstart(n_seqs + 1);
    FIN_CUDA(cudaMemcpy(d_starts.p, seqs.offsets.data(), (n_seqs + 1) * 8, cudaMemcpyHostToDevice));
    FIN_CUDA(cudaMemcpy(d_kstart.p, K.kstart.data(), (n_seqs + 1) * 8, cudaMemcpyHostToDevice));
    kmer_keys_kernel<Key><<<grid_for(n_pos), 256>>>(d_text.p, d_starts.p, d_kstart.p, n_seqs, n_pos, k, K.keys.p);
    g_launches++;
    if (canonical) {
        canonical_keys_kernel<Key><<<grid_for(n_pos), 256>>>(K.keys.p, n_pos, k);
        g_launches++;
    }
    FIN_CUDA(cudaGetLastError());
    FIN_CUDA(cudaDeviceSynchronize());
    return K;
}

template <class Key>
static SbwtRows build_rows_device(const SeqBatch& seqs, int k) {
    using clock = std::chrono::steady_clock;
    auto ms_since = [](clock::time_point t0) { return (float)std::chrono::duration<double, std::milli>(clock::now() - t0).count(); };
    auto t0 = clock::now();
    DevBuf<uint8_t> tmp;
    KmerKeys<Key> K = make_kmer_keys<Key>(seqs, k, false);
    const uint64_t n_pos = K.n_pos;
    DevBuf<Key> a = std::move(K.keys), b(n_pos);
    g_sbwt_timing.keys_ms = ms_since(t0);

    // ---- 2. the k-mer nodes: sort, unique
    NvtxRange range_sort("fin:sbwt_build:sort_and_join");
    t0 = clock::now();
    sort_keys<Key>(tmp, a.p, b.p, n_pos, 2 * k);
    uint64_t n = 0;
    {
        DevBuf<uint64_t> d_n(1);
        size_t bytes = 0;
        FIN_CUDA(cub::DeviceSelect::Unique(nullptr, bytes, b.p, a.p, d_n.p, (int64_t)n_pos));
        tmp.ensure(bytes + 256);
        FIN_CUDA(cub::DeviceSelect::Unique(tmp.p, bytes, b.p, a.p, d_n.p, (int64_t)n_pos));
        g_launches += 2;
        FIN_CUDA(cudaMemcpy(&n, d_n.p, 8, cudaMemcpyDeviceToHost));
    }
    Key* keys = a.p;  // n distinct k-mers in colex order
    // ---- 3. rotated copies, sorted: k-mers by (k-1)-prefix
    DevBuf<Key> rot(n);
    rotate_keys_kernel<Key><<<grid_for(n), 256>>>(keys, n, k, b.p);
    g_launches++;
    sort_keys<Key>(tmp, b.p, rot.p, n, 2 * k);
    b.release();
    FIN_CUDA(cudaDeviceSynchronize());
    g_sbwt_timing.sort_ms = ms_since(t0);

    t0 = clock::now();
    DevBuf<uint8_t> kflags(n), has_pred(n);
    FIN_CUDA(cudaMemset(has_pred.p, 0, n));
    group_labels_kernel<Key><<<grid_for(n), 256>>>(keys, rot.p, n, kflags.p, has_pred.p);
    g_launches++;
    // ---- 4. dummy nodes
    uint64_t m_count = 0;
    DevBuf<Key> nopred;
    {
        DevBuf<uint8_t> sel(n);
        DevBuf<uint64_t> d_m(1);
        no_pred_flags_kernel<Key><<<grid_for(n), 256>>>(has_pred.p, n, sel.p);
        g_launches++;
        DevBuf<Key> out(n);
        size_t bytes = 0;
        FIN_CUDA(cub::DeviceSelect::Flagged(nullptr, bytes, rot.p, sel.p, out.p, d_m.p, (int64_t)n));
        tmp.ensure(bytes + 256);
        FIN_CUDA(cub::DeviceSelect::Flagged(tmp.p, bytes, rot.p, sel.p, out.p, d_m.p, (int64_t)n));
        g_launches += 2;
        FIN_CUDA(cudaMemcpy(&m_count, d_m.p, 8, cudaMemcpyDeviceToHost));
        nopred.alloc(std::max<uint64_t>(1, m_count));
        if (m_count) FIN_CUDA(cudaMemcpy(nopred.p, out.p, m_count * sizeof(Key), cudaMemcpyDeviceToDevice));
    }
    rot.release();
    has_pred.release();
    const uint64_t d_raw = m_count * (uint64_t)k + 1;
    if (d_raw >= 0xFFFFFFF0ull) throw limit_error("too many dummy nodes for this build");
    uint64_t n_dummies = 0;
    DevBuf<Key> ukey(d_raw);
    DevBuf<unsigned> uflag(d_raw);
    {
        DevBuf<Key> dkey(d_raw), gkey(d_raw), skey(d_raw);
        DevBuf<uint8_t> dlen(d_raw), dflag(d_raw), slen(d_raw), sflag(d_raw), len_sorted(d_raw), ulen(d_raw);
        DevBuf<uint32_t> idx0(d_raw), idx1(d_raw), idx2(d_raw), head(d_raw), slot1(d_raw);
        dummy_nodes_kernel<Key><<<grid_for(d_raw), 256>>>(nopred.p, m_count, k, dkey.p, dlen.p, dflag.p);
        iota_kernel<<<grid_for(d_raw), 256>>>(idx0.p, d_raw);
        g_launches += 2;
        // stable two-pass sort by (key, length): length first, then key
        sort_pairs<uint8_t, uint32_t>(tmp, dlen.p, len_sorted.p, idx0.p, idx1.p, d_raw, 8);
        gather_kernel<Key><<<grid_for(d_raw), 256>>>(dkey.p, idx1.p, d_raw, gkey.p);
        g_launches++;
        sort_pairs<Key, uint32_t>(tmp, gkey.p, skey.p, idx1.p, idx2.p, d_raw, 2 * k);
        gather_kernel<uint8_t><<<grid_for(d_raw), 256>>>(dlen.p, idx2.p, d_raw, slen.p);
        gather_kernel<uint8_t><<<grid_for(d_raw), 256>>>(dflag.p, idx2.p, d_raw, sflag.p);
        dummy_heads_kernel<Key><<<grid_for(d_raw), 256>>>(skey.p, slen.p, d_raw, head.p);
        g_launches += 3;
        size_t bytes = 0;
        FIN_CUDA(cub::DeviceScan::InclusiveSum(nullptr, bytes, head.p, slot1.p, (int64_t)d_raw));
        tmp.ensure(bytes + 256);
        FIN_CUDA(cub::DeviceScan::InclusiveSum(tmp.p, bytes, head.p, slot1.p, (int64_t)d_raw));
        g_launches += 2;
        FIN_CUDA(cudaMemset(uflag.p, 0, d_raw * 4));
        dummy_fold_kernel<Key><<<grid_for(d_raw), 256>>>(skey.p, slen.p, sflag.p, slot1.p, d_raw, ukey.p, ulen.p, uflag.p);
        g_launches++;
        uint32_t last = 0;
        FIN_CUDA(cudaMemcpy(&last, slot1.p + (d_raw - 1), 4, cudaMemcpyDeviceToHost));
        n_dummies = last;
    }
    // ---- 5. merge by rank, set the bits
    const uint64_t n_nodes = n + n_dummies;
    if (n_nodes >= kIndexLimit32) throw limit_error("SBWT too large for the 32-bit column positions of this build");
    SbwtRows R;
    R.n_nodes = n_nodes;
    R.n_kmers = n;
    R.row_words = (n_nodes + 63) / 64 + 1;
    DevBuf<uint64_t> d_rows(5 * R.row_words);
    FIN_CUDA(cudaMemset(d_rows.p, 0, d_rows.bytes()));
    place_kmers_kernel<Key><<<grid_for(n), 256>>>(keys, kflags.p, n, ukey.p, n_dummies, (unsigned*)d_rows.p, R.row_words);
    place_dummies_kernel<Key><<<grid_for(n_dummies), 256>>>(ukey.p, uflag.p, n_dummies, keys, n, (unsigned*)d_rows.p, R.row_words);
    g_launches += 2;
    FIN_CUDA(cudaGetLastError());
    R.words.resize(5 * R.row_words);
    FIN_CUDA(cudaMemcpy(R.words.data(), d_rows.p, d_rows.bytes(), cudaMemcpyDeviceToHost));
    g_sbwt_timing.nodes_ms = ms_since(t0);
    return R;
}

void sbwt_build_device(const std::string& unitig_path, int64_t k, const std::string& out_path, int64_t precalc_k, int device) {
    using clock = std::chrono::steady_clock;
    auto ms_since = [](clock::time_point t0) { return (float)std::chrono::duration<double, std::milli>(clock::now() - t0).count(); };
    const auto t_all = clock::now();
    g_sbwt_timing = fin_sbwt_build_timing{};
    if (k < 2 || k > 64) throw limit_error("fin_sbwt_build supports 2 <= k <= 64");
    if (precalc_k < 0 || precalc_k > 12) throw arg_error("precalc_k must be in [0, 12]");
    if (precalc_k > k) precalc_k = k;  // sbwt_build.cpp:150-153
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) throw cuda_error("no CUDA device is visible");
    if (device < 0 || device >= n_dev) throw arg_error("CUDA device ordinal out of range");
    FIN_CUDA(cudaSetDevice(device));

    auto t0 = clock::now();
    SeqBatch seqs;
    for (const std::string& f : has_suffix(unitig_path, ".txt") ? read_lines(unitig_path) : std::vector<std::string>{unitig_path})
        append_sequences(f, seqs);
    if (seqs.seqs.size() >= kIndexLimit32) throw limit_error("input too large for the 32-bit positions of this build");
    g_sbwt_timing.read_ms = ms_since(t0);
    g_sbwt_timing.bases = (int64_t)seqs.seqs.size();

    const SbwtRows R = k <= 32 ? build_rows_device<uint64_t>(seqs, (int)k) : build_rows_device<u128>(seqs, (int)k);

    // ---- 6. host: C array, rank blobs, precalc, file
    t0 = clock::now();
    const uint64_t N = R.n_nodes, nw = (N + 63) / 64;
    const uint64_t* row[5];
    for (int c = 0; c < 5; c++) row[c] = R.words.data() + (uint64_t)c * R.row_words;
    int64_t Carr[4];
    std::vector<uint32_t> cum[4];
    {
        int64_t run = 1;  // SBWT.hh:277-282: C[0] = 1 (the root), then cumulative label counts
        for (int c = 0; c < 4; c++) {
            Carr[c] = run;
            cum[c].assign(nw + 1, 0);
            uint32_t s = 0;
            for (uint64_t w = 0; w < nw; w++) {
                cum[c][w] = s;
                s += (uint32_t)__builtin_popcountll(row[c][w]);
            }
            cum[c][nw] = s;
            run += s;
        }
    }
    std::vector<int64_t> precalc;
    if (precalc_k > 0) {  // SBWT.hh:529-555: entry d = interval of the precalc_k characters of d, 2 bits each, first char lowest
        auto rank = [&](int c, int64_t pos) -> int64_t {
            const uint64_t w = (uint64_t)pos >> 6, off = (uint64_t)pos & 63;
            return cum[c][w] + (off ? __builtin_popcountll(row[c][w] & ((1ULL << off) - 1)) : 0);
        };
        const uint64_t np = 1ULL << (2 * precalc_k);
        precalc.assign(2 * np, -1);
        for (uint64_t d = 0; d < np; d++) {
            int64_t l = 0, r = (int64_t)N - 1;
            for (int64_t i = 0; i < precalc_k && l != -1; i++) {
                const int c = (int)((d >> (2 * i)) & 3);
                const int64_t nl = Carr[c] + rank(c, l), nr = Carr[c] + rank(c, r + 1) - 1;
                if (nl > nr) l = r = -1;
                else { l = nl; r = nr; }
            }
            precalc[2 * d] = l;
            precalc[2 * d + 1] = r;
        }
    }
    FILE* f = fopen(out_path.c_str(), "wb");
    if (!f) throw io_error("cannot write " + out_path);
    try {
        auto put64 = [&](int64_t v) { write_all(f, &v, 8, out_path); };
        auto put_bits = [&](const uint64_t* words) {
            const uint64_t bits = N;
            write_all(f, &bits, 8, out_path);
            write_all(f, words, nw * 8, out_path);
        };
        put64(12); write_all(f, "plain-matrix", 12, out_path);  // sbwt_build.cpp: serialize_string(variant)
        put64(4);  write_all(f, "v0.1", 4, out_path);           // SBWT.hh:398
        for (int c = 0; c < 4; c++) put_bits(row[c]);
        for (int c = 0; c < 4; c++) {
            const std::vector<uint64_t> bb = rank_v5_blob(row[c], N);
            const uint64_t bits = (uint64_t)bb.size() * 64;
            write_all(f, &bits, 8, out_path);
            write_all(f, bb.data(), bb.size() * 8, out_path);
        }
        put_bits(row[4]);
        put64(32);
        write_all(f, Carr, 32, out_path);
        put64((int64_t)precalc.size() * 8);
        write_all(f, precalc.data(), precalc.size() * 8, out_path);
        put64(precalc_k); put64((int64_t)N); put64((int64_t)R.n_kmers); put64(k);
    } catch (...) {
        fclose(f);
        throw;
    }
    if (fclose(f) != 0) throw io_error("cannot close " + out_path);
    g_sbwt_timing.write_ms = ms_since(t0);
    g_sbwt_timing.total_ms = ms_since(t_all);
    g_sbwt_timing.n_nodes = (int64_t)N;
    g_sbwt_timing.n_kmers = (int64_t)R.n_kmers;
}

// ---------------------------------------------------------------------------------------------------------------
// Disjoint spectrum-preserving string set on the device: the stand-in for the unitig tools
// the reference's README chains in front of `sbwt build` (ggcat + unitig_flipper).  Every canonical k-mer (a k-mer and
// its reverse complement count as one, as in ggcat) is kept at its first occurrence in the input; the output strings
// are the maximal stretches of consecutive kept k-mers ("pieces", each >= k long, every k-mer exactly once in either
// orientation), optionally reverse-complemented at random ("flipped").  Not maximal unitigs - FinimizerIndex only
// needs a disjoint SPSS.
// ---------------------------------------------------------------------------------------------------------------
static __global__ void first_occurrence_kernel_u32(const uint32_t* __restrict__ pos, uint64_t n, uint8_t* __restrict__ keep, const void* keys,
                                                   int key_bytes) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bool head = i == 0;
    if (!head) {
        if (key_bytes == 8) head = ((const uint64_t*)keys)[i] != ((const uint64_t*)keys)[i - 1];
        else head = ((const u128*)keys)[i] != ((const u128*)keys)[i - 1];
    }
    keep[pos[i]] = head ? 1 : 0;  // the sort is stable: the head of a run of equal keys has the smallest input position
}
static __global__ void mark_sequence_bounds_kernel(const uint64_t* __restrict__ kstart, uint64_t n_seqs, uint8_t* __restrict__ keep) {
    const uint64_t u = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_seqs || kstart[u + 1] == kstart[u]) return;
    keep[kstart[u]] |= 2;          // first k-mer of its sequence
    keep[kstart[u + 1] - 1] |= 4;  // last k-mer of its sequence
}
// 1 = a piece starts at k-mer j, 2 = a piece ends at k-mer j
static __global__ void piece_bounds_kernel(const uint8_t* __restrict__ keep, uint64_t n, uint8_t* __restrict__ is_start, uint8_t* __restrict__ is_end) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const unsigned f = keep[j];
    const bool kept = f & 1;
    is_start[j] = kept && ((f & 2) || !(keep[j - 1] & 1));  // j > 0 whenever bit 2 is clear
    is_end[j] = kept && ((f & 4) || !(keep[j + 1] & 1));    // j + 1 < n whenever bit 4 is clear
}

template <class Key>
static void spss_pieces_device(const SeqBatch& seqs, int k, std::vector<uint64_t>& starts, std::vector<uint64_t>& ends) {
    KmerKeys<Key> K = make_kmer_keys<Key>(seqs, k, true);
    const uint64_t n = K.n_pos, n_seqs = seqs.size();
    if (n >= 0xFFFFFFF0ull) throw limit_error("too many k-mer positions for this build");
    DevBuf<uint8_t> tmp, keep(n + 1);
    {
        DevBuf<Key> sorted(n);
        DevBuf<uint32_t> pos(n), pos_sorted(n);
        iota_kernel<<<grid_for(n), 256>>>(pos.p, n);
        g_launches++;
        sort_pairs<Key, uint32_t>(tmp, K.keys.p, sorted.p, pos.p, pos_sorted.p, n, 2 * k);
        K.keys.release();
        first_occurrence_kernel_u32<<<grid_for(n), 256>>>(pos_sorted.p, n, keep.p, sorted.p, (int)sizeof(Key));
        g_launches++;
        FIN_CUDA(cudaGetLastError());
        FIN_CUDA(cudaDeviceSynchronize());
    }
    tmp.release();
    DevBuf<uint64_t> d_kstart(n_seqs + 1);
    FIN_CUDA(cudaMemcpy(d_kstart.p, K.kstart.data(), (n_seqs + 1) * 8, cudaMemcpyHostToDevice));
    mark_sequence_bounds_kernel<<<grid_for(n_seqs), 256>>>(d_kstart.p, n_seqs, keep.p);
    DevBuf<uint8_t> is_start(n), is_end(n);
    piece_bounds_kernel<<<grid_for(n), 256>>>(keep.p, n, is_start.p, is_end.p);
    g_launches += 2;
    DevBuf<uint32_t> sel(n);
    DevBuf<uint64_t> d_count(1);
    auto select = [&](const uint8_t* flags, std::vector<uint64_t>& out) {
        size_t bytes = 0;
        cub::CountingInputIterator<uint32_t> it(0);
        FIN_CUDA(cub::DeviceSelect::Flagged(nullptr, bytes, it, flags, sel.p, d_count.p, (int64_t)n));
        tmp.ensure(bytes + 256);
        FIN_CUDA(cub::DeviceSelect::Flagged(tmp.p, bytes, it, flags, sel.p, d_count.p, (int64_t)n));
        g_launches += 2;
        uint64_t c = 0;
        FIN_CUDA(cudaMemcpy(&c, d_count.p, 8, cudaMemcpyDeviceToHost));
        std::vector<uint32_t> h(c);
        if (c) FIN_CUDA(cudaMemcpy(h.data(), sel.p, c * 4, cudaMemcpyDeviceToHost));
        out.assign(h.begin(), h.end());
    };
    select(is_start.p, starts);
    select(is_end.p, ends);
    if (starts.size() != ends.size()) throw std::runtime_error("internal error: piece bounds do not pair up");
    // k-mer index -> text position
    uint64_t u = 0;
    for (size_t i = 0; i < starts.size(); i++) {
        while (K.kstart[u + 1] <= starts[i]) u++;
        const uint64_t shift = seqs.offsets[u] - K.kstart[u];
        if (ends[i] < starts[i] || ends[i] >= K.kstart[u + 1]) throw std::runtime_error("internal error: a piece leaves its sequence");
        starts[i] += shift;
        ends[i] += shift + (uint64_t)k;  // exclusive end in bases
    }
}

static uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

fin_spss_info spss_build_device(const std::string& in_path, int64_t k, const std::string& out_path, int flip, uint64_t seed, int device) {
    if (k < 2 || k > 64) throw limit_error("fin_spss_build supports 2 <= k <= 64");
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) throw cuda_error("no CUDA device is visible");
    if (device < 0 || device >= n_dev) throw arg_error("CUDA device ordinal out of range");
    FIN_CUDA(cudaSetDevice(device));
    SeqBatch seqs;
    for (const std::string& f : has_suffix(in_path, ".txt") ? read_lines(in_path) : std::vector<std::string>{in_path}) append_sequences(f, seqs);
    if (seqs.seqs.size() >= kIndexLimit32) throw limit_error("input too large for the 32-bit positions of this build");
    std::vector<uint64_t> starts, ends;
    if (k <= 32) spss_pieces_device<uint64_t>(seqs, (int)k, starts, ends);
    else spss_pieces_device<u128>(seqs, (int)k, starts, ends);
    FILE* f = fopen(out_path.c_str(), "wb");
    if (!f) throw io_error("cannot write " + out_path);
    fin_spss_info info{};
    try {
        static const char comp[4] = {'T', 'G', 'C', 'A'};
        std::vector<char> buf;
        buf.reserve(1 << 24);
        for (size_t i = 0; i < starts.size(); i++) {
            char head[32];
            const int hl = snprintf(head, sizeof head, ">%zu\n", i);
            buf.insert(buf.end(), head, head + hl);
            const char* src = seqs.seqs.data() + starts[i];
            const uint64_t len = ends[i] - starts[i];
            const size_t at = buf.size();
            buf.resize(at + len + 1);
            if (flip && (splitmix64(seed ^ (uint64_t)i) & 1)) {
                for (uint64_t q = 0; q < len; q++) {
                    const unsigned ch = (unsigned char)src[len - 1 - q];
                    buf[at + q] = comp[((ch >> 1) ^ (ch >> 2)) & 3u];
                }
            } else {
                memcpy(&buf[at], src, len);
            }
            buf[at + len] = '\n';
            info.bases += (int64_t)len;
            if (buf.size() > (1u << 24) - 4096 || len > (1u << 22)) {
                write_all(f, buf.data(), buf.size(), out_path);
                buf.clear();
            }
        }
        write_all(f, buf.data(), buf.size(), out_path);
    } catch (...) {
        fclose(f);
        throw;
    }
    if (fclose(f) != 0) throw io_error("cannot close " + out_path);
    info.n_strings = (int64_t)starts.size();
    info.n_kmers = info.bases - info.n_strings * (k - 1);
    info.input_bases = (int64_t)seqs.seqs.size();
    return info;
}

}  // namespace fin
