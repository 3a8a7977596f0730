// build-fmin on the device: host driver around the kernels of build_kernels.cuh.  Reproduces
// FinimizerIndexBuilder's constructor (include/FinimizerIndex.hh:273-319), the build_fmin driver
// (include/build_fmin.hh:302-401) and FinimizerIndex::serialize (:187-207).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>

#include <ctime>

#include "build_host.hpp"
#include <thread>

#include "engine.hpp"
#include "build_kernels.cuh"

namespace fin {

static void select_device(int device) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) throw cuda_error("no CUDA device is visible");
    if (device < 0 || device >= n_dev) throw arg_error("CUDA device ordinal out of range");
    FIN_CUDA(cudaSetDevice(device));
}

template <class T> static DevBuf<T> to_device(const std::vector<T>& v, size_t pad = 0) {
    DevBuf<T> d(v.size() + pad);
    if (pad) FIN_CUDA(cudaMemset(d.p, 0, d.bytes()));
    if (!v.empty()) FIN_CUDA(cudaMemcpy(d.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}

// ---- the SBWT rows on the device, with the rank of every 64-column word (shared by the LCS rounds and the lines) ----
struct DeviceSbwt {
    DevBuf<uint64_t> rows;       // rows[c * nw + w], bits past N cleared
    DevBuf<uint32_t> word_dst;   // rank_c(64 w)
    uint64_t nw = 0, N = 0;
    LcsArgs args{};
};

static DeviceSbwt upload_sbwt(const SbwtHost& s) {
    DeviceSbwt d;
    d.N = (uint64_t)s.n_nodes;
    d.nw = (d.N + 63) / 64;
    const uint64_t nw = d.nw;
    d.rows.alloc(4 * nw);
    for (int c = 0; c < 4; c++) {
        if (s.bits[c].words.size() < nw) throw format_error("SBWT row shorter than the number of columns");
        FIN_CUDA(cudaMemcpy(d.rows.p + (uint64_t)c * nw, s.bits[c].words.data(), nw * 8, cudaMemcpyHostToDevice));
    }
    mask_row_tails_kernel<<<1, 32>>>(d.rows.p, nw, d.N);
    d.word_dst.alloc(4 * nw);
    const unsigned T = 256;
    row_popcount_kernel<<<(unsigned)((4 * nw + T - 1) / T), T>>>(d.rows.p, nw, d.word_dst.p);
    g_launches += 2;
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d.word_dst.p, d.word_dst.p, (int64_t)nw);
    DevBuf<uint8_t> tmp(tmp_bytes + 256);
    for (int c = 0; c < 4; c++) {
        size_t tb = tmp_bytes;
        uint32_t* row = d.word_dst.p + (uint64_t)c * nw;
        FIN_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, row, row, (int64_t)nw));
        g_launches += 2;
    }
    d.args.rows = d.rows.p;
    d.args.word_dst = d.word_dst.p;
    d.args.n = d.N;
    d.args.n_words = nw;
    for (int c = 0; c < 4; c++) d.args.C[c] = (uint32_t)s.C[c];
    d.args.k = (uint8_t)s.k;
    FIN_CUDA(cudaGetLastError());
    return d;
}

// ---- B1: LCS bytes on the device (k launches of lcs_round_kernel), padded with zeros to 64 * nw + 64 bytes ----------
static DevBuf<uint8_t> lcs_on_device(const DeviceSbwt& d) {
    NvtxRange range("fin:build:lcs_rounds");
    const uint64_t padded = d.nw * 64 + 64;
    DevBuf<uint8_t> last(padded), next(padded), lcs(padded);
    const unsigned T = 256;
    FIN_CUDA(cudaMemset(last.p, 0, padded));
    FIN_CUDA(cudaMemset(next.p, 0, padded));
    FIN_CUDA(cudaMemset(lcs.p, 0, padded));
    lcs_init_kernel<<<(unsigned)((d.N + T - 1) / T), T>>>(d.args, last.p, lcs.p);
    g_launches++;
    uint8_t* cur = last.p;
    uint8_t* nxt = next.p;
    for (int round = 0; round < (int)d.args.k; round++) {
        const uint64_t words_per_block = 32ull * kLcsWarpsPerBlock;
        lcs_round_warp_kernel<<<(unsigned)((d.nw + words_per_block - 1) / words_per_block), 32 * kLcsWarpsPerBlock>>>(d.args, cur, nxt, lcs.p,
                                                                                                             (uint8_t)round);
        g_launches++;
        std::swap(cur, nxt);
    }
    FIN_CUDA(cudaGetLastError());
    FIN_CUDA(cudaDeviceSynchronize());
    return lcs;
}

std::vector<uint8_t> lcs_device(const SbwtHost& s, int device) {
    select_device(device);
    check_sbwt_limits(s);
    const DeviceSbwt ds = upload_sbwt(s);
    DevBuf<uint8_t> d = lcs_on_device(ds);
    std::vector<uint8_t> h((size_t)s.n_nodes);
    FIN_CUDA(cudaMemcpy(h.data(), d.p, h.size(), cudaMemcpyDeviceToHost));
    return h;
}

// device array -> sdsl int_vector of the given width (bit packing on the device, one D2H of the packed words)
template <class T> static PackedVec pack_on_device(const T* d_in, uint64_t n, int width) {
    PackedVec v;
    v.assign(n, (uint8_t)width);
    const uint64_t nw = v.disk_words();
    if (nw) {
        DevBuf<uint64_t> d_words(nw);
        pack_bits_kernel<T><<<(unsigned)((nw + 255) / 256), 256>>>(d_in, n, (uint32_t)width, d_words.p, nw);
        g_launches++;
        FIN_CUDA(cudaGetLastError());
        FIN_CUDA(cudaMemcpy(v.words.data(), d_words.p, nw * 8, cudaMemcpyDeviceToHost));
    }
    return v;
}

// LCS file -> padded bytes on the device
static DevBuf<uint8_t> lcs_bytes_from_packed(const PackedVec& lcs, uint64_t nw, int64_t k) {
    if (lcs.width > 8) throw format_error("LCS wider than 8 bits (k <= 255)");
    const uint64_t n = lcs.size(), padded = nw * 64 + 64;
    DevBuf<uint8_t> bytes(padded);
    FIN_CUDA(cudaMemset(bytes.p, 0, padded));
    DevBuf<uint64_t> words = to_device(lcs.words);
    if (n) {
        DevBuf<unsigned int> d_bad(1);
        FIN_CUDA(cudaMemset(d_bad.p, 0, 4));
        unpack_bits_to_bytes_kernel<<<(unsigned)((n + 255) / 256), 256>>>(words.p, n, lcs.width, bytes.p, (uint32_t)k, d_bad.p);
        g_launches++;
        FIN_CUDA(cudaGetLastError());
        unsigned bad = 0;
        FIN_CUDA(cudaMemcpy(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost));
        if (bad) throw format_error("the LCS file does not belong to this SBWT: LCS[0] must be 0 and every value below k");
    }
    return bytes;
}

// the 128-byte lines (device_index.cuh) from the rows and the LCS bytes; Ustart / fmin words (device) or null
static DevBuf<uint64_t> make_lines_device(const DeviceSbwt& d, const uint8_t* d_lcs_bytes, const uint64_t* d_ustart = nullptr,
                                          const uint64_t* d_fmin = nullptr) {
    const uint64_t n_lines = (d.N >> 6) + 2;
    DevBuf<uint64_t> lines(n_lines * 16);
    make_lines_kernel<<<(unsigned)((n_lines + 127) / 128), 128>>>(d.args, d_lcs_bytes, d_ustart, d_fmin, n_lines,
                                                                  reinterpret_cast<uint4*>(lines.p));
    g_launches++;
    FIN_CUDA(cudaGetLastError());
    return lines;
}

// FinimizerIndex::load: the lines of a complete index, laid out on the device (the host only reads the files)
DevBuf<uint64_t> index_lines_device(const IndexHost& h) {
    const uint64_t N = (uint64_t)h.sbwt.n_nodes;
    if (N >= kIndexLimit32) throw limit_error("index too large for the 32-bit column/text positions of this build");
    if (N < 2) throw domain_error("SBWT with fewer than 2 columns");
    const DeviceSbwt ds = upload_sbwt(h.sbwt);
    DevBuf<uint8_t> lcs = lcs_bytes_from_packed(h.lcs, ds.nw, h.sbwt.k);
    if (h.ustart.words.size() < ds.nw || h.fmin.words.size() < ds.nw) throw format_error("Ustart / fmin bit vectors shorter than the SBWT");
    DevBuf<uint64_t> us(ds.nw), fm(ds.nw);
    FIN_CUDA(cudaMemcpy(us.p, h.ustart.words.data(), ds.nw * 8, cudaMemcpyHostToDevice));
    FIN_CUDA(cudaMemcpy(fm.p, h.fmin.words.data(), ds.nw * 8, cudaMemcpyHostToDevice));
    DevBuf<uint64_t> lines = make_lines_device(ds, lcs.p, us.p, fm.p);
    FIN_CUDA(cudaDeviceSynchronize());
    return lines;
}

static int window_capacity(int k) { return k <= 32 ? 32 : k <= 64 ? 64 : k <= 128 ? 128 : 256; }

// ---- B2-B4: FinimizerIndexBuilder, staged so that the stream can be sharded over GPUs ----------------------------
// world == 1: open -> stream -> merge (nothing to fold) -> finish.
// world  > 1: every rank opens the same inputs; stream() covers this rank's chunks, export_tuples() hands out the
// (colex, key) pairs sorted by colex with the number that belongs to every owner, the caller moves them with ONE
// all-to-all (finito_b200/distributed.py: torch.distributed over NCCL), merge() folds what arrived into this rank's
// column slice, slice() returns the slice's fmin words and offsets in column order, and rank 0 — after gathering the
// slices — writes the files in finish().
struct TableNonZero {
    const unsigned long long* table;
    __device__ bool operator()(const uint32_t& i) const { return table[i] != 0ULL; }
};

static __global__ void tuple_gather_kernel(const unsigned long long* __restrict__ table, const uint32_t* __restrict__ idx, uint64_t n,
                                           uint64_t* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t colex = idx[i];
    reinterpret_cast<ulonglong2*>(out)[i] = make_ulonglong2((unsigned long long)colex, table[colex]);
}

// bounds[g] = first tuple whose colex >= g * slice_cols (idx ascending), g = 0 .. world
static __global__ void owner_bounds_kernel(const uint32_t* __restrict__ idx, uint64_t n, uint64_t slice_cols, int world,
                                           uint64_t* __restrict__ bounds) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g > world) return;
    const uint64_t want = (uint64_t)g * slice_cols;
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if ((uint64_t)idx[mid] >= want) hi = mid;
        else lo = mid + 1;
    }
    bounds[g] = g == world ? n : lo;
}

static __global__ void tuple_merge_kernel(const uint64_t* __restrict__ tuples, uint64_t n, uint64_t slice_begin, uint64_t slice_cols,
                                          unsigned long long* __restrict__ slice_table, unsigned int* __restrict__ error_flags) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const ulonglong2 t = reinterpret_cast<const ulonglong2*>(tuples)[i];
    const uint64_t local = t.x - slice_begin;
    if (t.x < slice_begin || local >= slice_cols) {
        atomicOr(error_flags, kErrLookupRange);
        return;
    }
    atomicMax(slice_table + local, t.y);
}

class RarestBuilder {
   public:
    // d_lcs_bytes: the LCS as padded bytes on the device (lcs_on_device / lcs_bytes_from_packed); consumed here
    RarestBuilder(SbwtHost&& sbwt, PackedVec&& lcs_file, const DeviceSbwt& ds, DevBuf<uint8_t>&& d_lcs_bytes, const SeqBatch& unitigs,
                  int rank, int world)
        : sbwt_(std::move(sbwt)), lcs_file_(std::move(lcs_file)), rank_(rank), world_(world) {
        if (world < 1 || rank < 0 || rank >= world) throw arg_error("build: rank / world out of range");
        k_ = (int)sbwt_.k;
        N_ = (uint64_t)sbwt_.n_nodes;
        const uint64_t U = unitigs.size();
        if (lcs_file_.size() != N_) throw format_error("LCS length differs from the number of SBWT columns");
        check_unitigs(unitigs, k_);
        d_lines_ = make_lines_device(ds, d_lcs_bytes.p);
        FIN_CUDA(cudaDeviceSynchronize());
        d_lcs_bytes.release();
        ix_.lines = reinterpret_cast<const uint4*>(d_lines_.p);
        ix_.n_nodes = (idx_t)N_;
        ix_.k = k_;
        d_err_.alloc(1);
        FIN_CUDA(cudaMemset(d_err_.p, 0, 4));

        // B2 on the device: first k-mer columns straight from the ASCII unitigs, host sort of the columns, then the
        // permuted 2-bit text in one pass
        const uint64_t T = unitigs.seqs.size();
        DevBuf<char> d_ascii(T + 1);
        FIN_CUDA(cudaMemcpy(d_ascii.p, unitigs.seqs.data(), T, cudaMemcpyHostToDevice));
        DevBuf<uint64_t> d_offsets = to_device(unitigs.offsets);
        DevBuf<uint32_t> d_cols(U);
        first_kmer_ascii_kernel<<<(unsigned)((U + 127) / 128), 128>>>(ix_, d_ascii.p, d_offsets.p, (uint32_t)U, d_cols.p, d_err_.p);
        g_launches++;
        FIN_CUDA(cudaGetLastError());
        std::vector<uint32_t> cols(U);
        FIN_CUDA(cudaMemcpy(cols.data(), d_cols.p, U * 4, cudaMemcpyDeviceToHost));
        check_bases();
        const std::vector<uint64_t> order = unitig_order(cols, N_, ustart_);
        std::vector<uint64_t> starts(U), src(U);
        ends_.assign(U, 64);  // width stays 64 (PackedStrings.hh:44 passes the bit length as the default VALUE)
        uint64_t pos = 0;
        for (uint64_t j = 0; j < U; j++) {
            const uint64_t u = order[j];
            starts[j] = pos;
            src[j] = unitigs.offsets[u];
            pos += unitigs.offsets[u + 1] - unitigs.offsets[u];
            ends_.words[j] = pos;
        }
        text_.assign(T, 2);
        const uint64_t n_text_words = text_.words.size();  // disk words + 1 pad
        d_text_.alloc(n_text_words + 2);
        FIN_CUDA(cudaMemset(d_text_.p, 0, d_text_.bytes()));
        {
            DevBuf<uint64_t> d_starts = to_device(starts), d_src = to_device(src);
            pack_text_kernel<<<(unsigned)((n_text_words + 255) / 256), 256>>>(d_ascii.p, d_starts.p, d_src.p, (uint32_t)U, T, d_text_.p,
                                                                            n_text_words, d_err_.p);
            g_launches++;
            FIN_CUDA(cudaGetLastError());
            FIN_CUDA(cudaMemcpy(text_.words.data(), d_text_.p, n_text_words * 8, cudaMemcpyDeviceToHost));
        }
        check_bases();

        chunks_ = plan_build_chunks(starts, ends_, k_, 0);
        const std::pair<size_t, size_t> mine = shard_chunks(chunks_, rank, world);
        chunk_begin_ = mine.first;
        chunk_end_ = mine.second;
        slice_cols_ = world == 1 ? N_ : slice_columns(N_, world);
        slice_begin_ = (uint64_t)rank * slice_cols_;
        for (size_t c = chunk_begin_; c < chunk_end_; c++) bases_streamed_ += chunks_[c].n_stream;
    }

    uint64_t n_nodes() const { return N_; }
    uint64_t slice_cols() const { return slice_cols_; }
    uint64_t text_len() const { return text_.size(); }
    uint64_t n_unitigs() const { return ends_.size(); }
    uint64_t n_chunks() const { return chunks_.size(); }
    uint64_t chunk_begin() const { return chunk_begin_; }
    uint64_t chunk_end() const { return chunk_end_; }
    uint64_t bases_streamed() const { return bases_streamed_; }
    int world() const { return world_; }
    int rank() const { return rank_; }
    float stream_ms() const { return stream_ms_; }

    // add_sequence over this rank's chunks; overwrite rule as a max over table[colex] (B3)
    uint64_t stream() {
        NvtxRange range("fin:build:stream");
        const size_t nc = chunk_end_ - chunk_begin_;
        d_table_.alloc(N_);
        FIN_CUDA(cudaMemset(d_table_.p, 0, d_table_.bytes()));
        cudaEvent_t e0, e1;
        FIN_CUDA(cudaEventCreate(&e0));
        FIN_CUDA(cudaEventCreate(&e1));
        FIN_CUDA(cudaEventRecord(e0));
        if (nc) {
            DevBuf<BuildChunk> d_chunks(nc);
            FIN_CUDA(cudaMemcpy(d_chunks.p, chunks_.data() + chunk_begin_, nc * sizeof(BuildChunk), cudaMemcpyHostToDevice));
            const unsigned blocks = (unsigned)((nc + 127) / 128);
            const uint32_t n32 = (uint32_t)nc;
            switch (window_capacity(k_)) {
                case 32: build_stream_kernel<32><<<blocks, 128>>>(ix_, d_text_.p, d_chunks.p, n32, d_table_.p, d_err_.p); break;
                case 64: build_stream_kernel<64><<<blocks, 128>>>(ix_, d_text_.p, d_chunks.p, n32, d_table_.p, d_err_.p); break;
                case 128: build_stream_kernel<128><<<blocks, 128>>>(ix_, d_text_.p, d_chunks.p, n32, d_table_.p, d_err_.p); break;
                default: build_stream_kernel<256><<<blocks, 128>>>(ix_, d_text_.p, d_chunks.p, n32, d_table_.p, d_err_.p); break;
            }
            g_launches++;
            FIN_CUDA(cudaGetLastError());
            FIN_CUDA(cudaEventRecord(e1));
            FIN_CUDA(cudaEventSynchronize(e1));
            FIN_CUDA(cudaEventElapsedTime(&stream_ms_, e0, e1));
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        check_errors();
        n_tuples_ = 0;
        if (world_ > 1) {  // the columns this rank has something to say about, ascending
            d_idx_.alloc(N_);
            DevBuf<uint64_t> d_n(1);
            cub::CountingInputIterator<uint32_t> first(0);
            size_t tb = 0;
            cub::DeviceSelect::If(nullptr, tb, first, d_idx_.p, d_n.p, (int64_t)N_, TableNonZero{d_table_.p});
            DevBuf<uint8_t> tmp(tb + 256);
            FIN_CUDA(cub::DeviceSelect::If(tmp.p, tb, first, d_idx_.p, d_n.p, (int64_t)N_, TableNonZero{d_table_.p}));
            g_launches += 2;
            FIN_CUDA(cudaMemcpy(&n_tuples_, d_n.p, 8, cudaMemcpyDeviceToHost));
        }
        streamed_ = true;
        return n_tuples_;
    }

    // {u64 colex, u64 key} per tuple into d_out (device), counts[g] = tuples owned by rank g (host, world entries)
    void export_tuples(uint64_t* d_out, uint64_t capacity, uint64_t* counts) {
        if (!streamed_) throw arg_error("build: export before stream");
        if (capacity < n_tuples_) throw arg_error("build: tuple buffer too small");
        if (world_ == 1) {
            counts[0] = 0;
            return;
        }
        if (n_tuples_) {
            tuple_gather_kernel<<<(unsigned)((n_tuples_ + 255) / 256), 256>>>(d_table_.p, d_idx_.p, n_tuples_, d_out);
            g_launches++;
        }
        DevBuf<uint64_t> d_bounds((size_t)world_ + 1);
        owner_bounds_kernel<<<(unsigned)((world_ + 1 + 63) / 64), 64>>>(d_idx_.p, n_tuples_, slice_cols_, world_, d_bounds.p);
        g_launches++;
        FIN_CUDA(cudaGetLastError());
        std::vector<uint64_t> bounds((size_t)world_ + 1);
        FIN_CUDA(cudaMemcpy(bounds.data(), d_bounds.p, bounds.size() * 8, cudaMemcpyDeviceToHost));
        for (int g = 0; g < world_; g++) counts[g] = bounds[(size_t)g + 1] - bounds[(size_t)g];
    }

    // fold the tuples that arrived for this rank's slice; then fmin bits and offsets of the slice in column order
    // (FinimizerIndex.hh:301-306).  Returns the number of finimizers in the slice.
    uint64_t merge(const uint64_t* d_tuples, uint64_t n) {
        if (!streamed_) throw arg_error("build: merge before stream");
        const unsigned long long* slice_table = d_table_.p;
        DevBuf<unsigned long long> d_slice;
        if (world_ > 1) {
            d_slice.alloc(slice_cols_);
            FIN_CUDA(cudaMemset(d_slice.p, 0, d_slice.bytes()));
            if (n) {
                tuple_merge_kernel<<<(unsigned)((n + 255) / 256), 256>>>(d_tuples, n, slice_begin_, slice_cols_, d_slice.p, d_err_.p);
                g_launches++;
                FIN_CUDA(cudaGetLastError());
            }
            check_errors();
            slice_table = d_slice.p;
            d_table_.release();
            d_idx_.release();
        }
        const uint64_t n_valid = slice_begin_ >= N_ ? 0 : std::min<uint64_t>(slice_cols_, N_ - slice_begin_);
        const uint64_t n_words32 = (slice_cols_ + 31) / 32;
        d_bits_.alloc(n_words32 + 2);
        FIN_CUDA(cudaMemset(d_bits_.p, 0, d_bits_.bytes()));
        n_slice_fmin_ = 0;
        if (n_valid) {
            DevBuf<uint32_t> d_offs(n_valid);
            d_sel_.alloc(n_valid);
            DevBuf<uint64_t> d_nsel(1);
            fmin_bits_kernel<<<(unsigned)((((n_valid + 31) / 32) * 32 + 255) / 256), 256>>>(slice_table, n_valid, d_bits_.p, d_offs.p);
            g_launches++;
            FIN_CUDA(cudaGetLastError());
            size_t tb = 0;
            cub::DeviceSelect::If(nullptr, tb, d_offs.p, d_sel_.p, d_nsel.p, (int64_t)n_valid, NotSentinel());
            DevBuf<uint8_t> tmp(tb + 256);
            FIN_CUDA(cub::DeviceSelect::If(tmp.p, tb, d_offs.p, d_sel_.p, d_nsel.p, (int64_t)n_valid, NotSentinel()));
            g_launches += 2;
            FIN_CUDA(cudaMemcpy(&n_slice_fmin_, d_nsel.p, 8, cudaMemcpyDeviceToHost));
        }
        merged_ = true;
        return n_slice_fmin_;
    }

    uint64_t slice_words() const { return (slice_cols_ + 63) / 64; }

    // device -> caller's DEVICE buffers: slice_words() u64 of fmin bits, n_slice_fmin u32 offsets
    void slice(uint64_t* d_bits_out, uint32_t* d_offsets_out) const {
        if (!merged_) throw arg_error("build: slice before merge");
        FIN_CUDA(cudaMemcpy(d_bits_out, d_bits_.p, slice_words() * 8, cudaMemcpyDeviceToDevice));
        if (n_slice_fmin_) FIN_CUDA(cudaMemcpy(d_offsets_out, d_sel_.p, n_slice_fmin_ * 4, cudaMemcpyDeviceToDevice));
    }
    void slice_to_host(std::vector<uint64_t>& bits, std::vector<uint32_t>& offs) const {
        if (!merged_) throw arg_error("build: slice before merge");
        bits.assign(slice_words(), 0);
        offs.resize(n_slice_fmin_);
        // d_bits_ holds 32-bit words; an odd count leaves the upper half of the last u64 zero (allocated + 2, zeroed)
        FIN_CUDA(cudaMemcpy(bits.data(), d_bits_.p, slice_words() * 8, cudaMemcpyDeviceToHost));
        if (n_slice_fmin_) FIN_CUDA(cudaMemcpy(offs.data(), d_sel_.p, n_slice_fmin_ * 4, cudaMemcpyDeviceToHost));
    }

    // all slices in rank order (HOST): world * slice_words() u64 of fmin bits, the offsets back to back
    void finish(const uint64_t* bits_all, uint64_t n_bit_words, const uint32_t* offsets_all, uint64_t n_offsets, const std::string& out_prefix) {
        NvtxRange range("fin:build:write_files");
        save_index_from_slices(std::move(sbwt_), std::move(lcs_file_), std::move(text_), std::move(ends_), std::move(ustart_), bits_all,
                               n_bit_words, offsets_all, n_offsets, out_prefix);
    }

   private:
    void check_bases() {
        unsigned err = 0;
        FIN_CUDA(cudaMemcpy(&err, d_err_.p, 4, cudaMemcpyDeviceToHost));
        if (err & kErrBadBase) throw domain_error("Invalid character");  // PackedStrings.hh:57
    }
    void check_errors() {
        unsigned err = 0;
        FIN_CUDA(cudaMemcpy(&err, d_err_.p, 4, cudaMemcpyDeviceToHost));
        if (err)
            throw domain_error("the unitigs are not a disjoint spectrum-preserving string set of this SBWT "
                               "(a unitig k-mer is missing from the SBWT or a window has no unique substring)");
    }

    SbwtHost sbwt_;
    PackedVec lcs_file_, text_, ends_, ustart_;
    int rank_, world_, k_ = 0;
    uint64_t N_ = 0, slice_cols_ = 0, slice_begin_ = 0, bases_streamed_ = 0, n_tuples_ = 0, n_slice_fmin_ = 0;
    size_t chunk_begin_ = 0, chunk_end_ = 0;
    float stream_ms_ = 0.f;
    bool streamed_ = false, merged_ = false;
    std::vector<BuildChunk> chunks_;
    IndexView ix_{};
    DevBuf<uint64_t> d_lines_, d_text_;
    DevBuf<unsigned long long> d_table_;
    DevBuf<uint32_t> d_idx_, d_bits_, d_sel_;
    DevBuf<unsigned int> d_err_;
};

extern thread_local fin_build_timing g_build_timing;
static double wall_ms();

static void build_rarest(SbwtHost&& sbwt, PackedVec&& lcs_file, const DeviceSbwt& ds, DevBuf<uint8_t>&& d_lcs_bytes, const SeqBatch& unitigs,
                         const std::string& out_prefix) {
    const double t0 = wall_ms();
    RarestBuilder b(std::move(sbwt), std::move(lcs_file), ds, std::move(d_lcs_bytes), unitigs, 0, 1);
    const double t1 = wall_ms();
    b.stream();
    b.merge(nullptr, 0);
    std::vector<uint64_t> bits;
    std::vector<uint32_t> offs;
    b.slice_to_host(bits, offs);
    const double t2 = wall_ms();
    b.finish(bits.data(), bits.size(), offs.data(), offs.size(), out_prefix);
    g_build_timing.prepare_ms = (float)(t1 - t0);
    g_build_timing.stream_merge_ms = (float)(t2 - t1);
    g_build_timing.stream_kernel_ms = b.stream_ms();
    g_build_timing.write_ms = (float)(wall_ms() - t2);
}

// ---- B5 / B6: --type shortest | verify, statistics only (build_fmin.hh:95-214,252-267) ---------------------------
// Unitigs stay in FILE order here (no permutation): the shared fmin_found bit vector of B5 makes the result depend on it.
// build-fmin --type shortest|verify in stages, shardable over GPUs.  Rank r of `world`:
//   shortest: streams a contiguous share of the chunks.  Pass 0 folds "first window per column whose finimizer ends at
//             or after position k-1 of its unitig" into table[colex] with atomicMin - a min over ALL ranks' tables gives
//             the table of the single pass (the closed form of build_fmin.hh:182-187 does not care who streamed what) -,
//             pass 1 emits this rank's keys against the reduced table.
//   verify  : a contiguous range of text positions; the per-position records are recomputed for the k positions past
//             the range (a window reads the records of its k starts), the keys of the range's windows are this rank's.
// The union of all ranks' keys, sorted and made unique, is the reference's std::set (common.hh:188-206).
class StatsBuilder {
   public:
    StatsBuilder(const SbwtHost& sbwt, const DeviceSbwt& ds, const uint8_t* d_lcs_bytes, const SeqBatch& unitigs, const std::string& type,
                 int64_t t, int rank, int world)
        : shortest_(type == "shortest"), t_((uint32_t)t), k_((int)sbwt.k), N_((uint64_t)sbwt.n_nodes), n_kmers_(sbwt.n_kmers) {
        if (type != "shortest" && type != "verify") throw arg_error("stats: type must be shortest or verify");
        if (world < 1 || rank < 0 || rank >= world) throw arg_error("stats: rank / world out of range");
        const uint64_t U = unitigs.size();
        check_unitigs(unitigs, k_);
        check_stats_t(t, sbwt.n_nodes);
        d_lines_ = make_lines_device(ds, d_lcs_bytes);
        ix_.lines = reinterpret_cast<const uint4*>(d_lines_.p);
        ix_.n_nodes = (idx_t)N_;
        ix_.k = k_;
        std::vector<uint64_t> order(U), starts;  // FILE order (build_fmin.hh:202-214), not the colex permutation
        std::iota(order.begin(), order.end(), 0);
        PackedVec text;
        pack_unitigs(unitigs, order, text, starts);
        const PackedVec ends = unitig_ends(unitigs, order, starts);
        T_ = text.size();
        d_text_ = to_device(text.words, 2);
        d_err_.alloc(1);
        FIN_CUDA(cudaMemset(d_err_.p, 0, 4));
        if (shortest_) {
            const std::vector<BuildChunk> chunks = plan_build_chunks(starts, ends, k_, 0);
            const std::pair<size_t, size_t> mine = shard_chunks(chunks, rank, world);
            nc_ = (uint32_t)(mine.second - mine.first);
            d_chunks_.alloc((size_t)nc_ + 1);
            if (nc_) FIN_CUDA(cudaMemcpy(d_chunks_.p, chunks.data() + mine.first, (size_t)nc_ * sizeof(BuildChunk), cudaMemcpyHostToDevice));
            d_first_.alloc(N_);
            d_runs_.alloc((size_t)nc_ + 1);
            FIN_CUDA(cudaMemset(d_first_.p, 0xFF, d_first_.bytes()));
            FIN_CUDA(cudaMemset(d_runs_.p, 0, d_runs_.bytes()));
        } else {
            std::vector<uint32_t> ends32(U);
            for (uint64_t u = 0; u < U; u++) ends32[u] = (uint32_t)ends.words[u];
            d_ends_ = to_device(ends32);
            n_unitigs_ = (uint32_t)U;
            p0_ = T_ * (uint64_t)rank / (uint64_t)world;
            p1_ = T_ * (uint64_t)(rank + 1) / (uint64_t)world;
        }
    }

    uint64_t table_entries() const { return shortest_ ? N_ : 0; }
    uint32_t* table() { return d_first_.p; }

    void first_pass() {
        if (shortest_) launch_shortest(std::integral_constant<int, 0>());
        passed0_ = true;
    }

    uint64_t second_pass() {
        if (!passed0_) throw arg_error("stats: second pass before the first");
        if (shortest_) {
            {  // run counts -> offsets; the element past the last chunk becomes the total
                size_t tb = 0;
                cub::DeviceScan::ExclusiveSum(nullptr, tb, d_runs_.p, d_runs_.p, (int64_t)nc_ + 1);
                DevBuf<uint8_t> tmp(tb + 256);
                FIN_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, d_runs_.p, d_runs_.p, (int64_t)nc_ + 1));
                g_launches += 2;
            }
            uint32_t total = 0;
            FIN_CUDA(cudaMemcpy(&total, d_runs_.p + nc_, 4, cudaMemcpyDeviceToHost));
            n_keys_ = total;
            d_keys_.alloc(n_keys_ + 1);
            launch_shortest(std::integral_constant<int, 1>());
        } else {
            const uint64_t r1 = std::min<uint64_t>(T_, p1_ + (uint64_t)k_);  // records of [p0, r1), keys of the windows that start in [p0, p1)
            n_keys_ = p1_ - p0_;
            DevBuf<uint64_t> d_rec(r1 - p0_ + 1);
            d_keys_.alloc(n_keys_ + 1);
            if (r1 > p0_) {
                rare_prefix_kernel<<<(unsigned)((r1 - p0_ + 255) / 256), 256>>>(ix_, d_text_.p, d_ends_.p, n_unitigs_, (uint32_t)p0_, (uint32_t)r1, t_,
                                                                              d_rec.p, d_err_.p);
                g_launches++;
            }
            if (n_keys_) {
                verify_window_kernel<<<(unsigned)((n_keys_ + 255) / 256), 256>>>(d_rec.p, d_ends_.p, n_unitigs_, (uint32_t)p0_, (uint32_t)p1_, k_,
                                                                               d_keys_.p, d_err_.p);
                g_launches++;
            }
            FIN_CUDA(cudaGetLastError());
            FIN_CUDA(cudaDeviceSynchronize());
        }
        unsigned err = 0;
        FIN_CUDA(cudaMemcpy(&err, d_err_.p, 4, cudaMemcpyDeviceToHost));
        if (err)
            throw domain_error("the unitigs are not a disjoint spectrum-preserving string set of this SBWT "
                               "(a unitig k-mer is missing from the SBWT or a window has no substring with at most t occurrences)");
        return n_keys_;
    }

    uint64_t n_keys() const { return n_keys_; }
    const uint64_t* keys() const { return d_keys_.p; }

    // the std::set of the reference over the keys of ALL ranks (device): sort, unique, sum -> the stats string
    std::string finish(const uint64_t* d_keys_all, uint64_t n_keys) const {
        unsigned long long sums[3] = {0, 0, 0};
        if (n_keys) {
            DevBuf<uint64_t> d_sorted(n_keys), d_unique(n_keys);
            DevBuf<uint64_t> d_nuniq(1);
            size_t tb1 = 0, tb2 = 0;
            cub::DeviceRadixSort::SortKeys(nullptr, tb1, d_keys_all, d_sorted.p, (int64_t)n_keys);
            cub::DeviceSelect::Unique(nullptr, tb2, d_sorted.p, d_unique.p, d_nuniq.p, (int64_t)n_keys);
            DevBuf<uint8_t> tmp(std::max(tb1, tb2) + 256);
            FIN_CUDA(cub::DeviceRadixSort::SortKeys(tmp.p, tb1, d_keys_all, d_sorted.p, (int64_t)n_keys));
            FIN_CUDA(cub::DeviceSelect::Unique(tmp.p, tb2, d_sorted.p, d_unique.p, d_nuniq.p, (int64_t)n_keys));
            g_launches += 10;
            uint64_t n_unique = 0;
            FIN_CUDA(cudaMemcpy(&n_unique, d_nuniq.p, 8, cudaMemcpyDeviceToHost));
            DevBuf<unsigned long long> d_sums(3);
            FIN_CUDA(cudaMemset(d_sums.p, 0, 24));
            if (n_unique) {
                stat_sum_kernel<<<(unsigned)((n_unique + 255) / 256), 256>>>(d_unique.p, n_unique, d_sums.p);
                g_launches++;
                FIN_CUDA(cudaGetLastError());
            }
            FIN_CUDA(cudaMemcpy(sums, d_sums.p, 24, cudaMemcpyDeviceToHost));
        }
        return stats_string(sums[0], sums[1], sums[2], n_kmers_);
    }

    int t() const { return (int)t_; }

   private:
    template <class Pass> void launch_shortest(Pass) {
        constexpr int P = Pass::value;
        if (nc_ == 0) return;
        const unsigned blocks = (nc_ + 127) / 128;
        switch (window_capacity(k_)) {
            case 32: shortest_stream_kernel<32, P><<<blocks, 128>>>(ix_, d_text_.p, d_chunks_.p, nc_, t_, d_first_.p, d_runs_.p, d_keys_.p, d_err_.p); break;
            case 64: shortest_stream_kernel<64, P><<<blocks, 128>>>(ix_, d_text_.p, d_chunks_.p, nc_, t_, d_first_.p, d_runs_.p, d_keys_.p, d_err_.p); break;
            case 128: shortest_stream_kernel<128, P><<<blocks, 128>>>(ix_, d_text_.p, d_chunks_.p, nc_, t_, d_first_.p, d_runs_.p, d_keys_.p, d_err_.p); break;
            default: shortest_stream_kernel<256, P><<<blocks, 128>>>(ix_, d_text_.p, d_chunks_.p, nc_, t_, d_first_.p, d_runs_.p, d_keys_.p, d_err_.p); break;
        }
        g_launches++;
        FIN_CUDA(cudaGetLastError());
        FIN_CUDA(cudaDeviceSynchronize());
    }

    bool shortest_, passed0_ = false;
    uint32_t t_;
    int k_;
    uint64_t N_, T_ = 0, n_keys_ = 0, p0_ = 0, p1_ = 0;
    int64_t n_kmers_;
    uint32_t nc_ = 0, n_unitigs_ = 0;
    IndexView ix_{};
    DevBuf<uint64_t> d_lines_, d_text_, d_keys_;
    DevBuf<BuildChunk> d_chunks_;
    DevBuf<uint32_t> d_first_, d_runs_, d_ends_;
    DevBuf<unsigned int> d_err_;
};

static std::string build_stats(const SbwtHost& sbwt, const DeviceSbwt& ds, const uint8_t* d_lcs_bytes, const SeqBatch& unitigs,
                               const std::string& type, int64_t t) {
    StatsBuilder b(sbwt, ds, d_lcs_bytes, unitigs, type, t, 0, 1);
    b.first_pass();
    const uint64_t n = b.second_pass();
    return b.finish(b.keys(), n);
}

// ---- staged builder behind the fin_builder_* entry points ---------------------------------------------------------
// LCS of a build: given as a file (--lcs) or computed on the device and saved (build_fmin.hh:373-380; the reference then
// loads the file back, which yields the same vector).  Returns the file form and leaves the padded bytes on the device.
static PackedVec lcs_for_build(const SbwtHost& sbwt, const DeviceSbwt& ds, const std::string& lcs_path, const std::string& out_prefix,
                               bool write_file, DevBuf<uint8_t>& d_bytes) {
    if (!lcs_path.empty()) {
        PackedVec lcs = load_packed_file(lcs_path, 0);
        if (lcs.size() != (uint64_t)sbwt.n_nodes) throw format_error("LCS length differs from the number of SBWT columns");
        d_bytes = lcs_bytes_from_packed(lcs, ds.nw, sbwt.k);
        return lcs;
    }
    d_bytes = lcs_on_device(ds);
    const int w = bit_length((uint64_t)(sbwt.k - 1));  // lcs_basic_parallel_algorithm.hpp:115
    if (w == 0) throw domain_error("k = 1 is outside the parity domain");
    PackedVec lcs = pack_on_device<uint8_t>(d_bytes.p, (uint64_t)sbwt.n_nodes, w);
    if (write_file) save_packed_file(out_prefix + ".LCS.sdsl", lcs, true);
    return lcs;
}

RarestBuilder* builder_open(const std::string& sbwt_path, const std::string& unitig_path, const std::string& lcs_path,
                            const std::string& out_prefix, int device, int rank, int world) {
    const std::vector<std::string> files = unitig_file_list(unitig_path);
    SeqBatch unitigs;  // read by a second host thread while this one loads the SBWT and computes the LCS
    std::string unitig_error;
    std::thread unitig_reader([&] {
        try {
            append_sequences(files[0], unitigs);
        } catch (const std::exception& e) {
            unitig_error = e.what();
            if (unitig_error.empty()) unitig_error = "cannot read the unitig file";
        }
    });
    struct Joiner {
        std::thread& t;
        ~Joiner() { if (t.joinable()) t.join(); }
    } joiner{unitig_reader};
    SbwtHost sbwt = load_sbwt_file(sbwt_path, true);
    select_device(device);
    check_sbwt_limits(sbwt);
    const DeviceSbwt ds = upload_sbwt(sbwt);
    DevBuf<uint8_t> d_lcs;
    PackedVec lcs = lcs_for_build(sbwt, ds, lcs_path, out_prefix, false, d_lcs);  // rank 0 writes it with the other files (finish)
    unitig_reader.join();
    if (!unitig_error.empty()) throw io_error(unitig_error);
    return new RarestBuilder(std::move(sbwt), std::move(lcs), ds, std::move(d_lcs), unitigs, rank, world);
}
void builder_free(RarestBuilder* b) { delete b; }
fin_builder_info builder_info(const RarestBuilder& b) {
    fin_builder_info i{};
    i.n_nodes = (int64_t)b.n_nodes();
    i.slice_columns = (int64_t)b.slice_cols();
    i.slice_words = (int64_t)b.slice_words();
    i.text_len = (int64_t)b.text_len();
    i.n_unitigs = (int64_t)b.n_unitigs();
    i.n_chunks = (int64_t)b.n_chunks();
    i.chunk_begin = (int64_t)b.chunk_begin();
    i.chunk_end = (int64_t)b.chunk_end();
    i.bases_streamed = (int64_t)b.bases_streamed();
    i.rank = b.rank();
    i.world = b.world();
    i.stream_ms = b.stream_ms();
    return i;
}
uint64_t builder_stream(RarestBuilder& b) { return b.stream(); }
void builder_export(RarestBuilder& b, uint64_t* d_out, uint64_t capacity, uint64_t* counts) { b.export_tuples(d_out, capacity, counts); }
uint64_t builder_merge(RarestBuilder& b, const uint64_t* d_tuples, uint64_t n) { return b.merge(d_tuples, n); }
void builder_slice(const RarestBuilder& b, uint64_t* d_bits, uint32_t* d_offsets) { b.slice(d_bits, d_offsets); }
void builder_finish(RarestBuilder& b, const uint64_t* bits_all, uint64_t n_bit_words, const uint32_t* offsets_all, uint64_t n_offsets,
                    const std::string& out_prefix) {
    b.finish(bits_all, n_bit_words, offsets_all, n_offsets, out_prefix);
    append_stats_line(out_prefix, 1, "");
}

thread_local fin_build_timing g_build_timing{};
fin_build_timing last_build_timing() { return g_build_timing; }

// ---- staged stats modes behind the fin_stats_* entry points --------------------------------------------------------
struct StatsHandle {
    SbwtHost sbwt;
    DevBuf<uint8_t> d_lcs;
    std::unique_ptr<StatsBuilder> b;
    std::string out_prefix;
    int rank = 0;
};

StatsHandle* stats_open(const std::string& sbwt_path, const std::string& unitig_path, const std::string& lcs_path, const std::string& out_prefix,
                        const std::string& type, int64_t t, int device, int rank, int world) {
    const std::vector<std::string> files = unitig_file_list(unitig_path);
    std::unique_ptr<StatsHandle> h(new StatsHandle());
    h->sbwt = load_sbwt_file(sbwt_path, true);
    if (type != "shortest" && type != "verify") throw arg_error("Error: unknown type: " + type);
    select_device(device);
    check_sbwt_limits(h->sbwt);
    const DeviceSbwt ds = upload_sbwt(h->sbwt);
    // the reference computes and saves the LCS for every type (build_fmin.hh:373-380): rank 0 writes it
    lcs_for_build(h->sbwt, ds, lcs_path, out_prefix, rank == 0 && !out_prefix.empty(), h->d_lcs);
    SeqBatch unitigs;
    append_sequences(files[0], unitigs);
    h->b.reset(new StatsBuilder(h->sbwt, ds, h->d_lcs.p, unitigs, type, (int64_t)(signed char)t, rank, world));
    h->out_prefix = out_prefix;
    h->rank = rank;
    return h.release();
}
void stats_free(StatsHandle* h) { delete h; }
uint64_t stats_first_pass(StatsHandle& h) {
    h.b->first_pass();
    return h.b->table_entries();
}
void stats_table(StatsHandle& h, uint32_t* d_table, bool set) {
    const uint64_t n = h.b->table_entries();
    if (!n) return;
    if (set) FIN_CUDA(cudaMemcpy(h.b->table(), d_table, n * 4, cudaMemcpyDeviceToDevice));
    else FIN_CUDA(cudaMemcpy(d_table, h.b->table(), n * 4, cudaMemcpyDeviceToDevice));
}
uint64_t stats_second_pass(StatsHandle& h) { return h.b->second_pass(); }
void stats_keys(const StatsHandle& h, uint64_t* d_keys) {
    if (h.b->n_keys()) FIN_CUDA(cudaMemcpy(d_keys, h.b->keys(), h.b->n_keys() * 8, cudaMemcpyDeviceToDevice));
}
std::string stats_finish(StatsHandle& h, const uint64_t* d_keys_all, uint64_t n_keys_all, bool write_line) {
    const std::string r = h.b->finish(d_keys_all, n_keys_all);
    if (write_line && !h.out_prefix.empty()) append_stats_line(h.out_prefix, h.b->t(), r);
    return r;
}

static double wall_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec / 1e6;
}

std::string build_fmin_device(const std::string& sbwt_path, const std::string& unitig_path, const std::string& out_prefix,
                              const std::string& type, int64_t t_arg, const std::string& lcs_path, int device) {
    const signed char t = (signed char)t_arg;  // build_fmin.hh:333 narrows to char
    g_build_timing = fin_build_timing{};
    const double t_begin = wall_ms();
    const std::vector<std::string> files = unitig_file_list(unitig_path);
    // the unitig file is read and parsed by a second host thread while this one loads the SBWT and computes the LCS
    SeqBatch unitigs;
    std::string unitig_error;
    std::thread unitig_reader([&] {
        try {
            append_sequences(files[0], unitigs);
        } catch (const std::exception& e) {
            unitig_error = e.what();
            if (unitig_error.empty()) unitig_error = "cannot read the unitig file";
        }
    });
    struct Joiner {
        std::thread& t;
        ~Joiner() { if (t.joinable()) t.join(); }
    } joiner{unitig_reader};
    SbwtHost sbwt = load_sbwt_file(sbwt_path, true);
    if (type != "rarest" && type != "shortest" && type != "verify") throw arg_error("Error: unknown type: " + type);
    select_device(device);
    check_sbwt_limits(sbwt);

    const double t_loaded = wall_ms();
    const DeviceSbwt ds = upload_sbwt(sbwt);
    DevBuf<uint8_t> d_lcs;
    // rarest writes <prefix>.LCS.sdsl with the other six files at the end; the stats types only write this one
    PackedVec lcs = lcs_for_build(sbwt, ds, lcs_path, out_prefix, type != "rarest", d_lcs);
    const double t_lcs = wall_ms();

    std::string result;
    unitig_reader.join();
    if (!unitig_error.empty()) throw io_error(unitig_error);
    const double t_unitigs = wall_ms();
    g_build_timing.load_sbwt_ms = (float)(t_loaded - t_begin);
    g_build_timing.lcs_ms = (float)(t_lcs - t_loaded);
    g_build_timing.read_unitigs_ms = (float)(t_unitigs - t_lcs);
    g_build_timing.unitig_bases = (int64_t)unitigs.seqs.size();
    if (type == "rarest") {
        if (t != 1) throw domain_error("t != 1 does not make sense with rarest type");  // :245-247
        build_rarest(std::move(sbwt), std::move(lcs), ds, std::move(d_lcs), unitigs, out_prefix);
    } else {
        result = build_stats(sbwt, ds, d_lcs.p, unitigs, type, t);
    }
    append_stats_line(out_prefix, (int)t, result);
    g_build_timing.total_ms = (float)(wall_ms() - t_begin);
    return result;
}

}  // namespace fin
