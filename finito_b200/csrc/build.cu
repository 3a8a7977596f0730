// build-fmin on the device: host driver around the kernels of build_kernels.cuh.  Reproduces
// FinimizerIndexBuilder's constructor (include/FinimizerIndex.hh:273-319), the build_fmin driver
// (include/build_fmin.hh:302-401) and FinimizerIndex::serialize (:187-207).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>

#include "build_host.hpp"
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

// ---- B1: LCS bytes on the device (k launches of lcs_round_kernel) --------------------------------------
static DevBuf<uint8_t> lcs_on_device(const SbwtHost& s) {
    uint64_t nw = 0;
    const std::vector<uint64_t> h_rows = sbwt_row_words(s, nw);
    DevBuf<uint64_t> rows = to_device(h_rows);
    const uint64_t N = (uint64_t)s.n_nodes;
    const uint64_t padded = nw * 64 + 64;
    DevBuf<uint8_t> last(padded), next(padded), lcs(padded);
    DevBuf<uint32_t> word_dst(4 * nw);
    const unsigned T = 256;
    row_popcount_kernel<<<(unsigned)((4 * nw + T - 1) / T), T>>>(rows.p, nw, word_dst.p);
    g_launches++;
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, word_dst.p, word_dst.p, (int64_t)nw);
    DevBuf<uint8_t> tmp(tmp_bytes + 256);
    for (int c = 0; c < 4; c++) {
        size_t tb = tmp_bytes;
        uint32_t* row = word_dst.p + (uint64_t)c * nw;
        FIN_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, row, row, (int64_t)nw));
        g_launches += 2;
    }
    LcsArgs a{};
    a.rows = rows.p;
    a.word_dst = word_dst.p;
    a.n = N;
    a.n_words = nw;
    for (int c = 0; c < 4; c++) a.C[c] = (uint32_t)s.C[c];
    a.k = (uint8_t)s.k;
    FIN_CUDA(cudaMemset(last.p, 0, padded));
    FIN_CUDA(cudaMemset(next.p, 0, padded));
    FIN_CUDA(cudaMemset(lcs.p, 0, padded));
    lcs_init_kernel<<<(unsigned)((N + T - 1) / T), T>>>(a, last.p, lcs.p);
    g_launches++;
    uint8_t* cur = last.p;
    uint8_t* nxt = next.p;
    for (int64_t round = 0; round < s.k; round++) {
        lcs_round_kernel<<<(unsigned)((nw + 127) / 128), 128>>>(a, cur, nxt, lcs.p, (uint8_t)round);
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
    DevBuf<uint8_t> d = lcs_on_device(s);
    std::vector<uint8_t> h((size_t)s.n_nodes);
    FIN_CUDA(cudaMemcpy(h.data(), d.p, h.size(), cudaMemcpyDeviceToHost));
    return h;
}

static int window_capacity(int k) { return k <= 32 ? 32 : k <= 64 ? 64 : k <= 128 ? 128 : 256; }

// ---- B2-B4: FinimizerIndexBuilder -------------------------------------------------------------------------
static void build_rarest(SbwtHost&& sbwt, PackedVec&& lcs_file, const SeqBatch& unitigs, const std::string& out_prefix) {
    const int k = (int)sbwt.k;
    const uint64_t N = (uint64_t)sbwt.n_nodes, U = unitigs.size();
    if (lcs_file.size() != N) throw format_error("LCS length differs from the number of SBWT columns");
    check_unitigs(unitigs, k);

    DevBuf<uint64_t> d_lines = to_device(make_lines(sbwt, &lcs_file, nullptr, nullptr));
    IndexView ix{};
    ix.lines = reinterpret_cast<const uint4*>(d_lines.p);
    ix.n_nodes = (idx_t)N;
    ix.k = k;

    // first k-mer columns on the file-order text, then the permuted text
    std::vector<uint64_t> order(U), starts;
    std::iota(order.begin(), order.end(), 0);
    PackedVec text;
    pack_unitigs(unitigs, order, text, starts);
    DevBuf<uint64_t> d_text = to_device(text.words, 2);
    DevBuf<uint64_t> d_starts = to_device(starts);
    DevBuf<uint32_t> d_cols(U);
    first_kmer_kernel<<<(unsigned)((U + 127) / 128), 128>>>(ix, d_text.p, d_starts.p, (uint32_t)U, d_cols.p);
    g_launches++;
    FIN_CUDA(cudaGetLastError());
    std::vector<uint32_t> cols(U);
    FIN_CUDA(cudaMemcpy(cols.data(), d_cols.p, U * 4, cudaMemcpyDeviceToHost));

    PackedVec ustart;
    order = unitig_order(cols, N, ustart);
    pack_unitigs(unitigs, order, text, starts);
    FIN_CUDA(cudaMemcpy(d_text.p, text.words.data(), text.words.size() * 8, cudaMemcpyHostToDevice));
    PackedVec ends = unitig_ends(unitigs, order, starts);

    // add_sequence over chunks; overwrite rule as a max over table[colex]
    const std::vector<BuildChunk> chunks = plan_build_chunks(starts, ends, k, 0);
    DevBuf<BuildChunk> d_chunks = to_device(chunks);
    DevBuf<unsigned long long> d_table(N);
    FIN_CUDA(cudaMemset(d_table.p, 0, d_table.bytes()));
    DevBuf<unsigned int> d_err(1);
    FIN_CUDA(cudaMemset(d_err.p, 0, 4));
    {
        const unsigned blocks = (unsigned)((chunks.size() + 127) / 128);
        const uint32_t nc = (uint32_t)chunks.size();
        switch (window_capacity(k)) {
            case 32: build_stream_kernel<32><<<blocks, 128>>>(ix, d_text.p, d_chunks.p, nc, d_table.p, d_err.p); break;
            case 64: build_stream_kernel<64><<<blocks, 128>>>(ix, d_text.p, d_chunks.p, nc, d_table.p, d_err.p); break;
            case 128: build_stream_kernel<128><<<blocks, 128>>>(ix, d_text.p, d_chunks.p, nc, d_table.p, d_err.p); break;
            default: build_stream_kernel<256><<<blocks, 128>>>(ix, d_text.p, d_chunks.p, nc, d_table.p, d_err.p); break;
        }
        g_launches++;
        FIN_CUDA(cudaGetLastError());
    }
    unsigned err = 0;
    FIN_CUDA(cudaMemcpy(&err, d_err.p, 4, cudaMemcpyDeviceToHost));
    if (err)
        throw domain_error("the unitigs are not a disjoint spectrum-preserving string set of this SBWT "
                           "(a unitig k-mer is missing from the SBWT or a window has no unique substring)");

    // fmin bit vector and offsets in column order (FinimizerIndex.hh:301-306)
    const uint64_t n_words32 = (N + 31) / 32;
    DevBuf<uint32_t> d_bits(n_words32 + 2), d_offs(N), d_sel(N);
    DevBuf<uint64_t> d_nsel(1);
    FIN_CUDA(cudaMemset(d_bits.p, 0, d_bits.bytes()));
    fmin_bits_kernel<<<(unsigned)((n_words32 * 32 + 255) / 256), 256>>>(d_table.p, N, d_bits.p, d_offs.p);
    g_launches++;
    FIN_CUDA(cudaGetLastError());
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

    IndexHost out;
    out.fmin.assign(N, 1);
    {
        std::vector<uint32_t> b32(n_words32 + 2, 0);
        FIN_CUDA(cudaMemcpy(b32.data(), d_bits.p, n_words32 * 4, cudaMemcpyDeviceToHost));
        memcpy(out.fmin.words.data(), b32.data(), std::min<size_t>(n_words32 * 4, out.fmin.words.size() * 8));
    }
    out.offsets = pack_offsets(offs);
    out.sbwt = std::move(sbwt);
    out.lcs = std::move(lcs_file);
    out.text = std::move(text);
    out.ends = std::move(ends);
    out.ustart = std::move(ustart);
    save_index_files(out, out_prefix);
}

// ---- B5 / B6: --type shortest | verify, statistics only (build_fmin.hh:95-214,252-267) ---------------------------
// Unitigs stay in FILE order here (no permutation): the shared fmin_found bit vector of B5 makes the result depend on it.
static std::string build_stats(const SbwtHost& sbwt, const PackedVec& lcs_file, const SeqBatch& unitigs, const std::string& type,
                               int64_t t) {
    const int k = (int)sbwt.k;
    const uint64_t N = (uint64_t)sbwt.n_nodes, U = unitigs.size();
    if (lcs_file.size() != N) throw format_error("LCS length differs from the number of SBWT columns");
    check_unitigs(unitigs, k);
    check_stats_t(t, sbwt.n_nodes);

    DevBuf<uint64_t> d_lines = to_device(make_lines(sbwt, &lcs_file, nullptr, nullptr));
    IndexView ix{};
    ix.lines = reinterpret_cast<const uint4*>(d_lines.p);
    ix.n_nodes = (idx_t)N;
    ix.k = k;
    std::vector<uint64_t> order(U), starts;
    std::iota(order.begin(), order.end(), 0);
    PackedVec text;
    pack_unitigs(unitigs, order, text, starts);
    const PackedVec ends = unitig_ends(unitigs, order, starts);
    const uint64_t T = text.size();
    DevBuf<uint64_t> d_text = to_device(text.words, 2);
    DevBuf<unsigned int> d_err(1);
    FIN_CUDA(cudaMemset(d_err.p, 0, 4));

    DevBuf<uint64_t> d_keys;
    uint64_t n_keys = 0;
    if (type == "shortest") {
        const std::vector<BuildChunk> chunks = plan_build_chunks(starts, ends, k, 0);
        const uint32_t nc = (uint32_t)chunks.size();
        DevBuf<BuildChunk> d_chunks = to_device(chunks);
        DevBuf<uint32_t> d_first(N), d_runs((size_t)nc + 1);
        FIN_CUDA(cudaMemset(d_first.p, 0xFF, d_first.bytes()));
        FIN_CUDA(cudaMemset(d_runs.p, 0, d_runs.bytes()));
        const unsigned blocks = (nc + 127) / 128;
        auto launch = [&](auto pass) {
            constexpr int P = decltype(pass)::value;
            switch (window_capacity(k)) {
                case 32: shortest_stream_kernel<32, P><<<blocks, 128>>>(ix, d_text.p, d_chunks.p, nc, (uint32_t)t, d_first.p, d_runs.p, d_keys.p, d_err.p); break;
                case 64: shortest_stream_kernel<64, P><<<blocks, 128>>>(ix, d_text.p, d_chunks.p, nc, (uint32_t)t, d_first.p, d_runs.p, d_keys.p, d_err.p); break;
                case 128: shortest_stream_kernel<128, P><<<blocks, 128>>>(ix, d_text.p, d_chunks.p, nc, (uint32_t)t, d_first.p, d_runs.p, d_keys.p, d_err.p); break;
                default: shortest_stream_kernel<256, P><<<blocks, 128>>>(ix, d_text.p, d_chunks.p, nc, (uint32_t)t, d_first.p, d_runs.p, d_keys.p, d_err.p); break;
            }
            g_launches++;
            FIN_CUDA(cudaGetLastError());
        };
        launch(std::integral_constant<int, 0>());
        {  // run counts -> offsets; the element past the last chunk becomes the total
            size_t tb = 0;
            cub::DeviceScan::ExclusiveSum(nullptr, tb, d_runs.p, d_runs.p, (int64_t)nc + 1);
            DevBuf<uint8_t> tmp(tb + 256);
            FIN_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tb, d_runs.p, d_runs.p, (int64_t)nc + 1));
            g_launches += 2;
        }
        uint32_t total = 0;
        FIN_CUDA(cudaMemcpy(&total, d_runs.p + nc, 4, cudaMemcpyDeviceToHost));
        n_keys = total;
        d_keys.alloc(n_keys + 1);
        launch(std::integral_constant<int, 1>());
    } else {  // verify
        std::vector<uint32_t> ends32(U);
        for (uint64_t u = 0; u < U; u++) ends32[u] = (uint32_t)ends.words[u];
        DevBuf<uint32_t> d_ends = to_device(ends32);
        DevBuf<uint64_t> d_rec(T + 1);
        n_keys = T;
        d_keys.alloc(T + 1);
        const unsigned blocks = (unsigned)((T + 255) / 256);
        rare_prefix_kernel<<<blocks, 256>>>(ix, d_text.p, d_ends.p, (uint32_t)U, (uint32_t)T, (uint32_t)t, d_rec.p, d_err.p);
        verify_window_kernel<<<blocks, 256>>>(d_rec.p, d_ends.p, (uint32_t)U, (uint32_t)T, k, d_keys.p, d_err.p);
        g_launches += 2;
        FIN_CUDA(cudaGetLastError());
    }
    unsigned err = 0;
    FIN_CUDA(cudaMemcpy(&err, d_err.p, 4, cudaMemcpyDeviceToHost));
    if (err)
        throw domain_error("the unitigs are not a disjoint spectrum-preserving string set of this SBWT "
                           "(a unitig k-mer is missing from the SBWT or a window has no substring with at most t occurrences)");

    // the std::set of the reference: sort, unique, sum
    unsigned long long sums[3] = {0, 0, 0};
    if (n_keys) {
        DevBuf<uint64_t> d_sorted(n_keys), d_unique(n_keys);
        DevBuf<uint64_t> d_nuniq(1);
        size_t tb1 = 0, tb2 = 0;
        cub::DeviceRadixSort::SortKeys(nullptr, tb1, d_keys.p, d_sorted.p, (int64_t)n_keys);
        cub::DeviceSelect::Unique(nullptr, tb2, d_sorted.p, d_unique.p, d_nuniq.p, (int64_t)n_keys);
        DevBuf<uint8_t> tmp(std::max(tb1, tb2) + 256);
        FIN_CUDA(cub::DeviceRadixSort::SortKeys(tmp.p, tb1, d_keys.p, d_sorted.p, (int64_t)n_keys));
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
    return stats_string(sums[0], sums[1], sums[2], sbwt.n_kmers);
}

std::string build_fmin_device(const std::string& sbwt_path, const std::string& unitig_path, const std::string& out_prefix,
                              const std::string& type, int64_t t_arg, const std::string& lcs_path, int device) {
    const signed char t = (signed char)t_arg;  // build_fmin.hh:333 narrows to char
    const std::vector<std::string> files = unitig_file_list(unitig_path);
    SbwtHost sbwt = load_sbwt_file(sbwt_path, true);
    if (type != "rarest" && type != "shortest" && type != "verify") throw arg_error("Error: unknown type: " + type);
    select_device(device);
    check_sbwt_limits(sbwt);

    std::string lcs_file = lcs_path;
    if (lcs_file.empty()) {  // :373-380: compute, save, then load it back like the reference does
        lcs_file = out_prefix + ".LCS.sdsl";
        DevBuf<uint8_t> d = lcs_on_device(sbwt);
        std::vector<uint8_t> h((size_t)sbwt.n_nodes);
        FIN_CUDA(cudaMemcpy(h.data(), d.p, h.size(), cudaMemcpyDeviceToHost));
        save_packed_file(lcs_file, lcs_bytes_to_packed(h, h.size(), sbwt.k), true);
    }
    PackedVec lcs = load_packed_file(lcs_file, 0);

    std::string result;
    SeqBatch unitigs;
    append_sequences(files[0], unitigs);
    if (type == "rarest") {
        if (t != 1) throw domain_error("t != 1 does not make sense with rarest type");  // :245-247
        build_rarest(std::move(sbwt), std::move(lcs), unitigs, out_prefix);
    } else {
        result = build_stats(sbwt, lcs, unitigs, type, t);
    }
    append_stats_line(out_prefix, (int)t, result);
    return result;
}

}  // namespace fin
