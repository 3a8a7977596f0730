// Pure host code: the seven FinimizerIndex files in memory and their re-layout into the arrays the
// device kernels read (device_index.cuh).  No CUDA here, so tests can build the same layout on the CPU.
#pragma once
#include <functional>
#include <thread>
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "formats.hpp"

namespace fin {

constexpr uint64_t kRankBlockBitsHost = 192;  // must equal kRankBlockBits (device_index.cuh)
constexpr uint64_t kIndexLimit32 = 0xFFFFFFFFull - 256;

// 32-byte rank sectors {ones before block, 3 words} over a bit vector; `n_rows` interleaved rows per block.
// rows[r] points at the words of row r (at least ceil(n_bits/64) words).  4*n_rows u64 per block for
// n_bits/192 + 1 blocks, so rank(n_bits) is addressable.
inline std::vector<uint64_t> make_rank_sectors(const uint64_t* const* rows, int n_rows, uint64_t n_bits) {
    const uint64_t n_blocks = n_bits / kRankBlockBitsHost + 1;
    const uint64_t n_words = (n_bits + 63) >> 6;
    const uint64_t tail_bits = n_bits & 63;
    std::vector<uint64_t> s(n_blocks * 4 * (uint64_t)n_rows, 0);
    for (int r = 0; r < n_rows; r++) {
        uint64_t cum = 0;
        const uint64_t* w = rows[r];
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
    }
    return s;
}

// The seven files of a FinimizerIndex as the reference holds them in memory (FinimizerIndex.hh:108-115)
struct IndexHost {
    SbwtHost sbwt;
    PackedVec lcs, fmin, offsets, text, ends, ustart;
};

inline IndexHost load_index_files(const std::string& prefix) {  // FinimizerIndex.hh:209-241
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

inline void save_index_files(const IndexHost& ix, const std::string& prefix) {  // FinimizerIndex.hh:187-207
    // seven independent files: one writer thread each (at 3 Gbp they add up to 7 GB)
    auto write_sbwt = [&] {
        const std::string path = prefix + ".sbwt";
        FILE* f = fopen(path.c_str(), "wb");
        if (!f) throw io_error("cannot write " + path);
        const size_t n = ix.sbwt.body_len;
        const bool ok = fwrite(ix.sbwt.body(), 1, n, f) == n;
        if (fclose(f) != 0 || !ok) throw io_error("short write on " + path);
    };
    std::vector<std::function<void()>> jobs = {
        write_sbwt,
        [&] { save_packed_file(prefix + ".LCS.sdsl", ix.lcs, true); },
        [&] { save_packed_file(prefix + ".O.sdsl", ix.offsets, true); },
        [&] { save_packed_file(prefix + ".packed_unitigs.sdsl", ix.text, false); },
        [&] { save_packed_file(prefix + ".FBV.sdsl", ix.fmin, false); },
        [&] { save_packed_file(prefix + ".Ustart.sdsl", ix.ustart, false); },
        [&] { save_packed_file(prefix + ".unitig_endpoints.sdsl", ix.ends, true); },
    };
    std::vector<std::string> errors(jobs.size());
    std::vector<std::thread> threads;
    for (size_t j = 0; j < jobs.size(); j++)
        threads.emplace_back([&, j] {
            try {
                jobs[j]();
            } catch (const std::exception& e) {
                errors[j] = e.what();
            }
        });
    for (auto& t : threads) t.join();
    for (const std::string& e : errors)
        if (!e.empty()) throw io_error(e);
}

// The 128-byte lines of device_index.cuh: per 64 columns four rows {u64 bits, u32 C[c] + ones before, u32 extra}
// followed by 64 LCS bytes.  lcs / ustart / fmin may be null (build-fmin streams before they exist): zeros then.
// (N >> 6) + 2 lines, so that column N (rank of the whole vector, LCS pad = 0) is addressable.
inline std::vector<uint64_t> make_lines(const SbwtHost& s, const PackedVec* lcs, const PackedVec* ustart, const PackedVec* fmin) {
    const uint64_t N = (uint64_t)s.n_nodes;
    const uint64_t n_lines = (N >> 6) + 2;
    const uint64_t n_words = (N + 63) >> 6;
    const uint64_t tail_bits = N & 63;
    std::vector<uint64_t> L(n_lines * 16, 0);
    uint64_t cum[4] = {(uint64_t)s.C[0], (uint64_t)s.C[1], (uint64_t)s.C[2], (uint64_t)s.C[3]};
    for (uint64_t b = 0; b < n_lines; b++) {
        uint64_t* line = &L[b * 16];
        const uint64_t us = ustart && b < n_words ? ustart->words[b] : 0;
        const uint64_t fm = fmin && b < n_words ? fmin->words[b] : 0;
        const uint32_t extra[4] = {(uint32_t)us, (uint32_t)(us >> 32), (uint32_t)fm, (uint32_t)(fm >> 32)};
        for (int c = 0; c < 4; c++) {
            uint64_t x = b < n_words ? s.bits[c].words[b] : 0;
            if (tail_bits && b == n_words - 1) x &= (1ULL << tail_bits) - 1;  // bits past the end never count
            line[2 * c] = x;
            line[2 * c + 1] = (cum[c] & 0xFFFFFFFFull) | ((uint64_t)extra[c] << 32);
            cum[c] += (uint64_t)__builtin_popcountll(x);
        }
        if (lcs) {
            uint8_t* lb = reinterpret_cast<uint8_t*>(line + 8);
            const uint64_t i0 = b * 64, i1 = std::min(N, i0 + 64);
            for (uint64_t i = i0; i < i1; i++) lb[i - i0] = (uint8_t)lcs->get(i);
        }
    }
    return L;
}

// Arrays in exactly the form the kernels read them (see device_index.cuh for the meaning of each).
struct HostLayout {
    std::vector<uint64_t> lines, ustart, fmin, tstart, text;
    std::vector<uint32_t> offsets, ends;
    uint64_t n_nodes = 0, text_len = 0, n_unitigs = 0, n_fmin = 0, max_unitig_len = 0;
    int64_t C[4] = {0, 0, 0, 0};
    int64_t k = 0, n_kmers = 0;
};

inline HostLayout build_host_layout(const IndexHost& h, bool with_lines = true) {
    HostLayout L;
    const uint64_t N = (uint64_t)h.sbwt.n_nodes, T = h.text.size(), U = h.ends.size(), F = h.offsets.size();
    if (N >= kIndexLimit32 || T >= kIndexLimit32)
        throw limit_error("index too large for the 32-bit column/text positions of this build");
    if (N < 2) throw domain_error("SBWT with fewer than 2 columns");
    if (U == 0 || h.ends.get(U - 1) != T) throw format_error("unitig end points do not match the packed text");
    L.n_nodes = N; L.text_len = T; L.n_unitigs = U; L.n_fmin = F;
    L.k = h.sbwt.k; L.n_kmers = h.sbwt.n_kmers;
    for (int c = 0; c < 4; c++) L.C[c] = h.sbwt.C[c];
    if (with_lines) L.lines = make_lines(h.sbwt, &h.lcs, &h.ustart, &h.fmin);  // the device path builds them with make_lines_kernel
    {
        const uint64_t* r[1] = {h.ustart.words.data()};
        L.ustart = make_rank_sectors(r, 1, N);
    }
    {
        const uint64_t* r[1] = {h.fmin.words.data()};
        L.fmin = make_rank_sectors(r, 1, N);
    }
    L.ends.resize(U);
    {
        std::vector<uint64_t> ts((T >> 6) + 2, 0);
        uint64_t prev = 0;
        for (uint64_t u = 0; u < U; u++) {
            const uint64_t e = h.ends.get(u);
            if (e <= prev || e > T) throw format_error("unitig end points are not strictly increasing");
            ts[prev >> 6] |= 1ULL << (prev & 63);
            L.ends[u] = (uint32_t)e;
            L.max_unitig_len = std::max(L.max_unitig_len, e - prev);
            prev = e;
        }
        const uint64_t* r[1] = {ts.data()};
        L.tstart = make_rank_sectors(r, 1, T);
    }
    L.offsets.assign(F + 1, 0);
    for (uint64_t i = 0; i < F; i++) {
        const uint64_t v = h.offsets.get(i);
        if (v >= kIndexLimit32) throw limit_error("finimizer offset beyond 32 bits");
        L.offsets[i] = (uint32_t)v;
    }
    const uint64_t nw = (T + 31) / 32 + 2;
    L.text.assign(nw, 0);
    memcpy(L.text.data(), h.text.words.data(), std::min<uint64_t>(nw, h.text.words.size()) * 8);
    return L;
}

}  // namespace fin
