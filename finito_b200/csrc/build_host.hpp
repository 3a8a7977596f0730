// Pure host logic of build-fmin shared by the CUDA driver (build.cu) and the CPU emulation used by the
// test-suite: unitig packing and permutation (PackedStrings.hh:36-64,104-137), the work plan of the
// build stream, and the assembly of the index files (FinimizerIndex.hh:301-306, :187-207).
#pragma once
#include <thread>
#include <algorithm>
#include <numeric>

#include "layout_host.hpp"
#include "seqio.hpp"

namespace fin {

struct BuildChunk {
    uint32_t text_begin;    // global text position of the first streamed base (inside the unitig)
    uint32_t first_pos;     // position of that base in its unitig
    uint32_t n_stream;      // bases to stream
    uint32_t first_emit;    // unitig-local END position of the first k-mer window this chunk reports
    uint32_t unitig_start;  // global text position of the unitig
};

// the four rows as plain words, row-major rows[c * n_words + w], bits past n_nodes cleared
inline std::vector<uint64_t> sbwt_row_words(const SbwtHost& s, uint64_t& n_words) {
    const uint64_t N = (uint64_t)s.n_nodes;
    n_words = (N + 63) / 64;
    std::vector<uint64_t> h(4 * n_words, 0);
    const uint64_t tail = N & 63;
    for (int c = 0; c < 4; c++) {
        memcpy(&h[(uint64_t)c * n_words], s.bits[c].words.data(), n_words * 8);
        if (tail) h[(uint64_t)c * n_words + n_words - 1] &= (1ULL << tail) - 1;
    }
    return h;
}

inline void check_sbwt_limits(const SbwtHost& s) {
    if ((uint64_t)s.n_nodes >= kIndexLimit32) throw limit_error("SBWT too large for the 32-bit columns of this build");
    if (s.n_nodes < 2) throw domain_error("SBWT with fewer than 2 columns");
}

inline PackedVec lcs_bytes_to_packed(const std::vector<uint8_t>& bytes, uint64_t n, int64_t k) {
    PackedVec v;
    const int w = bit_length((uint64_t)(k - 1));  // lcs_basic_parallel_algorithm.hpp:115
    if (w == 0) throw domain_error("k = 1 is outside the parity domain");
    v.assign(n, (uint8_t)w);
    for (uint64_t i = 0; i < n; i++) v.put(i, bytes[i]);
    return v;
}

inline void check_unitigs(const SeqBatch& unitigs, int64_t k) {
    if (unitigs.size() == 0) throw domain_error("no unitigs in the input");
    if (unitigs.seqs.size() >= kIndexLimit32) throw limit_error("unitig text too large for the 32-bit positions of this build");
    for (uint64_t u = 0; u < unitigs.size(); u++)
        if (unitigs.offsets[u + 1] - unitigs.offsets[u] < (uint64_t)k)
            throw domain_error("a unitig is shorter than k (the reference reads out of bounds there, PackedStrings.hh:116)");
}

// concatenate the unitigs in the given order, 2 bits per base (PackedStrings.hh:36-64)
inline void pack_unitigs(const SeqBatch& unitigs, const std::vector<uint64_t>& order, PackedVec& text,
                         std::vector<uint64_t>& starts) {
    const uint64_t U = unitigs.size();
    text.assign(unitigs.seqs.size(), 2);
    starts.resize(U);
    uint64_t pos = 0;
    for (uint64_t j = 0; j < U; j++) {
        const uint64_t u = order[j];
        const uint64_t b = unitigs.offsets[u], e = unitigs.offsets[u + 1];
        starts[j] = pos;
        for (uint64_t i = b; i < e; i++) {
            uint64_t code;
            switch (unitigs.seqs[i]) {
                case 'A': code = 0; break;
                case 'C': code = 1; break;
                case 'G': code = 2; break;
                case 'T': code = 3; break;
                default: throw domain_error("Invalid character");  // PackedStrings.hh:57
            }
            text.words[pos >> 5] |= code << (2 * (pos & 31));
            pos++;
        }
    }
}

// permute_unitigs (PackedStrings.hh:104-137): the colex order of the first k-mers is the order of their SBWT
// columns; ties (impossible in a DSPSS) fall back to the file index like the reference's (Kmer, index) pairs.
inline std::vector<uint64_t> unitig_order(const std::vector<uint32_t>& cols, uint64_t n_nodes, PackedVec& ustart) {
    const uint64_t U = cols.size();
    ustart.assign(n_nodes, 1);
    for (uint64_t u = 0; u < U; u++) {
        if (cols[u] == 0xFFFFFFFFu) throw domain_error("the first k-mer of a unitig is not in the SBWT (PackedStrings.hh:129-131)");
        ustart.words[cols[u] >> 6] |= 1ULL << (cols[u] & 63);
    }
    std::vector<uint64_t> order(U);
    std::iota(order.begin(), order.end(), 0);
    std::sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) { return cols[a] != cols[b] ? cols[a] < cols[b] : a < b; });
    return order;
}

inline PackedVec unitig_ends(const SeqBatch& unitigs, const std::vector<uint64_t>& order, const std::vector<uint64_t>& starts) {
    PackedVec ends;
    ends.assign(unitigs.size(), 64);  // width stays 64 (PackedStrings.hh:44 passes the bit length as the default VALUE)
    for (uint64_t j = 0; j < unitigs.size(); j++)
        ends.words[j] = starts[j] + (unitigs.offsets[order[j] + 1] - unitigs.offsets[order[j]]);
    return ends;
}

// Chunks of k-mer windows.  A chunk streams `warm` extra bases to the left of its first window; after k-1 bases
// the finimizer interval and after 2k-2 the candidate window no longer depend on where the stream began
// (DESIGN.md "chunked streams"), so every chunk reports exactly what one pass over the whole unitig would.
inline std::vector<BuildChunk> plan_build_chunks(const std::vector<uint64_t>& starts, const PackedVec& ends, int64_t k,
                                                 uint64_t per_chunk /*0 = choose*/) {
    const uint64_t U = starts.size();
    uint64_t total_windows = 0;
    for (uint64_t j = 0; j < U; j++) total_windows += ends.words[j] - starts[j] - (uint64_t)k + 1;
    if (per_chunk == 0) per_chunk = std::min<uint64_t>(1024, std::max<uint64_t>(64, total_windows / (148ull * 1024)));
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
    return chunks;
}

// ---- build-fmin sharded over GPUs -------------------------------------------------------
// Every rank knows the whole plan (SBWT, LCS, permutation, chunk list are replicated: they are cheap next to the
// stream).  Rank r streams the chunks [first, last) below — contiguous in permuted-unitig order, balanced by streamed
// bases — and OWNS the SBWT columns [r * S, (r + 1) * S), S a multiple of 64 so that the fmin words of the slices
// concatenate.  The (colex, key) tuples a rank produced for columns it does not own travel to their owner in one
// all-to-all; the owner folds them into its slice with the same max as the single-GPU table.
inline uint64_t slice_columns(uint64_t n_nodes, int world) {
    const uint64_t per = (n_nodes + (uint64_t)world - 1) / (uint64_t)world;
    return std::max<uint64_t>(64, (per + 63) / 64 * 64);
}

inline std::pair<size_t, size_t> shard_chunks(const std::vector<BuildChunk>& chunks, int rank, int world) {
    uint64_t total = 0;
    for (const BuildChunk& c : chunks) total += c.n_stream;
    auto boundary = [&](int r) -> size_t {  // first chunk whose prefix sum reaches r/world of the work
        if (r <= 0) return 0;
        if (r >= world) return chunks.size();
        const uint64_t want = total / (uint64_t)world * (uint64_t)r;
        uint64_t acc = 0;
        size_t i = 0;
        while (i < chunks.size() && acc < want) acc += chunks[i++].n_stream;
        return i;
    };
    return {boundary(rank), boundary(rank + 1)};
}

// tuples sorted by colex -> how many go to each owner
inline std::vector<uint64_t> tuples_per_owner(const uint64_t* tuples /* {colex, key} pairs */, uint64_t n, uint64_t slice_cols, int world) {
    std::vector<uint64_t> counts((size_t)world, 0);
    for (uint64_t i = 0; i < n; i++) {
        const uint64_t owner = tuples[2 * i] / slice_cols;
        if (owner >= (uint64_t)world) throw arg_error("internal: tuple beyond the last slice");
        counts[owner]++;
    }
    return counts;
}

// global_offsets of the marked columns, packed at the width of the largest one (FinimizerIndex.hh:301-306)
inline PackedVec pack_offsets(const uint32_t* offs, uint64_t n) {
    // blocks of 64 entries start on a word boundary whatever the width: the blocks are packed by several threads
    const unsigned T = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(16, n / (1u << 20)));
    const uint64_t per = ((n + T - 1) / T + 63) / 64 * 64;
    std::vector<uint64_t> maxes(T, 0);
    auto each = [&](auto&& fn) {
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < T; t++) pool.emplace_back(fn, t);
        fn(0u);
        for (auto& th : pool) th.join();
    };
    each([&](unsigned t) {
        uint64_t m = 0;
        for (uint64_t i = std::min(n, t * per), e = std::min(n, (t + 1) * per); i < e; i++) m = std::max<uint64_t>(m, offs[i]);
        maxes[t] = m;
    });
    const uint64_t max_off = *std::max_element(maxes.begin(), maxes.end());
    PackedVec p;
    const int w = bit_length(max_off);
    p.assign(n, (uint8_t)(w ? w : 1));  // sdsl cannot hold width 0 either; only reachable with a single base of text
    each([&](unsigned t) {
        for (uint64_t i = std::min(n, t * per), e = std::min(n, (t + 1) * per); i < e; i++) p.put(i, offs[i]);
    });
    return p;
}

// print_finimizer_stats (common.hh:188-197): "<#finimizers>,<sum freq>,<mean freq>,<mean len>,<#k-mers>", the means in
// float and printed by std::to_string(float)
inline std::string stats_string(uint64_t n_fmin, uint64_t sum_freq, uint64_t sum_len, int64_t n_kmers) {
    return std::to_string((int64_t)n_fmin) + "," + std::to_string((int64_t)sum_freq) + "," +
           std::to_string(static_cast<float>((int64_t)sum_freq) / static_cast<float>((int64_t)n_fmin)) + "," +
           std::to_string(static_cast<float>((int64_t)sum_len) / static_cast<float>((int64_t)n_fmin)) + "," + std::to_string(n_kmers);
}

// the parity domain of the statistics modes: 1 <= t <= 127 (the reference narrows t to char, build_fmin.hh:333)
// and t < n_nodes (with t >= n_nodes its shrink loop never ends, build_fmin.hh:155-166)
inline void check_stats_t(int64_t t, int64_t n_nodes) {
    if (t < 1 || t > 127) throw domain_error("-t outside 1..127 (the reference narrows it to char)");
    if (t >= n_nodes) throw domain_error("-t is not smaller than the number of SBWT columns (the reference does not terminate)");
}

// all column slices in rank order -> the seven index files (FinimizerIndex.hh:301-306, :187-207)
inline void save_index_from_slices(SbwtHost&& sbwt, PackedVec&& lcs, PackedVec&& text, PackedVec&& ends, PackedVec&& ustart,
                                   const uint64_t* bits_all, uint64_t n_bit_words, const uint32_t* offsets_all, uint64_t n_offsets,
                                   const std::string& out_prefix) {
    IndexHost out;
    out.fmin.assign((uint64_t)sbwt.n_nodes, 1);
    const uint64_t need = out.fmin.disk_words();
    if (n_bit_words < need) throw arg_error("build: fmin words missing");
    memcpy(out.fmin.words.data(), bits_all, need * 8);
    uint64_t ones = 0;
    for (uint64_t w = 0; w < need; w++) ones += (uint64_t)__builtin_popcountll(out.fmin.words[w]);
    if (ones != n_offsets) throw arg_error("build: the slices disagree on the number of finimizers");
    out.offsets = pack_offsets(offsets_all, n_offsets);
    out.sbwt = std::move(sbwt);
    out.lcs = std::move(lcs);
    out.text = std::move(text);
    out.ends = std::move(ends);
    out.ustart = std::move(ustart);
    save_index_files(out, out_prefix);
}

inline void append_stats_line(const std::string& out_prefix, int t, const std::string& result) {  // build_fmin.hh:386-395
    const std::string stats_path = out_prefix + "_stats.txt";
    FILE* sf = fopen(stats_path.c_str(), "ab");
    if (!sf) throw io_error("Error: Unable to open file.");
    const std::string line = std::to_string(t) + "," + result + "\n";
    fwrite(line.data(), 1, line.size(), sf);
    fclose(sf);
}

inline std::vector<std::string> unitig_file_list(const std::string& unitig_path) {  // build_fmin.hh:336-343
    std::vector<std::string> files;
    if (has_suffix(unitig_path, ".txt")) files = read_lines(unitig_path);
    else files.push_back(unitig_path);
    for (const std::string& f : files) {
        FILE* fp = fopen(f.c_str(), "rb");
        if (!fp) throw io_error("Error: cannot read file " + f);
        fclose(fp);
    }
    if (files.empty()) throw arg_error("no unitig file given");
    if (files.size() > 1)
        throw domain_error("more than one unitig file: the reference moves its SBWT into the first builder (build_fmin.hh:281-300)");
    return files;
}

}  // namespace fin
