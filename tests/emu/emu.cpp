// TEST INFRASTRUCTURE ONLY.  Compiles the per-thread bodies of the CUDA kernels (finito_b200/csrc/*.cuh,
// built with -DFIN_EMU) for the host, one "thread" after the other, so that the kernel logic can be diffed
// against the oracle in the CPU-only container before any GPU time is spent.  Nothing in the product links
// or loads this file; the shipped library has no CPU path.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

struct ulonglong2 { unsigned long long x, y; };
struct uint4 { unsigned int x, y, z, w; };
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
static inline int __clzll(long long x) { return x ? __builtin_clzll((unsigned long long)x) : 64; }
static inline int __ffsll(long long x) { return __builtin_ffsll(x); }
using std::max;
using std::min;

#define FIN_EMU 1
#include "../../finito_b200/csrc/build_kernels.cuh"
#include "../../finito_b200/csrc/runs_host.hpp"

using namespace fin;

static thread_local std::string g_err;
template <class F> static int guarded(F f) {
    try { f(); return 0; } catch (const std::exception& e) { g_err = e.what(); return 1; }
}

static IndexView view_of(const HostLayout& L) {
    IndexView v{};
    v.lines = reinterpret_cast<const uint4*>(L.lines.data());
    v.ustart = reinterpret_cast<const ulonglong2*>(L.ustart.data());
    v.fmin = reinterpret_cast<const ulonglong2*>(L.fmin.data());
    v.tstart = reinterpret_cast<const ulonglong2*>(L.tstart.data());
    v.offsets = L.offsets.data();
    v.ends = L.ends.data();
    v.text = L.text.data();
    v.n_nodes = (idx_t)L.n_nodes;
    v.text_len = (idx_t)L.text_len;
    v.n_unitigs = (uint32_t)L.n_unitigs;
    v.n_fmin = (uint32_t)L.n_fmin;
    v.k = (int32_t)L.k;
    return v;
}

template <int KCAP, class HitT, bool COUNT, bool RUNS>
static void run_chain(const IndexView& v, const uint64_t* packed, uint64_t b0, uint64_t len, uint64_t read, const SearchOut<HitT>& out,
                      ReadCounters& rc, int n_strands) {
    Chain<KCAP, HitT, COUNT, RUNS> ch;
    uint64_t dq[KCAP];
    if (!ch.begin_read(v, b0, len, read, out, rc, n_strands)) return;
    while (ch.st != ST_IDLE) ch.step(v, packed, rc, n_strands, dq, out);
}

template <class HitT, bool COUNT, bool RUNS>
static void run_read(const IndexView& v, const uint64_t* packed, uint64_t b0, uint64_t len, uint64_t read, const SearchOut<HitT>& out,
                     ReadCounters& rc, int n_strands) {
    const int k = v.k;
    if (k <= 32) run_chain<32, HitT, COUNT, RUNS>(v, packed, b0, len, read, out, rc, n_strands);
    else if (k <= 64) run_chain<64, HitT, COUNT, RUNS>(v, packed, b0, len, read, out, rc, n_strands);
    else if (k <= 128) run_chain<128, HitT, COUNT, RUNS>(v, packed, b0, len, read, out, rc, n_strands);
    else run_chain<256, HitT, COUNT, RUNS>(v, packed, b0, len, read, out, rc, n_strands);
}

static std::vector<uint8_t> emu_lcs_bytes(const SbwtHost& s) {
    uint64_t nw = 0;
    std::vector<uint64_t> rows = sbwt_row_words(s, nw);
    std::vector<uint32_t> word_dst(4 * nw);
    for (int c = 0; c < 4; c++) {
        uint32_t cum = 0;
        for (uint64_t w = 0; w < nw; w++) {
            word_dst[c * nw + w] = cum;
            cum += (uint32_t)__builtin_popcountll(rows[c * nw + w]);
        }
    }
    const uint64_t padded = nw * 64 + 64;
    std::vector<uint8_t> last(padded, 0), next(padded, 0), lcs(padded, 0);
    LcsArgs a{};
    a.rows = rows.data();
    a.word_dst = word_dst.data();
    a.n = (uint64_t)s.n_nodes;
    a.n_words = nw;
    for (int c = 0; c < 4; c++) a.C[c] = (uint32_t)s.C[c];
    a.k = (uint8_t)s.k;
    for (uint64_t i = 0; i < a.n; i++) { last[i] = lcs_first_column(a, i); lcs[i] = a.k; }
    for (int64_t r = 0; r < s.k; r++) {
        for (uint64_t w = 0; w < nw; w++) lcs_round_word(a, last.data(), next.data(), lcs.data(), w, (uint8_t)r);
        last.swap(next);
    }
    lcs.resize((size_t)s.n_nodes);
    return lcs;
}

// build-fmin --type shortest|verify with the kernel bodies (one "thread" after the other), staged like StatsBuilder in
// build.cu so that the gloo tests can shard it: rank r of `world` streams a share of the chunks (shortest) or of the text
// positions (verify)
struct EmuStats {
    SbwtHost sbwt;
    PackedVec lcs_file, text, ends;
    std::vector<uint64_t> lines;
    IndexView ix{};
    std::string type;
    uint32_t t = 1;
    int k = 0;
    uint64_t N = 0, T = 0, p0 = 0, p1 = 0;
    std::vector<BuildChunk> chunks;  // this rank's
    std::vector<uint32_t> first, runs;
    std::vector<uint64_t> keys;
    unsigned err = 0;

    EmuStats(const char* sbwt_path, const char* unitig_path, const std::string& type_, int64_t t_, uint64_t per_chunk, int rank, int world)
        : type(type_), t((uint32_t)t_) {
        if (type != "shortest" && type != "verify") throw arg_error("emu_stats: type");
        const std::vector<std::string> files = unitig_file_list(unitig_path);
        sbwt = load_sbwt_file(sbwt_path, true);
        check_sbwt_limits(sbwt);
        N = (uint64_t)sbwt.n_nodes;
        k = (int)sbwt.k;
        check_stats_t(t_, sbwt.n_nodes);
        lcs_file = lcs_bytes_to_packed(emu_lcs_bytes(sbwt), N, k);
        SeqBatch unitigs;
        append_sequences(files[0], unitigs);
        check_unitigs(unitigs, k);
        const uint64_t U = unitigs.size();
        lines = make_lines(sbwt, &lcs_file, nullptr, nullptr);
        ix.lines = reinterpret_cast<const uint4*>(lines.data());
        ix.n_nodes = (idx_t)N;
        ix.k = k;
        std::vector<uint64_t> order(U), starts;
        for (uint64_t u = 0; u < U; u++) order[u] = u;
        pack_unitigs(unitigs, order, text, starts);
        text.words.resize(text.words.size() + 2, 0);
        ends = unitig_ends(unitigs, order, starts);
        T = text.size();
        if (type == "shortest") {
            const std::vector<BuildChunk> all = plan_build_chunks(starts, ends, k, per_chunk);
            const std::pair<size_t, size_t> mine = shard_chunks(all, rank, world);
            chunks.assign(all.begin() + (long)mine.first, all.begin() + (long)mine.second);
            first.assign(N, 0xFFFFFFFFu);
            runs.assign(chunks.size() + 1, 0);
        } else {
            p0 = T * (uint64_t)rank / (uint64_t)world;
            p1 = T * (uint64_t)(rank + 1) / (uint64_t)world;
        }
    }

    template <class Pass> uint32_t run(Pass, size_t c, uint64_t* out) {
        constexpr int P = Pass::value;
        if (k <= 32) return shortest_stream_chunk<32, P>(ix, text.words.data(), chunks[c], t, first.data(), out, &err);
        if (k <= 64) return shortest_stream_chunk<64, P>(ix, text.words.data(), chunks[c], t, first.data(), out, &err);
        if (k <= 128) return shortest_stream_chunk<128, P>(ix, text.words.data(), chunks[c], t, first.data(), out, &err);
        return shortest_stream_chunk<256, P>(ix, text.words.data(), chunks[c], t, first.data(), out, &err);
    }

    uint64_t first_pass() {
        if (type != "shortest") return 0;
        // the chunks in REVERSE order: the result must not depend on the order the "threads" run in
        for (size_t c = chunks.size(); c-- > 0;) runs[c] = run(std::integral_constant<int, 0>(), c, nullptr);
        return N;
    }

    uint64_t second_pass() {
        keys.clear();
        if (type == "shortest") {
            uint64_t total = 0;
            for (size_t c = 0; c < chunks.size(); c++) { const uint32_t n = runs[c]; runs[c] = (uint32_t)total; total += n; }
            keys.assign(total + 1, kNoStatKey);
            for (size_t c = chunks.size(); c-- > 0;) run(std::integral_constant<int, 1>(), c, keys.data() + runs[c]);
            keys.resize(total);
        } else {
            const uint64_t r1 = std::min<uint64_t>(T, p1 + (uint64_t)k);
            std::vector<uint64_t> rec(r1 - p0 + 1, kNoStatKey);
            uint64_t u = 0;
            for (uint64_t p = p0; p < r1; p++) {
                while (ends.words[u] <= p) u++;
                rec[p - p0] = shortest_rare_prefix(ix, text.words.data(), p, (int)std::min<uint64_t>((uint64_t)k, ends.words[u] - p), t, err);
            }
            u = 0;
            for (uint64_t p = p0; p < p1; p++) {
                while (ends.words[u] <= p) u++;
                if (p + (uint64_t)k <= ends.words[u]) {
                    const uint64_t key = verify_window_min(rec.data(), p - p0, k);
                    if (key == kNoStatKey) err |= kErrWindowUnderflow;
                    keys.push_back(key);
                }
            }
        }
        if (err) throw domain_error("emu: stats stream error flags " + std::to_string(err));
        return keys.size();
    }

    std::string finish(const uint64_t* all, uint64_t n_all) const {
        std::vector<uint64_t> v(all, all + n_all);
        std::sort(v.begin(), v.end());
        v.erase(std::unique(v.begin(), v.end()), v.end());
        uint64_t n = 0, sf = 0, sl = 0;
        for (uint64_t key : v)
            if (key != kNoStatKey) { n++; sf += (key >> 32) & 0xFF; sl += key >> 40; }
        return stats_string(n, sf, sl, sbwt.n_kmers);
    }
};

extern "C" {

const char* emu_last_error() { return g_err.c_str(); }

// counts4: found_fwd, found_rc, positive, n_queries; access8: the eight access counters; flags: error bits
int emu_search(const char* prefix, const char* seqs, const uint64_t* offs, uint64_t n_reads, int64_t* out, int64_t* counts4,
               int64_t* access8, uint32_t* flags, int n_strands, int runs_mode, uint64_t* n_runs_out, int prefilter_s,
               uint64_t* strands_streamed) {
    return guarded([&] {
        IndexHost h = load_index_files(prefix);
        HostLayout L = build_host_layout(h);
        IndexView v = view_of(L);
        const uint64_t nb = offs[n_reads];
        std::vector<uint64_t> packed(nb / 32 + 2, 0);
        *flags = 0;
        for (uint64_t i = 0; i < nb; i++) {
            const uint32_t ch = (uint8_t)seqs[i] & ~32u;
            if (!(ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T')) *flags |= kErrBadBase;
            packed[i >> 5] |= (uint64_t)(((ch >> 1) ^ (ch >> 2)) & 3u) << (2 * (i & 31));
        }
        for (int i = 0; i < 4; i++) counts4[i] = 0;
        for (int i = 0; i < 8; i++) access8[i] = 0;
        // what the device pipeline prepares: output offsets (exclusive scan of the k-mer counts) and the found-mask scratch
        std::vector<uint64_t> out_offsets(n_reads + 1, 0);
        for (uint64_t r = 0; r < n_reads; r++) {
            const uint64_t len = offs[r + 1] - offs[r];
            out_offsets[r + 1] = out_offsets[r] + (len >= (uint64_t)v.k ? len - v.k + 1 : 0);
        }
        const uint64_t total = out_offsets[n_reads];
        std::vector<uint32_t> fmask((total >> 5) + n_reads + 2, 0xDEADBEEFu);
        std::vector<uint64_t> keys(total + 1);
        std::vector<RunRec> recs(total + 1);
        unsigned long long n_runs = 0;
        // strand prefilter (DeviceIndex's constructor + strand_filter_kernel): s-mer table of the text, flags per read
        std::vector<uint32_t> smer;
        std::vector<uint8_t> sflags(n_reads + 1, 0);
        if (prefilter_s > 0) {
            const int s = std::min<int>(prefilter_s, (int)v.k);
            smer.assign(std::max<uint64_t>(1, (1ull << (2 * s)) / 32), 0);
            for (uint64_t p = 0; p + (uint64_t)s <= L.text_len; p++) {
                const uint64_t w0 = L.text[p >> 5], w1 = L.text[(p >> 5) + 1];
                const int sh = 2 * (int)(p & 31);
                const uint64_t x = sh ? (w0 >> sh) | (w1 << (64 - sh)) : w0;
                const uint32_t code = (uint32_t)x & smer_mask(s);
                smer[code >> 5] |= 1u << (code & 31);
            }
            v.smer = smer.data();
            v.smer_s = s;
            for (uint64_t r = 0; r < n_reads; r++) {
                unsigned lookups = 0;
                unsigned f = strand_candidates(v, packed.data(), offs[r], offs[r + 1] - offs[r], lookups);
                if (n_strands < 2) f &= kStrandFwd;
                sflags[r] = (uint8_t)f;
                if (strands_streamed) *strands_streamed += (f & 1) + ((f >> 1) & 1);
                if (!runs_mode && f == 0)
                    for (uint64_t i = out_offsets[r]; i < out_offsets[r + 1]; i++) out[2 * i] = out[2 * i + 1] = -1;
            }
        }
        SearchOut<int64_t> so{reinterpret_cast<HitPair<int64_t>*>(out), out_offsets.data(), fmask.data(),
                              RunSink{keys.data(), recs.data(), &n_runs, total}, prefilter_s > 0 ? sflags.data() : nullptr};
        for (uint64_t r = 0; r < n_reads; r++) {
            ReadCounters rc;
            const uint64_t len = offs[r + 1] - offs[r];
            if (runs_mode) run_read<int64_t, true, true>(v, packed.data(), offs[r], len, r, so, rc, n_strands);
            else run_read<int64_t, true, false>(v, packed.data(), offs[r], len, r, so, rc, n_strands);
            counts4[0] += rc.found[0]; counts4[1] += rc.found[1]; counts4[2] += rc.positive; counts4[3] += rc.n_q;
            const unsigned a[8] = {rc.c_rank, rc.c_lcs, rc.c_ustart, rc.c_anchor, rc.c_branch, rc.c_walk, rc.c_bases, rc.c_kmers};
            for (int i = 0; i < 8; i++) access8[i] += a[i];
            *flags |= rc.err;
        }
        if (runs_mode) {  // the device pipeline's sort by (read, first k-mer) and CSR offsets, then the product's host expander
            if (n_runs > total) throw arg_error("more runs than k-mer positions");
            std::vector<uint64_t> order(n_runs);
            for (uint64_t t = 0; t < n_runs; t++) order[t] = t;
            std::sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) { return keys[a] < keys[b]; });
            std::vector<fin_run> sorted(n_runs);
            std::vector<uint64_t> run_offsets(n_reads + 1, 0);
            for (uint64_t t = 0; t < n_runs; t++) {
                const RunRec& q = recs[order[t]];
                sorted[t] = fin_run{q.i0, q.len_rc, q.unitig, q.pos0};
                run_offsets[(keys[order[t]] >> kRunKeyPosBits) + 1]++;
            }
            for (uint64_t r = 0; r < n_reads; r++) run_offsets[r + 1] += run_offsets[r];
            expand_runs_host(offs, n_reads, v.k, run_offsets.data(), sorted.data(), out, total, 8, 3);
            if (n_runs_out) *n_runs_out = n_runs;
        }
    });
}

// the incremental FASTA/FASTQ(.gz) reader of the CLI and of build-fmin (seqio.hpp), batch by batch with the given limits and
// refill size; returns the number of reads, the bases back to back in `seqs` (capacity cap) and n + 1 offsets
int64_t emu_read_sequences(const char* path, uint64_t chunk_bytes, uint64_t max_reads, uint64_t max_bases, char* seqs, uint64_t cap,
                           uint64_t* offsets, uint64_t offsets_cap, uint64_t* n_batches) {
    int64_t n = -1;
    guarded([&] {
        SeqReader reader(path, (size_t)chunk_bytes);
        uint64_t total = 0, reads = 0;
        *n_batches = 0;
        offsets[0] = 0;
        for (;;) {
            SeqBatch b;
            if (!reader.next(b, max_reads, max_bases)) break;
            (*n_batches)++;
            if (b.size() > max_reads) throw arg_error("a batch holds more reads than asked for");
            for (uint64_t i = 0; i < b.size(); i++) {
                const uint64_t len = b.offsets[i + 1] - b.offsets[i];
                if (total + len > cap || reads + 1 >= offsets_cap) throw arg_error("buffer too small");
                memcpy(seqs + total, b.seqs.data() + b.offsets[i], len);
                total += len;
                offsets[++reads] = total;
            }
        }
        n = (int64_t)reads;
    });
    return n;
}

int emu_lcs(const char* sbwt_path, int with_variant, uint8_t* out, uint64_t cap) {
    return guarded([&] {
        SbwtHost s = load_sbwt_file(sbwt_path, with_variant != 0);
        if ((uint64_t)s.n_nodes > cap) throw arg_error("buffer too small");
        std::vector<uint8_t> l = emu_lcs_bytes(s);
        memcpy(out, l.data(), l.size());
    });
}

// build-fmin --type rarest with the kernel bodies; per_chunk = 0 lets the planner choose
int emu_build(const char* sbwt_path, const char* unitig_path, const char* out_prefix, uint64_t per_chunk) {
    return guarded([&] {
        const std::vector<std::string> files = unitig_file_list(unitig_path);
        SbwtHost sbwt = load_sbwt_file(sbwt_path, true);
        check_sbwt_limits(sbwt);
        const uint64_t N = (uint64_t)sbwt.n_nodes;
        const int k = (int)sbwt.k;
        PackedVec lcs_file = lcs_bytes_to_packed(emu_lcs_bytes(sbwt), N, k);
        SeqBatch unitigs;
        append_sequences(files[0], unitigs);
        check_unitigs(unitigs, k);
        const uint64_t U = unitigs.size();
        std::vector<uint64_t> lines = make_lines(sbwt, &lcs_file, nullptr, nullptr);
        IndexView ix{};
        ix.lines = reinterpret_cast<const uint4*>(lines.data());
        ix.n_nodes = (idx_t)N;
        ix.k = k;
        std::vector<uint64_t> order(U), starts;
        for (uint64_t u = 0; u < U; u++) order[u] = u;
        PackedVec text;
        pack_unitigs(unitigs, order, text, starts);
        text.words.resize(text.words.size() + 2, 0);
        std::vector<uint32_t> cols(U);
        for (uint64_t u = 0; u < U; u++) cols[u] = first_kmer_column(ix, text.words.data(), starts[u]);
        PackedVec ustart;
        order = unitig_order(cols, N, ustart);
        pack_unitigs(unitigs, order, text, starts);
        text.words.resize(text.words.size() + 2, 0);
        PackedVec ends = unitig_ends(unitigs, order, starts);
        std::vector<BuildChunk> chunks = plan_build_chunks(starts, ends, k, per_chunk);
        std::vector<unsigned long long> table(N, 0);
        unsigned err = 0;
        for (const BuildChunk& ch : chunks) {
            if (k <= 32) build_stream_chunk<32>(ix, text.words.data(), ch, table.data(), &err);
            else if (k <= 64) build_stream_chunk<64>(ix, text.words.data(), ch, table.data(), &err);
            else if (k <= 128) build_stream_chunk<128>(ix, text.words.data(), ch, table.data(), &err);
            else build_stream_chunk<256>(ix, text.words.data(), ch, table.data(), &err);
        }
        if (err) throw domain_error("emu: build stream error flags " + std::to_string(err));
        IndexHost out;
        out.fmin.assign(N, 1);
        std::vector<uint32_t> offs;
        for (uint64_t i = 0; i < N; i++)
            if (table[i]) {
                out.fmin.words[i >> 6] |= 1ULL << (i & 63);
                offs.push_back(fmin_key_offset(table[i]));
            }
        out.offsets = pack_offsets(offs.data(), offs.size());
        text.words.resize(text.disk_words() + 1);
        out.sbwt = std::move(sbwt);
        out.lcs = std::move(lcs_file);
        out.text = std::move(text);
        out.ends = std::move(ends);
        out.ustart = std::move(ustart);
        save_index_files(out, out_prefix);
        append_stats_line(out_prefix, 1, "");
    });
}


void* emu_stats_open(const char* sbwt_path, const char* unitig_path, const char* type, int64_t t, int rank, int world) {
    EmuStats* h = nullptr;
    if (guarded([&] { h = new EmuStats(sbwt_path, unitig_path, type, t, 0, rank, world); })) return nullptr;
    return h;
}
void emu_stats_free(void* h) { delete static_cast<EmuStats*>(h); }
int emu_stats_first_pass(void* h, uint64_t* n_table) {
    return guarded([&] { *n_table = static_cast<EmuStats*>(h)->first_pass(); });
}
int emu_stats_table(void* h, uint32_t* table, int set) {
    return guarded([&] {
        EmuStats& e = *static_cast<EmuStats*>(h);
        if (set) memcpy(e.first.data(), table, e.first.size() * 4);
        else memcpy(table, e.first.data(), e.first.size() * 4);
    });
}
int emu_stats_second_pass(void* h, uint64_t* n_keys) {
    return guarded([&] { *n_keys = static_cast<EmuStats*>(h)->second_pass(); });
}
int emu_stats_keys(void* h, uint64_t* out) {
    return guarded([&] {
        const EmuStats& e = *static_cast<EmuStats*>(h);
        if (!e.keys.empty()) memcpy(out, e.keys.data(), e.keys.size() * 8);
    });
}
int emu_stats_finish(void* h, const uint64_t* keys, uint64_t n, char* result, uint64_t cap) {
    return guarded([&] {
        const std::string r = static_cast<EmuStats*>(h)->finish(keys, n);
        if (r.size() + 1 > cap) throw arg_error("buffer too small");
        memcpy(result, r.c_str(), r.size() + 1);
    });
}

int emu_stats(const char* sbwt_path, const char* unitig_path, const char* type_c, int64_t t, uint64_t per_chunk, char* result,
              uint64_t cap) {
    return guarded([&] {
        EmuStats e(sbwt_path, unitig_path, type_c, t, per_chunk, 0, 1);
        e.first_pass();
        e.second_pass();
        const std::string r = e.finish(e.keys.data(), e.keys.size());
        if (r.size() + 1 > cap) throw arg_error("buffer too small");
        memcpy(result, r.c_str(), r.size() + 1);
    });
}

// ---- one rank of the sharded build over HOST memory: same stages as fin_builder_* (include/finito_b200.h), the stream
//      run by the kernel bodies.  Lets the gloo world_size-2 tests drive finito_b200/distributed.py without a GPU.
struct EmuBuilder {
    SbwtHost sbwt;
    PackedVec lcs_file, text, ends, ustart;
    std::vector<uint64_t> lines;
    IndexView ix{};
    std::vector<BuildChunk> chunks;
    size_t c0 = 0, c1 = 0;
    int rank = 0, world = 1, k = 0;
    uint64_t N = 0, S = 0, bases = 0;
    std::vector<unsigned long long> table, slice_table;
    std::vector<uint64_t> tuples;   // {colex, key}
    std::vector<uint64_t> bits;
    std::vector<uint32_t> offs;
    std::string prefix;
};

void* emu_builder_open(const char* sbwt_path, const char* unitig_path, const char* out_prefix, int rank, int world) {
    EmuBuilder* b = nullptr;
    const int rc = guarded([&] {
        std::unique_ptr<EmuBuilder> e(new EmuBuilder());
        const std::vector<std::string> files = unitig_file_list(unitig_path);
        e->sbwt = load_sbwt_file(sbwt_path, true);
        check_sbwt_limits(e->sbwt);
        e->N = (uint64_t)e->sbwt.n_nodes;
        e->k = (int)e->sbwt.k;
        e->rank = rank; e->world = world; e->prefix = out_prefix;
        e->lcs_file = lcs_bytes_to_packed(emu_lcs_bytes(e->sbwt), e->N, e->k);
        if (rank == 0) save_packed_file(e->prefix + ".LCS.sdsl", e->lcs_file, true);
        SeqBatch unitigs;
        append_sequences(files[0], unitigs);
        check_unitigs(unitigs, e->k);
        const uint64_t U = unitigs.size();
        e->lines = make_lines(e->sbwt, &e->lcs_file, nullptr, nullptr);
        e->ix.lines = reinterpret_cast<const uint4*>(e->lines.data());
        e->ix.n_nodes = (idx_t)e->N;
        e->ix.k = e->k;
        std::vector<uint64_t> order(U), starts;
        for (uint64_t u = 0; u < U; u++) order[u] = u;
        pack_unitigs(unitigs, order, e->text, starts);
        e->text.words.resize(e->text.words.size() + 2, 0);
        std::vector<uint32_t> cols(U);
        for (uint64_t u = 0; u < U; u++) cols[u] = first_kmer_column(e->ix, e->text.words.data(), starts[u]);
        order = unitig_order(cols, e->N, e->ustart);
        pack_unitigs(unitigs, order, e->text, starts);
        e->text.words.resize(e->text.words.size() + 2, 0);
        e->ends = unitig_ends(unitigs, order, starts);
        e->chunks = plan_build_chunks(starts, e->ends, e->k, 0);
        const std::pair<size_t, size_t> mine = shard_chunks(e->chunks, rank, world);
        e->c0 = mine.first; e->c1 = mine.second;
        e->S = world == 1 ? e->N : slice_columns(e->N, world);
        for (size_t c = e->c0; c < e->c1; c++) e->bases += e->chunks[c].n_stream;
        b = e.release();
    });
    return rc ? nullptr : b;
}
void emu_builder_free(void* h) { delete static_cast<EmuBuilder*>(h); }
// info9: n_nodes, slice_columns, slice_words, text_len, n_unitigs, n_chunks, chunk_begin, chunk_end, bases_streamed
void emu_builder_info(void* h, int64_t* info9) {
    EmuBuilder& e = *static_cast<EmuBuilder*>(h);
    const int64_t v[9] = {(int64_t)e.N, (int64_t)e.S, (int64_t)((e.S + 63) / 64), (int64_t)e.text.size(), (int64_t)e.ends.size(),
                          (int64_t)e.chunks.size(), (int64_t)e.c0, (int64_t)e.c1, (int64_t)e.bases};
    memcpy(info9, v, sizeof v);
}
int emu_builder_stream(void* h, uint64_t* n_tuples) {
    return guarded([&] {
        EmuBuilder& e = *static_cast<EmuBuilder*>(h);
        e.table.assign(e.N, 0);
        unsigned err = 0;
        for (size_t c = e.c0; c < e.c1; c++) {
            const BuildChunk& ch = e.chunks[c];
            if (e.k <= 32) build_stream_chunk<32>(e.ix, e.text.words.data(), ch, e.table.data(), &err);
            else if (e.k <= 64) build_stream_chunk<64>(e.ix, e.text.words.data(), ch, e.table.data(), &err);
            else if (e.k <= 128) build_stream_chunk<128>(e.ix, e.text.words.data(), ch, e.table.data(), &err);
            else build_stream_chunk<256>(e.ix, e.text.words.data(), ch, e.table.data(), &err);
        }
        if (err) throw domain_error("emu: build stream error flags " + std::to_string(err));
        e.tuples.clear();
        if (e.world > 1)
            for (uint64_t i = 0; i < e.N; i++)
                if (e.table[i]) { e.tuples.push_back(i); e.tuples.push_back(e.table[i]); }
        *n_tuples = e.tuples.size() / 2;
    });
}
int emu_builder_export(void* h, uint64_t* out, uint64_t capacity, uint64_t* counts) {
    return guarded([&] {
        EmuBuilder& e = *static_cast<EmuBuilder*>(h);
        const uint64_t n = e.tuples.size() / 2;
        if (capacity < n) throw arg_error("tuple buffer too small");
        if (n) memcpy(out, e.tuples.data(), n * 16);
        const std::vector<uint64_t> c = tuples_per_owner(e.tuples.data(), n, e.S, e.world);
        for (int g = 0; g < e.world; g++) counts[g] = e.world == 1 ? 0 : c[(size_t)g];
    });
}
int emu_builder_merge(void* h, const uint64_t* tuples, uint64_t n, uint64_t* n_slice) {
    return guarded([&] {
        EmuBuilder& e = *static_cast<EmuBuilder*>(h);
        const uint64_t begin = (uint64_t)e.rank * e.S;
        if (e.world == 1) {
            e.slice_table = e.table;
        } else {
            e.slice_table.assign(e.S, 0);
            for (uint64_t i = 0; i < n; i++) {
                const uint64_t local = tuples[2 * i] - begin;
                if (tuples[2 * i] < begin || local >= e.S) throw arg_error("tuple outside the slice");
                fin_atomic_max(&e.slice_table[local], tuples[2 * i + 1]);
            }
        }
        const uint64_t n_valid = begin >= e.N ? 0 : std::min<uint64_t>(e.S, e.N - begin);
        e.bits.assign((e.S + 63) / 64, 0);
        e.offs.clear();
        for (uint64_t i = 0; i < n_valid; i++)
            if (e.slice_table[i]) {
                e.bits[i >> 6] |= 1ULL << (i & 63);
                e.offs.push_back(fmin_key_offset(e.slice_table[i]));
            }
        *n_slice = e.offs.size();
    });
}
void emu_builder_slice(void* h, uint64_t* bits, uint32_t* offs) {
    EmuBuilder& e = *static_cast<EmuBuilder*>(h);
    memcpy(bits, e.bits.data(), e.bits.size() * 8);
    if (!e.offs.empty()) memcpy(offs, e.offs.data(), e.offs.size() * 4);
}
int emu_builder_finish(void* h, const uint64_t* bits_all, uint64_t n_words, const uint32_t* offs_all, uint64_t n_offs) {
    return guarded([&] {
        EmuBuilder& e = *static_cast<EmuBuilder*>(h);
        e.text.words.resize(e.text.disk_words() + 1);
        save_index_from_slices(std::move(e.sbwt), std::move(e.lcs_file), std::move(e.text), std::move(e.ends), std::move(e.ustart), bits_all,
                               n_words, offs_all, n_offs, e.prefix);
        append_stats_line(e.prefix, 1, "");
    });
}

}  // extern "C"
