// fin_datagen — synthetic inputs for tests and bench.py (not part of the product path, not the oracle).
//
//   fin_datagen genome  <out.fna> <length> <seed> [families copies replen divergence]
//   fin_datagen bigenome <out.fna> <length> <seed> [chromosomes=24 repeat_fraction=0.45 families=2000 min_div=0.05 max_div=0.20]
//                                                              large genomes: iid bases generated
//                                                              in parallel, repeat families of 300 bp - 6 kb with Zipf copy numbers
//                                                              and per-copy divergence, cut into `chromosomes` records
//   fin_datagen unitigs <genome.fna> <k> <out.fna> <seed>      maximal non-branching paths of the node-centric
//                                                              de Bruijn graph of the forward k-mers, each unitig
//                                                              reverse-complemented with probability 1/2
//   fin_datagen sbwt    <unitigs.fna> <k> <out.sbwt> [precalc_k=8]
//                                                              plain-matrix SBWT file in the format `sbwt build` writes
//                                                              (SBWT/include/sbwt/SBWT.hh:395-424; node set and order as in
//                                                              NodeBOSSInMemoryConstructor.hh:98-165; rank_support_v5
//                                                              blobs as sdsl/rank_support_v5.hpp:65-109)
//   fin_datagen reads   <genome.fna> <n> <len> <error> <seed> <out.bin|out.fna>
//                                                              uniform start, strand 1/2, iid substitutions; .bin = n*len
//                                                              ASCII bases back to back (no headers)
// k <= 64 (2 bits per base in one 128-bit word).
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <parallel/algorithm>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

typedef unsigned __int128 u128;

static int code_of(char c) {
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return -1;
    }
}

static std::vector<std::string> read_fasta(const std::string& path) {
    std::ifstream in(path);
    if (!in) throw std::runtime_error("cannot open " + path);
    std::vector<std::string> seqs;
    std::string line;
    while (std::getline(in, line)) {
        if (!line.empty() && line[0] == '>') seqs.emplace_back();
        else if (!seqs.empty()) seqs.back() += line;
    }
    return seqs;
}

static std::string revcomp(const std::string& s) {
    std::string t(s.rbegin(), s.rend());
    for (char& c : t) c = "TGCA"[code_of(c)];
    return t;
}

static u128 low_mask(int bits) { return bits >= 128 ? ~(u128)0 : (((u128)1 << bits) - 1); }

// colex key: char i of the k-mer at bits [2i, 2i+2): the last char is the most significant
static void collect_kmers(const std::vector<std::string>& seqs, int k, std::vector<u128>& out) {
    for (const std::string& s : seqs) {
        if ((int)s.size() < k) continue;
        u128 key = 0;
        int valid = 0;
        for (size_t i = 0; i < s.size(); i++) {
            const int c = code_of(s[i]);
            if (c < 0) { valid = 0; key = 0; continue; }
            key = (key >> 2) | ((u128)c << (2 * (k - 1)));
            if (++valid >= k) out.push_back(key);
        }
    }
}

static void sort_unique(std::vector<u128>& v) {
    __gnu_parallel::sort(v.begin(), v.end());
    v.erase(std::unique(v.begin(), v.end()), v.end());
}

// ------------------------------------------------------------------------------------------------ genome
static int cmd_genome(int argc, char** argv) {
    if (argc < 5) return 2;
    const std::string out = argv[2];
    const uint64_t len = std::stoull(argv[3]);
    const uint64_t seed = std::stoull(argv[4]);
    const int families = argc > 5 ? atoi(argv[5]) : 50, copies = argc > 6 ? atoi(argv[6]) : 5;
    const uint64_t replen = argc > 7 ? std::stoull(argv[7]) : 2000;
    const double div = argc > 8 ? atof(argv[8]) : 0.01;
    std::mt19937_64 rng(seed);
    std::string g(len, 'A');
    for (uint64_t i = 0; i < len; i++) g[i] = "ACGT"[rng() & 3];
    if (replen < len) {
        for (int f = 0; f < families; f++) {
            std::string rep(replen, 'A');
            for (char& c : rep) c = "ACGT"[rng() & 3];
            for (int c = 0; c < copies; c++) {
                std::string r = rep;
                for (char& ch : r)
                    if ((double)(rng() >> 11) / 9007199254740992.0 < div) ch = "ACGT"[(code_of(ch) + 1 + rng() % 3) & 3];
                const uint64_t pos = rng() % (len - replen);
                memcpy(&g[pos], r.data(), replen);
            }
        }
    }
    std::ofstream o(out);
    o << ">genome\n" << g << "\n";
    return 0;
}

// ------------------------------------------------------------------------------------------------ bigenome
static uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

static int cmd_bigenome(int argc, char** argv) {
    if (argc < 5) return 2;
    const std::string out = argv[2];
    const uint64_t len = std::stoull(argv[3]);
    const uint64_t seed = std::stoull(argv[4]);
    const int chromosomes = argc > 5 ? atoi(argv[5]) : 24;
    const double rep_frac = argc > 6 ? atof(argv[6]) : 0.45;
    const int families = argc > 7 ? atoi(argv[7]) : 2000;
    const double min_div = argc > 8 ? atof(argv[8]) : 0.05, max_div = argc > 9 ? atof(argv[9]) : 0.20;
    std::string g(len, 'A');
    const uint64_t block = 1 << 20;
#pragma omp parallel for schedule(dynamic, 4)
    for (uint64_t b = 0; b < (len + block - 1) / block; b++) {
        uint64_t x = mix64(seed ^ (b * 0xD1342543DE82EF95ULL));
        const uint64_t end = std::min(len, (b + 1) * block);
        for (uint64_t i = b * block; i < end; i += 32) {
            x = mix64(x);
            uint64_t v = x;
            for (uint64_t j = i; j < std::min(end, i + 32); j++, v >>= 2) g[j] = "ACGT"[v & 3];
        }
    }
    // repeat families: copy number ~ 1 / rank (Zipf), total repeat bases = rep_frac * len; planted one after the other
    // (later copies may overwrite earlier ones, as transposons do)
    if (rep_frac > 0 && families > 0) {
        std::mt19937_64 rng(seed + 17);
        double harmonic = 0;
        for (int f = 1; f <= families; f++) harmonic += 1.0 / f;
        uint64_t planted = 0, n_copies = 0;
        for (int f = 1; f <= families; f++) {
            const uint64_t L = 300 + rng() % 5701;
            if (L * 4 > len) continue;
            std::string rep(L, 'A');
            for (char& c : rep) c = "ACGT"[rng() & 3];
            const double share = rep_frac * (double)len / harmonic / f;  // bases of this family
            const uint64_t copies = std::max<uint64_t>(2, (uint64_t)(share / (double)L));
            for (uint64_t c = 0; c < copies; c++) {
                const double div = min_div + (max_div - min_div) * ((double)(rng() >> 11) / 9007199254740992.0);
                const uint64_t thr = (uint64_t)(div * 4294967296.0);
                const uint64_t pos = rng() % (len - L);
                for (uint64_t i = 0; i < L; i++) {
                    char ch = rep[i];
                    const uint64_t r = rng();
                    if ((r & 0xFFFFFFFFull) < thr) ch = "ACGT"[(code_of(ch) + 1 + (int)((r >> 32) % 3)) & 3];
                    g[pos + i] = ch;
                }
                planted += L;
                n_copies++;
            }
        }
        std::cerr << "bigenome: " << n_copies << " repeat copies, " << planted << " bases planted\n";
    }
    FILE* o = fopen(out.c_str(), "wb");
    if (!o) throw std::runtime_error("cannot write " + out);
    const uint64_t per = (len + chromosomes - 1) / chromosomes;
    for (int c = 0; c < chromosomes; c++) {
        const uint64_t a = (uint64_t)c * per, b = std::min(len, a + per);
        if (a >= b) break;
        fprintf(o, ">chr%d\n", c + 1);
        fwrite(g.data() + a, 1, b - a, o);
        fputc('\n', o);
    }
    fclose(o);
    return 0;
}

// ------------------------------------------------------------------------------------------------ unitigs
static int64_t find_key(const std::vector<u128>& v, u128 x) {
    auto it = std::lower_bound(v.begin(), v.end(), x);
    return (it != v.end() && *it == x) ? (int64_t)(it - v.begin()) : -1;
}

static int cmd_unitigs(int argc, char** argv) {
    if (argc < 6) return 2;
    const std::vector<std::string> genome = read_fasta(argv[2]);
    const int k = atoi(argv[3]);
    if (k < 2 || k > 64) throw std::runtime_error("k must be in [2,64]");
    const std::string out = argv[4];
    std::mt19937_64 rng(std::stoull(argv[5]));
    std::vector<u128> km;
    collect_kmers(genome, k, km);
    sort_unique(km);
    const int64_t n = (int64_t)km.size();
    const u128 pre_mask = low_mask(2 * (k - 1));
    // successor / predecessor lists as counts and (when unique) the neighbour's index
    std::vector<int64_t> succ(n, -1), pred(n, -1);
    std::vector<uint8_t> outdeg(n, 0), indeg(n, 0);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        for (int c = 0; c < 4; c++) {
            const int64_t j = find_key(km, (km[i] >> 2) | ((u128)c << (2 * (k - 1))));
            if (j >= 0) { outdeg[i]++; succ[i] = j; }
            const int64_t p = find_key(km, ((km[i] & pre_mask) << 2) | (u128)c);
            if (p >= 0) { indeg[i]++; pred[i] = p; }
        }
    }
    auto key_to_string = [&](u128 key) {
        std::string s(k, 'A');
        for (int i = 0; i < k; i++) s[i] = "ACGT"[(int)((key >> (2 * i)) & 3)];
        return s;
    };
    std::vector<uint8_t> used(n, 0);
    std::vector<std::string> unitigs;
    auto walk_from = [&](int64_t i) {
        std::string u = key_to_string(km[i]);
        used[i] = 1;
        int64_t cur = i;
        while (outdeg[cur] == 1) {
            const int64_t nx = succ[cur];
            if (indeg[nx] != 1 || used[nx]) break;
            u.push_back("ACGT"[(int)((km[nx] >> (2 * (k - 1))) & 3)]);
            used[nx] = 1;
            cur = nx;
        }
        unitigs.push_back(u);
    };
    for (int64_t i = 0; i < n; i++) {  // unitig starts: not the unique continuation of a unique predecessor
        const bool is_start = !(indeg[i] == 1 && outdeg[pred[i]] == 1);
        if (is_start && !used[i]) walk_from(i);
    }
    for (int64_t i = 0; i < n; i++)  // isolated cycles
        if (!used[i]) walk_from(i);
    std::ofstream o(out);
    uint64_t total = 0;
    for (size_t i = 0; i < unitigs.size(); i++) {
        const std::string& u = (rng() & 1) ? unitigs[i] : revcomp(unitigs[i]);
        o << ">" << i << "\n" << u << "\n";
        total += u.size();
    }
    std::cerr << "unitigs " << unitigs.size() << " kmers " << n << " bases " << total << "\n";
    return 0;
}

// ------------------------------------------------------------------------------------------------ sbwt
struct Node {
    u128 key;  // aligned: last char in the top 2 bits of the 2k-bit field, missing left chars read as 0
    uint8_t len, flags;
};
static bool node_less(const Node& a, const Node& b) { return a.key != b.key ? a.key < b.key : a.len < b.len; }

template <class T> static void put(std::ostream& o, const T& v) { o.write((const char*)&v, sizeof(T)); }

static void write_bitvector(std::ostream& o, const std::vector<uint64_t>& words, uint64_t n_bits) {
    put<uint64_t>(o, n_bits);
    o.write((const char*)words.data(), (std::streamsize)(((n_bits + 63) / 64) * 8));
}

// sdsl::rank_support_v5<1> over a bit vector (rank_support_v5.hpp:65-109), serialized as int_vector<64>
static void write_rank_v5(std::ostream& o, const std::vector<uint64_t>& words, uint64_t n_bits) {
    const uint64_t capacity = ((n_bits + 63) >> 6) << 6;
    std::vector<uint64_t> bb;
    if (n_bits == 0) {
        bb.assign(2, 0);
    } else {
        bb.assign(((capacity >> 11) + 1) << 1, 0);
        uint64_t j = 0, sum = (uint64_t)__builtin_popcountll(words[0]), second = 0, cnt_words = 1;
        const uint64_t n_words = capacity >> 6;
        for (uint64_t i = 1; i < n_words; ++i, ++cnt_words) {
            if (cnt_words == 32) {
                j += 2;
                bb[j - 1] = second;
                bb[j] = bb[j - 2] + sum;
                second = sum = cnt_words = 0;
            } else if ((cnt_words % 6) == 0) {
                second |= sum << (60 - 12 * (cnt_words / 6));
            }
            sum += (uint64_t)__builtin_popcountll(words[i]);
        }
        if ((cnt_words % 6) == 0) second |= sum << (60 - 12 * (cnt_words / 6));
        if (cnt_words == 32) {
            j += 2;
            bb[j - 1] = second;
            bb[j] = bb[j - 2] + sum;
            bb[j + 1] = 0;
        } else {
            bb[j + 1] = second;
        }
    }
    put<uint64_t>(o, (uint64_t)bb.size() * 64);
    o.write((const char*)bb.data(), (std::streamsize)(bb.size() * 8));
}

static int cmd_sbwt(int argc, char** argv) {
    if (argc < 5) return 2;
    const std::vector<std::string> seqs = read_fasta(argv[2]);
    const int k = atoi(argv[3]);
    if (k < 2 || k > 64) throw std::runtime_error("k must be in [2,64]");
    const std::string out = argv[4];
    int64_t precalc_k = argc > 5 ? atoll(argv[5]) : 8;
    if (precalc_k > k) precalc_k = k;

    std::vector<u128> km;
    collect_kmers(seqs, k, km);
    const size_t with_dups = km.size();
    sort_unique(km);
    if (km.size() != with_dups)
        std::cerr << "warning: " << (with_dups - km.size()) << " repeated k-mers (not a disjoint SPSS)\n";
    const int64_t nk = (int64_t)km.size();
    const u128 pre_mask = low_mask(2 * (k - 1));

    // (k-1)-suffix groups: consecutive k-mers with equal key >> 2; group start carries the outgoing edges
    std::vector<int64_t> group_start;  // index of the first k-mer of every group
    std::vector<u128> group_key;
    for (int64_t i = 0; i < nk; i++)
        if (i == 0 || (km[i] >> 2) != (km[i - 1] >> 2)) {
            group_start.push_back(i);
            group_key.push_back(km[i] >> 2);
        }
    std::vector<uint8_t> kflags(nk, 0);
    std::vector<uint8_t> has_pred(nk, 0);
    std::vector<int64_t> pred_group(nk, -1);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nk; i++) {  // the k-mer y = km[i] has an incoming edge iff its (k-1)-prefix is some group's suffix
        const u128 p = km[i] & pre_mask;
        auto it = std::lower_bound(group_key.begin(), group_key.end(), p);
        if (it != group_key.end() && *it == p) {
            has_pred[i] = 1;
            pred_group[i] = (int64_t)(it - group_key.begin());
        }
    }
    for (int64_t i = 0; i < nk; i++)
        if (has_pred[i]) kflags[group_start[pred_group[i]]] |= (uint8_t)(1u << (int)((km[i] >> (2 * (k - 1))) & 3));

    // dummy nodes: every proper prefix of a k-mer without an incoming edge; the root always exists
    std::vector<Node> dummies;
    dummies.push_back(Node{0, 0, 0});
    for (int64_t i = 0; i < nk; i++) {
        if (has_pred[i]) continue;
        for (int L = 0; L < k; L++) {
            Node nd;
            nd.len = (uint8_t)L;
            nd.key = L ? ((km[i] & low_mask(2 * L)) << (2 * (k - L))) : (u128)0;
            nd.flags = (uint8_t)(1u << (int)((km[i] >> (2 * L)) & 3));
            dummies.push_back(nd);
        }
    }
    std::sort(dummies.begin(), dummies.end(), node_less);
    {
        size_t j = 0;
        for (size_t i = 0; i < dummies.size(); i++) {
            if (j && dummies[j - 1].key == dummies[i].key && dummies[j - 1].len == dummies[i].len) dummies[j - 1].flags |= dummies[i].flags;
            else dummies[j++] = dummies[i];
        }
        dummies.resize(j);
    }
    // merge the two sorted node lists
    const int64_t n_nodes = nk + (int64_t)dummies.size();
    const uint64_t n_words = ((uint64_t)n_nodes + 63) / 64;
    std::vector<uint64_t> rows[4], ssup(n_words + 1, 0);
    for (auto& r : rows) r.assign(n_words + 1, 0);
    {
        int64_t a = 0;
        size_t b = 0;
        Node prev{0, 0, 0};
        for (int64_t i = 0; i < n_nodes; i++) {
            Node cur;
            const bool take_dummy = b < dummies.size() && (a >= nk || node_less(dummies[b], Node{km[a], (uint8_t)k, 0}));
            if (take_dummy) cur = dummies[b++];
            else { cur = Node{km[a], (uint8_t)k, kflags[a]}; a++; }
            for (int c = 0; c < 4; c++)
                if (cur.flags & (1u << c)) rows[c][(uint64_t)i >> 6] |= 1ULL << (i & 63);
            // suffix group starts (NodeBOSSInMemoryConstructor.hh:166-178): compare after dropping the left char of full k-mers
            Node norm = cur;
            if (norm.len == k) { norm.key &= ~(u128)3; norm.len = (uint8_t)(k - 1); }
            if (i == 0 || norm.key != prev.key || norm.len != prev.len) ssup[(uint64_t)i >> 6] |= 1ULL << (i & 63);
            prev = norm;
        }
    }
    // C array (SBWT.hh:277-282) and per-word cumulative counts for the precalc table
    int64_t Carr[4];
    std::vector<uint32_t> cum[4];
    {
        int64_t run = 1;
        for (int c = 0; c < 4; c++) {
            Carr[c] = run;
            cum[c].assign(n_words + 1, 0);
            uint32_t s = 0;
            for (uint64_t w = 0; w < n_words; w++) { cum[c][w] = s; s += (uint32_t)__builtin_popcountll(rows[c][w]); }
            cum[c][n_words] = s;
            run += s;
        }
    }
    auto rank = [&](int c, int64_t pos) -> int64_t {
        const uint64_t w = (uint64_t)pos >> 6, off = (uint64_t)pos & 63;
        return cum[c][w] + (off ? __builtin_popcountll(rows[c][w] & ((1ULL << off) - 1)) : 0);
    };
    std::vector<int64_t> precalc;
    if (precalc_k > 0) {  // SBWT.hh:529-555: index = chars LSB first
        const uint64_t np = 1ULL << (2 * precalc_k);
        precalc.assign(2 * np, -1);
        for (uint64_t d = 0; d < np; d++) {
            int64_t l = 0, r = n_nodes - 1;
            for (int64_t i = 0; i < precalc_k && l != -1; i++) {
                const int c = (int)((d >> (2 * i)) & 3);
                const int64_t nl = Carr[c] + rank(c, l), nr = Carr[c] + rank(c, r + 1) - 1;
                if (nl > nr) { l = r = -1; } else { l = nl; r = nr; }
            }
            precalc[2 * d] = l;
            precalc[2 * d + 1] = r;
        }
    }
    std::ofstream o(out, std::ios::binary);
    if (!o) throw std::runtime_error("cannot write " + out);
    put<int64_t>(o, 12); o.write("plain-matrix", 12);
    put<int64_t>(o, 4);  o.write("v0.1", 4);
    for (int c = 0; c < 4; c++) write_bitvector(o, rows[c], (uint64_t)n_nodes);
    for (int c = 0; c < 4; c++) write_rank_v5(o, rows[c], (uint64_t)n_nodes);
    write_bitvector(o, ssup, (uint64_t)n_nodes);
    put<int64_t>(o, 32);
    o.write((const char*)Carr, 32);
    put<int64_t>(o, (int64_t)precalc.size() * 8);
    o.write((const char*)precalc.data(), (std::streamsize)(precalc.size() * 8));
    put<int64_t>(o, precalc_k);
    put<int64_t>(o, n_nodes);
    put<int64_t>(o, nk);
    put<int64_t>(o, (int64_t)k);
    std::cerr << "n_nodes " << n_nodes << " n_kmers " << nk << "\n";
    return 0;
}

// ------------------------------------------------------------------------------------------------ reads
static int cmd_reads(int argc, char** argv) {
    if (argc < 8) return 2;
    const std::vector<std::string> genome = read_fasta(argv[2]);
    const uint64_t n = std::stoull(argv[3]), len = std::stoull(argv[4]);
    const double err = atof(argv[5]);
    const uint64_t seed = std::stoull(argv[6]);
    const std::string out = argv[7];
    const std::string& g = genome.at(0);
    if (g.size() < len) throw std::runtime_error("genome shorter than the read length");
    const bool fasta = out.size() > 4 && out.substr(out.size() - 4) != ".bin";
    std::vector<char> buf(n * len);
    const uint64_t thr = (uint64_t)(err * 4294967296.0);
#pragma omp parallel
    {
#pragma omp for schedule(static)
        for (uint64_t blk = 0; blk < (n + 4095) / 4096; blk++) {
            std::mt19937_64 rng(seed * 0x9E3779B97F4A7C15ULL + blk);
            for (uint64_t r = blk * 4096; r < std::min(n, (blk + 1) * 4096); r++) {
                const uint64_t pos = rng() % (g.size() - len + 1);
                const bool rc = rng() & 1;
                char* dst = &buf[r * len];
                for (uint64_t i = 0; i < len; i++) {
                    int c = code_of(rc ? g[pos + len - 1 - i] : g[pos + i]);
                    if (rc) c = 3 - c;
                    if ((rng() & 0xFFFFFFFFull) < thr) c = (c + 1 + (int)(rng() % 3)) & 3;
                    dst[i] = "ACGT"[c];
                }
            }
        }
    }
    std::ofstream o(out, std::ios::binary);
    if (!o) throw std::runtime_error("cannot write " + out);
    if (!fasta) {
        o.write(buf.data(), (std::streamsize)buf.size());
    } else {
        std::string rec;
        for (uint64_t r = 0; r < n; r++) {
            rec = ">" + std::to_string(r) + "\n";
            rec.append(&buf[r * len], len);
            rec.push_back('\n');
            o.write(rec.data(), (std::streamsize)rec.size());
        }
    }
    return 0;
}

int main(int argc, char** argv) {
    try {
        const std::string cmd = argc > 1 ? argv[1] : "";
        int rc = 2;
        if (cmd == "genome") rc = cmd_genome(argc, argv);
        else if (cmd == "bigenome") rc = cmd_bigenome(argc, argv);
        else if (cmd == "unitigs") rc = cmd_unitigs(argc, argv);
        else if (cmd == "sbwt") rc = cmd_sbwt(argc, argv);
        else if (cmd == "reads") rc = cmd_reads(argc, argv);
        if (rc == 2) std::cerr << "usage: fin_datagen genome|bigenome|unitigs|sbwt|reads ... (see the header of tools/fin_datagen.cpp)\n";
        return rc;
    } catch (const std::exception& e) {
        std::cerr << "fin_datagen: " << e.what() << "\n";
        return 1;
    }
}
