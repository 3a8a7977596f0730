// Host-side readers / writers for the on-disk formats of the finimizer path.
// Byte layouts follow what the reference writes through sdsl-lite:
//   int_vector<w>  : u64 size_in_bits ; u8 width only for run-time width (w == 0) ; ceil(bits/64) LE words
//                    (SBWT/sdsl-lite/include/sdsl/int_vector.hpp:624-640 header, :1577-1628 body)
//   .sbwt          : [i64 len,"plain-matrix"] i64 len,"v0.1" ; 4 bit_vectors ; 4 rank_support_v5 ;
//                    suffix_group_starts ; C ; precalc ; precalc_k,n_nodes,n_kmers,k
//                    (SBWT/include/sbwt/SBWT.hh:395-424, SubsetMatrixRank.hh serialize, build_fmin.hh:352-364)
// Nothing here touches the GPU.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace fin {

struct io_error : std::runtime_error { using std::runtime_error::runtime_error; };
struct format_error : std::runtime_error { using std::runtime_error::runtime_error; };
struct domain_error : std::runtime_error { using std::runtime_error::runtime_error; };
struct limit_error : std::runtime_error { using std::runtime_error::runtime_error; };
struct arg_error : std::runtime_error { using std::runtime_error::runtime_error; };

inline int bit_length(uint64_t x) { return x ? 64 - __builtin_clzll(x) : 0; }

// A packed vector of fixed-width unsigned integers, LSB-first across little-endian 64-bit words.
struct PackedVec {
    uint64_t n_bits = 0;
    uint8_t width = 1;
    std::vector<uint64_t> words;  // at least ceil(n_bits/64) + 1 (one pad word so get() may straddle)

    uint64_t size() const { return width ? n_bits / width : 0; }
    uint64_t disk_words() const { return (n_bits + 63) >> 6; }

    void assign(uint64_t n, uint8_t w) {
        width = w;
        n_bits = n * (uint64_t)w;
        words.assign(disk_words() + 1, 0);
    }
    uint64_t get(uint64_t i) const {
        const uint64_t bit = i * width, wi = bit >> 6, off = bit & 63;
        uint64_t v = words[wi] >> off;
        if (off + width > 64) v |= words[wi + 1] << (64 - off);
        return width < 64 ? v & ((1ULL << width) - 1) : v;
    }
    void put(uint64_t i, uint64_t v) {
        const uint64_t bit = i * width, wi = bit >> 6, off = bit & 63;
        const uint64_t mask = width < 64 ? (1ULL << width) - 1 : ~0ULL;
        v &= mask;
        words[wi] = (words[wi] & ~(mask << off)) | (v << off);
        if (off + width > 64) {
            const unsigned hi = 64 - (unsigned)off;
            words[wi + 1] = (words[wi + 1] & ~(mask >> hi)) | (v >> hi);
        }
    }
    bool bit(uint64_t i) const { return (words[i >> 6] >> (i & 63)) & 1; }
};

// ---- byte cursor over a whole file held in memory ---------------------------------------------------
struct Cursor {
    const uint8_t* p;
    const uint8_t* end;
    const std::string* name;
    void need(size_t n) const {
        if ((size_t)(end - p) < n) throw io_error("truncated file: " + *name);
    }
    template <class T> T pod() {
        need(sizeof(T));
        T v;
        memcpy(&v, p, sizeof(T));
        p += sizeof(T);
        return v;
    }
    void bytes(void* dst, size_t n) {
        need(n);
        memcpy(dst, p, n);
        p += n;
    }
    void skip(size_t n) {
        need(n);
        p += n;
    }
};

inline std::vector<uint8_t> slurp(const std::string& path) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw io_error("cannot open " + path);
    std::vector<uint8_t> buf;
    if (fseek(f, 0, SEEK_END) == 0) {
        long n = ftell(f);
        if (n > 0) buf.resize((size_t)n);
        fseek(f, 0, SEEK_SET);
    }
    size_t got = buf.empty() ? 0 : fread(buf.data(), 1, buf.size(), f);
    fclose(f);
    if (got != buf.size()) throw io_error("short read on " + path);
    return buf;
}

// fixed_width == 0 means the width is a run-time property stored in the file (sdsl int_vector<0>)
inline PackedVec read_packed(Cursor& c, uint8_t fixed_width) {
    PackedVec v;
    v.n_bits = c.pod<uint64_t>();
    v.width = fixed_width ? fixed_width : c.pod<uint8_t>();
    if (v.width == 0 || v.width > 64) throw format_error("bad int_vector width in " + *c.name);
    const uint64_t nw = v.disk_words();
    c.need(nw * 8);
    v.words.assign(nw + 1, 0);
    if (nw) c.bytes(v.words.data(), nw * 8);
    return v;
}

inline void skip_packed(Cursor& c, uint8_t fixed_width) {
    const uint64_t n_bits = c.pod<uint64_t>();
    if (!fixed_width) c.skip(1);
    c.skip(((n_bits + 63) >> 6) * 8);
}

inline PackedVec load_packed_file(const std::string& path, uint8_t fixed_width) {
    std::vector<uint8_t> buf = slurp(path);
    Cursor c{buf.data(), buf.data() + buf.size(), &path};
    return read_packed(c, fixed_width);
}

inline void write_all(FILE* f, const void* p, size_t n, const std::string& path) {
    if (n && fwrite(p, 1, n, f) != n) throw io_error("short write on " + path);
}

inline void save_packed_file(const std::string& path, const PackedVec& v, bool runtime_width) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw io_error("cannot write " + path);
    try {
        write_all(f, &v.n_bits, 8, path);
        if (runtime_width) write_all(f, &v.width, 1, path);
        write_all(f, v.words.data(), v.disk_words() * 8, path);
    } catch (...) {
        fclose(f);
        throw;
    }
    if (fclose(f) != 0) throw io_error("cannot close " + path);
}

// ---- plain-matrix SBWT as stored on disk -----------------------------------------------------------------
struct SbwtHost {
    int64_t n_nodes = 0, n_kmers = 0, k = 0, precalc_k = 0;
    int64_t C[4] = {0, 0, 0, 0};
    PackedVec bits[4];                 // A, C, G, T rows of the subset matrix
    std::vector<int64_t> precalc;      // 2 * 4^precalc_k entries: inclusive interval per prefix, {-1,-1} when absent
    std::vector<uint8_t> file;         // the whole file as read; the serialized SBWT without the variant string is
    size_t body_off = 0, body_len = 0;  //   file[body_off, body_off + body_len), written verbatim as <prefix>.sbwt
    const uint8_t* body() const { return file.data() + body_off; }
};

inline std::string read_len_string(Cursor& c) {  // SBWT/src/globals.cpp:56-62 load_string
    const int64_t n = c.pod<int64_t>();
    if (n < 0 || n > 4096) throw format_error("bad string length in " + *c.name);
    std::string s((size_t)n, '\0');
    if (n) c.bytes(&s[0], (size_t)n);
    return s;
}

inline SbwtHost load_sbwt_file(const std::string& path, bool with_variant) {
    std::vector<uint8_t> buf = slurp(path);
    Cursor c{buf.data(), buf.data() + buf.size(), &path};
    SbwtHost s;
    if (with_variant) {
        if (read_len_string(c) != "plain-matrix")  // build_fmin.hh:355-358
            throw format_error("Error loading index from file: unrecognized variant specified in the file");
    }
    const uint8_t* body_begin = c.p;
    if (read_len_string(c) != "v0.1")  // SBWT.hh:429-432
        throw format_error("Error: Corrupt index file, or the index was constructed with an incompatible version of SBWT.");
    for (int ch = 0; ch < 4; ch++) s.bits[ch] = read_packed(c, 1);
    for (int ch = 0; ch < 4; ch++) skip_packed(c, 64);  // rank_support_v5 directories: the device layout carries its own counts
    skip_packed(c, 1);                                   // suffix_group_starts: not read by this path
    if (c.pod<int64_t>() != 32) throw format_error("unexpected C array size in " + path);
    c.bytes(s.C, 32);
    const int64_t pre_bytes = c.pod<int64_t>();
    if (pre_bytes < 0 || pre_bytes % 16) throw format_error("bad precalc table in " + path);
    s.precalc.resize((size_t)pre_bytes / 8);
    if (pre_bytes) c.bytes(s.precalc.data(), (size_t)pre_bytes);
    s.precalc_k = c.pod<int64_t>();
    s.n_nodes = c.pod<int64_t>();
    s.n_kmers = c.pod<int64_t>();
    s.k = c.pod<int64_t>();
    for (int ch = 0; ch < 4; ch++)
        if ((int64_t)s.bits[ch].n_bits != s.n_nodes) throw format_error("bit vector length differs from n_nodes in " + path);
    if (s.k < 1 || s.k > 255) throw format_error("k out of range in " + path);
    s.body_off = (size_t)(body_begin - buf.data());
    s.body_len = (size_t)(c.p - body_begin);
    s.file = std::move(buf);  // no second copy of the (up to 2 GB) body
    return s;
}

inline bool has_suffix(const std::string& s, const std::string& suf) {
    return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
}

}  // namespace fin
