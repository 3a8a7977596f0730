// FASTA / FASTQ (optionally gzip) input with the behaviour of the reference reader
// (SBWT/include/sbwt/SeqIO.hh:190-292, SBWT/src/SeqIO.cpp:10-52): format from the file extension,
// headers dropped, FASTA sequences may span lines, FASTQ is strictly four lines per record,
// every base upper-cased, '\r' is NOT stripped, an empty sequence is an error.
// Sequences are returned back to back with an offsets array — the batch layout of fin_search_batch.
#pragma once
#include <cctype>
#include <string>
#include <vector>
#include <zlib.h>

#include "formats.hpp"

namespace fin {

struct SeqBatch {
    std::vector<char> seqs;          // concatenated, upper case
    std::vector<uint64_t> offsets;   // n + 1 entries
    uint64_t size() const { return offsets.empty() ? 0 : offsets.size() - 1; }
};

enum class SeqFormat { Fasta, Fastq };

inline SeqFormat sequence_file_format(std::string name) {  // SBWT/src/SeqIO.cpp:19-52
    const std::string full = name;
    if (has_suffix(name, ".gz")) name.resize(name.size() - 3);
    const size_t dot = name.rfind('.');
    if (dot != std::string::npos) {
        const std::string ext = name.substr(dot);
        for (const char* s : {".fasta", ".fna", ".ffn", ".faa", ".frn", ".fa"})
            if (ext == s) return SeqFormat::Fasta;
        for (const char* s : {".fastq", ".fq"})
            if (ext == s) return SeqFormat::Fastq;
    }
    throw io_error("Unknown file format: " + full);
}

inline std::vector<char> read_maybe_gz(const std::string& path) {
    if (!has_suffix(path, ".gz")) {  // the reference decides by the extension too (SeqIO.cpp:19-52): plain file, one read
        FILE* pf = fopen(path.c_str(), "rb");
        if (!pf) throw io_error("Error opening file " + path);
        std::vector<char> whole;
        if (fseek(pf, 0, SEEK_END) == 0) {
            const long n = ftell(pf);
            if (n > 0) whole.resize((size_t)n);
            fseek(pf, 0, SEEK_SET);
        }
        const size_t got = whole.empty() ? 0 : fread(whole.data(), 1, whole.size(), pf);
        fclose(pf);
        if (got != whole.size()) throw io_error("Error reading file " + path);
        return whole;
    }
    gzFile f = gzopen(path.c_str(), "rb");
    if (!f) throw io_error("Error opening file " + path);
    gzbuffer(f, 1 << 20);
    std::vector<char> buf;
    size_t used = 0;
    for (;;) {
        if (buf.size() - used < (1u << 22)) buf.resize(buf.size() * 2 + (1u << 23));
        const int got = gzread(f, buf.data() + used, (unsigned)std::min<size_t>(buf.size() - used, 1u << 30));
        if (got < 0) {
            gzclose(f);
            throw io_error("Error reading file " + path);
        }
        if (got == 0) break;
        used += (size_t)got;
    }
    gzclose(f);
    buf.resize(used);
    return buf;
}

inline void append_sequences(const std::string& path, SeqBatch& out) {
    const SeqFormat fmt = sequence_file_format(path);
    const std::vector<char> buf = read_maybe_gz(path);
    const char* p = buf.data();
    const char* const end = p + buf.size();
    if (out.offsets.empty()) out.offsets.push_back(0);
    if (p == end || *p != (fmt == SeqFormat::Fasta ? '>' : '@'))  // SeqIO.hh read_first_char_and_sanity_check
        throw io_error(std::string("ERROR: ") + (fmt == SeqFormat::Fasta ? "FASTA" : "FASTQ") + " file " + path +
                       " does not start with '" + (fmt == SeqFormat::Fasta ? ">" : "@") + "'");
    // worst case every byte is a base; whole lines are copied with memcpy and upper-cased afterwards in one pass
    const size_t base0 = out.seqs.size();
    out.seqs.resize(base0 + buf.size());
    char* dst = out.seqs.data() + base0;
    auto skip_line = [&]() {
        const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
        p = nl ? nl + 1 : end;
    };
    auto copy_line = [&]() {  // copies up to (not including) the newline, consumes the newline
        const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
        const char* stop = nl ? nl : end;
        memcpy(dst, p, (size_t)(stop - p));
        dst += stop - p;
        p = nl ? nl + 1 : end;
    };
    if (fmt == SeqFormat::Fasta) {
        while (p < end) {
            skip_line();  // header
            const char* before = dst;
            while (p < end && *p != '>') copy_line();
            if (dst == before) throw domain_error("Error: empty sequence in FASTA file.");
            out.offsets.push_back((uint64_t)(dst - out.seqs.data()));
        }
    } else {
        while (p < end) {
            skip_line();  // header
            const char* before = dst;
            copy_line();
            skip_line();  // '+'
            skip_line();  // qualities
            if (dst == before) throw domain_error("Error: empty sequence in FASTQ file.");
            out.offsets.push_back((uint64_t)(dst - out.seqs.data()));
        }
    }
    const size_t n_new = (size_t)(dst - (out.seqs.data() + base0));
    out.seqs.resize(base0 + n_new);
    char* q = out.seqs.data() + base0;
    for (size_t i = 0; i < n_new; i++) {  // toupper of the reference's reader (SeqIO.hh:232), branch-free so it vectorises
        const char c = q[i];
        q[i] = (char)(c - ((c >= 'a' && c <= 'z') ? 32 : 0));
    }
}

inline std::vector<std::string> read_lines(const std::string& path) {  // SBWT/src/globals.cpp:23-31
    const std::vector<uint8_t> buf = slurp(path);
    std::vector<std::string> lines;
    std::string cur;
    for (uint8_t c : buf) {
        if (c == '\n') {
            lines.push_back(cur);
            cur.clear();
        } else {
            cur.push_back((char)c);
        }
    }
    if (!cur.empty()) lines.push_back(cur);
    return lines;
}

}  // namespace fin
