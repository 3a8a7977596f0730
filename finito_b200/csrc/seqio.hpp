// FASTA / FASTQ (optionally gzip) input with the behaviour of the reference reader
// (SBWT/include/sbwt/SeqIO.hh:190-292, SBWT/src/SeqIO.cpp:10-52): format from the file extension,
// headers dropped, FASTA sequences may span lines, FASTQ is strictly four lines per record,
// every base upper-cased, '\r' is NOT stripped, an empty sequence is an error.
// Sequences are returned back to back with an offsets array — the batch layout of fin_search_batch.
#pragma once
#include <cctype>
#include <cstdio>
#include <cstring>
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

// Incremental reader: records are handed out in batches of bounded size, the file is read (or inflated) in chunks, so
// memory stays O(batch) however large the input is (the reference streams one read at a time, search_fmin.hh:43-72).
class SeqReader {
   public:
    // chunk_bytes: how much is read (or inflated) per refill; tests use tiny chunks to put every boundary case on a record
    explicit SeqReader(const std::string& path, size_t chunk_bytes = 16u << 20)
        : path_(path), fmt_(sequence_file_format(path)), kChunk(std::max<size_t>(chunk_bytes, 16)) {
        if (has_suffix(path, ".gz")) {  // the reference decides by the extension too (SeqIO.cpp:19-52)
            gz_ = gzopen(path.c_str(), "rb");
            if (!gz_) throw io_error("Error opening file " + path);
            gzbuffer(gz_, 1 << 20);
        } else {
            f_ = fopen(path.c_str(), "rb");
            if (!f_) throw io_error("Error opening file " + path);
        }
        buf_.resize(kChunk);
        fill();
        const char first = fmt_ == SeqFormat::Fasta ? '>' : '@';
        if (end_ == 0 || buf_[0] != first)  // SeqIO.hh read_first_char_and_sanity_check
            throw io_error(std::string("ERROR: ") + (fmt_ == SeqFormat::Fasta ? "FASTA" : "FASTQ") + " file " + path + " does not start with '" +
                           first + "'");
    }
    SeqReader(const SeqReader&) = delete;
    SeqReader& operator=(const SeqReader&) = delete;
    ~SeqReader() {
        if (gz_) gzclose(gz_);
        if (f_) fclose(f_);
    }

    // Appends whole records to `out` until it holds max_reads reads or max_bases bases (always at least one record when
    // one is left).  Returns false when the input is exhausted and nothing was appended.
    bool next(SeqBatch& out, uint64_t max_reads, uint64_t max_bases) {
        if (out.offsets.empty()) out.offsets.push_back(0);
        const uint64_t reads0 = out.size();
        for (;;) {
            if (out.size() - reads0 >= max_reads || (out.size() > reads0 && out.seqs.size() >= max_bases)) break;
            size_t rec_end = 0;
            if (!find_record(rec_end)) {
                if (eof_) {
                    if (pos_ >= end_) break;
                    rec_end = end_;  // the last record has no terminator
                } else {
                    fill();
                    continue;
                }
            }
            append_record(out, rec_end);
            pos_ = rec_end;
        }
        return out.size() > reads0;
    }

   private:

    void fill() {  // keeps [pos_, end_), reads more behind it
        if (pos_ > 0) {
            memmove(buf_.data(), buf_.data() + pos_, end_ - pos_);
            end_ -= pos_;
            scan_ -= std::min(scan_, pos_);
            pos_ = 0;
        }
        if (buf_.size() - end_ < kChunk / 2) buf_.resize(buf_.size() * 2);
        while (!eof_ && end_ < buf_.size()) {
            const size_t want = std::min<size_t>(buf_.size() - end_, 1u << 30);
            long got;
            if (gz_) {
                got = gzread(gz_, buf_.data() + end_, (unsigned)want);
                if (got < 0) throw io_error("Error reading file " + path_);
            } else {
                got = (long)fread(buf_.data() + end_, 1, want, f_);
                if (got == 0 && ferror(f_)) throw io_error("Error reading file " + path_);
            }
            if (got == 0) eof_ = true;
            end_ += (size_t)got;
            if (end_ - pos_ >= kChunk) break;
        }
    }

    // end (exclusive) of the record that starts at pos_, if it is complete in the buffer
    bool find_record(size_t& rec_end) {
        if (pos_ >= end_) return false;
        const char* b = buf_.data();
        if (fmt_ == SeqFormat::Fasta) {  // up to the next '>' at the start of a line
            size_t q = std::max(scan_, pos_ + 1);  // scan_: no record start before it (kept across refills)
            while (q < end_) {
                const char* nl = static_cast<const char*>(memchr(b + q, '\n', end_ - q));
                if (!nl) {
                    scan_ = end_;
                    return false;
                }
                const size_t after = (size_t)(nl - b) + 1;
                if (after >= end_) {  // the first character of the next line is not in the buffer yet: look at this newline again
                    scan_ = (size_t)(nl - b);
                    return false;
                }
                if (b[after] == '>') {
                    rec_end = after;
                    scan_ = after;
                    return true;
                }
                q = after;
            }
            scan_ = std::max(scan_, std::min(q, end_));
            return false;
        }
        size_t q = pos_;  // FASTQ: four lines
        for (int line = 0; line < 4; line++) {
            const char* nl = static_cast<const char*>(memchr(b + q, '\n', end_ - q));
            if (!nl) return false;
            q = (size_t)(nl - b) + 1;
        }
        rec_end = q;
        return true;
    }

    void append_record(SeqBatch& out, size_t rec_end) {
        const char* p = buf_.data() + pos_;
        const char* const end = buf_.data() + rec_end;
        const size_t base0 = out.seqs.size();
        out.seqs.resize(base0 + (size_t)(end - p));
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
        skip_line();  // header
        if (fmt_ == SeqFormat::Fasta) {
            while (p < end) copy_line();
        } else {
            copy_line();  // the '+' line and the qualities are dropped
        }
        const size_t n_new = (size_t)(dst - (out.seqs.data() + base0));
        if (n_new == 0) throw domain_error(std::string("Error: empty sequence in ") + (fmt_ == SeqFormat::Fasta ? "FASTA" : "FASTQ") + " file.");
        out.seqs.resize(base0 + n_new);
        char* q = out.seqs.data() + base0;
        for (size_t i = 0; i < n_new; i++) {  // toupper of the reference's reader (SeqIO.hh:232), branch-free so it vectorises
            const char c = q[i];
            q[i] = (char)(c - ((c >= 'a' && c <= 'z') ? 32 : 0));
        }
        out.offsets.push_back((uint64_t)out.seqs.size());
    }

    std::string path_;
    SeqFormat fmt_;
    size_t kChunk;
    gzFile gz_ = nullptr;
    FILE* f_ = nullptr;
    std::vector<char> buf_;
    size_t pos_ = 0, end_ = 0, scan_ = 0;
    bool eof_ = false;
};

// the whole file (build-fmin reads its unitigs this way)
inline void append_sequences(const std::string& path, SeqBatch& out) {
    SeqReader reader(path);
    while (reader.next(out, ~0ull, ~0ull)) {
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
