// finito_b200 — drop-in for the reference's `benchmark build-fmin` / `benchmark search-fmin` commands
// (src/main.cpp:21-59, include/build_fmin.hh:302-401, include/search_fmin.hh:130-212): same flags, same files,
// same text output.  Host code only: parsing, batching and text formatting; all computation goes through the
// C ABI of libfinito_b200.so (include/finito_b200.h) to the GPU.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include <unistd.h>


#include "../../include/finito_b200.h"
#include "seqio.hpp"

namespace {

int64_t now_micros() {
    return std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
const int64_t g_start = now_micros();

void log_line(const std::string& msg) {  // SBWT/src/globals.cpp:93-105 write_log, stderr only
    std::fprintf(stderr, "%.4f %s\n", (double)(now_micros() - g_start) / 1e6, msg.c_str());
}

struct Options {
    std::map<std::string, std::string> values;
    bool help = false;
    bool has(const std::string& k) const { return values.count(k) != 0; }
    std::string get(const std::string& k, const std::string& dflt = "") const {
        auto it = values.find(k);
        return it == values.end() ? dflt : it->second;
    }
};

// long name -> short letter; every option takes a value except help
Options parse_options(int argc, char** argv, const std::map<std::string, char>& spec) {
    Options o;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i], name, value;
        bool have_value = false;
        if (a == "-h" || a == "--help") {
            o.help = true;
            continue;
        }
        if (a.rfind("--", 0) == 0) {
            const size_t eq = a.find('=');
            name = a.substr(2, eq == std::string::npos ? std::string::npos : eq - 2);
            if (eq != std::string::npos) {
                value = a.substr(eq + 1);
                have_value = true;
            }
        } else if (a.size() >= 2 && a[0] == '-') {
            for (const auto& kv : spec)
                if (kv.second == a[1]) name = kv.first;
            if (name.empty()) throw std::runtime_error("Option '" + a.substr(1, 1) + "' does not exist");
            if (a.size() > 2) {
                value = a.substr(2);
                have_value = true;
            }
        } else {
            continue;  // positional arguments are ignored like cxxopts does without a positional spec
        }
        if (!spec.count(name)) throw std::runtime_error("Option '" + name + "' does not exist");
        if (!have_value) {
            if (i + 1 >= argc) throw std::runtime_error("Option '" + name + "' is missing an argument");
            value = argv[++i];
        }
        o.values[name] = value;
    }
    return o;
}

void check(fin_status st) {
    if (st != FIN_OK) throw std::runtime_error(fin_last_error());
}

void check_readable(const std::string& path) {  // SBWT/src/globals.cpp check_readable
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("Error reading file: " + path);
    std::fclose(f);
}

void check_writable(const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "ab");
    if (!f) throw std::runtime_error("Error writing to file: " + path);
    std::fclose(f);
}

// ---------------------------------------------------------------------------------------------------------
int build_fmin_command(int argc, char** argv) {
    const std::map<std::string, char> spec = {{"out-file", 'o'}, {"index-file", 'i'}, {"in-file", 'u'}, {"type", 0}, {"t", 't'}, {"lcs", 0}};
    const Options opts = parse_options(argc, argv, spec);
    if (argc == 1 || opts.help) {
        std::cerr << "Find all Finimizers of all input reads.\nUsage:\n  " << argv[0]
                  << " [OPTION...]\n\n  -o, --out-file arg    Output index filename prefix.\n  -i, --index-file arg  SBWT file. This has to be a binary matrix.\n"
                     "  -u, --in-file arg     The SPSS in FASTA or FASTQ format, possibly gzipped. If the file extension is .txt, this is a list of files.\n"
                     "      --type arg        Available types: rarest shortest verify (default: rarest)\n"
                     "      --t arg           Maximum finimizer frequency (default: 1)\n      --lcs arg         Provide in input the LCS file if available.\n  -h, --help            Print usage\n";
        return 1;
    }
    for (const char* need : {"in-file", "out-file", "index-file"})
        if (!opts.has(need)) throw std::runtime_error(std::string("Option '") + need + "' has no value");
    const int64_t t = std::stoll(opts.get("t", "1"));
    char stats[512];
    log_line("Building the finimizer index on CUDA device 0");
    check(fin_build(opts.get("index-file").c_str(), opts.get("in-file").c_str(), opts.get("out-file").c_str(),
                    opts.get("type", "rarest").c_str(), t, opts.get("lcs").c_str(), 0, stats, sizeof stats));
    std::cout << "String appended to the file successfully." << std::endl;  // build_fmin.hh:396
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// "(u,p) (u,p) ...\n" per read (search_fmin.hh:54-65), straight from the run-length answers of fin_search_batch_runs
inline char* put_uint(char* p, uint64_t v) {
    char tmp[24];
    int n = 0;
    do {
        tmp[n++] = (char)('0' + v % 10);
        v /= 10;
    } while (v);
    while (n) *p++ = tmp[--n];
    return p;
}

struct TextBuf {
    std::unique_ptr<char[]> p;
    size_t cap = 0, len = 0;
    void reserve(size_t n) {
        if (n > cap) {
            p.reset(new char[n]);  // never zero-filled
            cap = n;
        }
    }
};

// Text of reads [r0, r1) of a batch.  Worst case per k-mer position: "(" + digits + "," + digits + ")" + " ", bounded by the
// widest pair the index can report (or "(-1,-1) ").
void format_reads(const fin_run* runs, const uint64_t* run_offsets, const uint64_t* offsets, uint64_t r0, uint64_t r1, int64_t k,
                  size_t bytes_per_hit, TextBuf& out) {
    uint64_t n_hits = 0;
    for (uint64_t r = r0; r < r1; r++) {
        const uint64_t len = offsets[r + 1] - offsets[r];
        if (len >= (uint64_t)k) n_hits += len - k + 1;
    }
    out.reserve(n_hits * bytes_per_hit + (r1 - r0) + 16);
    char* p = out.p.get();
    static const char kMiss[] = " (-1,-1)";
    for (uint64_t r = r0; r < r1; r++) {
        const uint64_t len = offsets[r + 1] - offsets[r];
        const uint64_t n = len >= (uint64_t)k ? len - k + 1 : 0;
        uint64_t i = 0;
        auto misses_until = [&](uint64_t stop) {
            for (; i < stop; i++) {
                memcpy(p, kMiss + (i ? 0 : 1), 8);  // without the leading blank at the start of the line
                p += i ? 8 : 7;
            }
        };
        for (uint64_t t = run_offsets[r]; t < run_offsets[r + 1]; t++) {
            const fin_run& q = runs[t];
            misses_until(q.first_kmer);
            const uint64_t cnt = q.len_rc & ~FIN_RUN_RC;
            // "(unitig," is the same for the whole run
            char head[24];
            char* h = head;
            *h++ = ' ';
            *h++ = '(';
            h = put_uint(h, q.unitig);
            *h++ = ',';
            const size_t hl = (size_t)(h - head);
            uint64_t off = q.first_offset;
            const bool rc = q.len_rc & FIN_RUN_RC;
            for (uint64_t d = 0; d < cnt; d++, i++) {
                memcpy(p, head + (i ? 0 : 1), hl);
                p += i ? hl : hl - 1;
                p = put_uint(p, off);
                *p++ = ')';
                off = rc ? off - 1 : off + 1;
            }
        }
        misses_until(n);
        *p++ = '\n';
    }
    out.len = (size_t)(p - out.p.get());
}

int decimal_digits(int64_t v) {
    int n = 1;
    while (v >= 10) {
        v /= 10;
        n++;
    }
    return n;
}

struct QueryTotals {
    int64_t found_fwd = 0, found_rc = 0, positive = 0, n_queries = 0;
};

// run_fmin_queries_streaming (search_fmin.hh:33-84) as a three-stage pipeline over a ring of batches: a parser thread
// reads the next records of the file (bounded memory, whatever the file size), the main thread searches a batch on
// the GPU(s) - the answers come back as run-length records, ~50 bytes per read instead of 8 bytes per k-mer -, a writer
// thread formats the text on all host cores and writes it in read order (regular files: one pwrite per core).
struct BatchSlot {
    enum State { Empty, Parsed, Searched, Formatted } state = Empty;
    fin::SeqBatch reads;
    bool last = false;  // set by the parser: no batch follows (this one may be empty)
    struct Block {      // one contiguous block of the batch's reads per GPU
        uint64_t r0 = 0, r1 = 0;
        std::vector<uint64_t> offs, run_offsets;
        fin_run* runs = nullptr;  // pinned
        uint64_t cap = 0;
        fin_search_counts counts{};
    };
    std::vector<Block> blocks;
    std::vector<TextBuf> parts;  // the text of the batch: blocks x formatting threads, in read order
};

void search_block(fin_index* index, const fin::SeqBatch& reads, BatchSlot::Block& b) {
    const uint64_t n = b.r1 - b.r0, base = reads.offsets[b.r0];
    b.offs.resize(n + 1);
    for (uint64_t i = 0; i <= n; i++) b.offs[i] = reads.offsets[b.r0 + i] - base;
    b.run_offsets.resize(n + 1);
    b.counts = fin_search_counts{};
    for (int attempt = 0;; attempt++) {
        const uint64_t want = std::max<uint64_t>(4096, 8 * n);
        if (b.cap < want) {
            fin_host_free(b.runs);
            b.runs = nullptr;
            void* p = nullptr;
            check(fin_host_alloc(want * sizeof(fin_run), &p));
            b.runs = static_cast<fin_run*>(p);
            b.cap = want;
        }
        uint64_t n_runs = 0;
        const fin_status st = fin_search_batch_runs(index, reads.seqs.data() + base, b.offs.data(), n, b.run_offsets.data(), b.runs, b.cap,
                                                    &n_runs, &b.counts);
        if (st == FIN_ERR_ARG && n_runs > b.cap && attempt == 0) {  // more runs than expected: the call says how many
            fin_host_free(b.runs);
            b.runs = nullptr;
            b.cap = 0;
            void* p = nullptr;
            check(fin_host_alloc(n_runs * sizeof(fin_run), &p));
            b.runs = static_cast<fin_run*>(p);
            b.cap = n_runs;
            continue;
        }
        check(st);
        return;
    }
}

void run_query_file(const std::vector<fin_index*>& indexes, const fin_index_info& info, const std::string& infile, FILE* out,
                    QueryTotals& tot) {
    log_line("Running streaming queries from input file " + infile);
    const int64_t k = info.k;
    // "(u,p) ": the widest pair this index can report, or "(-1,-1) "
    const size_t bytes_per_hit = (size_t)std::max(8, decimal_digits(info.n_unitigs) + decimal_digits(info.text_len) + 4);
    const uint64_t max_reads = 1u << 19, max_bases = 128ull << 20;
    const unsigned n_threads = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
    constexpr int kRing = 3;
    BatchSlot ring[kRing];
    std::mutex mu;
    std::condition_variable cv;
    std::string parse_error, write_error, search_error;
    bool abort = false;
    int64_t us_parse = 0, us_format = 0, us_write = 0, us_search = 0, us_wait_parser = 0, us_wait_writer = 0;
    uint64_t n_batches = 0;

    std::thread parser([&] {
        try {
            fin::SeqReader reader(infile);
            for (size_t b = 0;; b++) {
                BatchSlot& s = ring[b % kRing];
                {
                    std::unique_lock<std::mutex> lk(mu);
                    cv.wait(lk, [&] { return s.state == BatchSlot::Empty || abort; });
                    if (abort) return;
                }
                const int64_t t0 = now_micros();
                s.reads.seqs.clear();
                s.reads.offsets.clear();
                const bool more = reader.next(s.reads, max_reads, max_bases);
                us_parse += now_micros() - t0;
                {
                    std::lock_guard<std::mutex> lk(mu);
                    s.last = !more;
                    s.state = BatchSlot::Parsed;
                }
                cv.notify_all();
                if (!more) return;
            }
        } catch (const std::exception& e) {
            std::lock_guard<std::mutex> lk(mu);
            parse_error = e.what();
            abort = true;
            cv.notify_all();
        }
    });

    // formatting on all cores, then one sequential write per part (buffered writes to ONE file serialise on the inode
    // lock, and filling a shared mapping from many threads stalls the search thread's pinning of pageable memory on the
    // process's mmap lock: measured, profiles/r2_cli_timing.log); the two stages run on consecutive batches
    std::thread formatter([&] {
        for (size_t b = 0;; b++) {
            BatchSlot& s = ring[b % kRing];
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return s.state == BatchSlot::Searched || abort; });
                if (abort) return;
            }
            const int64_t t0 = now_micros();
            s.parts.resize(s.blocks.size() * n_threads);
            std::vector<std::thread> workers;
            for (size_t d = 0; d < s.blocks.size(); d++) {
                BatchSlot::Block& blk = s.blocks[d];
                const uint64_t nr = blk.r1 - blk.r0, per = (nr + n_threads - 1) / n_threads;
                for (unsigned t = 0; t < n_threads; t++) {
                    const uint64_t a = std::min<uint64_t>(nr, t * per), e = std::min<uint64_t>(nr, (t + 1) * per);
                    TextBuf& part = s.parts[d * n_threads + t];
                    workers.emplace_back([&blk, &part, a, e, k, bytes_per_hit] {
                        format_reads(blk.runs, blk.run_offsets.data(), blk.offs.data(), a, e, k, bytes_per_hit, part);
                    });
                }
            }
            for (auto& w : workers) w.join();
            us_format += now_micros() - t0;
            const bool last = s.last;
            {
                std::lock_guard<std::mutex> lk(mu);
                s.state = BatchSlot::Formatted;
            }
            cv.notify_all();
            if (last) return;
        }
    });

    std::thread writer([&] {
        for (size_t b = 0;; b++) {
            BatchSlot& s = ring[b % kRing];
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return s.state == BatchSlot::Formatted || abort; });
                if (abort) return;
            }
            const int64_t t0 = now_micros();
            for (const TextBuf& part : s.parts)
                if (write_error.empty() && part.len && std::fwrite(part.p.get(), 1, part.len, out) != part.len)
                    write_error = "Error writing the output file";
            us_write += now_micros() - t0;
            const bool last = s.last;
            {
                std::lock_guard<std::mutex> lk(mu);
                s.state = BatchSlot::Empty;
                if (!write_error.empty()) abort = true;
            }
            cv.notify_all();
            if (last || !write_error.empty()) return;
        }
    });

    for (size_t b = 0;; b++) {
        BatchSlot& s = ring[b % kRing];
        const int64_t tw = now_micros();
        {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return s.state == BatchSlot::Parsed || abort; });
            if (abort) break;
        }
        const int64_t ts = now_micros();
        us_wait_parser += ts - tw;
        const uint64_t n = s.reads.size();
        // reads are independent (search_fmin.hh:43): the batch is cut into one contiguous block per GPU, every GPU holds
        // the whole index, one host thread drives each handle
        const size_t n_dev = std::max<size_t>(1, std::min<size_t>(indexes.size(), (size_t)std::max<uint64_t>(1, n)));
        if (s.blocks.size() < n_dev) s.blocks.resize(n_dev);
        std::vector<std::string> errors(n_dev);
        std::vector<std::thread> gpu_threads;
        for (size_t d = 0; d < s.blocks.size(); d++) {
            BatchSlot::Block& blk = s.blocks[d];
            blk.r0 = d < n_dev ? n * d / n_dev : n;
            blk.r1 = d < n_dev ? n * (d + 1) / n_dev : n;
            if (d >= n_dev) continue;
            auto work = [&, d] {
                try {
                    search_block(indexes[d], s.reads, s.blocks[d]);
                } catch (const std::exception& e) {
                    errors[d] = e.what();
                }
            };
            if (n_dev == 1) work();
            else gpu_threads.emplace_back(work);
        }
        for (auto& t : gpu_threads) t.join();
        for (size_t d = 0; d < n_dev; d++) {
            if (!errors[d].empty()) search_error = errors[d];
            const fin_search_counts& c = s.blocks[d].counts;
            tot.found_fwd += c.found_fwd;
            tot.found_rc += c.found_rc;
            tot.positive += c.positive;
            tot.n_queries += c.n_queries;
        }
        us_search += now_micros() - ts;
        n_batches++;
        const bool last = s.last;
        {
            std::lock_guard<std::mutex> lk(mu);
            if (!search_error.empty()) abort = true;
            else s.state = BatchSlot::Searched;
        }
        cv.notify_all();
        if (last || !search_error.empty()) break;
        const int64_t tq = now_micros();
        {   // the slot the next batch will use must have been written out: that wait is the writer's backlog
            BatchSlot& nx = ring[(b + 1) % kRing];
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return (nx.state != BatchSlot::Searched && nx.state != BatchSlot::Formatted) || abort; });
        }
        us_wait_writer += now_micros() - tq;
    }
    parser.join();
    formatter.join();
    writer.join();
    for (BatchSlot& s : ring)
        for (BatchSlot::Block& blk : s.blocks) fin_host_free(blk.runs);
    log_line("pipeline: parser " + std::to_string(us_parse / 1000) + " ms | GPU search " + std::to_string(us_search / 1000) +
             " ms (waited " + std::to_string(us_wait_parser / 1000) + " ms for the parser, " + std::to_string(us_wait_writer / 1000) +
             " ms for the writer) | writer: formatting " + std::to_string(us_format / 1000) + " ms, write " + std::to_string(us_write / 1000) +
             " ms (" + std::to_string(n_batches) + " batches)");
    if (!parse_error.empty()) throw std::runtime_error(parse_error);
    if (!search_error.empty()) throw std::runtime_error(search_error);
    if (!write_error.empty()) throw std::runtime_error(write_error);
}

// Everything is written and closed when a command returns 0; what is left is giving back pinned buffers, device memory
// and the CUDA context, which costs 0.5-1 s of a 2 s run and which the operating system does anyway.  FIN_CLEAN_EXIT=1
// keeps the orderly teardown (leak checkers, debugging).
static bool clean_exit() {
    const char* e = std::getenv("FIN_CLEAN_EXIT");
    return e && *e && *e != '0';
}

int search_fmin_command(int argc, char** argv) {
    const std::map<std::string, char> spec = {{"out-file", 'o'}, {"index-file", 'i'}, {"query-file", 'q'}};
    const Options opts = parse_options(argc, argv, spec);
    if (argc == 1 || opts.help) {
        std::cerr << "Query all Finimizers of all input reads.\nUsage:\n  " << argv[0]
                  << " [OPTION...]\n\n  -o, --out-file arg    Output filename, or stdout if not given.\n  -i, --index-file arg  Index filename prefix.\n"
                     "  -q, --query-file arg  The query in FASTA or FASTQ format, possibly gzipped. If the file extension is .txt, this is a list of query files,\n"
                     "                        one per line, and --out-file is a list of output files in the same manner.\n  -h, --help            Print usage\n";
        std::exit(1);
    }
    for (const char* need : {"query-file", "index-file"})
        if (!opts.has(need)) throw std::runtime_error(std::string("Option '") + need + "' has no value");
    const std::string queryfile = opts.get("query-file");
    const bool multi = fin::has_suffix(queryfile, ".txt");
    std::vector<std::string> query_files = multi ? fin::read_lines(queryfile) : std::vector<std::string>{queryfile};
    for (const std::string& f : query_files) check_readable(f);
    std::vector<std::string> out_files;
    const bool to_stdout = !opts.has("out-file");
    if (!to_stdout) {
        out_files = multi ? fin::read_lines(opts.get("out-file")) : std::vector<std::string>{opts.get("out-file")};
        for (const std::string& f : out_files) check_writable(f);
        if (out_files.size() != query_files.size())
            throw std::runtime_error("Number of input and output files does not match (" + std::to_string(query_files.size()) + " vs " +
                                     std::to_string(out_files.size()) + ")");
    } else {
        log_line("No output file given, writing to stdout");
    }
    const std::string prefix = opts.get("index-file");
    std::cerr << "Loading index..." << std::endl;
    // FIN_DEVICES = number of GPUs to shard the reads over (default: all visible ones)
    int n_devices = std::max(1, fin_device_count());
    if (const char* e = std::getenv("FIN_DEVICES")) n_devices = std::max(1, std::min(n_devices, std::atoi(e)));
    std::vector<fin_index*> indexes((size_t)n_devices, nullptr);
    auto free_all = [&] {
        for (fin_index* ix : indexes) fin_index_free(ix);
    };
    try {
        for (int d = 0; d < n_devices; d++) check(fin_index_load(prefix.c_str(), d, &indexes[(size_t)d]));
    } catch (...) {
        free_all();
        throw;
    }
    fin_index* index = indexes[0];
    std::cerr << "Index loaded" << (n_devices > 1 ? " on " + std::to_string(n_devices) + " GPUs" : std::string()) << std::endl;
    fin_index_info info{};
    check(fin_index_info_get(index, &info));
    std::cerr << "k = " << info.k << " SBWT nodes: " << info.n_nodes << " kmers: " << info.n_kmers << std::endl;

    int64_t total_queries = 0;
    for (size_t i = 0; i < query_files.size(); i++) {
        FILE* out = to_stdout ? stdout : std::fopen(out_files[i].c_str(), "wb");
        if (!out) throw std::runtime_error("Error writing to file: " + out_files[i]);
        QueryTotals tot;
        const int64_t t0 = now_micros();
        try {
            run_query_file(indexes, info, query_files[i], out, tot);
        } catch (...) {
            if (!to_stdout) std::fclose(out);
            free_all();
            throw;
        }
        if (!to_stdout) std::fclose(out);
        else std::fflush(stdout);
        const int64_t micros = now_micros() - t0;
        log_line("k " + std::to_string(info.k));
        log_line("us/query: " + std::to_string((double)micros / (double)std::max<int64_t>(1, tot.n_queries)) + " (including parsing and text output)");
        log_line("Found kmers: " + std::to_string(tot.found_fwd));
        log_line("Found kmers reverse : " + std::to_string(tot.found_rc));
        log_line("Total found kmers: " + std::to_string(tot.positive));
        FILE* sf = std::fopen((prefix + ".stats").c_str(), "ab");  // search_fmin.hh:79-82 (no newline)
        if (sf) {
            const std::string s = std::to_string(info.k) + "," + std::to_string(tot.found_fwd + tot.found_rc) + "," + std::to_string(tot.n_queries);
            std::fwrite(s.data(), 1, s.size(), sf);
            std::fclose(sf);
        }
        total_queries += tot.n_queries;
    }
    const int64_t total_micros = now_micros() - g_start;
    const double us_per_query = (double)total_micros / (double)std::max<int64_t>(1, total_queries);
    log_line("us/query end-to-end: " + std::to_string(us_per_query));
    log_line("total number of queries: " + std::to_string(total_queries));
    FILE* sf2 = std::fopen((prefix + "stats.txt").c_str(), "ab");  // search_fmin.hh:196-206; bytes = device layout here
    if (sf2) {
        const std::string s = "," + std::to_string(us_per_query) + "," + std::to_string(info.device_bytes) + "," +
                              std::to_string((double)(info.device_bytes * 8) / (double)std::max<int64_t>(1, info.n_kmers)) + "\n," +
                              std::to_string(info.n_kmers) + "\n";
        std::fwrite(s.data(), 1, s.size(), sf2);
        std::fclose(sf2);
    }
    log_line("bytes: " + std::to_string(info.device_bytes));
    if (clean_exit()) free_all();  // otherwise main() leaves with _exit: see there
    return 0;
}

void print_help(char** argv) {
    std::cerr << "Available commands: \n   " << argv[0] << " build-fmin\n   " << argv[0]
              << " search-fmin\nRunning a command without arguments prints the usage instructions for the command.\n";
}

}  // namespace

int main(int argc, char** argv) {
    if (argc == 1) {
        print_help(argv);
        return 1;
    }
    const std::string command = argv[1];
    if (command == "--help" || command == "-h") {
        print_help(argv);
        return 1;
    }
    for (int i = 1; i < argc; i++) argv[i - 1] = argv[i];
    argc--;
    try {
        int rc = -1;
        if (command == "build-fmin") rc = build_fmin_command(argc, argv);
        else if (command == "search-fmin") rc = search_fmin_command(argc, argv);
        else throw std::runtime_error("Invalid command: " + command);
        if (rc == 0 && !clean_exit()) {
            std::cout.flush();
            std::cerr.flush();
            std::fflush(nullptr);
            _exit(0);
        }
        return rc;
    } catch (const std::runtime_error& e) {
        std::cerr << "Runtime error: " << e.what() << '\n';
        return 1;
    } catch (const std::exception& e) {
        std::cerr << "Error: " << e.what() << '\n';
        return 1;
    }
}
