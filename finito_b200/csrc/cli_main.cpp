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
// "(u,p) (u,p) ...\n" per read (search_fmin.hh:54-65)
inline char* put_int(char* p, int64_t v) {
    if (v < 0) {
        *p++ = '-';
        v = -v;
    }
    char tmp[24];
    int n = 0;
    do {
        tmp[n++] = (char)('0' + v % 10);
        v /= 10;
    } while (v);
    while (n) *p++ = tmp[--n];
    return p;
}

// Text of reads [r0, r1) of a batch into `out` (grown as needed, never zero-filled).  Worst case per hit: "(" + 20 + "," +
// 20 + ")" + " " = 44 bytes; the bound actually used comes from the largest values the index can report.
struct TextBuf {
    std::unique_ptr<char[]> p;
    size_t cap = 0, len = 0;
    void reserve(size_t n) {
        if (n > cap) {
            p.reset(new char[n]);
            cap = n;
        }
    }
};

template <class HitT>
void format_reads(const HitT* hits, const uint64_t* offsets, uint64_t r0, uint64_t r1, uint64_t hit0, int64_t k, size_t bytes_per_hit,
                  TextBuf& out) {
    uint64_t n_hits = 0;
    for (uint64_t r = r0; r < r1; r++) {
        const uint64_t len = offsets[r + 1] - offsets[r];
        if (len >= (uint64_t)k) n_hits += len - k + 1;
    }
    out.reserve(n_hits * bytes_per_hit + (r1 - r0) + 16);
    char* p = out.p.get();
    const HitT* h = hits + 2 * hit0;
    for (uint64_t r = r0; r < r1; r++) {
        const uint64_t len = offsets[r + 1] - offsets[r];
        const uint64_t n = len >= (uint64_t)k ? len - k + 1 : 0;
        for (uint64_t i = 0; i < n; i++, h += 2) {
            if (i) *p++ = ' ';
            *p++ = '(';
            p = put_int(p, (int64_t)h[0]);
            *p++ = ',';
            p = put_int(p, (int64_t)h[1]);
            *p++ = ')';
        }
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

// run_fmin_queries_streaming (search_fmin.hh:33-84) as a two-stage pipeline: the main thread searches batch i+1 on the
// GPU(s) while a writer thread formats batch i on all host cores and writes it, in read order.  Hits travel as 2 x int32
// whenever the index allows it, through pinned buffers.
template <class HitT>
void run_query_batches(const std::vector<fin_index*>& indexes, int64_t k, const fin::SeqBatch& all, FILE* out, QueryTotals& tot,
                       size_t bytes_per_hit) {
    const uint64_t n = all.size();
    const uint64_t max_reads = 1u << 19, max_bases = 128ull << 20;
    const unsigned n_threads = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
    // batch boundaries and the largest batch (sizes the two pinned hit buffers)
    std::vector<uint64_t> cuts{0};
    uint64_t max_hits = 1;
    for (uint64_t r0 = 0; r0 < n;) {
        uint64_t r1 = r0 + 1;
        while (r1 < n && r1 - r0 < max_reads && all.offsets[r1 + 1] - all.offsets[r0] <= max_bases) r1++;
        max_hits = std::max(max_hits, fin_count_kmers(all.offsets.data() + r0, r1 - r0, k));
        cuts.push_back(r1);
        r0 = r1;
    }
    struct Slot {
        HitT* hits = nullptr;
        std::vector<uint64_t> offs;
        uint64_t n_hits = 0;
        bool full = false;
    } slots[2];
    for (Slot& s : slots) {
        void* p = nullptr;
        check(fin_host_alloc(max_hits * 2 * sizeof(HitT), &p));
        s.hits = static_cast<HitT*>(p);
    }
    std::mutex mu;
    std::condition_variable cv;
    bool done = false;
    std::string writer_error;
    int64_t us_format = 0, us_write = 0, us_search = 0, us_wait_slot = 0;

    std::thread writer([&] {
        std::vector<TextBuf> parts(n_threads);
        for (size_t b = 0;; b++) {
            Slot& s = slots[b & 1];
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return s.full || done; });
                if (!s.full) return;
            }
            if (writer_error.empty()) {
                const uint64_t nr = s.offs.size() - 1, per = (nr + n_threads - 1) / n_threads;
                std::vector<uint64_t> hit_start(n_threads + 1, 0);
                for (unsigned t = 0; t < n_threads; t++) {
                    const uint64_t a = std::min<uint64_t>(nr, t * per), e = std::min<uint64_t>(nr, (t + 1) * per);
                    hit_start[t + 1] = hit_start[t] + fin_count_kmers(s.offs.data() + a, e - a, k);
                }
                const int64_t t0 = now_micros();
                std::vector<std::thread> workers;
                for (unsigned t = 0; t < n_threads; t++) {
                    const uint64_t a = std::min<uint64_t>(nr, t * per), e = std::min<uint64_t>(nr, (t + 1) * per);
                    workers.emplace_back([&, t, a, e] { format_reads<HitT>(s.hits, s.offs.data(), a, e, hit_start[t], k, bytes_per_hit, parts[t]); });
                }
                for (auto& w : workers) w.join();
                const int64_t t1 = now_micros();
                for (unsigned t = 0; t < n_threads; t++)
                    if (parts[t].len && std::fwrite(parts[t].p.get(), 1, parts[t].len, out) != parts[t].len)
                        writer_error = "Error writing the output file";
                us_format += t1 - t0;
                us_write += now_micros() - t1;
            }
            {
                std::lock_guard<std::mutex> lk(mu);
                s.full = false;
            }
            cv.notify_all();
        }
    });

    std::string search_error;
    for (size_t b = 0; b + 1 < cuts.size() && search_error.empty(); b++) {
        const uint64_t r0 = cuts[b], r1 = cuts[b + 1];
        Slot& s = slots[b & 1];
        const int64_t tw = now_micros();
        {
            std::unique_lock<std::mutex> lk(mu);
            cv.wait(lk, [&] { return !s.full; });
        }
        const int64_t ts = now_micros();
        us_wait_slot += ts - tw;
        s.offs.assign(r1 - r0 + 1, 0);
        for (uint64_t i = r0; i <= r1; i++) s.offs[i - r0] = all.offsets[i] - all.offsets[r0];
        s.n_hits = fin_count_kmers(s.offs.data(), r1 - r0, k);
        // reads are independent (search_fmin.hh:43): the batch is cut into one contiguous block per GPU, every GPU holds
        // the whole index, one host thread drives each handle; the hits land in disjoint ranges of the same buffer
        const size_t n_dev = std::max<size_t>(1, std::min<size_t>(indexes.size(), (size_t)(r1 - r0)));
        std::vector<fin_search_counts> cs(n_dev);
        std::vector<std::string> errors(n_dev);
        std::vector<std::thread> gpu_threads;
        const uint64_t nb_reads = r1 - r0;
        for (size_t d = 0; d < n_dev; d++) {
            const uint64_t a = nb_reads * d / n_dev, e = nb_reads * (d + 1) / n_dev;
            const uint64_t hit_a = fin_count_kmers(s.offs.data(), a, k), hit_n = fin_count_kmers(s.offs.data() + a, e - a, k);
            auto work = [&, d, a, e, hit_a, hit_n] {
                std::vector<uint64_t> sub(s.offs.begin() + (long)a, s.offs.begin() + (long)e + 1);
                for (uint64_t& o : sub) o -= s.offs[a];
                cs[d] = fin_search_counts{};
                const char* seqs = all.seqs.data() + all.offsets[r0] + s.offs[a];
                fin_status st;
                if (sizeof(HitT) == 4)
                    st = fin_search_batch32(indexes[d], seqs, sub.data(), e - a, reinterpret_cast<int32_t*>(s.hits) + 2 * hit_a, hit_n, &cs[d]);
                else
                    st = fin_search_batch(indexes[d], seqs, sub.data(), e - a, reinterpret_cast<int64_t*>(s.hits) + 2 * hit_a, hit_n, &cs[d]);
                if (st != FIN_OK) errors[d] = fin_last_error();  // thread-local message: read it on this thread
            };
            if (n_dev == 1) work();
            else gpu_threads.emplace_back(work);
        }
        for (auto& t : gpu_threads) t.join();
        for (size_t d = 0; d < n_dev; d++) {
            if (!errors[d].empty()) search_error = errors[d];
            tot.found_fwd += cs[d].found_fwd;
            tot.found_rc += cs[d].found_rc;
            tot.positive += cs[d].positive;
            tot.n_queries += cs[d].n_queries;
        }
        us_search += now_micros() - ts;
        if (!search_error.empty()) break;
        {
            std::lock_guard<std::mutex> lk(mu);
            s.full = true;
        }
        cv.notify_all();
    }
    {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return (!slots[0].full && !slots[1].full) || !search_error.empty(); });
        done = true;
    }
    cv.notify_all();
    writer.join();
    for (Slot& s : slots) fin_host_free(s.hits);
    log_line("pipeline: GPU search " + std::to_string(us_search / 1000) + " ms, waiting for the writer " + std::to_string(us_wait_slot / 1000) +
             " ms | writer: formatting " + std::to_string(us_format / 1000) + " ms, fwrite " + std::to_string(us_write / 1000) + " ms (" +
             std::to_string(cuts.size() - 1) + " batches)");
    if (!search_error.empty()) throw std::runtime_error(search_error);
    if (!writer_error.empty()) throw std::runtime_error(writer_error);
}

void run_query_file(const std::vector<fin_index*>& indexes, const fin_index_info& info, const std::string& infile, FILE* out,
                    QueryTotals& tot) {
    log_line("Running streaming queries from input file " + infile);
    fin::SeqBatch all;
    fin::append_sequences(infile, all);
    // "(u,p) ": the widest pair this index can report, or "(-1,-1) "
    const size_t bytes_per_hit = (size_t)std::max(8, decimal_digits(info.n_unitigs) + decimal_digits(info.text_len) + 4);
    const bool fits32 = info.n_unitigs < 0x7FFFFFFF && info.text_len < 0x7FFFFFFF;
    if (fits32) run_query_batches<int32_t>(indexes, info.k, all, out, tot, bytes_per_hit);
    else run_query_batches<int64_t>(indexes, info.k, all, out, tot, bytes_per_hit);
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
    free_all();
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
        if (command == "build-fmin") return build_fmin_command(argc, argv);
        if (command == "search-fmin") return search_fmin_command(argc, argv);
        throw std::runtime_error("Invalid command: " + command);
    } catch (const std::runtime_error& e) {
        std::cerr << "Runtime error: " << e.what() << '\n';
        return 1;
    } catch (const std::exception& e) {
        std::cerr << "Error: " << e.what() << '\n';
        return 1;
    }
}
