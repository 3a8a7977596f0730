// Host side of the run-length hits (include/finito_b200.h fin_run): runs -> the dense (unitig, offset) table that
// search-fmin prints (include/search_fmin.hh:54-65).  Pure host arithmetic, shared by the library (fin_expand_runs)
// and the CLI's formatter.
#pragma once
#include <algorithm>
#include <cstdint>
#include <thread>
#include <vector>

#include "formats.hpp"
#include "../../include/finito_b200.h"

namespace fin {

// -------------------------------------------------------------------------------------------------------
// runs -> the dense (unitig, offset) table, on the host
// -------------------------------------------------------------------------------------------------------
template <class HitT>
static bool expand_range(const uint64_t* read_offsets, uint64_t ra, uint64_t rb, uint64_t k, const uint64_t* run_offsets, const fin_run* runs,
                         HitT* out) {
    for (uint64_t r = ra; r < rb; r++) {
        const uint64_t len = read_offsets[r + 1] - read_offsets[r];
        const uint64_t nk = len >= k ? len - k + 1 : 0;
        uint64_t i = 0;
        for (uint64_t t = run_offsets[r]; t < run_offsets[r + 1]; t++) {
            const fin_run& q = runs[t];
            const uint64_t n = q.len_rc & ~FIN_RUN_RC;
            if (q.first_kmer < i || (uint64_t)q.first_kmer + n > nk) return false;  // unsorted, overlapping or out of the read
            for (; i < q.first_kmer; i++) out[2 * i] = out[2 * i + 1] = (HitT)-1;
            const HitT u = (HitT)q.unitig;
            if (q.len_rc & FIN_RUN_RC) {
                if (n > (uint64_t)q.first_offset + 1) return false;
                for (uint64_t d = 0; d < n; d++, i++) { out[2 * i] = u; out[2 * i + 1] = (HitT)(q.first_offset - d); }
            } else {
                for (uint64_t d = 0; d < n; d++, i++) { out[2 * i] = u; out[2 * i + 1] = (HitT)(q.first_offset + d); }
            }
        }
        for (; i < nk; i++) out[2 * i] = out[2 * i + 1] = (HitT)-1;
        out += 2 * nk;
    }
    return true;
}

inline void expand_runs_host(const uint64_t* read_offsets, uint64_t n_reads, int64_t k, const uint64_t* run_offsets, const fin_run* runs,
                      void* out_pairs, uint64_t out_capacity, int hit_bytes, int n_threads) {
    if (hit_bytes != 4 && hit_bytes != 8) throw arg_error("hit_bytes must be 4 or 8");
    if (k < 1) throw arg_error("k must be positive");
    const int T = (int)std::max<int64_t>(1, std::min<int64_t>(n_threads, (int64_t)(n_reads / 1024 + 1)));
    // contiguous blocks of reads per thread; the first output slot of each block from one pass over the offsets
    std::vector<uint64_t> first_read(T + 1), first_out(T + 1);
    uint64_t total = 0;
    for (int t = 0, r = 0; t <= T; t++) {
        const uint64_t stop = n_reads * (uint64_t)t / (uint64_t)T;
        for (; (uint64_t)r < stop; r++) {
            if (read_offsets[r + 1] < read_offsets[r]) throw arg_error("read_offsets must be non-decreasing");
            const uint64_t len = read_offsets[r + 1] - read_offsets[r];
            total += len >= (uint64_t)k ? len - (uint64_t)k + 1 : 0;
        }
        first_read[t] = stop;
        first_out[t] = total;
    }
    if (total > out_capacity) throw arg_error("out_capacity is smaller than the number of k-mer positions of the batch");
    std::vector<char> ok(T, 1);
    auto work = [&](int t) {
        if (hit_bytes == 4)
            ok[t] = expand_range<int32_t>(read_offsets, first_read[t], first_read[t + 1], (uint64_t)k, run_offsets, runs,
                                          (int32_t*)out_pairs + 2 * first_out[t]);
        else
            ok[t] = expand_range<int64_t>(read_offsets, first_read[t], first_read[t + 1], (uint64_t)k, run_offsets, runs,
                                          (int64_t*)out_pairs + 2 * first_out[t]);
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < T; t++) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
    for (int t = 0; t < T; t++)
        if (!ok[t]) throw arg_error("runs are not sorted, overlap, or leave their read");
}

}  // namespace fin
