// extern "C" surface of libfinito_b200.so (include/finito_b200.h).  Exceptions stop here.
#include <cstring>
#include <new>

#include "engine.hpp"
#include "runs_host.hpp"

struct fin_index {
    std::unique_ptr<fin::DeviceIndex> dev;
};

struct fin_stats {
    fin::StatsHandle* h = nullptr;
    ~fin_stats() { fin::stats_free(h); }
};

struct fin_builder {
    fin::RarestBuilder* b = nullptr;
    ~fin_builder() { fin::builder_free(b); }
};

namespace {

thread_local std::string g_error;

template <class F> fin_status guarded(F&& f) {
    try {
        f();
        g_error.clear();
        return FIN_OK;
    } catch (const fin::arg_error& e) { g_error = e.what(); return FIN_ERR_ARG;
    } catch (const fin::io_error& e) { g_error = e.what(); return FIN_ERR_IO;
    } catch (const fin::format_error& e) { g_error = e.what(); return FIN_ERR_FORMAT;
    } catch (const fin::cuda_error& e) { g_error = e.what(); cudaGetLastError(); return FIN_ERR_CUDA;
    } catch (const fin::domain_error& e) { g_error = e.what(); return FIN_ERR_DOMAIN;
    } catch (const fin::limit_error& e) { g_error = e.what(); return FIN_ERR_LIMIT;
    } catch (const std::bad_alloc&) { g_error = "out of host memory"; return FIN_ERR_IO;
    } catch (const std::exception& e) { g_error = e.what(); return FIN_ERR_ARG; }
}

void require(bool ok, const char* what) {
    if (!ok) throw fin::arg_error(what);
}

}  // namespace

extern "C" {

const char* fin_last_error(void) { return g_error.c_str(); }

int fin_abi_version(void) { return 3; }

int fin_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

fin_status fin_index_load(const char* prefix, int device, fin_index** out) {
    return guarded([&] {
        require(prefix && out, "fin_index_load: null argument");
        *out = nullptr;
        fin::IndexHost host = fin::load_index_files(prefix);
        std::unique_ptr<fin_index> h(new fin_index());
        h->dev.reset(new fin::DeviceIndex(host, device));
        *out = h.release();
    });
}

void fin_index_free(fin_index* index) { delete index; }

fin_status fin_index_info_get(const fin_index* index, fin_index_info* out) {
    return guarded([&] {
        require(index && out, "fin_index_info_get: null argument");
        *out = index->dev->info();
    });
}

uint64_t fin_count_kmers(const uint64_t* read_offsets, uint64_t n_reads, int64_t k) {
    uint64_t total = 0;
    if (!read_offsets || k < 1) return 0;
    for (uint64_t i = 0; i < n_reads; i++) {
        const uint64_t len = read_offsets[i + 1] - read_offsets[i];
        if (len >= (uint64_t)k) total += len - (uint64_t)k + 1;
    }
    return total;
}

fin_status fin_search_batch(fin_index* index, const char* seqs, const uint64_t* read_offsets, uint64_t n_reads,
                            int64_t* out_pairs, uint64_t out_capacity, fin_search_counts* counts) {
    return guarded([&] {
        require(index && read_offsets && (seqs || n_reads == 0) && (out_pairs || out_capacity == 0), "fin_search_batch: null argument");
        index->dev->search_host<int64_t>(seqs, read_offsets, n_reads, out_pairs, out_capacity, counts);
    });
}

fin_status fin_search_forward(fin_index* index, const char* seqs, const uint64_t* read_offsets, uint64_t n_reads,
                              int64_t* out_pairs, uint64_t out_capacity, fin_search_counts* counts) {
    return guarded([&] {
        require(index && read_offsets && (seqs || n_reads == 0) && (out_pairs || out_capacity == 0), "fin_search_forward: null argument");
        index->dev->search_host<int64_t>(seqs, read_offsets, n_reads, out_pairs, out_capacity, counts, 1);
    });
}

fin_status fin_search_batch32(fin_index* index, const char* seqs, const uint64_t* read_offsets, uint64_t n_reads,
                              int32_t* out_pairs, uint64_t out_capacity, fin_search_counts* counts) {
    return guarded([&] {
        require(index && read_offsets && (seqs || n_reads == 0) && (out_pairs || out_capacity == 0), "fin_search_batch32: null argument");
        index->dev->search_host<int32_t>(seqs, read_offsets, n_reads, out_pairs, out_capacity, counts);
    });
}

fin_status fin_search_device(fin_index* index, const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads,
                             uint64_t n_bases, uint64_t total_kmers, void* d_out, int hit_bytes, fin_search_counts* counts,
                             float* kernel_ms) {
    return guarded([&] {
        require(index && d_read_offsets && (d_seqs || n_bases == 0) && (d_out || total_kmers == 0), "fin_search_device: null argument");
        index->dev->search_device(d_seqs, d_read_offsets, n_reads, n_bases, total_kmers, d_out, hit_bytes, counts, kernel_ms);
    });
}

fin_status fin_search_device_noalloc(fin_index* index, const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads,
                                     uint64_t n_bases, uint64_t total_kmers, void* d_out, int hit_bytes,
                                     fin_search_counts* counts) {
    return guarded([&] {
        require(index && d_read_offsets && (d_seqs || n_bases == 0) && (d_out || total_kmers == 0),
                "fin_search_device_noalloc: null argument");
        index->dev->search_device_noalloc(d_seqs, d_read_offsets, n_reads, n_bases, total_kmers, d_out, hit_bytes, counts);
    });
}

fin_status fin_search_batch_runs(fin_index* index, const char* seqs, const uint64_t* read_offsets, uint64_t n_reads,
                                 uint64_t* run_offsets, fin_run* runs, uint64_t runs_capacity, uint64_t* n_runs,
                                 fin_search_counts* counts) {
    return guarded([&] {
        require(index && read_offsets && run_offsets && n_runs && (seqs || n_reads == 0) && (runs || runs_capacity == 0),
                "fin_search_batch_runs: null argument");
        *n_runs = index->dev->search_host_runs(seqs, read_offsets, n_reads, run_offsets, runs, runs_capacity, counts);
        if (*n_runs > runs_capacity) throw fin::arg_error("runs_capacity is smaller than the number of runs of the batch (see *n_runs)");
    });
}

fin_status fin_search_device_runs(fin_index* index, const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads,
                                  uint64_t n_bases, uint64_t* d_run_offsets, fin_run* d_runs, uint64_t runs_capacity,
                                  uint64_t* n_runs, fin_search_counts* counts, float* kernel_ms) {
    return guarded([&] {
        require(index && d_read_offsets && d_run_offsets && n_runs && (d_seqs || n_bases == 0) && (d_runs || runs_capacity == 0),
                "fin_search_device_runs: null argument");
        *n_runs = index->dev->search_device_runs(d_seqs, d_read_offsets, n_reads, n_bases, d_run_offsets, d_runs, runs_capacity, counts,
                                                 kernel_ms);
        if (*n_runs > runs_capacity) throw fin::arg_error("runs_capacity is smaller than the number of runs of the batch (see *n_runs)");
    });
}

fin_status fin_expand_runs(const uint64_t* read_offsets, uint64_t n_reads, int64_t k, const uint64_t* run_offsets,
                           const fin_run* runs, void* out_pairs, uint64_t out_capacity, int hit_bytes, int n_threads) {
    return guarded([&] {
        require(read_offsets && run_offsets && (runs || run_offsets[n_reads] == 0) && (out_pairs || out_capacity == 0),
                "fin_expand_runs: null argument");
        fin::expand_runs_host(read_offsets, n_reads, k, run_offsets, runs, out_pairs, out_capacity, hit_bytes, n_threads);
    });
}

fin_status fin_sbwt_build(const char* unitig_path, int64_t k, const char* out_sbwt_path, int64_t precalc_k, int device) {
    return guarded([&] {
        require(unitig_path && out_sbwt_path, "fin_sbwt_build: null argument");
        fin::sbwt_build_device(unitig_path, k, out_sbwt_path, precalc_k, device);
    });
}

fin_status fin_spss_build(const char* in_path, int64_t k, const char* out_fasta_path, int flip, uint64_t seed, int device,
                          fin_spss_info* info) {
    return guarded([&] {
        require(in_path && out_fasta_path, "fin_spss_build: null argument");
        const fin_spss_info i = fin::spss_build_device(in_path, k, out_fasta_path, flip, seed, device);
        if (info) *info = i;
    });
}

fin_status fin_sbwt_build_last_timing(fin_sbwt_build_timing* out) {
    return guarded([&] {
        require(out != nullptr, "fin_sbwt_build_last_timing: null argument");
        *out = fin::last_sbwt_build_timing();
    });
}

fin_status fin_search_access_counters(const fin_index* index, fin_access_counters* out) {
    return guarded([&] {
        require(index && out, "fin_search_access_counters: null argument");
        *out = index->dev->last_access;
    });
}

fin_status fin_search_set_counting(fin_index* index, int enabled) {
    return guarded([&] {
        require(index != nullptr, "fin_search_set_counting: null argument");
        index->dev->count_accesses = enabled != 0;
    });
}

fin_status fin_search_last_timing(const fin_index* index, fin_search_timing* out) {
    return guarded([&] {
        require(index && out, "fin_search_last_timing: null argument");
        *out = index->dev->last_timing;
    });
}

uint64_t fin_launch_count(void) { return fin::g_launches; }

fin_status fin_probe_random_sectors(int device, uint64_t table_bytes, int repeats, int bytes_per_access, double* gb_per_s) {
    return guarded([&] {
        require(gb_per_s != nullptr, "fin_probe_random_sectors: null argument");
        *gb_per_s = fin::probe_random_sectors(device, table_bytes, repeats, bytes_per_access);
    });
}

fin_status fin_host_alloc(size_t bytes, void** out) {
    return guarded([&] {
        require(out != nullptr, "fin_host_alloc: null argument");
        *out = nullptr;
        FIN_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    });
}

void fin_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

fin_status fin_build(const char* sbwt_path, const char* unitig_path, const char* out_prefix, const char* type, int64_t t,
                     const char* lcs_path, int device, char* stats_out, size_t stats_cap) {
    return guarded([&] {
        require(sbwt_path && unitig_path && out_prefix && type, "fin_build: null argument");
        const std::string stats =
            fin::build_fmin_device(sbwt_path, unitig_path, out_prefix, type, t, lcs_path ? lcs_path : "", device);
        if (stats_out && stats_cap) {
            strncpy(stats_out, stats.c_str(), stats_cap - 1);
            stats_out[stats_cap - 1] = 0;
        }
    });
}

fin_status fin_build_last_timing(fin_build_timing* out) {
    return guarded([&] {
        require(out != nullptr, "fin_build_last_timing: null argument");
        *out = fin::last_build_timing();
    });
}

fin_status fin_lcs(const char* sbwt_path, int with_variant_header, int device, uint8_t* out, uint64_t out_capacity,
                   uint64_t* n_nodes) {
    return guarded([&] {
        require(sbwt_path && out, "fin_lcs: null argument");
        fin::SbwtHost s = fin::load_sbwt_file(sbwt_path, with_variant_header != 0);
        if (n_nodes) *n_nodes = (uint64_t)s.n_nodes;
        require((uint64_t)s.n_nodes <= out_capacity, "fin_lcs: output buffer too small");
        std::vector<uint8_t> l = fin::lcs_device(s, device);
        memcpy(out, l.data(), (size_t)s.n_nodes);
    });
}

fin_status fin_builder_open(const char* sbwt_path, const char* unitig_path, const char* lcs_path, const char* out_prefix,
                            int device, int rank, int world, fin_builder** out) {
    return guarded([&] {
        require(sbwt_path && unitig_path && out_prefix && out, "fin_builder_open: null argument");
        *out = nullptr;
        std::unique_ptr<fin_builder> h(new fin_builder());
        h->b = fin::builder_open(sbwt_path, unitig_path, lcs_path ? lcs_path : "", out_prefix, device, rank, world);
        *out = h.release();
    });
}

void fin_builder_free(fin_builder* b) { delete b; }

fin_status fin_builder_info_get(const fin_builder* b, fin_builder_info* out) {
    return guarded([&] {
        require(b && out, "fin_builder_info_get: null argument");
        *out = fin::builder_info(*b->b);
    });
}

fin_status fin_builder_stream(fin_builder* b, uint64_t* n_tuples) {
    return guarded([&] {
        require(b && n_tuples, "fin_builder_stream: null argument");
        *n_tuples = fin::builder_stream(*b->b);
    });
}

fin_status fin_builder_export(fin_builder* b, void* d_tuples, uint64_t capacity_tuples, uint64_t* counts_per_rank) {
    return guarded([&] {
        require(b && counts_per_rank && (d_tuples || capacity_tuples == 0), "fin_builder_export: null argument");
        fin::builder_export(*b->b, static_cast<uint64_t*>(d_tuples), capacity_tuples, counts_per_rank);
    });
}

fin_status fin_builder_merge(fin_builder* b, const void* d_tuples, uint64_t n_tuples, uint64_t* n_slice_fmin) {
    return guarded([&] {
        require(b && n_slice_fmin && (d_tuples || n_tuples == 0), "fin_builder_merge: null argument");
        *n_slice_fmin = fin::builder_merge(*b->b, static_cast<const uint64_t*>(d_tuples), n_tuples);
    });
}

fin_status fin_builder_slice(const fin_builder* b, void* d_bits, void* d_offsets) {
    return guarded([&] {
        require(b && d_bits, "fin_builder_slice: null argument");
        fin::builder_slice(*b->b, static_cast<uint64_t*>(d_bits), static_cast<uint32_t*>(d_offsets));
    });
}

fin_status fin_builder_finish(fin_builder* b, const uint64_t* bits_all, uint64_t n_bit_words, const uint32_t* offsets_all,
                              uint64_t n_offsets, const char* out_prefix) {
    return guarded([&] {
        require(b && bits_all && out_prefix && (offsets_all || n_offsets == 0), "fin_builder_finish: null argument");
        fin::builder_finish(*b->b, bits_all, n_bit_words, offsets_all, n_offsets, out_prefix);
    });
}


fin_status fin_stats_open(const char* sbwt_path, const char* unitig_path, const char* lcs_path, const char* out_prefix, const char* type,
                          int64_t t, int device, int rank, int world, fin_stats** out) {
    return guarded([&] {
        require(sbwt_path && unitig_path && type && out, "fin_stats_open: null argument");
        *out = nullptr;
        std::unique_ptr<fin_stats> s(new fin_stats());
        s->h = fin::stats_open(sbwt_path, unitig_path, lcs_path ? lcs_path : "", out_prefix ? out_prefix : "", type, t, device, rank, world);
        *out = s.release();
    });
}
void fin_stats_free(fin_stats* s) { delete s; }
fin_status fin_stats_first_pass(fin_stats* s, uint64_t* n_table) {
    return guarded([&] {
        require(s && n_table, "fin_stats_first_pass: null argument");
        *n_table = fin::stats_first_pass(*s->h);
    });
}
fin_status fin_stats_table(fin_stats* s, void* d_table, int set) {
    return guarded([&] {
        require(s && d_table, "fin_stats_table: null argument");
        fin::stats_table(*s->h, static_cast<uint32_t*>(d_table), set != 0);
    });
}
fin_status fin_stats_second_pass(fin_stats* s, uint64_t* n_keys) {
    return guarded([&] {
        require(s && n_keys, "fin_stats_second_pass: null argument");
        *n_keys = fin::stats_second_pass(*s->h);
    });
}
fin_status fin_stats_keys(const fin_stats* s, void* d_keys) {
    return guarded([&] {
        require(s && d_keys, "fin_stats_keys: null argument");
        fin::stats_keys(*s->h, static_cast<uint64_t*>(d_keys));
    });
}
fin_status fin_stats_finish(fin_stats* s, const void* d_keys_all, uint64_t n_keys_all, int write_line, char* stats_out, size_t stats_cap) {
    return guarded([&] {
        require(s && (d_keys_all || n_keys_all == 0), "fin_stats_finish: null argument");
        const std::string r = fin::stats_finish(*s->h, static_cast<const uint64_t*>(d_keys_all), n_keys_all, write_line != 0);
        if (stats_out) {
            require(r.size() + 1 <= stats_cap, "fin_stats_finish: stats buffer too small");
            memcpy(stats_out, r.c_str(), r.size() + 1);
        }
    });
}
}  // extern "C"
