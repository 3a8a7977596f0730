// Host-side engine objects behind the C ABI (include/finito_b200.h).
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

#include "device_index.cuh"
#include "formats.hpp"
#include "layout_host.hpp"
#include "../../include/finito_b200.h"

namespace fin {

struct cuda_error : std::runtime_error { using std::runtime_error::runtime_error; };

#define FIN_CUDA(expr)                                                                                         \
    do {                                                                                                       \
        cudaError_t _e = (expr);                                                                               \
        if (_e != cudaSuccess)                                                                                 \
            throw ::fin::cuda_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + ":" + \
                                    std::to_string(__LINE__) + ")");                                           \
    } while (0)

extern thread_local uint64_t g_launches;  // kernels launched by this library on this thread

// NVTX range per pipeline stage (header-only nvtx3: a no-op unless a profiler is attached)
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

// RAII device allocation
template <class T> struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    DevBuf() {}
    explicit DevBuf(size_t count) { alloc(count); }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
        return *this;
    }
    ~DevBuf() { release(); }
    void alloc(size_t count) {
        release();
        n = count;
        if (count) FIN_CUDA(cudaMalloc((void**)&p, count * sizeof(T)));
    }
    void ensure(size_t count) { if (count > n) alloc(count); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    size_t bytes() const { return n * sizeof(T); }
};

struct SearchSlot;  // per-stream staging buffers of the batch pipeline
struct SearchTotals;
struct EventTimer;

class DeviceIndex {
   public:
    DeviceIndex(const IndexHost& host, int device);
    ~DeviceIndex();
    const IndexView& view() const { return view_; }
    int device() const { return device_; }
    fin_index_info info() const { return info_; }
    bool fits_hit32() const { return fits32_; }

    // search_fmin.hh:43-72 for a batch in host memory / in device memory
    template <class HitT>
    void search_host(const char* seqs, const uint64_t* read_offsets, uint64_t n_reads, HitT* out_pairs, uint64_t out_capacity,
                     fin_search_counts* counts, int n_strands = 2);
    void search_device(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases,
                       uint64_t total_kmers, void* d_out, int hit_bytes, fin_search_counts* counts, float* kernel_ms);
    // the same answers as run-length records (include/finito_b200.h fin_run); returns the number of runs
    uint64_t search_host_runs(const char* seqs, const uint64_t* read_offsets, uint64_t n_reads, uint64_t* run_offsets, fin_run* runs,
                              uint64_t runs_capacity, fin_search_counts* counts);
    uint64_t search_device_runs(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases,
                                uint64_t* d_run_offsets, fin_run* d_runs, uint64_t runs_capacity, fin_search_counts* counts,
                                float* kernel_ms);
    // search_device (dense hits) with every scratch array inside a fixed file-scope __device__ array: no allocation at all
    void search_device_noalloc(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases,
                               uint64_t total_kmers, void* d_out, int hit_bytes, fin_search_counts* counts);
    fin_access_counters last_access{};
    bool count_accesses = false;
    fin_search_timing last_timing{};

   private:
    uint64_t prepare_device_batch(const void* d_seqs, const uint64_t* d_read_offsets, uint64_t n_reads, uint64_t n_bases);
    void finish_device_batch(EventTimer& t, fin_search_counts* counts, float* kernel_ms, SearchTotals& tot);
    void ensure_slots();
    int device_;
    IndexView view_{};
    fin_index_info info_{};
    bool fits32_ = true;
    DevBuf<uint64_t> lines_, ustart_, fmin_, tstart_, text_;
    DevBuf<uint32_t> offsets_, ends_, smer_;
    std::vector<std::unique_ptr<SearchSlot>> slots_;
    // scratch of search_device
    DevBuf<uint64_t> dev_packed_, dev_out_offsets_, dev_keys_a_, dev_keys_b_;
    DevBuf<uint32_t> dev_fmask_;
    DevBuf<uint8_t> dev_strand_flags_;
    DevBuf<uint8_t> dev_scan_tmp_, dev_sort_tmp_, dev_recs_a_;
    DevBuf<uint8_t> dev_totals_;
};

// build-fmin on the device (build.cu)
std::string build_fmin_device(const std::string& sbwt_path, const std::string& unitig_path, const std::string& out_prefix,
                              const std::string& type, int64_t t, const std::string& lcs_path, int device);
std::vector<uint8_t> lcs_device(const SbwtHost& sbwt, int device);
fin_build_timing last_build_timing();
DevBuf<uint64_t> index_lines_device(const IndexHost& h);  // build.cu: the 128-byte lines of a loaded index, built on the device
double probe_random_sectors(int device, uint64_t table_bytes, int repeats, int bytes_per_access);  // probe.cu

// plain-matrix SBWT construction on the device (sbwt_build.cu)
void sbwt_build_device(const std::string& unitig_path, int64_t k, const std::string& out_path, int64_t precalc_k, int device);
fin_sbwt_build_timing last_sbwt_build_timing();
fin_spss_info spss_build_device(const std::string& in_path, int64_t k, const std::string& out_path, int flip, uint64_t seed, int device);

// staged build-fmin --type rarest, shardable over GPUs (build.cu)
class RarestBuilder;
RarestBuilder* builder_open(const std::string& sbwt_path, const std::string& unitig_path, const std::string& lcs_path,
                            const std::string& out_prefix, int device, int rank, int world);
void builder_free(RarestBuilder* b);
fin_builder_info builder_info(const RarestBuilder& b);
uint64_t builder_stream(RarestBuilder& b);
void builder_export(RarestBuilder& b, uint64_t* d_out, uint64_t capacity, uint64_t* counts);
uint64_t builder_merge(RarestBuilder& b, const uint64_t* d_tuples, uint64_t n);
void builder_slice(const RarestBuilder& b, uint64_t* d_bits, uint32_t* d_offsets);
void builder_finish(RarestBuilder& b, const uint64_t* bits_all, uint64_t n_bit_words, const uint32_t* offsets_all, uint64_t n_offsets,
                    const std::string& out_prefix);

// staged build-fmin --type shortest|verify, shardable over GPUs (build.cu)
struct StatsHandle;
StatsHandle* stats_open(const std::string& sbwt_path, const std::string& unitig_path, const std::string& lcs_path, const std::string& out_prefix,
                        const std::string& type, int64_t t, int device, int rank, int world);
void stats_free(StatsHandle* h);
uint64_t stats_first_pass(StatsHandle& h);
void stats_table(StatsHandle& h, uint32_t* d_table, bool set);
uint64_t stats_second_pass(StatsHandle& h);
void stats_keys(const StatsHandle& h, uint64_t* d_keys);
std::string stats_finish(StatsHandle& h, const uint64_t* d_keys_all, uint64_t n_keys_all, bool write_line);

}  // namespace fin
