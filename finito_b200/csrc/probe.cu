// Random-sector probe: the "HBM random-access roofline" SURVEY.md section 8(d) asks for beside the copy peak.
// Every thread issues independent 32-byte loads (one DRAM sector each) at pseudo-random sector addresses of a table
// far larger than the 126 MB L2, eight in flight per thread; the figure is sectors * 32 B / CUDA-event time.  This is
// what a kernel made ONLY of data-independent random sector reads reaches on this GPU — the ceiling for the rank /
// LCS accesses of search_reads_kernel once the index no longer fits the L2.
#include "engine.hpp"

namespace fin {

// SECTORS consecutive 32-byte sectors per random location (1 = one sector, 4 = a whole 128-byte line)
template <int SECTORS>
static __global__ void __launch_bounds__(256) random_sector_kernel(const uint4* __restrict__ table, uint64_t n_locations, uint32_t iters,
                                                                  unsigned long long* __restrict__ sink) {
    uint64_t s = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ULL + 0x632BE59BD9B4E019ULL;
    uint32_t acc = 0;
    constexpr int INFLIGHT = 8 / SECTORS > 0 ? 8 / SECTORS : 1;
    for (uint32_t i = 0; i < iters; i++) {
        uint4 v[INFLIGHT][2 * SECTORS];
#pragma unroll
        for (int j = 0; j < INFLIGHT; j++) {
            s = s * 6364136223846793005ULL + 1442695040888963407ULL;
            const uint64_t loc = __umul64hi(s, n_locations);  // uniform in [0, n_locations) without a division
#pragma unroll
            for (int q = 0; q < 2 * SECTORS; q++) v[j][q] = __ldg(table + 2 * SECTORS * loc + q);
        }
#pragma unroll
        for (int j = 0; j < INFLIGHT; j++)
#pragma unroll
            for (int q = 0; q < 2 * SECTORS; q++) acc ^= v[j][q].x ^ v[j][q].w;
    }
    if (acc == 0x12345678u) atomicAdd(sink, 1ULL);  // keeps the loads alive
}

double probe_random_sectors(int device, uint64_t table_bytes, int repeats, int bytes_per_access) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) throw cuda_error("no CUDA device is visible");
    if (device < 0 || device >= n_dev) throw arg_error("CUDA device ordinal out of range");
    FIN_CUDA(cudaSetDevice(device));
    if (table_bytes < (1ull << 20)) throw arg_error("probe table too small");
    if (bytes_per_access != 32 && bytes_per_access != 64 && bytes_per_access != 128) throw arg_error("probe: bytes per access must be 32, 64 or 128");
    const uint64_t n_sectors = table_bytes / 32;
    const int sectors = bytes_per_access / 32;
    const uint64_t n_loc = n_sectors / (uint64_t)sectors;
    const int inflight = 8 / sectors;
    DevBuf<uint4> table(n_sectors * 2);
    FIN_CUDA(cudaMemset(table.p, 0x5A, table.bytes()));
    DevBuf<unsigned long long> sink(1);
    FIN_CUDA(cudaMemset(sink.p, 0, 8));
    int sms = 0;
    FIN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const unsigned blocks = (unsigned)sms * 8, threads = 256;
    const uint32_t iters = 64;
    cudaEvent_t e0, e1;
    FIN_CUDA(cudaEventCreate(&e0));
    FIN_CUDA(cudaEventCreate(&e1));
    auto launch = [&](uint32_t it) {
        if (sectors == 1) random_sector_kernel<1><<<blocks, threads>>>(table.p, n_loc, it, sink.p);
        else if (sectors == 2) random_sector_kernel<2><<<blocks, threads>>>(table.p, n_loc, it, sink.p);
        else random_sector_kernel<4><<<blocks, threads>>>(table.p, n_loc, it, sink.p);
    };
    launch(4);  // warm-up
    g_launches++;
    float best = 1e30f;
    for (int r = 0; r < std::max(1, repeats); r++) {
        FIN_CUDA(cudaEventRecord(e0));
        launch(iters);
        g_launches++;
        FIN_CUDA(cudaEventRecord(e1));
        FIN_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        FIN_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        best = std::min(best, ms);
    }
    FIN_CUDA(cudaGetLastError());
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double accesses = (double)blocks * threads * iters * (double)inflight;
    return accesses * (double)bytes_per_access / (best * 1e-3) / 1e9;
}

}  // namespace fin
