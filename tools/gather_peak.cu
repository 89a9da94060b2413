// gather_peak.cu -- measures the B200's random-gather ceilings that bound the hash-grid encoder (SURVEY.md section 8d:
// "the builder must measure an L2 gather peak").  Stand-alone executable, not part of libngp_b200.so.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/gather_peak tools/gather_peak.cu && tools/gather_peak
//
// Every thread issues ILP independent loads per round from a table of `bytes` bytes at pseudo-random element indices
// (xorshift per thread -> no two lanes of a warp share a 128 B line except by chance).  Variants:
//   elem 4/8/16 B   LDG.32 / LDG.64 / LDG.128 per gather
//   warp_lines      how many distinct 128 B lines one warp-wide load touches (32 = fully random, 8, 1 = coalesced)
// Prints one JSON object per configuration: G gathers/s, algorithmic GB/s (elem bytes), sector GB/s (32 B per gather
// when random) and cycles per warp-load per SM.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t xorshift(uint32_t &s) { s ^= s << 13; s ^= s >> 17; s ^= s << 5; return s; }

template <typename T> __device__ __forceinline__ uint32_t fold(T v);
template <> __device__ __forceinline__ uint32_t fold<uint32_t>(uint32_t v) { return v; }
template <> __device__ __forceinline__ uint32_t fold<uint2>(uint2 v) { return v.x ^ v.y; }
template <> __device__ __forceinline__ uint32_t fold<uint4>(uint4 v) { return v.x ^ v.y ^ v.z ^ v.w; }

// warp_lines: lanes are grouped so that 32/warp_lines consecutive lanes read the same 128 B line (different elements)
template <typename T, int ILP>
__global__ void __launch_bounds__(256) k_gather(const T *__restrict__ table, uint32_t n_elem_mask, uint32_t rounds, uint32_t warp_lines,
                                                uint32_t *__restrict__ sink) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t group = lane / (32 / warp_lines);            // which of the warp's lines this lane reads
    uint32_t s = (blockIdx.x * blockDim.x + (threadIdx.x & ~31u) + group) * 2654435761u + 12345u;
    const uint32_t per_line = 128 / sizeof(T);
    const uint32_t within = lane % per_line;
    uint32_t acc = 0;
    for (uint32_t r = 0; r < rounds; r++) {
        T v[ILP];
        uint32_t idx[ILP];
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            const uint32_t line = xorshift(s);
            idx[i] = warp_lines == 32 ? (line & n_elem_mask) : (((line * per_line) + within) & n_elem_mask);
        }
#pragma unroll
        for (int i = 0; i < ILP; i++) v[i] = __ldg(table + idx[i]);
#pragma unroll
        for (int i = 0; i < ILP; i++) acc ^= fold<T>(v[i]);
    }
    if (acc == 0x12345678u) sink[0] = acc;
}

template <typename T, int ILP>
static void run(const void *table, size_t bytes, uint32_t warp_lines, int sm_count, int clock_khz, uint32_t *sink, const char *label) {
    const uint32_t n_elem = (uint32_t)(bytes / sizeof(T));
    uint32_t pow2 = 1;
    while (pow2 * 2 <= n_elem) pow2 *= 2;
    const uint32_t rounds = 256;
    const int blocks = sm_count * 8, threads = 256;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        CK(cudaEventRecord(e0));
        k_gather<T, ILP><<<blocks, threads>>>(reinterpret_cast<const T *>(table), pow2 - 1, rounds, warp_lines, sink);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    const double gathers = (double)blocks * threads * rounds * ILP;
    const double gps = gathers / (best * 1e-3);
    const double warp_loads_per_sm = gathers / 32 / sm_count;
    const double cyc = best * 1e-3 * clock_khz * 1e3 / warp_loads_per_sm;
    printf("{\"test\": \"%s\", \"elem_bytes\": %zu, \"table_mb\": %.1f, \"warp_lines\": %u, \"ilp\": %d, \"ms\": %.3f, \"ggathers_per_s\": %.2f, "
           "\"alg_gbs\": %.1f, \"sector_gbs\": %.1f, \"cyc_per_warp_load_per_sm_at_max_clock\": %.2f}\n",
           label, sizeof(T), pow2 * sizeof(T) / 1048576.0, warp_lines, ILP, best, gps / 1e9, gps * sizeof(T) / 1e9,
           gps * (warp_lines == 32 ? 32.0 : 128.0 * warp_lines / 32) / 1e9, cyc);
    fflush(stdout);
}

int main() {
    int dev = 0, sm = 0, khz = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    const size_t max_bytes = 256ull << 20;
    void *table;
    uint32_t *sink;
    CK(cudaMalloc(&table, max_bytes));
    CK(cudaMemset(table, 1, max_bytes));
    CK(cudaMalloc(&sink, 4));
    const size_t sizes[] = {2ull << 20, 16ull << 20, 128ull << 20, 256ull << 20};   // one hashed level, ~the 2^19 table, > L2 (2^22 table)
    for (size_t b : sizes) {
        run<uint32_t, 8>(table, b, 32, sm, khz, sink, "random4");
        run<uint32_t, 16>(table, b, 32, sm, khz, sink, "random4");
        run<uint2, 8>(table, b, 32, sm, khz, sink, "random8");
        run<uint4, 8>(table, b, 32, sm, khz, sink, "random16");
    }
    for (uint32_t wl : {16u, 8u, 4u, 1u}) run<uint32_t, 8>(table, 16ull << 20, wl, sm, khz, sink, "grouped4");
    return 0;
}
