// tc_mlp.cuh -- the fully fused 64-wide density + colour MLP on Blackwell's 5th-generation tensor cores (tcgen05).
//
// Replaces, for the fused paths, the wmma formulation of ffmlp/src/ffmlp.cu:331-407 (kernel_mlp_fused, 16x16x16 fragments):
//   * a tile is 128 sample rows = the 128 TMEM lanes; thread = row; a warp owns the 32 rows of its lane quadrant (warp % 4);
//   * every layer is `tcgen05.mma.cta_group::1.kind::f16` M=128, N=64|16, K=16 per instruction, issued by ONE elected thread:
//     A (activations) is read from TENSOR MEMORY (fp16, two halves per 32-bit column, row = lane: what tcgen05.st.32x32b writes),
//     B (weights) from shared memory through a matrix descriptor, D (fp32 accumulators) lives in tensor memory;
//   * with an output-lane mask a warp runs the layer for ITS 32 rows only, so the four quadrants of a tile are independent chains:
//     tcgen05.ld D -> ReLU / fp16 -> tcgen05.st next A -> issue next layer -> wait for its commit, no other warp involved;
//   * no accumulator, A fragment or B fragment ever lives in the register file across a layer, and no activation in shared memory.
//
// Shared-memory layout of a weight matrix [N x K] (K-major, no swizzle):  uint4 w[K / 8][N]
//     element (n, k) lives in w[k / 8][n], half k % 8.  In the descriptor's terms (cute/arch/mma_sm100_desc.hpp, "Major-K,
//     INTERLEAVE: ((8,n),2):((1,SBO),LBO)" in 16-byte units): 8 consecutive rows of one 16-byte K-chunk are the 128-byte core
//     matrix, SBO = 128 B between 8-row groups, LBO = N * 16 B between K-chunks.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "device_math.cuh"

namespace ngp {
namespace tc {

constexpr int kRows = 128;                // rows per tile = threads per warpgroup
constexpr uint32_t kTmemColsPerTile = 64; // fp32 accumulator columns of one tile

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- shared-memory matrix descriptor (SmemDescriptor, version 1, SWIZZLE_NONE, K-major)
__host__ __device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// ---- instruction descriptor (InstrDescriptor): D = F32, A = B = F16, both K-major, M = 128
__host__ __device__ __forceinline__ constexpr uint32_t instr_desc(uint32_t N) { return (1u << 4) | ((N >> 3) << 17) | ((128u >> 4) << 24); }

__device__ __forceinline__ void mma_f16_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mma_commit(uint32_t mbar_saddr) {   // implies tcgen05.fence::before_thread_sync
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar_saddr) : "memory");
}
// One lane of a converged warp (the lowest active one), as a predicate the compiler understands: a tcgen05.mma under it is issued
// straight, without the per-operand waterfall loop that an `if (lane == 0)` costs.
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void group_sync(uint32_t bar_id) { asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory"); }

__device__ __forceinline__ void mbar_init(uint32_t mbar_saddr, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar_saddr), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t mbar_saddr) {   // release at CTA scope
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbar_saddr) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Development builds (-DNGP_WATCHDOG): a wait that lasts longer than ~2 s raises a device-wide abort flag, reports who waited for
// what (`info` is the caller's tag) and returns; every other wait then returns too, so the kernel ends and the report is flushed.
#ifdef NGP_WATCHDOG
__device__ int g_wd_abort;
#endif
__device__ __forceinline__ bool mbar_try(uint32_t mbar_saddr, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"   // %3: suspend-time hint
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(done)
        : "r"(mbar_saddr), "r"(parity), "r"(0x989680u)
        : "memory");
    return done != 0;
}
// kSleepNs: how long to stay off the issue port between two looks at the barrier.  The hardware's own suspension (the hint above) is
// woken by ANY barrier traffic of the CTA -- with sixteen pipelines per SM that was ~40 wake-ups per wait, 10 % of all instructions
// the kernel executed -- so a waiter that expects a long wait sleeps on the timer instead.
template <uint32_t kSleepNs = 0>
__device__ __forceinline__ void mbar_wait(uint32_t mbar_saddr, uint32_t parity, uint32_t info = 0) {
#ifdef NGP_WATCHDOG
    uint32_t spins = 0;
    const long long t0 = clock64();
#endif
    while (!mbar_try(mbar_saddr, parity)) {
        if (kSleepNs) __nanosleep(kSleepNs);
#ifdef NGP_WATCHDOG
        if ((++spins & 255u) == 0u) {
            if (*reinterpret_cast<volatile int *>(&g_wd_abort)) break;
            if (clock64() - t0 > 4000000000ll) {
                atomicExch(&g_wd_abort, 1);
                printf("WATCHDOG block %d warp %d lane %d: barrier 0x%x parity %u info 0x%x\n", (int)blockIdx.x, (int)(threadIdx.x >> 5),
                       (int)(threadIdx.x & 31), mbar_saddr, parity, info);
                break;
            }
        }
#else
        (void)info;
#endif
    }
}

// ---- tensor memory
__device__ __forceinline__ void tmem_alloc(uint32_t dst_saddr, uint32_t cols) {   // one full warp; cols = power of two >= 32
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_saddr), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {     // the same warp that allocated
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// this thread's TMEM lane, 16 / 4 consecutive fp32 columns starting at taddr's column
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float (&v)[4]) {
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 4; i++) v[i] = __uint_as_float(r[i]);
}

// ---- A operand in tensor memory (the renderer's pipeline): row r of A lives in TMEM lane r, halves (2c, 2c + 1) of the row in
//      32-bit column c -- the layout a row-per-thread tcgen05.st.32x32b writes (cute: tmem_frg_1sm<half, half>, M = 128).
//      One K = 16 MMA consumes 8 columns.
__device__ __forceinline__ void mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// The same with an output-lane mask: D lanes whose bit is set in {m0..m3} (lane 32 j + i = bit i of m_j) are NOT written.  One warp
// can thereby run the MMA for ITS 32 rows only -- the other 96 rows of A are read but their results are discarded -- which makes
// the four lane quadrants of a tile independent of each other: no rendezvous of four warps per layer.
__device__ __forceinline__ void mma_f16_ts_masked(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate,
                                                  uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %6, %7, %8}, p;\n\t"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(m0), "r"(m1), "r"(m2), "r"(m3)
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st2(uint32_t taddr, uint32_t a, uint32_t b) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(taddr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void tmem_st4(uint32_t taddr, const uint4 &v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]),
                 "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
// 16 accumulator columns -> ReLU -> 8 packed fp16 columns of the next A operand (same lane)
__device__ __forceinline__ void relu16_to_tmem(uint32_t t_src, uint32_t t_dst) {
    float v[16];
    tmem_ld16(t_src, v);
    uint32_t p[8];
#pragma unroll
    for (int i = 0; i < 8; i++) p[i] = pack_half2_relu(v[2 * i], v[2 * i + 1]);
    tmem_st8(t_dst, p);
}
// 32 accumulator columns -> ReLU -> 16 packed fp16 columns: half as many TMEM round trips per layer as the 16-column form
__device__ __forceinline__ void relu32_to_tmem(uint32_t t_src, uint32_t t_dst) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, "
        "%22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
          "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
          "=r"(r[31])
        : "r"(t_src)
        : "memory");
    tmem_ld_wait();
    uint32_t p[16];
#pragma unroll
    for (int i = 0; i < 16; i++) p[i] = pack_half2_relu(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(t_dst),
                 "r"(p[0]), "r"(p[1]), "r"(p[2]), "r"(p[3]), "r"(p[4]), "r"(p[5]), "r"(p[6]), "r"(p[7]), "r"(p[8]), "r"(p[9]), "r"(p[10]), "r"(p[11]),
                 "r"(p[12]), "r"(p[13]), "r"(p[14]), "r"(p[15])
                 : "memory");
}
// The K / 16 MMAs of one layer with A in tensor memory, issued by ONE thread.  tmem_d / tmem_a: lane 0, first column.
__device__ __forceinline__ void issue_layer_ts(uint32_t tmem_d, uint32_t tmem_a, uint32_t w_saddr, uint32_t N, uint32_t K) {
    const uint32_t idesc = instr_desc(N);
    for (uint32_t k = 0; k < K / 16; k++) {
        const uint64_t bd = smem_desc(w_saddr + k * 2 * (N * 16), N * 16, 128);
        mma_f16_ts(tmem_d, tmem_a + k * 8, bd, idesc, k > 0 ? 1u : 0u);
    }
}
// ... for the 32 rows of lane quadrant q only (see mma_f16_ts_masked)
__device__ __forceinline__ void issue_layer_ts_quadrant(uint32_t tmem_d, uint32_t tmem_a, uint32_t w_saddr, uint32_t N, uint32_t K, int q) {
    const uint32_t idesc = instr_desc(N);
    const uint32_t m0 = q == 0 ? 0u : 0xffffffffu, m1 = q == 1 ? 0u : 0xffffffffu, m2 = q == 2 ? 0u : 0xffffffffu, m3 = q == 3 ? 0u : 0xffffffffu;
    for (uint32_t k = 0; k < K / 16; k++) {
        const uint64_t bd = smem_desc(w_saddr + k * 2 * (N * 16), N * 16, 128);
        mma_f16_ts_masked(tmem_d, tmem_a + k * 8, bd, idesc, k > 0 ? 1u : 0u, m0, m1, m2, m3);
    }
}
// CTA-scope acq_rel arrival counter in shared memory: returns the value before the increment
__device__ __forceinline__ uint32_t smem_arrive_count(uint32_t saddr) {
    uint32_t old;
    asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(saddr) : "memory");
    return old;
}

// ---- packed weights: per layer uint4 [K / 8][N], layers in evaluation order
struct WeightMap {   // byte offsets into the packed blob
    uint32_t s0, sh, sl, c0, ch, cl, total;
};
__host__ __device__ __forceinline__ WeightMap weight_map(uint32_t nh_sigma, uint32_t nh_color) {
    WeightMap m;
    m.s0 = 0;                          // [32 -> 64]  4 chunks x 64 rows
    m.sh = m.s0 + 4 * 64 * 16;         // nh_sigma x [64 -> 64]  8 x 64
    m.sl = m.sh + nh_sigma * 8 * 64 * 16;   // [64 -> 16]  8 x 16
    m.c0 = m.sl + 8 * 16 * 16;         // [32 -> 64], input columns shifted (see k_pack_network_tc)
    m.ch = m.c0 + 4 * 64 * 16;
    m.cl = m.ch + nh_color * 8 * 64 * 16;
    m.total = m.cl + 8 * 16 * 16;
    return m;
}

// One [N x K] row-major half matrix -> uint4 [K / 8][N].  shift_cols_from16: packed column 16 <- zero, 17..31 <- source
// 16..30 (the colour net's input is [SH16 | sigma-net output 0..15 with column 0 zeroed], network_ff.py:67-68).
__device__ __forceinline__ void pack_weight_chunks(const __half *__restrict__ W, int N, int K, uint4 *__restrict__ dst, int tid, int nthreads,
                                                   bool shift_cols_from16) {
    for (int i = tid; i < (K / 8) * N; i += nthreads) {
        const int c = i / N, n = i % N;
        __half h[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int k = c * 8 + j;
            if (!shift_cols_from16) h[j] = W[n * K + k];
            else h[j] = k < 16 ? W[n * K + k] : (k == 16 ? __float2half(0.f) : W[n * K + k - 1]);
        }
        dst[i] = *reinterpret_cast<const uint4 *>(h);
    }
}

// ---- one warp, its 32 rows of a 128-row tile, activations in tensor memory (the stand-alone network kernels)
// TMEM columns of a tile, relative to its base: accumulators D, hidden activations H (64 halves = 32 columns), and the first
// layer's operand E (the 32-half hash-grid encoding = 16 columns).
constexpr uint32_t kColD = 0, kColH = 64, kColE = 96;
constexpr uint32_t kTmemColsTs = 128;   // power of two >= 112

struct Tile {
    uint32_t w_saddr;     // packed weights (shared-space address of the blob)
    uint32_t mbar_saddr;  // one mbarrier (count 1) owned by this WARP
    uint32_t tmem;        // TMEM address of the tile's first column, lane field = this warp's first lane
    int q;                // this warp's lane quadrant (warp index % 4)
    uint32_t parity;      // phase of the mbarrier
};

// One layer for this warp's 32 rows: D[32 x N] = A[32 x K] (TMEM columns a_col ..) * W[N x K]^T (smem), as an M = 128 MMA whose other
// three lane quadrants are masked out.  All 32 lanes call it after their tcgen05.st of the A rows; on return the accumulators are
// readable with tcgen05.ld.  No other warp is involved.
__device__ __forceinline__ void layer_ts(Tile &t, int lane, uint32_t a_col, uint32_t w_off, uint32_t N, uint32_t K) {
    tmem_st_wait();          // this warp's A-row stores have landed in tensor memory
    fence_before_sync();     // ... and are ordered, with its tcgen05.ld of the previous accumulators, before the hand-over to lane 0
    __syncwarp();
    (void)lane;
    if (elect_one()) {
        fence_after_sync();
        const uint32_t base = t.tmem & 0x0000FFFFu;   // lane 0 of the tile
        issue_layer_ts_quadrant(base + kColD, base + a_col, t.w_saddr + w_off, N, K, t.q);
        mma_commit(t.mbar_saddr);
    }
    mbar_wait(t.mbar_saddr, t.parity);
    t.parity ^= 1u;
    fence_after_sync();
}
__device__ __forceinline__ void relu_row_ts(const Tile &t) {
#pragma unroll
    for (int q = 0; q < 4; q++) relu16_to_tmem(t.tmem + kColD + q * 16, t.tmem + kColH + q * 8);
}
// sigma-net output row (16 accumulators) -> sigma, and the colour net's first operand [SH16 | 0, geo15] in H columns 0..15
__device__ __forceinline__ float sigma_and_color_input(uint32_t t_d, uint32_t t_h, float density_scale, const uint4 &sh_lo, const uint4 &sh_hi) {
    float v[16];
    tmem_ld16(t_d, v);
    uint32_t g0 = pack_half2(v[0], v[1]);
    const float sigma = expf(__half2float(__ushort_as_half((unsigned short)(g0 & 0xffffu)))) * density_scale;
    g0 &= 0xffff0000u;   // column 0 (log-density) is not a colour feature
    tmem_st4(t_h, sh_lo);
    tmem_st4(t_h + 4, sh_hi);
    tmem_st4(t_h + 8, make_uint4(g0, pack_half2(v[2], v[3]), pack_half2(v[4], v[5]), pack_half2(v[6], v[7])));
    tmem_st4(t_h + 12, make_uint4(pack_half2(v[8], v[9]), pack_half2(v[10], v[11]), pack_half2(v[12], v[13]), pack_half2(v[14], v[15])));
    return sigma;
}
__device__ __forceinline__ void rgb_from_acc(uint32_t t_d, float &r, float &g, float &b) {
    float v[4];
    tmem_ld4(t_d, v);
    const float h0 = __half2float(__float2half_rn(v[0])), h1 = __half2float(__float2half_rn(v[1])), h2 = __half2float(__float2half_rn(v[2]));
    r = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h0))));
    g = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h1))));
    b = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h2))));
}

// Both MLPs for the 128 rows of one warpgroup.  Precondition: every thread stored its row's hash-grid encoding (32 halves) to the
// 16 TMEM columns at kColE.  sh_lo / sh_hi: the row's 16 SH halves.  Returns this row's sigma and rgb with the dtype flow of the
// reference under autocast (network_ff.py:56-72): sigma = exp(float(half h0)) * density_scale, rgb = half(sigmoid(float(half h))).
__device__ __forceinline__ void mlp_tile128(Tile &t, int lane, uint32_t nh_sigma, uint32_t nh_color, float density_scale, const uint4 &sh_lo,
                                            const uint4 &sh_hi, float &sigma, float &r, float &g, float &b) {
    const WeightMap wm = weight_map(nh_sigma, nh_color);
    layer_ts(t, lane, kColE, wm.s0, 64, 32);
    relu_row_ts(t);
    for (uint32_t k = 0; k < nh_sigma; k++) {
        layer_ts(t, lane, kColH, wm.sh + k * 8 * 64 * 16, 64, 64);
        relu_row_ts(t);
    }
    layer_ts(t, lane, kColH, wm.sl, 16, 64);
    sigma = sigma_and_color_input(t.tmem + kColD, t.tmem + kColH, density_scale, sh_lo, sh_hi);
    layer_ts(t, lane, kColH, wm.c0, 64, 32);
    relu_row_ts(t);
    for (uint32_t k = 0; k < nh_color; k++) {
        layer_ts(t, lane, kColH, wm.ch + k * 8 * 64 * 16, 64, 64);
        relu_row_ts(t);
    }
    layer_ts(t, lane, kColH, wm.cl, 16, 64);
    rgb_from_acc(t.tmem + kColD, r, g, b);
}

// The density net alone for the 128 rows of one warpgroup (NeRFNetwork.density, network_ff.py:76-89).  Same precondition as
// mlp_tile128.  h[16]: the row's sigma-net outputs rounded to half as the reference's FFMLP stores them (h[0] = log-density,
// h[1..15] = geo_feat).
__device__ __forceinline__ void sigma_tile128(Tile &t, int lane, uint32_t nh_sigma, float (&h)[16]) {
    const WeightMap wm = weight_map(nh_sigma, 0);
    layer_ts(t, lane, kColE, wm.s0, 64, 32);
    relu_row_ts(t);
    for (uint32_t k = 0; k < nh_sigma; k++) {
        layer_ts(t, lane, kColH, wm.sh + k * 8 * 64 * 16, 64, 64);
        relu_row_ts(t);
    }
    layer_ts(t, lane, kColH, wm.sl, 16, 64);
    tmem_ld16(t.tmem + kColD, h);
#pragma unroll
    for (int i = 0; i < 16; i++) h[i] = __half2float(__float2half_rn(h[i]));
}

}  // namespace tc
}  // namespace ngp
