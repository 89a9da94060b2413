// tc_mlp.cuh -- the fully fused 64-wide density + colour MLP on Blackwell's 5th-generation tensor cores (tcgen05).
//
// Replaces, for the fused paths, the mma.sync formulation of ffmlp/src/ffmlp.cu:331-407 (kernel_mlp_fused, wmma 16x16x16):
//   * one 128-row tile per WARPGROUP (4 warps; thread = row = TMEM lane), D accumulators (128 x 64 fp32) in tensor memory;
//   * every layer is `tcgen05.mma.cta_group::1.kind::f16` M=128, N=64|16, K=16 issued by ONE thread, A (activations) and B (weights)
//     read straight from shared memory through matrix descriptors;
//   * between layers each thread pulls its own row back with `tcgen05.ld.32x32b`, applies ReLU + fp16 rounding in registers and
//     stores the row as the next layer's A operand.  No accumulator, A fragment or B fragment ever lives in the register file
//     across the tile, which is what lets the renderer keep twice as many warps resident as the mma.sync version.
//
// Shared-memory operand layout (both A and B), K-major, no swizzle:  uint4 tile[K / 8][rows]
//     element (row r, k) lives in tile[k / 8][r], half k % 8.  In the descriptor's terms (cute/arch/mma_sm100_desc.hpp,
//     "Major-K, INTERLEAVE: ((8,n),2):((1,SBO),LBO)" in 16-byte units): 8 consecutive rows of one 16-byte K-chunk are the 128-byte
//     core matrix, SBO = 128 B between 8-row groups, LBO = rows * 16 B between K-chunks.  A row-per-thread store of one chunk is
//     32 consecutive uint4 per warp: conflict-free.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

#include "device_math.cuh"

namespace ngp {
namespace tc {

constexpr int kRows = 128;                // rows per tile = threads per warpgroup
constexpr uint32_t kTmemColsPerTile = 64; // fp32 accumulator columns of one tile

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- shared-memory matrix descriptor (SmemDescriptor, version 1, SWIZZLE_NONE, K-major)
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// ---- instruction descriptor (InstrDescriptor): D = F32, A = B = F16, both K-major, M = 128
__device__ __forceinline__ constexpr uint32_t instr_desc(uint32_t N) { return (1u << 4) | ((N >> 3) << 17) | ((128u >> 4) << 24); }

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
__device__ __forceinline__ void mbar_wait(uint32_t mbar_saddr, uint32_t parity) {
    uint32_t done = 0;
#ifdef NGP_SPIN_TRAP
    uint32_t spins = 0;
#endif
    do {
#ifdef NGP_SPIN_TRAP   // development builds: a tensor-core commit that never arrives traps instead of hanging the device
        if (++spins == (1u << 28)) __trap();
#endif
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n"
            : "=r"(done)
            : "r"(mbar_saddr), "r"(parity)
            : "memory");
    } while (!done);
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

// ---- per-warpgroup tile resources
struct Tile {
    uint4 *a;             // A operand: uint4 [8][128] (16 KB), generic pointer
    uint32_t a_saddr;     // its shared-space address
    uint32_t w_saddr;     // packed weights (shared-space address of the blob)
    uint32_t mbar_saddr;  // one mbarrier (count 1) owned by this warpgroup
    uint32_t tmem;        // TMEM address of this tile's 64 accumulator columns, lane field = this WARP's first lane
    uint32_t bar_id;      // named barrier of this warpgroup (1..15)
    uint32_t parity;      // phase of the mbarrier
    // pipelined variant only (mlp_tile128_pipelined)
    uint4 *e;             // encoding staging: uint4 [4][128] (8 KB), the first layer's A operand
    uint32_t e_saddr;
    uint32_t aready_saddr;   // mbarrier (count 4): every warp of the group has stored its rows of the next A operand
    uint32_t parity_a;
};

// One layer: D[128 x N] (TMEM) = A[128 x K] (smem) * W[N x K]^T (smem).  All 128 threads of the warpgroup call it after having
// written their A rows with ordinary stores; on return the accumulators are readable with tcgen05.ld.
__device__ __forceinline__ void layer_mma(Tile &t, int row, uint32_t w_off, uint32_t N, uint32_t K) {
    fence_async_smem();      // this thread's A-row stores -> visible to the tensor core's (async proxy) reads
    fence_before_sync();     // this thread's tcgen05.ld of the previous accumulators are ordered before the barrier
    group_sync(t.bar_id);
    if (row == 0) {
        fence_after_sync();
        const uint32_t idesc = instr_desc(N);
        const uint32_t tmem_d = t.tmem & 0x0000FFFFu;   // lane 0 of the tile's columns
        for (uint32_t k = 0; k < K / 16; k++) {
            const uint64_t ad = smem_desc(t.a_saddr + k * 2 * (kRows * 16), kRows * 16, 128);
            const uint64_t bd = smem_desc(t.w_saddr + w_off + k * 2 * (N * 16), N * 16, 128);
            mma_f16_ss(tmem_d, ad, bd, idesc, k > 0 ? 1u : 0u);
        }
        mma_commit(t.mbar_saddr);
        mbar_wait(t.mbar_saddr, t.parity);   // one poller per warpgroup; the other 127 threads sleep in the hardware barrier below
    }
    t.parity ^= 1u;
    group_sync(t.bar_id);
    fence_after_sync();
}

// ReLU + fp16 rounding of this thread's 64 accumulators -> its row of the next A operand (8 chunks).
__device__ __forceinline__ void relu_row_to_a(const Tile &t, int row) {
#pragma unroll
    for (int q = 0; q < 4; q++) {
        float v[16];
        tmem_ld16(t.tmem + q * 16, v);
        t.a[(2 * q) * kRows + row] = make_uint4(pack_half2_relu(v[0], v[1]), pack_half2_relu(v[2], v[3]), pack_half2_relu(v[4], v[5]), pack_half2_relu(v[6], v[7]));
        t.a[(2 * q + 1) * kRows + row] =
            make_uint4(pack_half2_relu(v[8], v[9]), pack_half2_relu(v[10], v[11]), pack_half2_relu(v[12], v[13]), pack_half2_relu(v[14], v[15]));
    }
}

// Both MLPs for the 128 rows of one warpgroup.  Precondition: every thread stored the 4 chunks of its row's hash-grid encoding
// (32 halves) to t.a[c * 128 + row], c = 0..3.  sh_lo / sh_hi: the row's 16 SH halves.  Returns this row's sigma and rgb with the
// dtype flow of the reference under autocast (network_ff.py:56-72): sigma = exp(float(half h0)) * density_scale,
// rgb = half(sigmoid(float(half h))).
__device__ __forceinline__ void mlp_tile128(Tile &t, int row, uint32_t nh_sigma, uint32_t nh_color, float density_scale, const uint4 &sh_lo,
                                            const uint4 &sh_hi, float &sigma, float &r, float &g, float &b) {
    const WeightMap wm = weight_map(nh_sigma, nh_color);
    layer_mma(t, row, wm.s0, 64, 32);
    relu_row_to_a(t, row);
    for (uint32_t k = 0; k < nh_sigma; k++) {
        layer_mma(t, row, wm.sh + k * 8 * 64 * 16, 64, 64);
        relu_row_to_a(t, row);
    }
    layer_mma(t, row, wm.sl, 16, 64);
    {
        float v[16];
        tmem_ld16(t.tmem, v);
        uint32_t g0 = pack_half2(v[0], v[1]);
        sigma = expf(__half2float(__ushort_as_half((unsigned short)(g0 & 0xffffu)))) * density_scale;
        g0 &= 0xffff0000u;   // column 0 (log-density) is not a colour feature
        t.a[0 * kRows + row] = sh_lo;
        t.a[1 * kRows + row] = sh_hi;
        t.a[2 * kRows + row] = make_uint4(g0, pack_half2(v[2], v[3]), pack_half2(v[4], v[5]), pack_half2(v[6], v[7]));
        t.a[3 * kRows + row] = make_uint4(pack_half2(v[8], v[9]), pack_half2(v[10], v[11]), pack_half2(v[12], v[13]), pack_half2(v[14], v[15]));
    }
    layer_mma(t, row, wm.c0, 64, 32);
    relu_row_to_a(t, row);
    for (uint32_t k = 0; k < nh_color; k++) {
        layer_mma(t, row, wm.ch + k * 8 * 64 * 16, 64, 64);
        relu_row_to_a(t, row);
    }
    layer_mma(t, row, wm.cl, 16, 64);
    {
        float v[4];
        tmem_ld4(t.tmem, v);
        const float h0 = __half2float(__float2half_rn(v[0])), h1 = __half2float(__float2half_rn(v[1])), h2 = __half2float(__float2half_rn(v[2]));
        r = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h0))));
        g = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h1))));
        b = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h2))));
    }
}

// The density net alone for the 128 rows of one warpgroup (NeRFNetwork.density, network_ff.py:76-89).  Same precondition as
// mlp_tile128.  h[16]: the row's sigma-net outputs rounded to half as the reference's FFMLP stores them (h[0] = log-density,
// h[1..15] = geo_feat).
__device__ __forceinline__ void sigma_tile128(Tile &t, int row, uint32_t nh_sigma, float (&h)[16]) {
    const WeightMap wm = weight_map(nh_sigma, 0);
    layer_mma(t, row, wm.s0, 64, 32);
    relu_row_to_a(t, row);
    for (uint32_t k = 0; k < nh_sigma; k++) {
        layer_mma(t, row, wm.sh + k * 8 * 64 * 16, 64, 64);
        relu_row_to_a(t, row);
    }
    layer_mma(t, row, wm.sl, 16, 64);
    tmem_ld16(t.tmem, h);
#pragma unroll
    for (int i = 0; i < 16; i++) h[i] = __half2float(__float2half_rn(h[i]));
}

// ---- pipelined variant: issue and completion of a layer are split, so that the caller can do independent work (the next pass's
//      gather) while the tensor core runs and while the other warps of the group catch up.  No thread waits at a barrier:
//        every warp     : stores its rows, fence.proxy.async, then ONE arrive on `aready` (count 4)
//        thread 0       : waits for `aready`, issues the MMAs, commits to `mbar`
//        every thread   : ... independent work ..., then waits for `mbar` (normally already complete)
__device__ __forceinline__ void layer_issue(Tile &t, int row, int lane, uint32_t a_saddr, uint32_t w_off, uint32_t N, uint32_t K) {
    fence_async_smem();      // this thread's A-row stores -> visible to the tensor core's (async proxy) reads
    fence_before_sync();     // this thread's tcgen05.ld of the previous accumulators are ordered before the arrive
    __syncwarp();
#ifdef NGP_LAST_ARRIVER
    // the LAST of the four warps to arrive issues the layer: nobody waits for the operand, and the issue happens the moment the
    // operand is complete.  `aready` is a plain 32-bit arrival counter here (acq_rel at CTA scope: the other warps' stores and
    // proxy fences happen-before the issuer's MMAs).
    (void)row;
    if (lane == 0) {
        uint32_t old;
        asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(t.aready_saddr) : "memory");
        if ((old & 3u) == 3u) {
            fence_after_sync();
            const uint32_t idesc = instr_desc(N);
            const uint32_t tmem_d = t.tmem & 0x0000FFFFu;
            for (uint32_t k = 0; k < K / 16; k++) {
                const uint64_t ad = smem_desc(a_saddr + k * 2 * (kRows * 16), kRows * 16, 128);
                const uint64_t bd = smem_desc(t.w_saddr + w_off + k * 2 * (N * 16), N * 16, 128);
                mma_f16_ss(tmem_d, ad, bd, idesc, k > 0 ? 1u : 0u);
            }
            mma_commit(t.mbar_saddr);
        }
    }
#else
    if (lane == 0) mbar_arrive(t.aready_saddr);
    if (row == 0) {
        mbar_wait(t.aready_saddr, t.parity_a);
        fence_after_sync();
        const uint32_t idesc = instr_desc(N);
        const uint32_t tmem_d = t.tmem & 0x0000FFFFu;
        for (uint32_t k = 0; k < K / 16; k++) {
            const uint64_t ad = smem_desc(a_saddr + k * 2 * (kRows * 16), kRows * 16, 128);
            const uint64_t bd = smem_desc(t.w_saddr + w_off + k * 2 * (N * 16), N * 16, 128);
            mma_f16_ss(tmem_d, ad, bd, idesc, k > 0 ? 1u : 0u);
        }
        mma_commit(t.mbar_saddr);
    }
    t.parity_a ^= 1u;
#endif
    __syncwarp();
}
__device__ __forceinline__ void layer_wait(Tile &t) {
    mbar_wait(t.mbar_saddr, t.parity);
    t.parity ^= 1u;
    fence_after_sync();
}

// As mlp_tile128, with the first layer's A operand in t.e and `between(step)` called after the MMAs of layer `step` (0-based, in
// evaluation order) have been handed to the tensor core.  `between` must not touch t.a, and may write t.e only for step >= 1.
template <class Between>
__device__ __forceinline__ void mlp_tile128_pipelined(Tile &t, int row, int lane, uint32_t nh_sigma, uint32_t nh_color, float density_scale,
                                                      const uint4 &sh_lo, const uint4 &sh_hi, float &sigma, float &r, float &g, float &b,
                                                      Between &&between) {
    const WeightMap wm = weight_map(nh_sigma, nh_color);
    const int n_sig = 2 + (int)nh_sigma, n_all = n_sig + 2 + (int)nh_color;
    // one rolled loop over the layers of both nets: whatever `between` does (a gather trip) exists ONCE in the instruction stream
#pragma unroll 1
    for (int step = 0; step < n_all; step++) {
        const bool in_color = step >= n_sig;
        const int l = in_color ? step - n_sig : step;                           // layer within its net
        const bool first = l == 0, last = l == (int)(in_color ? nh_color : nh_sigma) + 1;
        const uint32_t w_off = (in_color ? wm.c0 : wm.s0) + (first ? 0u : 4096u + (uint32_t)(l - 1) * 8192u);   // [32->64] is 4 KB, [64->64] 8 KB
        layer_issue(t, row, lane, step == 0 ? t.e_saddr : t.a_saddr, w_off, last ? 16u : 64u, first ? 32u : 64u);
        between(step);
        layer_wait(t);
        if (!last) {
            relu_row_to_a(t, row);
        } else if (!in_color) {
            float v[16];
            tmem_ld16(t.tmem, v);
            uint32_t g0 = pack_half2(v[0], v[1]);
            sigma = expf(__half2float(__ushort_as_half((unsigned short)(g0 & 0xffffu)))) * density_scale;
            g0 &= 0xffff0000u;   // column 0 (log-density) is not a colour feature
            t.a[0 * kRows + row] = sh_lo;
            t.a[1 * kRows + row] = sh_hi;
            t.a[2 * kRows + row] = make_uint4(g0, pack_half2(v[2], v[3]), pack_half2(v[4], v[5]), pack_half2(v[6], v[7]));
            t.a[3 * kRows + row] = make_uint4(pack_half2(v[8], v[9]), pack_half2(v[10], v[11]), pack_half2(v[12], v[13]), pack_half2(v[14], v[15]));
        } else {
            float v[4];
            tmem_ld4(t.tmem, v);
            const float h0 = __half2float(__float2half_rn(v[0])), h1 = __half2float(__float2half_rn(v[1])), h2 = __half2float(__float2half_rn(v[2]));
            r = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h0))));
            g = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h1))));
            b = __half2float(__float2half_rn(__fdividef(1.0f, 1.0f + __expf(-h2))));
        }
    }
}

}  // namespace tc
}  // namespace ngp
