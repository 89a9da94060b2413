// fused.cu -- the B200 fast path (no upstream counterpart):
//   ngp_network_forward(_tc) : hash-grid encode -> sigma MLP -> SH -> colour MLP -> exp/sigmoid, one kernel, nothing but
//                         xyz/dir in and sigma/rgb out touches HBM        (nerf/network_ff.py:51-74); _tc = tcgen05 tiles
//   ngp_density_forward : the density half of it (nerf/network_ff.py:76-89)
//   ngp_render          : NeRFRenderer.run_cuda's whole inference loop (nerf/renderer.py:261-284,329-386) as ONE
//                         persistent cooperative launch, one CTA of 5 warpgroups per SM.  The reference's global schedule
//                         `n_step = clamp(N // n_alive, 1, 8)` decides where every ray's chunks end and is reproduced
//                         exactly, but in speculative EPOCHS of many reference iterations (two grid barriers per epoch):
//                         a lane-dynamic lockstep MARCH phase fills a sample buffer, an EVALUATE + COMPOSITE phase walks
//                         it in 128-row warpgroup tiles (hash-grid gather -> both MLPs on tcgen05 with TMEM accumulators
//                         -> compositing in registers).  See the comment above k_render and DESIGN.md section 4.1.
#include <cooperative_groups.h>
#include <float.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <initializer_list>

#include "common.cuh"
#include "device_math.cuh"
#include "scan.cuh"
#include "tc_mlp.cuh"

namespace cg = cooperative_groups;

namespace ngp {

constexpr int kTile = 128;        // sample rows per tile == threads per CTA
constexpr int kEncStride = 40;    // halves per staged encoding row (32 + 8 pad: conflict-free fragment loads)
constexpr int kShStride = 24;     // halves per staged SH row (16 + 8 pad)

struct NetConst {
    const __half2 *table;
    const uint2 *wpack;            // global, fragment order: sigma net then colour net
    uint32_t wpack_count;          // uint2 elements
    const uint4 *wtc;              // global, tcgen05 operand order (tc_mlp.cuh), sigma net then colour net
    uint32_t wtc_bytes;
    uint32_t nh_sigma, nh_color;   // hidden (64x64) matmuls per net = num_layers - 1
    float bound, two_bound, density_scale;
    LevelConst lv[kMaxLevels];
    const __half2 *lvbase[kMaxLevels];   // table + lv[l].offset: one 64-bit base per level, 32-bit row offsets in the gather
    // The renderer's MMAs, layer by layer in evaluation order (sigma net, then colour net), as kernel PARAMETERS: the issuing thread reads
    // them from the constant bank straight into uniform registers.  ldesc_b = shared-memory descriptor of the layer's first weight
    // K-chunk with the address field relative to the weight blob (the kernel adds its blob address); chunk k is + k * bstep.
    uint64_t ldesc_b[10];
    uint32_t ldesc_i[10];          // instruction descriptor (N = 64 | 16)
    uint32_t ldesc_nk_bstep[10];   // (K / 16) << 16 | bstep
};

// ---------------------------------------------------------------------------------------------- device pieces

__device__ __forceinline__ void copy_wpack_to_smem(const NetConst &nc, uint2 *s_w) {
    const uint4 *src = reinterpret_cast<const uint4 *>(nc.wpack);
    uint4 *dst = reinterpret_cast<uint4 *>(s_w);
    for (uint32_t i = threadIdx.x; i < nc.wpack_count / 2; i += blockDim.x) dst[i] = __ldg(src + i);
}

// grid.py:144  inputs = (inputs + bound) / (2 * bound); returns false when the point lies outside [0,1]^3 (gridencoder.cu:99-123:
// the encoding is all zeros then)
__device__ __forceinline__ bool encode_unit(const NetConst &nc, float x, float y, float z, float &u0, float &u1, float &u2) {
    u0 = __fdiv_rn(__fadd_rn(x, nc.bound), nc.two_bound);
    u1 = __fdiv_rn(__fadd_rn(y, nc.bound), nc.two_bound);
    u2 = __fdiv_rn(__fadd_rn(z, nc.bound), nc.two_bound);
    return !((u0 < 0 || u0 > 1) || (u1 < 0 || u1 > 1) || (u2 < 0 || u2 > 1));
}

// One trip of the gather: levels LPT trip .. LPT trip + LPT - 1 of the point (u0, u1, u2) -> LPT half2 (8 LPT independent 4-byte
// loads in flight).  LPT = 4 fills one 16-byte operand chunk per trip; LPT = 2 halves the registers a trip holds (builds with
// more than 16 warps per SM).
#ifndef NGP_LEVELS_PER_TRIP
#define NGP_LEVELS_PER_TRIP 2
#endif
constexpr int kRenderLevelsPerTrip = NGP_LEVELS_PER_TRIP;   // the renderer: 20 warps per SM at 96 registers (A/B in DESIGN.md section 8)

template <int LPT, bool kGeneral = true>
__device__ __forceinline__ void encode_trip(const NetConst &nc, float u0, float u1, float u2, int trip, uint32_t (&packed)[LPT]) {
    float f[LPT][3];
    __half2 v[LPT][8];
#pragma unroll
    for (int j = 0; j < LPT; j++) {
        const LevelConst &lc = nc.lv[trip * LPT + j];
        const float scale = lc.scale;
        float p0 = __fmaf_rn(u0, scale, 0.5f), p1 = __fmaf_rn(u1, scale, 0.5f), p2 = __fmaf_rn(u2, scale, 0.5f);
        const uint32_t g0 = (uint32_t)floorf(p0), g1 = (uint32_t)floorf(p1), g2 = (uint32_t)floorf(p2);
        f[j][0] = __fsub_rn(p0, (float)g0); f[j][1] = __fsub_rn(p1, (float)g1); f[j][2] = __fsub_rn(p2, (float)g2);
        const __half2 *tl = nc.lvbase[trip * LPT + j];
        uint32_t row[8];
        cell_rows<kGeneral>(lc, g0, g1, g2, row);
#pragma unroll
        for (int k = 0; k < 8; k++) v[j][k] = __ldg(tl + row[k]);
    }
#ifndef NGP_FP32_BLEND   // default: trilinear blend in packed fp16 (both channels of a corner in one HFMA2; intermediate results rounded
                        // to half like the reference's own accumulator, gridencoder.cu:146-171) instead of fp32 FFMA2 after 16
                        // conversions per level: -2 % frame time, and the image's distance from the reference extensions is unchanged
                        // (1.4e-4, set by the fp32- vs fp16-accumulated MLPs).  Written as the lerp tree x -> y -> z (14 packed
                        // operations per level, no per-corner weight products or broadcasts).
#pragma unroll
    for (int j = 0; j < LPT; j++) {
        const __half2 fx = __float2half2_rn(f[j][0]), fy = __float2half2_rn(f[j][1]), fz = __float2half2_rn(f[j][2]);
        const __half2 a0 = __hfma2(fx, __hsub2(v[j][1], v[j][0]), v[j][0]), a1 = __hfma2(fx, __hsub2(v[j][3], v[j][2]), v[j][2]);
        const __half2 a2 = __hfma2(fx, __hsub2(v[j][5], v[j][4]), v[j][4]), a3 = __hfma2(fx, __hsub2(v[j][7], v[j][6]), v[j][6]);
        const __half2 b0 = __hfma2(fy, __hsub2(a1, a0), a0), b1 = __hfma2(fy, __hsub2(a3, a2), a2);
        const __half2 c = __hfma2(fz, __hsub2(b1, b0), b0);
        packed[j] = *reinterpret_cast<const uint32_t *>(&c);
    }
    return;
#endif
#pragma unroll
    for (int j = 0; j < LPT; j++) {
        const float p0 = f[j][0], p1 = f[j][1], p2 = f[j][2];
        const float q0 = 1.0f - p0, q1 = 1.0f - p1, q2 = 1.0f - p2;
        // (w_x * w_y) * w_z, gridencoder.cu:148-160, as packed pairs: (corner 2m, corner 2m + 1) differ in x only
        const float2 wx = make_float2(q0, p0);
        const float2 w0 = fmul2(wx, make_float2(q1, q1)), w1 = fmul2(wx, make_float2(p1, p1));
        const float2 w[4] = {fmul2(w0, make_float2(q2, q2)), fmul2(w1, make_float2(q2, q2)), fmul2(w0, make_float2(p2, p2)),
                             fmul2(w1, make_float2(p2, p2))};
        float2 acc = make_float2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < 8; k++) {   // the reference's corner order; both channels in one FFMA2
            const float wk = (k & 1) ? w[k >> 1].y : w[k >> 1].x;
            acc = ffma2(make_float2(wk, wk), __half22float2(v[j][k]), acc);
        }
        packed[j] = pack_half2(acc.x, acc.y);
    }
}

// 16-level hash-grid encoding of one point into 16 half2: dst[g * stride], g = 0..3, receives levels 4g .. 4g+3 (8 halves).
template <int kLevelsPerTrip = 4, bool kGeneral = true>
__device__ __forceinline__ void encode_point(const NetConst &nc, float x, float y, float z, uint4 *dst, int stride = 1) {
    static_assert(kLevelsPerTrip == 4 || kLevelsPerTrip == 2 || kLevelsPerTrip == 1, "a trip covers a whole, a half or a quarter of a 16-byte operand chunk");
    float u0, u1, u2;
    if (!encode_unit(nc, x, y, z, u0, u1, u2)) {
        dst[0] = dst[stride] = dst[2 * stride] = dst[3 * stride] = make_uint4(0, 0, 0, 0);
        return;
    }
    // the trip loop is deliberately NOT unrolled: a small gather body keeps the renderer's warps, which sit in different
    // phases, inside the instruction cache
#pragma unroll 1
    for (int trip = 0; trip < 16 / kLevelsPerTrip; trip++) {
        uint32_t packed[kLevelsPerTrip];
        encode_trip<kLevelsPerTrip, kGeneral>(nc, u0, u1, u2, trip, packed);
        if (kLevelsPerTrip == 4) dst[trip * stride] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
        else if (kLevelsPerTrip == 2) reinterpret_cast<uint2 *>(dst + (trip >> 1) * stride)[trip & 1] = make_uint2(packed[0], packed[kLevelsPerTrip - 1]);
        else reinterpret_cast<uint32_t *>(dst + (trip >> 2) * stride)[trip & 3] = packed[0];
    }
}

// 16-level hash-grid encoding of one point -> the 16 TMEM columns at taddr (this thread's lane).  tcgen05.st is warp-collective:
// every lane calls this; lanes with !active store zeros without gathering (the reference's encoding of a point outside the unit
// cube is zero as well, gridencoder.cu:99-123).
template <int LPT, bool kGeneral = true>
__device__ __forceinline__ void encode_point_tmem(const NetConst &nc, float x, float y, float z, bool active, uint32_t taddr) {
    static_assert(LPT == 4 || LPT == 2, "a trip stores 4 or 2 columns");
    float u0 = 0.f, u1 = 0.f, u2 = 0.f;
    const bool inside = active && encode_unit(nc, x, y, z, u0, u1, u2);
    // the trip loop is deliberately NOT unrolled: a small gather body keeps the warps of both roles inside the instruction cache
#pragma unroll 1
    for (int trip = 0; trip < 16 / LPT; trip++) {
        uint32_t packed[LPT];
#pragma unroll
        for (int j = 0; j < LPT; j++) packed[j] = 0u;
        if (inside) encode_trip<LPT, kGeneral>(nc, u0, u1, u2, trip, packed);
        __syncwarp();
        if (LPT == 4) tc::tmem_st4(taddr + (uint32_t)trip * 4u, make_uint4(packed[0], packed[1], packed[2], packed[LPT - 1]));
        else tc::tmem_st2(taddr + (uint32_t)trip * 2u, packed[0], packed[1]);
    }
}

// degree-4 SH of a direction -> 16 halves (2 x uint4)
__device__ __forceinline__ void sh_to_half(float dx, float dy, float dz, uint4 &lo, uint4 &hi) {
    float o[16];
    sh_basis(dx, dy, dz, 4, o);
    lo = make_uint4(pack_half2(o[0], o[1]), pack_half2(o[2], o[3]), pack_half2(o[4], o[5]), pack_half2(o[6], o[7]));
    hi = make_uint4(pack_half2(o[8], o[9]), pack_half2(o[10], o[11]), pack_half2(o[12], o[13]), pack_half2(o[14], o[15]));
}

__device__ __forceinline__ float round_half(float v) { return __half2float(__float2half_rn(v)); }

// Both MLPs for the 32 rows [row0, row0+32) owned by this warp.
//   s_enc / s_sh : staged fp16 inputs.   emit(row, sigma, r, g, b) is called by the lane that owns the value pair.
// dtype flow as in the reference under autocast: sigma_net output is half; sigma = exp(float(h0)) * density_scale;
// colour input = [SH16 (half), geo15 (half), 0]; rgb = half(sigmoid(float(half(h))))   (network_ff.py:56-72).
//   sh_row(r)    : staged SH row (16 halves) of sample row r -- per row in the network kernel, per RAY in the renderer.
template <class ShRow, class EmitSigma, class EmitRgb>
__device__ __forceinline__ void mlp_rows32(const NetConst &nc, const uint2 *__restrict__ s_w, const __half *s_enc, ShRow &&sh_row,
                                           int row0, int lane, EmitSigma &&emit_sigma, EmitRgb &&emit_rgb) {
    const int g = lane >> 2, t = lane & 3;
    const uint2 *w_s0 = s_w;                                   // [64][32]
    const uint2 *w_sh = w_s0 + 512;                            // nh_sigma x [64][64]
    const uint2 *w_sl = w_sh + (size_t)nc.nh_sigma * 1024;     // [16][64]
    const uint2 *w_c0 = w_sl + 256;                            // [64][32] (columns shifted, see pack kernel)
    const uint2 *w_ch = w_c0 + 512;
    const uint2 *w_cl = w_ch + (size_t)nc.nh_color * 1024;

    uint32_t A0[2][2][4];
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        const uint32_t *lo = reinterpret_cast<const uint32_t *>(s_enc + (row0 + mt * 16 + g) * kEncStride);
        const uint32_t *hi = reinterpret_cast<const uint32_t *>(s_enc + (row0 + mt * 16 + g + 8) * kEncStride);
#pragma unroll
        for (int kt = 0; kt < 2; kt++) {
            A0[mt][kt][0] = lo[kt * 8 + t]; A0[mt][kt][1] = hi[kt * 8 + t];
            A0[mt][kt][2] = lo[kt * 8 + 4 + t]; A0[mt][kt][3] = hi[kt * 8 + 4 + t];
        }
    }
    float C[2][8][4];
    uint32_t A[2][4][4];
    mlp_layer<2, 2, 8>(A0, w_s0, lane, C);
    relu_to_frag<2, 8>(C, A);
    for (uint32_t k = 0; k < nc.nh_sigma; k++) {
        mlp_layer<2, 4, 8>(A, w_sh + (size_t)k * 1024, lane, C);
        relu_to_frag<2, 8>(C, A);
    }
    float O[2][2][4];
    mlp_layer<2, 4, 2>(A, w_sl, lane, O);

    // colour-net input: k-tile 0 = SH (staged), k-tile 1 = sigma_net output columns 0..15 with column 0 (log-density)
    // forced to zero; the packed first-layer weights carry the matching column shift.
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        const uint32_t *lo = reinterpret_cast<const uint32_t *>(sh_row(row0 + mt * 16 + g));
        const uint32_t *hi = reinterpret_cast<const uint32_t *>(sh_row(row0 + mt * 16 + g + 8));
        A0[mt][0][0] = lo[t]; A0[mt][0][1] = hi[t]; A0[mt][0][2] = lo[4 + t]; A0[mt][0][3] = hi[4 + t];
        uint32_t g0 = pack_half2(O[mt][0][0], O[mt][0][1]), g1 = pack_half2(O[mt][0][2], O[mt][0][3]);
        if (t == 0) {
            emit_sigma(row0 + mt * 16 + g, expf(__half2float(__ushort_as_half((unsigned short)(g0 & 0xffffu)))) * nc.density_scale);
            emit_sigma(row0 + mt * 16 + g + 8, expf(__half2float(__ushort_as_half((unsigned short)(g1 & 0xffffu)))) * nc.density_scale);
            g0 &= 0xffff0000u;
            g1 &= 0xffff0000u;
        }
        A0[mt][1][0] = g0; A0[mt][1][1] = g1;
        A0[mt][1][2] = pack_half2(O[mt][1][0], O[mt][1][1]); A0[mt][1][3] = pack_half2(O[mt][1][2], O[mt][1][3]);
    }
    mlp_layer<2, 2, 8>(A0, w_c0, lane, C);
    relu_to_frag<2, 8>(C, A);
    for (uint32_t k = 0; k < nc.nh_color; k++) {
        mlp_layer<2, 4, 8>(A, w_ch + (size_t)k * 1024, lane, C);
        relu_to_frag<2, 8>(C, A);
    }
    float Oc[2][1][4];
    mlp_layer<2, 4, 1>(A, w_cl, lane, Oc);   // only output columns 0..7 are needed (rgb = 0..2)
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        if (t < 2) {
            // t == 0 holds columns 0,1 (r,g); t == 1 holds column 2 (b) in element 0
            const float v0 = round_half(__fdividef(1.0f, 1.0f + __expf(-round_half(Oc[mt][0][0]))));
            const float v1 = round_half(__fdividef(1.0f, 1.0f + __expf(-round_half(Oc[mt][0][1]))));
            const float v2 = round_half(__fdividef(1.0f, 1.0f + __expf(-round_half(Oc[mt][0][2]))));
            const float v3 = round_half(__fdividef(1.0f, 1.0f + __expf(-round_half(Oc[mt][0][3]))));
            emit_rgb(row0 + mt * 16 + g, t, v0, v1);
            emit_rgb(row0 + mt * 16 + g + 8, t, v2, v3);
        }
    }
}

#include "diff_render.cuh"   // differentiable small-batch render + density backward (same building blocks)

// ---------------------------------------------------------------------------------------------- coarse occupancy
// bit j of the result <- any of the 64 cells with Morton index 64 j .. 64 j + 63 (an aligned 4x4x4 block) is occupied.
__global__ void k_coarse_occupancy(const unsigned long long *__restrict__ bitfield64, uint32_t n_blocks, uint32_t *__restrict__ coarse) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    const bool any = j < n_blocks && __ldg(bitfield64 + j) != 0ull;
    const uint32_t word = __ballot_sync(0xffffffffu, any);
    if ((threadIdx.x & 31) == 0 && j < n_blocks) coarse[j >> 5] = word;
}

// ---------------------------------------------------------------------------------------------- weight packing
// sigma: [64][32], nh x [64][64], [16][64]; colour: [64][32]*, nh x [64][64], [16][64]   (* input columns shifted:
// packed column 16 <- zero, packed columns 17..31 <- source columns 16..30; source column 31 multiplies the zero pad).
__global__ void k_pack_network(const __half *__restrict__ w_sigma, uint32_t nh_sigma, const __half *__restrict__ w_color,
                               uint32_t nh_color, uint2 *__restrict__ dst) {
    const int tid = threadIdx.x, nt = blockDim.x;
    uint2 *p = dst;
    pack_weight_frags(w_sigma, 64, 32, p, tid, nt, false); p += 512;
    for (uint32_t k = 0; k < nh_sigma; k++) { pack_weight_frags(w_sigma + 2048 + (size_t)k * 4096, 64, 64, p, tid, nt, false); p += 1024; }
    pack_weight_frags(w_sigma + 2048 + (size_t)nh_sigma * 4096, 16, 64, p, tid, nt, false); p += 256;
    pack_weight_frags(w_color, 64, 32, p, tid, nt, true); p += 512;
    for (uint32_t k = 0; k < nh_color; k++) { pack_weight_frags(w_color + 2048 + (size_t)k * 4096, 64, 64, p, tid, nt, false); p += 1024; }
    pack_weight_frags(w_color + 2048 + (size_t)nh_color * 4096, 16, 64, p, tid, nt, false);
}

// ---------------------------------------------------------------------------------------------- network forward
__global__ void __launch_bounds__(kTile, 4)
k_network_forward(NetConst nc, const float *__restrict__ xyzs, const float *__restrict__ dirs, uint32_t M, float *__restrict__ sigmas,
                  float *__restrict__ rgbs, uint32_t n_tiles) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint2 *s_w = reinterpret_cast<uint2 *>(smem_raw);
    __half *s_enc = reinterpret_cast<__half *>(s_w + nc.wpack_count);
    __half *s_sh = s_enc + kTile * kEncStride;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    copy_wpack_to_smem(nc, s_w);
    __syncthreads();
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint32_t row = tile * kTile + tid;
        uint4 *erow = reinterpret_cast<uint4 *>(s_enc + tid * kEncStride);
        uint4 *srow = reinterpret_cast<uint4 *>(s_sh + tid * kShStride);
        if (row < M) {
            encode_point(nc, __ldg(xyzs + (size_t)row * 3), __ldg(xyzs + (size_t)row * 3 + 1), __ldg(xyzs + (size_t)row * 3 + 2), erow);
            sh_to_half(__ldg(dirs + (size_t)row * 3), __ldg(dirs + (size_t)row * 3 + 1), __ldg(dirs + (size_t)row * 3 + 2), srow[0], srow[1]);
        } else {
            erow[0] = erow[1] = erow[2] = erow[3] = make_uint4(0, 0, 0, 0);
            srow[0] = srow[1] = make_uint4(0, 0, 0, 0);
        }
        __syncwarp();
        const uint32_t base = tile * kTile;
        mlp_rows32(nc, s_w, s_enc, [&](int r) { return s_sh + r * kShStride; }, warp * 32, lane,
                   [&](int r, float s) { if (base + r < M) sigmas[base + r] = s; },
                   [&](int r, int t, float a, float b) {
                       if (base + r < M) {
                           float *o = rgbs + (size_t)(base + r) * 3;
                           if (t == 0) { o[0] = a; o[1] = b; } else { o[2] = a; }
                       }
                   });
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------- tcgen05 variants
__global__ void k_pack_network_tc(const __half *__restrict__ w_sigma, uint32_t nh_sigma, const __half *__restrict__ w_color,
                                  uint32_t nh_color, uint4 *__restrict__ dst) {
    const int tid = threadIdx.x, nt = blockDim.x;
    const tc::WeightMap wm = tc::weight_map(nh_sigma, nh_color);
    unsigned char *base = reinterpret_cast<unsigned char *>(dst);
    tc::pack_weight_chunks(w_sigma, 64, 32, reinterpret_cast<uint4 *>(base + wm.s0), tid, nt, false);
    for (uint32_t k = 0; k < nh_sigma; k++)
        tc::pack_weight_chunks(w_sigma + 2048 + (size_t)k * 4096, 64, 64, reinterpret_cast<uint4 *>(base + wm.sh + k * 8192), tid, nt, false);
    tc::pack_weight_chunks(w_sigma + 2048 + (size_t)nh_sigma * 4096, 16, 64, reinterpret_cast<uint4 *>(base + wm.sl), tid, nt, false);
    tc::pack_weight_chunks(w_color, 64, 32, reinterpret_cast<uint4 *>(base + wm.c0), tid, nt, true);
    for (uint32_t k = 0; k < nh_color; k++)
        tc::pack_weight_chunks(w_color + 2048 + (size_t)k * 4096, 64, 64, reinterpret_cast<uint4 *>(base + wm.ch + k * 8192), tid, nt, false);
    tc::pack_weight_chunks(w_color + 2048 + (size_t)nh_color * 4096, 16, 64, reinterpret_cast<uint4 *>(base + wm.cl), tid, nt, false);
}

// network_ff.py:51-74 with both MLPs on tcgen05: one warpgroup per CTA, thread = row of a 128-row tile, activations in tensor memory.
//   kDensityOnly: NeRFNetwork.density (network_ff.py:76-89) -- encoder + sigma net; `rgbs` is then a half [M,15] geo_feat buffer or NULL.
template <bool kDensityOnly>
__global__ void __launch_bounds__(128) k_network_forward_tc(NetConst nc, const float *__restrict__ xyzs, const float *__restrict__ dirs, uint32_t M,
                                                            float *__restrict__ sigmas, void *__restrict__ rgbs_or_geo, uint32_t n_tiles) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint4 *s_w = reinterpret_cast<uint4 *>(smem_raw);
    uint64_t *s_mbar = reinterpret_cast<uint64_t *>(s_w + nc.wtc_bytes / 16);   // one per warp
    uint32_t *s_tmem = reinterpret_cast<uint32_t *>(s_mbar + 4);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (uint32_t i = tid; i < nc.wtc_bytes / 16; i += blockDim.x) s_w[i] = __ldg(nc.wtc + i);
    tc::fence_async_smem();
    if (warp == 0) tc::tmem_alloc(tc::smem_u32(s_tmem), tc::kTmemColsTs);
    if (tid == 0) {
        for (int i = 0; i < 4; i++) tc::mbar_init(tc::smem_u32(s_mbar + i), 1);
        tc::mbar_init_fence();
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(s_tmem);
    tc::Tile t;   // the four warps run independent chains on their own lane quadrant of the tile
    t.w_saddr = tc::smem_u32(s_w); t.mbar_saddr = tc::smem_u32(s_mbar + warp);
    t.tmem = tmem_base + ((uint32_t)(warp * 32) << 16); t.q = warp; t.parity = 0;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint32_t row = tile * tc::kRows + tid;
        uint4 sh_lo = make_uint4(0, 0, 0, 0), sh_hi = sh_lo;
        float x = 0.f, y = 0.f, z = 0.f;
        if (row < M) {
            x = __ldg(xyzs + (size_t)row * 3); y = __ldg(xyzs + (size_t)row * 3 + 1); z = __ldg(xyzs + (size_t)row * 3 + 2);
            if (!kDensityOnly)
                sh_to_half(__ldg(dirs + (size_t)row * 3), __ldg(dirs + (size_t)row * 3 + 1), __ldg(dirs + (size_t)row * 3 + 2), sh_lo, sh_hi);
        }
        encode_point_tmem<4>(nc, x, y, z, row < M, t.tmem + tc::kColE);
        if (kDensityOnly) {
            float h[16];
            tc::sigma_tile128(t, lane, nc.nh_sigma, h);
            if (row < M) {
                sigmas[row] = expf(h[0]) * nc.density_scale;   // trunc_exp of the half output, in fp32 (activation.py:5-18)
                if (rgbs_or_geo) {
                    __half *o = reinterpret_cast<__half *>(rgbs_or_geo) + (size_t)row * 15;
#pragma unroll
                    for (int i = 0; i < 15; i++) o[i] = __float2half_rn(h[i + 1]);
                }
            }
        } else {
            float sg, r, g, b;
            tc::mlp_tile128(t, lane, nc.nh_sigma, nc.nh_color, nc.density_scale, sh_lo, sh_hi, sg, r, g, b);
            if (row < M) {
                sigmas[row] = sg;
                float *o = reinterpret_cast<float *>(rgbs_or_geo) + (size_t)row * 3;
                o[0] = r; o[1] = g; o[2] = b;
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, tc::kTmemColsTs);
}

// ---------------------------------------------------------------------------------------------- persistent renderer
// NeRFRenderer.run_cuda's inference loop (nerf/renderer.py:329-376) as ONE persistent cooperative launch, one 32-warp CTA per
// SM.  The reference's global schedule `n_step = clamp(N // n_alive, 1, 8)` (renderer.py:357) decides where every ray's
// chunks end, and a chunk boundary is visible in the result (rays_t restarts from the compositor's sum of deltas), so the
// schedule is reproduced exactly -- but not by meeting once per iteration.  Two facts make that unnecessary:
//
//   (1) given the schedule, a ray is independent of every other ray, and the marcher does not need the network: the value the
//       compositor writes back to rays_t is rays_t + the fp32 sum of the chunk's deltas[:,1] (raymarching.cu:870-905), which
//       the marcher can accumulate itself.  So one march call can emit E chunks of s samples with the restart applied at
//       every chunk boundary, bit for bit what E reference iterations with n_step = s would have emitted.
//   (2) n_alive never grows, so n_step never shrinks: the schedule is seven thresholds.  An EPOCH runs E reference iterations
//       speculating that n_step stays s; every ray reports the iteration it died in (ran out of samples: known to the
//       marcher; T < 1e-4: found by the compositor) into a histogram, from which every CTA recomputes n_alive of each of the
//       E iterations after the epoch's barrier.  If n_step would have changed at iteration j < E, the epoch is REPLAYED from
//       the committed ray state with E = j (exact by construction); otherwise its results are committed.  E is chosen from
//       the distance to the next threshold and the last observed death rate, so replays are rare; they cost time, never
//       correctness.
//
// Per epoch, two phases separated by a grid barrier (2 barriers per epoch instead of 2 per reference iteration):
//   MARCH     all 32 warps.  Lane-level dynamic scheduling: every LANE owns one live ray, probes the occupancy bitfield one cell
//             per trip in lockstep with the other 31 lanes of its warp, and the moment its ray has its E*s samples (or left the
//             box) the lane takes the next live ray from the list.  A sample is 8 bytes in the streaming buffer -- (t, delta_t);
//             position and dt are pure functions of t -- laid out per 32-ray tile as [pass][ray][k] so the evaluator reads it
//             fully coalesced.
//   EVALUATE  WARP-SPECIALISED, four independent STREAMS per SM.  A stream walks 128-row tiles (4 warps x R = 32 / K
//   + COMPOSITE  consecutive live rays x K consecutive samples per pass; thread = row = TMEM lane) through a producer /
//             tensor-core / consumer pipeline that lives entirely in tensor memory (128 of the 512 TMEM columns per stream):
//               gather team (4 warps)    16-level hash-grid gather of the pass's rows -> fp16 encoding written with tcgen05.st
//                                        into one of two E slots (16 columns); the LAST warp to arrive issues layer 0
//                                        (tcgen05.mma, A from TMEM, weights from shared memory) once the accumulators are free.
//               epilogue team (4 warps)  per layer: tcgen05.ld of the accumulators D -> ReLU / fp16 -> tcgen05.st of the next
//                                        operand H; the LAST warp to arrive issues the next layer.  After the sigma net: SH of
//                                        the ray direction + geo features -> colour net; after the colour net: sigmoid and
//                                        front-to-back compositing of the pass's K samples per ray in registers
//                                        (raymarching.cu:828-913), early termination included.
//             The gather of pass p + 1 and p + 2 runs under the seven-layer chain of pass p; nobody polls for anybody else's
//             rows; no activation ever touches shared memory (which holds only the weights), so 200 KB of L1 serve the gather.
//             Ray state makes one 32-byte round trip per EPOCH; dirs / sigmas / rgbs never exist in memory.
constexpr int kStreams = 4;                      // independent tile pipelines per SM
constexpr int kGatherWarps = 4 * kStreams;       // warps 0 .. 15: stream = warp / 4, TMEM lane quadrant = warp % 4
constexpr int kEpiWarps = 4 * kStreams;          // warps 16 .. 31: stream = (warp - 16) / 4
constexpr int kRenderWarps = kGatherWarps + kEpiWarps;
constexpr int kRenderThreads = kRenderWarps * 32;
constexpr uint32_t kRenderTmemCols = 512;
constexpr uint32_t kStreamCols = kRenderTmemCols / kStreams;   // D 0..63 | H 64..95 | E slot 0 96..111 | E slot 1 112..127
static_assert(kStreamCols == 128 && tc::kColE + 32 == kStreamCols, "a stream owns 128 TMEM columns");
// register budgets of the two roles during EVALUATE (setmaxnreg; the kernel launches with 64 per thread).  0 = leave it alone,
// which is the default: moving 8 registers per thread from the epilogue to the gather warps measured 5 % SLOWER (12.36 vs 11.74 ms),
// and the profiling build of the kernel hung with it.
#ifndef NGP_GATHER_REGS
#define NGP_GATHER_REGS 0
#endif
#ifndef NGP_EPI_REGS
#define NGP_EPI_REGS 0
#endif
// nanoseconds a waiter sleeps between two looks at its barrier (tc::mbar_wait): the epilogue warp waiting for a layer's MMAs, the gather
// warp waiting for a free operand slot, either of them waiting for the next tile
#ifndef NGP_SLEEP_LAYER
#define NGP_SLEEP_LAYER 64
#endif
#ifndef NGP_SLEEP_SLOT
#define NGP_SLEEP_SLOT 256
#endif
#ifndef NGP_SLEEP_TICKET
#define NGP_SLEEP_TICKET 256
#endif
#ifndef NGP_EXTRA_SMEM
#define NGP_EXTRA_SMEM 0
#endif
// A/B switch (tools/ab_build.sh): NGP_COARSE_SMEM = 0 reads the coarse occupancy from global / L1 instead of shared memory.
#ifndef NGP_COARSE_SMEM
#define NGP_COARSE_SMEM 1
#endif
constexpr int kMaxStep = 8;                    // renderer.py:357
constexpr int kMaxEpoch = 128;                 // reference iterations per epoch (max_steps 1024 / n_step 8)
#ifndef NGP_MAX_SAMPLES_PER_EPOCH
#define NGP_MAX_SAMPLES_PER_EPOCH 128
#endif
constexpr int kMaxSamplesPerEpoch = NGP_MAX_SAMPLES_PER_EPOCH;      // E * s cap: bounds the longest tile of an epoch
constexpr int kScratchPerWarp = 32 * (2 + 3 + 2);   // float4 per epilogue warp: compositing scratch (2 per lane) + ray state (3) + SH row (2)
constexpr uint32_t kTilesPerTicket = 1;      // EVALUATE: consecutive 32-row warp tiles (adjacent rays: shared hash-grid cells) per claimed ticket
#ifndef NGP_REFILL_LANES
#define NGP_REFILL_LANES 8
#endif
constexpr int kRefillLanes = NGP_REFILL_LANES;                // a warp refills its idle lanes from the live list once this many are idle

struct EpochCtr {
    uint32_t cursor;            // MARCH phase: next unclaimed entry of the live list
    uint32_t ticket;            // EVALUATE phase: next unclaimed 128-row stream tile
    uint32_t n_out;             // survivors appended to the next live list
    uint32_t opaque;            // bits of the largest weights_sum any surviving ray ended this epoch with (a float in [0, 1): orders like its bits)
    uint32_t hist[kMaxEpoch];   // rays that died in iteration j of the epoch
    uint32_t mhist[kMaxEpoch];  // of those, the ones that ran out of samples -- known to the marcher, before anything is evaluated
};
struct RenderCtr {
    EpochCtr e[3];              // attempt a uses e[a % 3]; e[(a + 1) % 3] is zeroed during attempt a
    unsigned long long samples;
    uint32_t error;
};

// Tunables of the epoch policy (ngp_ctx_set_option); the defaults are what every committed number was measured with.
struct RenderOptions {
    uint32_t rows_per_ray = 64;   // sample-buffer rows per ray (one epoch's samples of every live ray must fit); 8 B each
    uint32_t max_epoch = 0;       // cap on E (0 = none); 1 reproduces one barrier pair per reference iteration
    uint32_t kshift = 4;          // >= 4: automatic; else log2 of the samples per ray per evaluation pass
    uint32_t tile_balance = 4;    // automatic K: the smallest K in {2, 4, 8} that still leaves tile_balance warp tiles per pipeline (0 = always 2)
    uint32_t t_div = 2;           // once rays die by transmittance: E = margin / (t_div * rate)
    uint32_t t_cap = 16;          // samples per epoch while transmittance deaths can still move n_step (0 = no cap)
    uint32_t e0_cap = 4;          // cap on E of the first epoch (0 = none)
    uint32_t rate0_div = 128;     // death-rate prior of the first epochs: N / rate0_div + 1 per iteration
    uint32_t no_coarse = 0;       // 1 = do not build / use the coarse occupancy grid
    uint32_t early_cap = 32;      // samples per epoch until every ray has had `early_steps` samples (0 = off, the default: the
    uint32_t early_steps = 0;     //   transmittance horizon below decides how long an epoch dares to be)
    uint32_t horizon_div = 2;     // an epoch spends at most 1 / horizon_div of the samples left before the first extrapolated T < 1e-4
};

struct RenderArgs {
    const float *rays_o, *rays_d;
    // rays from camera poses (ngp_render_poses*): when `poses` is set, the prologue computes the ray of every pixel itself (nerf/utils.py:
    // 52-116) and writes it to gen_o / gen_d (== rays_o / rays_d, the context's staging buffers) on its way to near / far -- no separate
    // ray-generation launch, no second pass over the rays
    const float *poses;
    float *gen_o, *gen_d;
    uint32_t img_w, img_hw;
    float fx, fy, cx, cy;
    uint32_t N;
    float min_near, bg_color;
    uint32_t max_steps;
    float aabb[6];
    MarchConst mc;
    float *nears, *fars;
    float4 *state[2];  // [2 N] per buffer: (t, weights_sum, depth, seq) , (r, g, b, samples); committed / tentative ping-pong
    int *alive[2];
    uint32_t *n_emit;  // [N] samples the marcher emitted for live-list entry n in this epoch
    const uint32_t *coarse;   // [coarse_words] 4x4x4-block occupancy bits (Morton order) or NULL
    uint32_t coarse_words;
    float2 *smp;       // [cap_rows] (t, t_new - last_t) per sample of the current epoch
    uint32_t cap_rows;
    uint32_t max_epoch;   // optional cap on E (0 = none); 1 reproduces one barrier pair per reference iteration
    uint32_t force_kshift;   // >= 4: automatic; else log2 of the samples per ray per evaluation pass (profiling A/B)
    uint32_t tile_balance;   // see RenderOptions
    uint32_t t_div;          // once rays die by transmittance: E = margin / (t_div * rate)
    // iterations-per-epoch caps per n_step (index s - 1), host-computed so the device does no division: cap_n = 128 / s bounds the
    // longest tile; cap_t = 16 / s applies once rays die by transmittance while n_step can still change (s < 8) and bounds what a
    // replay throws away
    uint32_t cap_n[kMaxStep], cap_t[kMaxStep];
    uint32_t e0_cap;         // cap on E of the first epoch (0 = none): a short probe epoch learns cheaply whether rays die by transmittance
    uint32_t rate0;          // death-rate prior for the first epochs: N / 128 + 1 deaths per iteration
    uint32_t early_cap, early_steps;   // E * s <= early_cap while step < early_steps (RenderOptions)
    float horizon_div;                 // epochs use at most 1 / horizon_div of the samples left before the first transmittance death
    RenderCtr *ctr;
    float *image, *depth, *weights_sum;
    int *samples_per_ray;
    long long *stats;
    long long *prof;   // optional [8]: warp-cycles per phase (diagnostics, ngp_ctx_phase_cycles)
    // ---- optional outputs for uncertainty quantification (ngp_render_uq; all NULL in ngp_render)
    float *agg[2];           // [N] per buffer: sum of w * sigma per ray, committed / tentative like `state`
    float *agg_out;          // [N] aggregated_density (nerf/renderer.py:244)
    float4 *dump;            // [dump_rows] (sigma, r, g, b) of every sample of the CURRENT epoch at its sample-buffer row
    uint32_t dump_rows;
    int4 *ray_info;          // [N] by ray id: (seq of the epoch that last evaluated the ray, live-list position, samples, T-death iteration)
    long long *dump_hdr;     // [8] of the last committed epoch: seq, s, iterations run, log2 K, row slots per ray, n_alive, samples per ray, 0
};

__device__ __forceinline__ uint32_t clamp_n_step(uint32_t N, uint32_t n_alive) {   // renderer.py:357
    const uint32_t q = N / n_alive;
    return q > (uint32_t)kMaxStep ? (uint32_t)kMaxStep : (q < 1u ? 1u : q);
}

// Tile geometry of one epoch: a warp tile is R = 32 / K consecutive live rays, one evaluation pass covers K consecutive samples
// of each (lane = r * K + k).  K = 1 is "32 adjacent pixels, one sample each"; K = 8 gives 8x more (and 8x shorter) tiles, which
// is what keeps the streams evenly loaded once few rays are left.  Sample rows of a tile are stored [pass][ray][k], so a pass
// reads 32 consecutive rows.
struct TileGeom {
    uint32_t kshift;   // log2 K
    uint32_t ESp;      // samples per ray rounded up to a multiple of K (row slots per ray)
};
__device__ __forceinline__ uint32_t sample_row0(const TileGeom &tg, uint32_t n) {   // < cap_rows < 2^31
    const uint32_t rshift = 5u - tg.kshift;
    return (n >> rshift) * ((32u >> tg.kshift) * tg.ESp) + ((n & ((1u << rshift) - 1u)) << tg.kshift);
}

// MARCH phase of one epoch for one warp (see the header comment): up to ES = E * s samples per live ray, in chunks of s.
__device__ __forceinline__ void march_phase(const RenderArgs &ra, const uint32_t *s_coarse, int lane, uint32_t lt_mask,
                                            const int *__restrict__ alive_in, const float4 *__restrict__ state_in, uint32_t n_alive,
                                            uint32_t s, uint32_t ES, const TileGeom tg, uint32_t *cursor, uint32_t *s_hist) {
    bool has = false, list_done = false;
    uint32_t n = 0, emitted = 0, in_chunk = 0;
    const uint32_t kmask = (1u << tg.kshift) - 1u;
    uint32_t row0 = 0;
    float ox = 0.f, oy = 0.f, oz = 0.f, tm = 0.f, last_t = 0.f, far = 0.f, acc_t = 0.f;
    RayDir rd = make_ray_dir(1.f, 1.f, 1.f);
    for (;;) {
        const uint32_t need = __ballot_sync(0xffffffffu, !has);
        if (!list_done && __popc(need) >= kRefillLanes) {
            const uint32_t cnt = (uint32_t)__popc(need);
            uint32_t start = 0;
            if (lane == 0) start = atomicAdd(cursor, cnt);
            start = __shfl_sync(0xffffffffu, start, 0);
            if (!has) {
                const uint32_t my = start + (uint32_t)__popc(need & lt_mask);
                if (my < n_alive) {
                    n = my;
                    const int index = __ldcg(alive_in + n);
                    const float *o = ra.rays_o + (size_t)index * 3, *d = ra.rays_d + (size_t)index * 3;
                    ox = __ldcg(o); oy = __ldcg(o + 1); oz = __ldcg(o + 2);   // L2: the prologue may have written the rays (poses)
                    rd = make_ray_dir(__ldcg(d), __ldcg(d + 1), __ldcg(d + 2));
                    tm = __ldcg(reinterpret_cast<const float *>(state_in + 2 * (size_t)index));   // rays_t
                    last_t = tm; acc_t = tm;
                    far = __ldcg(ra.fars + index);
                    emitted = 0; in_chunk = 0;
                    row0 = sample_row0(tg, n);
                    has = true;
                }
            }
            list_done = start + cnt >= n_alive;
        }
        if (!__any_sync(0xffffffffu, has)) {
            if (list_done) break;
            continue;
        }
        if (has) {   // one lockstep probe
            float x, y, z, dt, dl;
            const float t_at = tm;
            if (tm < far && march_probe(ra.mc, ox, oy, oz, rd, tm, last_t, x, y, z, dt, dl, s_coarse)) {
                const uint32_t row = row0 + (emitted >> tg.kshift) * 32u + (emitted & kmask);
                __stcs(ra.smp + row, make_float2(t_at, dl));
                emitted++;
                acc_t = __fadd_rn(acc_t, dl);           // what composite_rays will store to rays_t (raymarching.cu:870,905)
                if (++in_chunk == s) {                  // chunk boundary: the next march call starts from rays_t (:723-725)
                    in_chunk = 0;
                    tm = acc_t; last_t = acc_t;
                }
            }
            if (emitted == ES || !(tm < far)) {         // epoch complete (raymarching.cu:752) or ray exhausted
                __stcg(ra.n_emit + n, emitted);
                if (emitted < ES) atomicAdd(s_hist + emitted / s, 1u);   // dies in the chunk that came up short (:860-863)
                has = false;
            }
        }
        __syncwarp();
    }
}

// ---- EVALUATE: the pipeline state of one PAIR (gather warp + epilogue warp on the same lane quadrant of the same TMEM columns)
struct __align__(16) PairSync {
    uint64_t dfull;      // mbarrier, count 1: tcgen05.commit of every layer -- this quadrant's accumulators D are complete
    uint64_t empty[2];   // mbarrier, count 1: tcgen05.commit of layer 0 -- the E slot has been consumed (pre-armed once)
    uint64_t tfull[2];   // mbarrier, count 1: the gather warp has published the pair's next tile ...
    uint64_t tempty[2];  // mbarrier, count 1: ... and the epilogue warp has read it (pre-armed once)
    uint32_t cnt_e[2];   // per E slot: one arrival of the gather warp (the slot is filled) + one of the epilogue warp (the previous pass
                         // has let go of D); whoever makes the second issues layer 0 -- neither waits for the other
    uint32_t ticket[2];  // the pair's current / next warp tile, claimed from EpochCtr::ticket
    unsigned long long dead;   // (tile + 1) << 32 | lanes whose ray has terminated -- a hint for the gather warp
};
__device__ __forceinline__ void pair_sync_init(PairSync &ps) {   // one thread
    tc::mbar_init(tc::smem_u32(&ps.dfull), 1);
    for (int i = 0; i < 2; i++) {
        tc::mbar_init(tc::smem_u32(&ps.empty[i]), 1);
        tc::mbar_init(tc::smem_u32(&ps.tfull[i]), 1);
        tc::mbar_init(tc::smem_u32(&ps.tempty[i]), 1);
    }
    ps.cnt_e[0] = 1u;   // D is free before the first pass: the epilogue arrival of slot 0 is pre-paid
    ps.cnt_e[1] = 0u;
    ps.dead = 0ull;
}
__device__ __forceinline__ void pair_sync_prearm(PairSync &ps) {   // one thread, after the init fence: phase 0 of every "free" barrier is complete
    for (int i = 0; i < 2; i++) {
        tc::mbar_arrive(tc::smem_u32(&ps.tempty[i]));
        tc::mbar_arrive(tc::smem_u32(&ps.empty[i]));
    }
}

// Dynamic tile scheduling.  Both warps of a pair walk the same sequence of 32-row warp tiles; tile number T of that sequence (counted
// since the kernel started, like the pass counter) is claimed from the epoch's global ticket counter by the gather warp and handed to
// the epilogue warp through a two-deep ring.  All waits sleep in hardware, none polls.
__device__ __forceinline__ uint32_t ticket_publish(PairSync &ps, uint32_t *ticket_ctr, uint32_t T, int lane) {   // gather warp
    const uint32_t ts = T & 1u;
    tc::mbar_wait<NGP_SLEEP_TICKET>(tc::smem_u32(&ps.tempty[ts]), (T >> 1) & 1u, 0x40000000u | T);
    uint32_t t = 0;
    if (lane == 0) {
        t = atomicAdd(ticket_ctr, 1u);
        *reinterpret_cast<volatile uint32_t *>(&ps.ticket[ts]) = t;
        tc::mbar_arrive(tc::smem_u32(&ps.tfull[ts]));   // release: the ticket is visible to whoever sees the phase complete
    }
    return __shfl_sync(0xffffffffu, t, 0);
}
__device__ __forceinline__ uint32_t ticket_consume(PairSync &ps, uint32_t T, int lane) {   // epilogue warp
    const uint32_t ts = T & 1u;
    tc::mbar_wait<NGP_SLEEP_TICKET>(tc::smem_u32(&ps.tfull[ts]), (T >> 1) & 1u, 0x50000000u | T);
    const uint32_t t = *reinterpret_cast<volatile uint32_t *>(&ps.ticket[ts]);
    __syncwarp();
    if (lane == 0) tc::mbar_arrive(tc::smem_u32(&ps.tempty[ts]));
    return t;
}

constexpr int kMaxLayers = 10;   // 2 nets x (first + <= 3 hidden + last)
static void layer_table_init(NetConst &nc) {   // host, ngp_ctx_set_network
    const tc::WeightMap wm = tc::weight_map(nc.nh_sigma, nc.nh_color);
    const int n_sig = 2 + (int)nc.nh_sigma, n_all = n_sig + 2 + (int)nc.nh_color;
    for (int i = 0; i < n_all; i++) {
        const bool in_color = i >= n_sig;
        const int l = in_color ? i - n_sig : i;
        const bool first = l == 0, last = l == (int)(in_color ? nc.nh_color : nc.nh_sigma) + 1;
        const uint32_t w_off = (in_color ? wm.c0 : wm.s0) + (first ? 0u : 4096u + (uint32_t)(l - 1) * 8192u);   // [32->64] is 4 KB, [64->64] 8 KB
        const uint32_t N = last ? 16u : 64u, K = first ? 32u : 64u;
        nc.ldesc_b[i] = tc::smem_desc(w_off, N * 16, 128);
        nc.ldesc_i[i] = tc::instr_desc(N);
        nc.ldesc_nk_bstep[i] = ((K / 16) << 16) | (2u * N);   // 2 * (N * 16) bytes per K-chunk >> 4
    }
}
// One layer for lane quadrant q: D = A (TMEM columns from tmem_a) x W^T.  One thread.  All K / 16 MMAs sit in ONE asm block, and every
// operand is a warp-uniform value the compiler can prove uniform (kernel parameters, the warp index, literals), so that the uniform
// operands are set up once per layer (a masked tcgen05.mma issued alone, from per-lane registers, costs ~19 SASS instructions).
template <int NK>
__device__ __forceinline__ void issue_layer_nk(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc0, uint32_t idesc, uint32_t bstep, uint32_t m0,
                                               uint32_t m1, uint32_t m2, uint32_t m3) {
    static_assert(NK == 2 || NK == 4, "a layer has K = 32 or K = 64");
    if (NK == 2) {
        asm volatile(
            "{\n\t"
            ".reg .pred p0, p1;\n\t"
            ".reg .b32 a1;\n\t"
            ".reg .b64 b1, st;\n\t"
            "setp.ne.b32 p0, 0, 0;\n\t"
            "setp.eq.b32 p1, 0, 0;\n\t"
            "cvt.u64.u32 st, %4;\n\t"
            "add.u64 b1, %2, st;\n\t"
            "add.u32 a1, %1, 8;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %6, %7, %8}, p0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], b1, %3, {%5, %6, %7, %8}, p1;\n\t"
            "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc0), "r"(idesc), "r"(bstep), "r"(m0), "r"(m1), "r"(m2), "r"(m3)
            : "memory");
    } else {
        asm volatile(
            "{\n\t"
            ".reg .pred p0, p1;\n\t"
            ".reg .b32 a1, a2, a3;\n\t"
            ".reg .b64 b1, b2, b3, st;\n\t"
            "setp.ne.b32 p0, 0, 0;\n\t"
            "setp.eq.b32 p1, 0, 0;\n\t"
            "cvt.u64.u32 st, %4;\n\t"
            "add.u64 b1, %2, st; add.u64 b2, b1, st; add.u64 b3, b2, st;\n\t"
            "add.u32 a1, %1, 8; add.u32 a2, %1, 16; add.u32 a3, %1, 24;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %6, %7, %8}, p0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], b1, %3, {%5, %6, %7, %8}, p1;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], b2, %3, {%5, %6, %7, %8}, p1;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], b3, %3, {%5, %6, %7, %8}, p1;\n\t"
            "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc0), "r"(idesc), "r"(bstep), "r"(m0), "r"(m1), "r"(m2), "r"(m3)
            : "memory");
    }
}
// w_desc = (shared address of the weight blob) >> 4: added to the relative descriptor (the address field holds 14 bits; 48 KB blob + base fit)
__device__ __forceinline__ void issue_layer(const NetConst &nc, int layer, uint32_t w_desc, uint32_t tmem_d, uint32_t tmem_a, int q) {
    const uint64_t bdesc0 = nc.ldesc_b[layer] + w_desc;
    const uint32_t idesc = nc.ldesc_i[layer], nkb = nc.ldesc_nk_bstep[layer], bstep = nkb & 0xffffu;
    const uint32_t m0 = q == 0 ? 0u : 0xffffffffu, m1 = q == 1 ? 0u : 0xffffffffu, m2 = q == 2 ? 0u : 0xffffffffu, m3 = q == 3 ? 0u : 0xffffffffu;
    if ((nkb >> 16) == 4u) issue_layer_nk<4>(tmem_d, tmem_a, bdesc0, idesc, bstep, m0, m1, m2, m3);
    else issue_layer_nk<2>(tmem_d, tmem_a, bdesc0, idesc, bstep, m0, m1, m2, m3);
}

// One arrival of this (converged) warp on a two-party counter; true in every lane iff it was the second one.
__device__ __forceinline__ bool second_arrival(uint32_t cnt_saddr, int lane) {
    uint32_t old = 0;
    if (lane == 0) old = tc::smem_arrive_count(cnt_saddr);
    return (__shfl_sync(0xffffffffu, old, 0) & 1u) == 1u;
}

// Layer 0 of a pass (A = E slot, 32 -> 64) for the pair's lane quadrant, issued by whichever of the two warps made the second arrival
// on the slot's counter.
__device__ __forceinline__ void issue_layer0(const NetConst &nc, uint32_t tmem_col, uint32_t w_desc, uint32_t slot, int q, uint32_t empty_saddr, uint32_t dfull_saddr) {
    tc::fence_after_sync();
#ifndef NGP_ABLATE_MLP
    // two call sites, so that each sees a warp-uniform operand address (`slot` comes from a data-dependent pass counter)
    if (slot) issue_layer(nc, 0, w_desc, tmem_col + tc::kColD, tmem_col + tc::kColE + 16u, q);
    else issue_layer(nc, 0, w_desc, tmem_col + tc::kColD, tmem_col + tc::kColE, q);
#endif
    tc::mma_commit(empty_saddr);   // the slot may be refilled ...
    tc::mma_commit(dfull_saddr);   // ... and the epilogue warp may read D once these MMAs have completed
}

// GATHER role of one warp: every pass of every warp tile the pair claims.  A warp tile is R = 32 / K consecutive live rays; a pass is K
// consecutive samples of each (lane = r * K + k = TMEM lane within the quadrant).  P / T: passes / tiles this pair has issued since the
// kernel started (they select the E slot, the ticket slot and the barrier phases; identical in both warps of the pair).
template <bool kProf>
__device__ __forceinline__ void gather_role(const NetConst &nc, const RenderArgs &ra, PairSync &ps, uint32_t tmem_col, uint32_t w_desc, int lane,
                                            int q, uint32_t *ticket_ctr, const int *__restrict__ alive_in, uint32_t n_alive,
                                            uint32_t ES, const TileGeom tg, uint32_t &P, uint32_t &T, long long (&pc)[8]) {
    const uint32_t kshift = tg.kshift, K = 1u << kshift, R = 32u >> kshift;
    const uint32_t n_tiles = (n_alive + R - 1) / R;
    const uint32_t my_r = (uint32_t)lane >> kshift, my_k = (uint32_t)lane & (K - 1u);
    const uint32_t tile_rows = R * tg.ESp;
    const uint32_t ps_s = tc::smem_u32(&ps);   // one base register; the fields are immediate offsets
    constexpr uint32_t kEmpty = offsetof(PairSync, empty), kCnt = offsetof(PairSync, cnt_e), kDfull = offsetof(PairSync, dfull);
    const uint32_t tmem_lane = tmem_col + ((uint32_t)(q * 32) << 16);   // this warp's lanes of the stream's columns
    const volatile unsigned long long *dead_w = &ps.dead;
    long long c0 = kProf ? clock64() : 0, c1;
#define NGP_PROF(k) if (kProf) { c1 = clock64(); pc[k] += c1 - c0; c0 = c1; }
    const uint32_t n_tickets = (n_tiles + kTilesPerTicket - 1) / kTilesPerTicket;
    uint32_t ticket = ticket_publish(ps, ticket_ctr, T, lane);
    T++;
    while (ticket < n_tickets) {
      const uint32_t next_ticket = ticket_publish(ps, ticket_ctr, T, lane);   // one ahead: the claim's latency hides behind these tiles
      T++;
      for (uint32_t tile = ticket * kTilesPerTicket; tile < min(n_tiles, (ticket + 1u) * kTilesPerTicket); tile++) {
        const uint32_t n = tile * R + my_r;
        const bool has_ray = n < n_alive;
        uint32_t ne = 0;
        float ox = 0.f, oy = 0.f, oz = 0.f, dx = 0.f, dy = 0.f, dz = 0.f;
        if (has_ray) {
            ne = min(__ldcg(ra.n_emit + n), ES);   // ES may be shorter than what was marched (pre-validated epoch)
            const int index = __ldcg(alive_in + n);
            const float *o = ra.rays_o + (size_t)index * 3, *d = ra.rays_d + (size_t)index * 3;
            ox = __ldcg(o); oy = __ldcg(o + 1); oz = __ldcg(o + 2);
            dx = __ldcg(d); dy = __ldcg(d + 1); dz = __ldcg(d + 2);
        }
        uint32_t max_ne = ne;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) max_ne = max(max_ne, __shfl_xor_sync(0xffffffffu, max_ne, o));
        const uint32_t tbase = tile * tile_rows + (uint32_t)lane;
        float t_next = 0.f;
        if (my_k < ne) t_next = __ldcg(reinterpret_cast<const float *>(ra.smp + tbase));
        NGP_PROF(0)
        for (uint32_t i0 = 0; i0 < max_ne; i0 += K, P++) {
            const uint32_t slot = P & 1u;
            tc::mbar_wait<NGP_SLEEP_SLOT>(ps_s + kEmpty + slot * 8u, (P >> 1) & 1u, 0x10000000u | P);   // layer 0 of pass P - 2 has consumed this slot
            tc::fence_after_sync();
            NGP_PROF(5)
            const uint32_t i = i0 + my_k;
            const unsigned long long dw = *dead_w;
            const bool dead = ra.dump == nullptr && (uint32_t)(dw >> 32) == tile + 1u && (((uint32_t)dw >> lane) & 1u);   // a dump wants every marched sample
            const bool active = has_ray && i < ne && !dead;
            const float t = t_next;
            if (!dead && i + K < ne) t_next = __ldcg(reinterpret_cast<const float *>(ra.smp + tbase + (i0 + K) * R));
#ifdef NGP_ABLATE_GATHER   // ablation builds (tools/ab_build.sh): what does the frame cost without one of the two big phases?
            const bool gather_on = false;   // the operand is all zeros: deterministic, so the epoch speculation still converges
#else
            const bool gather_on = true;
#endif
            encode_point_tmem<kRenderLevelsPerTrip, false>(nc, sample_pos(ra.mc, ox, dx, t), sample_pos(ra.mc, oy, dy, t), sample_pos(ra.mc, oz, dz, t),
                                                    active && gather_on, tmem_lane + tc::kColE + slot * 16u);
            tc::tmem_st_wait();
            tc::fence_before_sync();
            __syncwarp();
            NGP_PROF(3)
            if (second_arrival(ps_s + kCnt + slot * 4u, lane) && tc::elect_one())   // slot filled AND D free: issue layer 0
                issue_layer0(nc, tmem_col, w_desc, slot, q, ps_s + kEmpty + slot * 8u, ps_s + kDfull);
            __syncwarp();
            NGP_PROF(5)
        }
      }
      ticket = next_ticket;
    }
#undef NGP_PROF
}

// EPILOGUE role of one warp: every layer's activation for the pair's 32 rows, then compositing, ray state and the next live list.
// par: phase parity of ps.dfull (one phase per layer), persistent like P and T.
template <bool kProf>
__device__ __forceinline__ void epilogue_role(const NetConst &nc, const RenderArgs &ra, PairSync &ps, uint32_t tmem_col, uint32_t w_desc,
                                              float4 *scratch, int lane, int q, uint32_t lt_mask,
                                              const int *__restrict__ alive_in, int *__restrict__ alive_out, const float4 *__restrict__ state_in,
                                              float4 *__restrict__ state_out, uint32_t n_alive, uint32_t s, uint32_t ES, const TileGeom tg,
                                              uint32_t seq, uint32_t agg_in, EpochCtr *ec, uint32_t *s_hist, uint32_t &P, uint32_t &T, uint32_t &par, long long (&pc)[8]) {
    const uint32_t kshift = tg.kshift, K = 1u << kshift, R = 32u >> kshift;
    const uint32_t n_tiles = (n_alive + R - 1) / R;
    const uint32_t my_r = (uint32_t)lane >> kshift, my_k = (uint32_t)lane & (K - 1u);
    const bool leader = my_k == 0;
    const int lead_lane = lane & ~(int)(K - 1u);
    const uint32_t tile_rows = R * tg.ESp;
    const uint32_t ps_s = tc::smem_u32(&ps);   // one base register; the fields are immediate offsets
    constexpr uint32_t kEmpty = offsetof(PairSync, empty), kCnt = offsetof(PairSync, cnt_e), kDfull = offsetof(PairSync, dfull);
    const int n_sig = 2 + (int)nc.nh_sigma, n_all = n_sig + 2 + (int)nc.nh_color;
    const uint32_t tmem_lane = tmem_col + ((uint32_t)(q * 32) << 16);   // this warp's lanes of the stream's columns
    const uint32_t t_d = tmem_lane + tc::kColD, t_h = tmem_lane + tc::kColH;
    float4 *scr = scratch + 2 * lane;   // this lane's 32 bytes of its warp's compositing scratch
    float4 *rs = scratch + 64 + 3 * lane;   // ... the 48 bytes of ray state behind it
    uint4 *shs = reinterpret_cast<uint4 *>(scratch + 160) + 2 * lane;   // ... and the 16 SH halves of this lane's ray
    long long c0 = kProf ? clock64() : 0, c1;
#define NGP_PROF(k) if (kProf) { c1 = clock64(); pc[k] += c1 - c0; c0 = c1; }
    const uint32_t n_tickets = (n_tiles + kTilesPerTicket - 1) / kTilesPerTicket;
    for (uint32_t ticket = ticket_consume(ps, T++, lane); ticket < n_tickets; ticket = ticket_consume(ps, T++, lane)) {
      NGP_PROF(6)
      for (uint32_t tile = ticket * kTilesPerTicket; tile < min(n_tiles, (ticket + 1u) * kTilesPerTicket); tile++) {
        const uint32_t n = tile * R + my_r;
        const bool has_ray = n < n_alive;
        uint32_t ne = 0;
        int index = 0;
        // The accumulators of a ray live in SHARED memory between passes (3 x 16 B in its leader lane's slot): only the leader lanes touch
        // them, for a few instructions per pass, and keeping them out of the register file is what keeps the layer loop free of spills.
        //   rs[0] = (t, weights_sum, depth, samples)   rs[1] = (r, g, b, death)   rs[2] = (chunk, in_chunk, sum of w * sigma, T-death iteration)
        if (has_ray) {
            ne = min(__ldcg(ra.n_emit + n), ES);
            index = __ldcg(alive_in + n);
            const float *d = ra.rays_d + (size_t)index * 3;
            sh_to_half(__ldcg(d), __ldcg(d + 1), __ldcg(d + 2), shs[0], shs[1]);   // the direction is constant along the ray: SH once per tile,
            if (leader) {                                                       // kept in this lane's shared-memory slot
                const float4 r0 = __ldcg(state_in + 2 * (size_t)index), r1 = __ldcg(state_in + 2 * (size_t)index + 1);
                rs[0] = make_float4(r0.x, r0.y, r0.z, r1.w);
                rs[1] = make_float4(r1.x, r1.y, r1.z, 0.f);
                rs[2] = make_float4(0.f, 0.f, ra.agg_out ? __ldcg(ra.agg[agg_in] + index) : 0.f, __uint_as_float(0x7fffffffu));
            }
        }
        uint32_t max_ne = ne;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) max_ne = max(max_ne, __shfl_xor_sync(0xffffffffu, max_ne, o));
        const uint32_t tbase = tile * tile_rows + (uint32_t)lane;
        float2 sn = make_float2(0.f, 0.f);
        if (my_k < ne) sn = __ldcg(ra.smp + tbase);
        bool dead = false;           // the ray of this lane (leader decides, its K lanes follow)
        NGP_PROF(0)
        for (uint32_t i0 = 0; i0 < max_ne; i0 += K, P++) {
            const uint32_t i = i0 + my_k;
            const bool active = has_ray && !dead && i < ne;
            const float2 sc = sn;
            if (!dead && i + K < ne) sn = __ldcg(ra.smp + tbase + (i0 + K) * R);
            float sg = 0.f, cr = 0.f, cg = 0.f, cbl = 0.f;
#ifndef NGP_ABLATE_MLP
#pragma unroll 1
            for (int step = 0; step < n_all; step++) {
                tc::mbar_wait<NGP_SLEEP_LAYER>(ps_s + kDfull, par, 0x30000000u | ((uint32_t)step << 24) | P);
                par ^= 1u;
                tc::fence_after_sync();
                NGP_PROF(6)
                const bool in_color = step >= n_sig;
                const int l = in_color ? step - n_sig : step;                           // layer within its net
                const bool last = l == (int)(in_color ? nc.nh_color : nc.nh_sigma) + 1;
                if (last && in_color) {
                    tc::rgb_from_acc(t_d, cr, cg, cbl);
                    tc::fence_before_sync();
                    __syncwarp();
                    // D is free: that is one of the two arrivals layer 0 of the NEXT pass needs; if its slot is already filled, issue it here
                    const uint32_t nslot = (P + 1u) & 1u;
                    if (second_arrival(ps_s + kCnt + nslot * 4u, lane) && tc::elect_one())
                        issue_layer0(nc, tmem_col, w_desc, nslot, q, ps_s + kEmpty + nslot * 8u, ps_s + kDfull);
                    break;
                }
                if (!last) {
                    tc::relu32_to_tmem(t_d, t_h);
                    tc::relu32_to_tmem(t_d + 32, t_h + 16);
                } else {   // end of the sigma net
                    sg = tc::sigma_and_color_input(t_d, t_h, nc.density_scale, shs[0], shs[1]);
                }
                tc::tmem_st_wait();
                tc::fence_before_sync();
                __syncwarp();
                if (tc::elect_one()) {   // this warp's rows of the operand are complete: run the next layer for them, nobody else is involved
                    tc::fence_after_sync();
                    issue_layer(nc, step + 1, w_desc, tmem_col + tc::kColD, tmem_col + tc::kColH, q);   // consumes what this step wrote
                    tc::mma_commit(ps_s + kDfull);
                }
                __syncwarp();
                NGP_PROF(4)
            }
#else
            {   // no MLPs: consume layer 0's (empty) commit, hand D back
                tc::mbar_wait(ps_s + kDfull, par);
                par ^= 1u;
                tc::fence_after_sync();
                sg = 1.0f + (float)(lane & 1); cr = cg = cbl = 0.5f;
                tc::fence_before_sync();
                __syncwarp();
                const uint32_t nslot = (P + 1u) & 1u;
                if (second_arrival(ps_s + kCnt + nslot * 4u, lane) && tc::elect_one())
                    issue_layer0(nc, tmem_col, w_desc, nslot, q, ps_s + kEmpty + nslot * 8u, ps_s + kDfull);
            }
#endif
            if (ra.dump != nullptr && has_ray && i < ne) {   // opt-in: the per-sample outputs uncertain.py-style callers read (renderer.py:381-384)
                const uint32_t row = tbase + i0 * R;
                if (row < ra.dump_rows) __stcs(ra.dump + row, make_float4(sg, cr, cg, cbl));
            }
            const float dt = sample_dt(ra.mc, sc.x);
            if (K > 1) {
                scr[0] = make_float4(sg, cr, cg, cbl);
                scr[1] = make_float4(dt, sc.y, 0.f, 0.f);
                __syncwarp();
            }
            if (leader && has_ray && !dead && i0 < ne) {   // composite this pass's <= K samples of the ray in order (raymarching.cu:866-894)
                const float4 s0 = rs[0], s1 = rs[1], s2 = rs[2];
                RayAcc a = {s0.x, s0.y, s0.z, s1.x, s1.y, s1.z};
                uint32_t samples = __float_as_uint(s0.w), death = __float_as_uint(s1.w), chunk = __float_as_uint(s2.x), in_chunk = __float_as_uint(s2.y);
                float agg = s2.z;
                uint32_t t_death = __float_as_uint(s2.w);
                const uint32_t cnt = min(K, ne - i0);
                for (uint32_t k = 0; k < cnt; k++) {
                    float4 o = make_float4(sg, cr, cg, cbl), dd = make_float4(dt, sc.y, 0.f, 0.f);
                    if (K > 1) { o = scr[2 * k]; dd = scr[2 * k + 1]; }
                    float w;
                    const bool done = composite_sample(a, o.x, o.y, o.z, o.w, dd.x, dd.y, &w);
                    agg = __fmaf_rn(w, o.x, agg);   // renderer.py:244: aggregated_density = sum(weights * sigma)
                    if (done) {
                        dead = true;
                        death = chunk;
                        t_death = chunk;
                        samples += min(ne, (chunk + 1) * s);   // the reference marched the whole chunk before compositing it
                        break;
                    }
                    if (++in_chunk == s) { in_chunk = 0; chunk++; }
                }
                rs[0] = make_float4(a.t, a.ws, a.d, __uint_as_float(samples));
                rs[1] = make_float4(a.r, a.g, a.b, __uint_as_float(death));
                rs[2] = make_float4(__uint_as_float(chunk), __uint_as_float(in_chunk), agg, __uint_as_float(t_death));
            }
            if (K > 1) dead = __shfl_sync(0xffffffffu, dead, lead_lane);
            {   // tell the gather warp which rows it may skip from now on (a hint: a stale or foreign word only costs work)
                const uint32_t dm = __ballot_sync(0xffffffffu, dead);
                if (lane == 0) *reinterpret_cast<volatile unsigned long long *>(&ps.dead) = ((unsigned long long)(tile + 1u) << 32) | dm;
            }
            NGP_PROF(4)
        }
        // ---- end of the epoch for these rays: who survives (raymarching.cu:896-911), tentative state, next live list
        bool survives = false;
        if (has_ray && leader) {
            const float4 s0 = rs[0], s1 = rs[1];
            RayAcc a = {s0.x, s0.y, s0.z, s1.x, s1.y, s1.z};
            uint32_t samples = __float_as_uint(s0.w), death = __float_as_uint(s1.w);
            if (!dead) {
                samples += ne;
                if (ne < ES) { dead = true; death = ne / s; }   // the chunk that came up short: delta == 0 sentinel (:860-863)
                else survives = true;
            }
            if (dead) atomicAdd(s_hist + death, 1u);
            else if (__float_as_uint(a.ws) > __ldcg(&ec->opaque)) atomicMax(&ec->opaque, __float_as_uint(a.ws));   // how far has the most opaque ray got
            float4 *so = state_out + 2 * (size_t)index;
            __stcg(so, make_float4(a.t, a.ws, a.d, __uint_as_float(seq)));
            __stcg(so + 1, make_float4(a.r, a.g, a.b, __uint_as_float(samples)));
            if (ra.agg_out) __stcg(ra.agg[agg_in ^ 1] + index, rs[2].z);
            if (ra.ray_info) __stcg(ra.ray_info + index, make_int4((int)seq, (int)n, (int)ne, (int)__float_as_uint(rs[2].w)));
        }
        const uint32_t ballot = __ballot_sync(0xffffffffu, survives);   // renderer.py:369
        if (ballot) {
            uint32_t b0 = 0;
            if (lane == 0) b0 = atomicAdd(&ec->n_out, (uint32_t)__popc(ballot));
            b0 = __shfl_sync(0xffffffffu, b0, 0);
            if (survives) __stcg(alive_out + b0 + __popc(ballot & lt_mask), index);
        }
        NGP_PROF(0)
      }
    }
#undef NGP_PROF
}

template <int kRegs>
__device__ __forceinline__ void set_max_regs_inc() {
    if (kRegs > 0) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegs > 0 ? kRegs : 64));
}
template <int kRegs>
__device__ __forceinline__ void set_max_regs_dec() {
    if (kRegs > 0) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegs > 0 ? kRegs : 64));
}

template <bool kProf>
__global__ void __launch_bounds__(kRenderThreads, 1) k_render(NetConst nc, RenderArgs ra) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ uint32_t s_hist[kMaxEpoch], s_val[kMaxEpoch], s_mval[kMaxEpoch];
    __shared__ PairSync s_sync[kGatherWarps];   // pair = (stream, lane quadrant) = gather warp index = epilogue warp index - 16
    __shared__ uint32_t s_tmem;
    // dynamic shared memory: [tcgen05 weights][compositing scratch + ray state + SH rows, 3.5 KB per epilogue warp][coarse occupancy, optional]
    uint4 *s_w = reinterpret_cast<uint4 *>(smem_raw);
    float4 *s_scratch = reinterpret_cast<float4 *>(s_w + nc.wtc_bytes / 16);
    const int tid = threadIdx.x, lane = tid & 31;
    // the warp index through a broadcast: the compiler then knows that everything derived from it (role, stream, quadrant, TMEM columns,
    // barrier addresses) is warp-uniform and keeps it in uniform registers -- which do not count against the 64 per thread
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const bool is_gather = warp < kGatherWarps;
    const int stream = (warp & (kGatherWarps - 1)) >> 2, q = warp & 3;
    // coarse occupancy: a shared-memory copy (NGP_COARSE_SMEM) or the global array itself (4 KB, stays in L1)
    const bool coarse_in_smem = NGP_COARSE_SMEM != 0 && ra.coarse != nullptr;
    uint32_t *s_coarse = coarse_in_smem ? reinterpret_cast<uint32_t *>(s_scratch + (size_t)kEpiWarps * kScratchPerWarp) : const_cast<uint32_t *>(ra.coarse);
    const uint32_t N = ra.N;
    const uint32_t lt_mask = (1u << lane) - 1u;
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long k_start = kProf ? clock64() : 0;
    for (uint32_t i = tid; i < nc.wtc_bytes / 16; i += blockDim.x) s_w[i] = __ldg(nc.wtc + i);
    tc::fence_async_smem();
    if (coarse_in_smem)
        for (uint32_t i = tid; i < ra.coarse_words; i += blockDim.x) s_coarse[i] = __ldg(ra.coarse + i);
    if (tid < kMaxEpoch) s_hist[tid] = 0;
    if (warp == 0) tc::tmem_alloc(tc::smem_u32(&s_tmem), kRenderTmemCols);
    if (tid == 0) {
        for (int g = 0; g < kGatherWarps; g++) pair_sync_init(s_sync[g]);
        tc::mbar_init_fence();
        for (int g = 0; g < kGatherWarps; g++) pair_sync_prearm(s_sync[g]);
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    // All 512 columns are ours, so the allocation starts at column 0, lane 0; relying on that keeps every TMEM address a warp-uniform
    // value the compiler can see (an address loaded from shared memory is a per-lane value to it, and every tcgen05.mma operand that is
    // not provably uniform costs a waterfall loop at the issue site).
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t *>(&s_tmem);
    if (tmem_base != 0u) __trap();
    const uint32_t tmem_col = (uint32_t)stream * kStreamCols;   // lane field 0: what the MMAs address
    const uint32_t w_desc = (tc::smem_u32(s_w) >> 4) & 0x3FFFu;
    PairSync &ss = s_sync[warp & (kGatherWarps - 1)];
    uint32_t passes = 0, tiles = 0, par = 0;   // the stream's pipeline position (see gather_role / epilogue_role): lives across epochs

    // ---- prologue: near/far, ray state, first live list (renderer.py:273-274, 335-343)
    for (uint32_t n = blockIdx.x * blockDim.x + tid; n < N; n += gridDim.x * blockDim.x) {
        float o[3], d[3];
        if (ra.poses) {
            pose_ray(ra.poses + (size_t)(n / ra.img_hw) * 16, n % ra.img_hw, ra.img_w, ra.fx, ra.fy, ra.cx, ra.cy, o, d);
#pragma unroll
            for (int k = 0; k < 3; k++) { ra.gen_o[(size_t)n * 3 + k] = o[k]; ra.gen_d[(size_t)n * 3 + k] = d[k]; }
        } else {
#pragma unroll
            for (int k = 0; k < 3; k++) { o[k] = __ldg(ra.rays_o + (size_t)n * 3 + k); d[k] = __ldg(ra.rays_d + (size_t)n * 3 + k); }
        }
        float near, far;
        if (!slab_near_far(o[0], o[1], o[2], d[0], d[1], d[2], ra.aabb, ra.min_near, near, far))
            near = far = FLT_MAX;
        ra.nears[n] = near; ra.fars[n] = far;
        ra.alive[0][n] = (int)n;
        ra.state[0][2 * (size_t)n] = make_float4(near, 0.f, 0.f, __uint_as_float(1u));
        ra.state[0][2 * (size_t)n + 1] = make_float4(0.f, 0.f, 0.f, __uint_as_float(0u));
        ra.state[1][2 * (size_t)n] = make_float4(0.f, 0.f, 0.f, __uint_as_float(0u));
        if (ra.agg_out) ra.agg[0][n] = 0.f;
        if (ra.ray_info) ra.ray_info[n] = make_int4(0, 0, 0, 0);
    }
    grid.sync();

    uint32_t n_alive = N, step = 0, iter = 0, attempt = 0, in = 0, last_rate = ra.rate0, E = 1, replays = 0, t_deaths = 0, horizon = 0xffffffffu;
    bool replay = false;
    while (step < ra.max_steps && n_alive > 0) {
        const uint32_t s = clamp_n_step(N, n_alive);                         // renderer.py:357
        if (!replay) {
            if (s >= (uint32_t)kMaxStep) E = kMaxEpoch;                      // n_alive only shrinks: n_step stays 8
            else {
                const uint32_t margin = n_alive - N / (s + 1);               // deaths until N // n_alive reaches s + 1
                const uint32_t rate = last_rate < 1u ? 1u : last_rate;
                // pre-validation makes over-shooting cheap (only marching is wasted) as long as rays die by running out of samples;
                // transmittance deaths are only found by the evaluator, so be conservative once they show up
                const uint32_t e = t_deaths ? margin / (ra.t_div * rate) : (margin + margin / 2u) / rate + 1u;
                E = e < 1u ? 1u : (e > (uint32_t)kMaxEpoch ? (uint32_t)kMaxEpoch : e);
            }
            if (ra.max_epoch && E > ra.max_epoch) E = ra.max_epoch;
            if (iter == 0 && ra.e0_cap && E > ra.e0_cap) E = ra.e0_cap;
            if (step < ra.early_steps && E * s > ra.early_cap) E = ra.early_cap / s > 1u ? ra.early_cap / s : 1u;
            if (E * s > horizon) {   // do not run into the first transmittance deaths blind: beyond 32 samples an epoch needs the horizon's consent
                const uint32_t lim = horizon > 32u ? horizon : 32u;
                if (E * s > lim) E = lim / s > 1u ? lim / s : 1u;
            }
        }
        {
            const uint32_t left = (ra.max_steps - step + s - 1) / s;         // iterations the `while step < max_steps` loop still allows
            if (E > left) E = left;
            const uint32_t cap = (t_deaths && s < (uint32_t)kMaxStep) ? ra.cap_t[s - 1u] : ra.cap_n[s - 1u];   // both >= 1
            if (E > cap) E = cap;
            const uint32_t g = s == 8 ? 1u : (s == 4 ? 2u : (s == 2 || s == 6 ? 4u : 8u));   // E * s a multiple of 8 once it can be
            if (E >= g) E -= E % g;
        }
        uint32_t ES;
        TileGeom tg;
        for (;;) {   // tile geometry; shrink the epoch until its samples fit the buffer (rarely more than one trip)
            ES = E * s;
            // K samples of R = 32 / K rays per evaluation pass.  A tile's fixed cost (ticket, ray state in and out, SH rows) is per RAY
            // GROUP, so few long tiles (K = 2: 16 rays) beat many short ones (K = 8: 4 rays) for as long as every one of the SM's 16
            // pipelines still gets several tiles; near the end of a frame K grows again to keep them evenly loaded.  K <= ES.
            uint32_t ks = ES >= 5u ? 3u : (ES >= 3u ? 2u : ES - 1u);
            if (ks > 1u) {
                const uint32_t want_tiles = ra.tile_balance * gridDim.x * (uint32_t)kGatherWarps;
                uint32_t k = 1u;
                while (k < ks && (n_alive >> (5u - k)) < want_tiles) k++;
                ks = k;
            }
            tg.kshift = ra.force_kshift < 4u ? ra.force_kshift : ks;
            tg.ESp = (ES + (1u << tg.kshift) - 1u) & ~((1u << tg.kshift) - 1u);
            const uint32_t R = 32u >> tg.kshift;
            unsigned long long rows = (unsigned long long)((n_alive + R - 1) / R) * R * tg.ESp;
            if (rows <= ra.cap_rows) break;
            if (E > 1) { E = E > 8u ? E - 8u : E - 1u; continue; }
            tg.kshift = 0;                // one iteration, minimal padding: (n_alive + 31) * (s + 1) <= 2 N + 279 rows always fits
            tg.ESp = ES;
            break;
        }
        EpochCtr *ec = &ra.ctr->e[attempt % 3];
        if (blockIdx.x == 0 && tid < (int)(sizeof(EpochCtr) / 4))   // nobody else touches this slot during this attempt
            reinterpret_cast<uint32_t *>(&ra.ctr->e[(attempt + 1) % 3])[tid] = 0;
        const int *alive_in = ra.alive[in];
        int *alive_out = ra.alive[in ^ 1];
        const float4 *state_in = ra.state[in];
        float4 *state_out = ra.state[in ^ 1];
        long long g0 = kProf ? clock64() : 0;
        march_phase(ra, s_coarse, lane, lt_mask, alive_in, state_in, n_alive, s, ES, tg, &ec->cursor, s_hist);
        if (kProf) { const long long g1 = clock64(); pc[1] += g1 - g0; g0 = g1; }
        __syncthreads();
        if (tid < kMaxEpoch) {
            const uint32_t v = s_hist[tid];
            if (v) atomicAdd(&ec->mhist[tid], v);
            s_hist[tid] = 0;
        }
        grid.sync();
        // ---- pre-validation: the rays that ran out of samples are known now.  If they alone move n_step at iteration j, only
        //      the first j iterations of what was marched are evaluated (the marched samples of those are exact).
        if (tid < kMaxEpoch) s_val[tid] = s_mval[tid] = tid < (int)E ? __ldcg(&ec->mhist[tid]) : 0u;
        __syncthreads();
        {
            uint32_t na = n_alive, ok_iters = 0;
            for (uint32_t j = 0; j < E; j++) {
                if (j > 0 && (na == 0 || clamp_n_step(N, na) != s)) break;
                na -= s_val[j];
                ok_iters++;
            }
            E = ok_iters;
            ES = E * s;
        }
        if (tid < kGatherWarps) s_sync[tid].dead = 0ull;   // the skip hints are keyed by tile number: none survives an epoch
        __syncthreads();   // s_val is rewritten below
        if (kProf) pc[2] += clock64() - g0;
        if (is_gather) {
            set_max_regs_inc<NGP_GATHER_REGS>();
            gather_role<kProf>(nc, ra, ss, tmem_col, w_desc, lane, q, &ec->ticket, alive_in, n_alive, ES, tg, passes, tiles, pc);
            set_max_regs_dec<NGP_GATHER_REGS ? 64 : 0>();
        } else {
            set_max_regs_dec<NGP_EPI_REGS>();
            epilogue_role<kProf>(nc, ra, ss, tmem_col, w_desc, s_scratch + (size_t)(warp - kGatherWarps) * kScratchPerWarp, lane, q, lt_mask, alive_in,
                                 alive_out, state_in, state_out, n_alive, s, ES, tg, attempt + 2, in, ec, s_hist, passes, tiles, par, pc);
            set_max_regs_inc<NGP_EPI_REGS ? 64 : 0>();
        }
        if (kProf) g0 = clock64();
        __syncthreads();
        if (tid < kMaxEpoch) {
            const uint32_t v = s_hist[tid];
            if (v) atomicAdd(&ec->hist[tid], v);
            s_hist[tid] = 0;
        }
        grid.sync();
        if (tid < kMaxEpoch) s_val[tid] = tid < (int)E ? __ldcg(&ec->hist[tid]) : 0u;
        __syncthreads();
        // ---- validate the speculation: replay the reference's schedule over the epoch's death histogram
        uint32_t na = n_alive, ran = 0, last = 0;
        bool ok = true, any_t = false;
        for (uint32_t j = 0; j < E; j++) {
            if (j > 0) {
                if (na == 0) break;                                         // renderer.py:352-353: the loop ended here
                if (clamp_n_step(N, na) != s) { ok = false; break; }        // n_step changed: iterations >= j ran with the wrong chunking
            }
            last = s_val[j];
            na -= last;
            ran++;
            if (last != s_mval[j]) any_t = true;   // somebody died here without running out of samples: transmittance
        }
        __syncthreads();   // s_val is rewritten after the next epoch
        if (kProf) pc[7] += clock64() - g0;
        attempt++;
        if (!ok) {          // replay from the committed state with the E that is known to be valid
            E = ran;
            replay = true;
            replays++;
            continue;
        }
        if (blockIdx.x == 0 && tid == 0 && __ldcg(&ec->n_out) != na) ra.ctr->error = 1;   // histogram and live list disagree
        if (ra.dump_hdr && blockIdx.x == 0 && tid == 0) {   // this epoch is committed: it is the last one so far
            ra.dump_hdr[0] = (long long)(attempt + 1u); ra.dump_hdr[1] = (long long)s; ra.dump_hdr[2] = (long long)ran;
            ra.dump_hdr[3] = (long long)tg.kshift; ra.dump_hdr[4] = (long long)tg.ESp; ra.dump_hdr[5] = (long long)n_alive;
            ra.dump_hdr[6] = (long long)ES; ra.dump_hdr[7] = 0;
        }
        replay = false;
        step += ran * s;
        iter += ran;
        // the rays that miss the box all die in iteration 0: that burst says nothing about the rate of the iterations after it
        last_rate = (iter == ran && ran == 1u) ? ra.rate0 : last;
        // be conservative as soon as rays die by transmittance -- or are about to (a ray has lost 90 % of its light)
        const float ws_max = __uint_as_float(__ldcg(&ec->opaque));
        t_deaths = (any_t || ws_max > 0.9f) ? 1u : 0u;
        // Transmittance horizon: the most opaque ray has lost ws_max of its light in `step` samples; at that rate it reaches T = 1e-4
        // after step * ln(1e-4) / ln(1 - ws_max) samples.  Epochs stay within a fraction of what is left (a heuristic: it costs a replay,
        // never correctness, when the density ahead is higher than the density behind).
        if (ws_max > 1e-6f && ws_max < 1.0f) {
            const float total = (float)step * __fdividef(-9.2103404f, __logf(1.0f - ws_max));
            const float left = total - (float)step;
            horizon = left < 32.0f ? 16u : (left > 4.0e6f ? 0xffffffffu : (uint32_t)(left * (1.0f / ra.horizon_div)));
        } else horizon = ws_max >= 1.0f ? 16u : 0xffffffffu;
        n_alive = na;
        in ^= 1;
    }

    // ---- epilogue: background + depth normalisation (renderer.py:375-376) from the newest record of every ray
    uint32_t my_samples = 0;
    const float poison = __ldcg(&ra.ctr->error) ? __uint_as_float(0x7fc00000u) : 0.0f;   // schedule check failed: no plausible-looking image
    for (uint32_t n = blockIdx.x * blockDim.x + tid; n < N; n += gridDim.x * blockDim.x) {
        float4 r0 = __ldcg(ra.state[0] + 2 * (size_t)n);
        const float4 q0 = __ldcg(ra.state[1] + 2 * (size_t)n);
        const bool second = __float_as_uint(q0.w) > __float_as_uint(r0.w);
        if (second) r0 = q0;
        const float4 r1 = __ldcg(ra.state[second ? 1 : 0] + 2 * (size_t)n + 1);
        const float ws_n = r0.y;
        const float k = __fadd_rn(__fmul_rn(__fsub_rn(1.0f, ws_n), ra.bg_color), poison);
        float *im = ra.image + (size_t)n * 3;
        im[0] = __fadd_rn(r1.x, k); im[1] = __fadd_rn(r1.y, k); im[2] = __fadd_rn(r1.z, k);
        ra.weights_sum[n] = ws_n;
        if (ra.agg_out) ra.agg_out[n] = __ldcg(ra.agg[second ? 1 : 0] + n);
        const float near = ra.nears[n], far = ra.fars[n];
        const float dd = __fsub_rn(r0.z, near);
        ra.depth[n] = __fdiv_rn(dd < 0.f ? 0.f : dd, __fsub_rn(far, near));   // 0/0 = NaN for misses, like the reference
        const uint32_t cnt = __float_as_uint(r1.w);
        if (ra.samples_per_ray) ra.samples_per_ray[n] = (int)cnt;
        my_samples += cnt;
    }
    if (kProf && lane == 0) {
        // slots: 0 tile set-up / tile end (both roles) + prologue, epilogue, loop control; 1 march; 2 barrier after march;
        //        3 gather (sample load + 16 levels + store); 4 epilogue (activations, compositing); 5 gather team waiting (slot, D);
        //        6 epilogue team waiting (layer commits); 7 barrier after evaluate
        const long long total = clock64() - k_start;
        pc[0] += total - (pc[0] + pc[1] + pc[2] + pc[3] + pc[4] + pc[5] + pc[6] + pc[7]);
        for (int k = 0; k < 8; k++) atomicAdd(reinterpret_cast<unsigned long long *>(ra.prof + k), (unsigned long long)pc[k]);
    }
    if (ra.stats) {
        const uint32_t wsum = warp_sum(my_samples);
        if (lane == 0) atomicAdd(&ra.ctr->samples, (unsigned long long)wsum);
        grid.sync();   // every warp's counters are in
        if (blockIdx.x == 0 && tid == 0) {
            ra.stats[0] = (long long)iter;
            ra.stats[1] = (long long)__ldcg(&ra.ctr->samples);
            ra.stats[2] = __ldcg(&ra.ctr->error) ? -1ll : (long long)attempt;   // epochs run (replays included); -1 = internal inconsistency
            ra.stats[3] = (long long)replays;
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem_base, kRenderTmemCols);
}

}  // namespace ngp

using namespace ngp;

// ---------------------------------------------------------------------------------------------- context
struct ngp_ctx {
    int device = 0, sm_count = 0;
    size_t max_smem_optin = 0;
    bool has_net = false, has_scene = false;
    bool fast_levels = false;   // every level is a plain dense grid or a power-of-two hash table (what the persistent renderer is built for)
    NetConst nc{};
    uint2 *wpack = nullptr;
    uint2 *wpack_t = nullptr;     // transposed pack for the input-gradient kernels (diff_render.cuh)
    uint4 *wtc = nullptr;
    size_t diff_smem[2] = {0, 0}; // dynamic shared memory of k_diff_render<false / true> (== k_density_backward for [1])
    // scene
    const uint8_t *bitfield = nullptr;
    uint32_t cascade = 1, grid_size = 128;
    float aabb[6] = {-1, -1, -1, 1, 1, 1};
    float min_near = 0.2f;
    // render workspace
    uint32_t capN = 0;
    uint32_t capRows = 0;
    float *nears = nullptr, *fars = nullptr;
    float2 *smp = nullptr;
    float4 *state[2] = {nullptr, nullptr};
    RenderOptions opt;
    float *agg[2] = {nullptr, nullptr};   // aggregated-density accumulators (ngp_render_uq), allocated on first use
    uint32_t capAgg = 0;
    uint32_t *h_error = nullptr;  // pinned host word: RenderCtr::error of the last *_host render
    uint32_t *n_emit = nullptr;
    int *alive[2] = {nullptr, nullptr};
    RenderCtr *ctr = nullptr;
    // staging for ngp_render_host
    uint32_t capHost = 0;
    float *d_o = nullptr, *d_d = nullptr, *d_img = nullptr, *d_depth = nullptr, *d_ws = nullptr;
    float *d_poses = nullptr;     // staging for ngp_render_poses_host
    uint32_t capPoses = 0;
    uint32_t *coarse = nullptr;   // 4x4x4-block occupancy bits, rebuilt from the borrowed bitfield at every ngp_render
    uint32_t coarse_words = 0;
    long long *prof = nullptr;    // device [8], allocated by ngp_ctx_enable_profile
    int render_blocks = 0, fwd_blocks = 0;
    size_t render_smem = 0, render_smem_max = 0, fwd_smem = 0, fwd_tc_smem = 0;
};

static size_t fwd_smem_bytes(const NetConst &nc) { return (size_t)nc.wpack_count * 8 + kTile * (kEncStride + kShStride) * 2; }
static size_t render_smem_bytes(const NetConst &nc) { return (size_t)nc.wtc_bytes + (size_t)kEpiWarps * kScratchPerWarp * sizeof(float4) + NGP_EXTRA_SMEM; }

// Frees a set of device buffers and nulls the pointers (a failed re-allocation must not leave anything dangling).
template <class... T>
static void free_all(T *&...p) {
    (void)std::initializer_list<int>{(cudaFree(p), p = nullptr, 0)...};
}

extern "C" {

int ngp_ctx_create(ngp_ctx **out) {
    NGP_REQUIRE(out != nullptr, "ctx_create: null out pointer");
    int n = 0;
    const cudaError_t e = cudaGetDeviceCount(&n);
    NGP_REQUIRE(e == cudaSuccess && n > 0, "ctx_create: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    ngp_ctx *c = new ngp_ctx();
    NGP_CUDA(cudaGetDevice(&c->device));
    NGP_CUDA(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, c->device));
    {
        int v = 0;
        NGP_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device));
        c->max_smem_optin = (size_t)v;
    }
    NGP_CUDA(cudaMalloc(&c->ctr, sizeof(RenderCtr)));
    *out = c;
    return 0;
}

int ngp_ctx_destroy(ngp_ctx *c) {
    if (!c) return 0;
    cudaFree(c->wpack); cudaFree(c->wpack_t); cudaFree(c->wtc); cudaFree(c->nears); cudaFree(c->fars); cudaFree(c->state[0]); cudaFree(c->state[1]); cudaFree(c->n_emit);
    cudaFree(c->alive[0]); cudaFree(c->alive[1]); cudaFree(c->smp); cudaFree(c->agg[0]); cudaFree(c->agg[1]);
    if (c->h_error) cudaFreeHost(c->h_error);
    cudaFree(c->ctr); cudaFree(c->prof); cudaFree(c->coarse); cudaFree(c->d_o); cudaFree(c->d_d); cudaFree(c->d_img); cudaFree(c->d_depth); cudaFree(c->d_ws); cudaFree(c->d_poses);
    delete c;
    return 0;
}

int ngp_ctx_set_network(ngp_ctx *c, const void *table_f16, const int32_t *offsets_dev, const int32_t *offsets_host, uint32_t L, float S,
                        uint32_t H, const void *w_sigma_f16, uint32_t nl_sigma, const void *w_color_f16, uint32_t nl_color, float bound,
                        float density_scale, void *stream) {
    NGP_REQUIRE(c != nullptr, "ctx_set_network: null ctx");
    NGP_REQUIRE(L == 16, "fused path: the hash grid must have 16 levels x 2 channels (sigma_net input 32); got L=%u", L);
    NGP_REQUIRE(nl_sigma >= 2 && nl_sigma <= 4 && nl_color >= 2 && nl_color <= 4, "fused path: num_layers must be in [2,4] for both nets");
    NGP_DEVPTR(table_f16); NGP_DEVPTR(offsets_dev); NGP_DEVPTR(w_sigma_f16); NGP_DEVPTR(w_color_f16);
    NGP_REQUIRE(offsets_host != nullptr, "ctx_set_network: offsets_host is null");
    cudaStream_t st = as_stream(stream);
    NetConst &nc = c->nc;
    nc.table = reinterpret_cast<const __half2 *>(table_f16);
    nc.nh_sigma = nl_sigma - 1; nc.nh_color = nl_color - 1;
    nc.bound = bound; nc.two_bound = 2.0f * bound; nc.density_scale = density_scale;
    // per-level scale exactly as the device evaluates exp2f (gridencoder.cu:126)
    float *d_scales = nullptr, h_scales[kMaxLevels];
    NGP_CUDA(cudaMalloc(&d_scales, sizeof(float) * L));
    if (int rc = ngp_grid_level_scales(L, S, H, d_scales, stream)) { cudaFree(d_scales); return rc; }
    NGP_CUDA(cudaMemcpyAsync(h_scales, d_scales, sizeof(float) * L, cudaMemcpyDeviceToHost, st));
    NGP_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_scales);
    for (uint32_t l = 0; l < L; l++) {
        LevelConst &lc = nc.lv[l];
        lc.offset = (uint32_t)offsets_host[l];
        nc.lvbase[l] = nc.table + lc.offset;
        lc.size = (uint32_t)(offsets_host[l + 1] - offsets_host[l]);
        lc.scale = h_scales[l];
        const uint32_t res = (uint32_t)ceilf(lc.scale) + 1;
        lc.stride = res + 1;
        // replay the reference's uint32 stride loop (gridencoder.cu:60-69) to decide dense vs hashed
        uint32_t stride = 1;
        for (int d = 0; d < 3 && stride <= lc.size; d++) stride *= lc.stride;
        const bool hashed = stride > lc.size;
        const unsigned long long cube = (unsigned long long)lc.stride * lc.stride * lc.stride;
        if (!hashed) lc.mode = cube <= lc.size ? 0u : 3u;
        else lc.mode = (lc.size & (lc.size - 1)) == 0 ? 1u : 2u;
    }
    c->fast_levels = true;
    for (uint32_t l = 0; l < L; l++) c->fast_levels = c->fast_levels && nc.lv[l].mode <= 1u;
    const uint32_t count = 512 + nc.nh_sigma * 1024 + 256 + 512 + nc.nh_color * 1024 + 256;
    if (c->wpack) { cudaFree(c->wpack); c->wpack = nullptr; }
    NGP_CUDA(cudaMalloc(&c->wpack, sizeof(uint2) * count));
    k_pack_network<<<1, 256, 0, st>>>(reinterpret_cast<const __half *>(w_sigma_f16), nc.nh_sigma, reinterpret_cast<const __half *>(w_color_f16),
                                      nc.nh_color, c->wpack);
    if (int rc = after_launch("pack_network")) return rc;
    nc.wpack = c->wpack;
    nc.wpack_count = count;
    {
        if (c->wpack_t) { cudaFree(c->wpack_t); c->wpack_t = nullptr; }
        NGP_CUDA(cudaMalloc(&c->wpack_t, sizeof(uint2) * count));
        k_pack_network_transposed<<<1, 256, 0, st>>>(reinterpret_cast<const __half *>(w_sigma_f16), nc.nh_sigma,
                                                     reinterpret_cast<const __half *>(w_color_f16), nc.nh_color, c->wpack_t);
        if (int rc = after_launch("pack_network_transposed")) return rc;
        for (int b = 0; b < 2; b++) c->diff_smem[b] = (size_t)count * 8 * (b + 1) + diff::kWarps * diff::warp_smem_bytes(b != 0);
        NGP_REQUIRE(c->diff_smem[1] <= c->max_smem_optin, "differentiable render: %zu B of shared memory do not fit an SM", c->diff_smem[1]);
        NGP_CUDA(cudaFuncSetAttribute(diff::k_diff_render<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->diff_smem[0]));
        NGP_CUDA(cudaFuncSetAttribute(diff::k_diff_render<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->diff_smem[0]));
        NGP_CUDA(cudaFuncSetAttribute(diff::k_diff_render<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->diff_smem[1]));
        NGP_CUDA(cudaFuncSetAttribute(diff::k_diff_render<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->diff_smem[1]));
        NGP_CUDA(cudaFuncSetAttribute(diff::k_density_backward<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->diff_smem[1]));
        NGP_CUDA(cudaFuncSetAttribute(diff::k_density_backward<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->diff_smem[1]));
    }
    {
        const tc::WeightMap wm = tc::weight_map(nc.nh_sigma, nc.nh_color);
        if (c->wtc) { cudaFree(c->wtc); c->wtc = nullptr; }
        NGP_CUDA(cudaMalloc(&c->wtc, wm.total));
        k_pack_network_tc<<<1, 256, 0, st>>>(reinterpret_cast<const __half *>(w_sigma_f16), nc.nh_sigma, reinterpret_cast<const __half *>(w_color_f16),
                                             nc.nh_color, c->wtc);
        if (int rc = after_launch("pack_network_tc")) return rc;
        nc.wtc = c->wtc;
        nc.wtc_bytes = wm.total;
        layer_table_init(nc);
        c->fwd_tc_smem = wm.total + 64;
        NGP_CUDA(cudaFuncSetAttribute(k_network_forward_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->fwd_tc_smem));
        NGP_CUDA(cudaFuncSetAttribute(k_network_forward_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->fwd_tc_smem));
    }
    c->fwd_smem = fwd_smem_bytes(nc);
    c->render_smem = render_smem_bytes(nc);
    NGP_CUDA(cudaFuncSetAttribute(k_network_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->fwd_smem));
    {
        cudaFuncAttributes fa;
        NGP_CUDA(cudaFuncGetAttributes(&fa, k_render<false>));
        c->render_smem_max = c->max_smem_optin - fa.sharedSizeBytes;   // dynamic part: opt-in limit minus the kernel's static arrays
        NGP_CUDA(cudaFuncSetAttribute(k_render<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->render_smem_max));
        NGP_CUDA(cudaFuncSetAttribute(k_render<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->render_smem_max));
    }
    int occ = 0;
    NGP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_network_forward, kTile, c->fwd_smem));
    NGP_REQUIRE(occ >= 1, "network_forward kernel does not fit on an SM (smem %zu B)", c->fwd_smem);
    c->fwd_blocks = occ * c->sm_count;
    NGP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_render<false>, kRenderThreads, c->render_smem));
    NGP_REQUIRE(occ >= 1, "render kernel does not fit on an SM (smem %zu B)", c->render_smem);
    c->render_blocks = occ * c->sm_count;
    c->has_net = true;
    return 0;
}

int ngp_ctx_enable_profile(ngp_ctx *c, int enable) {
    NGP_REQUIRE(c != nullptr, "ctx_enable_profile: null ctx");
    if (enable && !c->prof) {
        NGP_CUDA(cudaMalloc(&c->prof, 8 * sizeof(long long)));
        NGP_CUDA(cudaMemset(c->prof, 0, 8 * sizeof(long long)));
    } else if (!enable && c->prof) {
        cudaFree(c->prof);
        c->prof = nullptr;
    }
    return 0;
}

int ngp_ctx_phase_cycles(ngp_ctx *c, int64_t *out_host, int reset) {
    NGP_REQUIRE(c != nullptr && c->prof != nullptr && out_host != nullptr, "ctx_phase_cycles: profiling is not enabled (ngp_ctx_enable_profile)");
    NGP_CUDA(cudaDeviceSynchronize());
    NGP_CUDA(cudaMemcpy(out_host, c->prof, 8 * sizeof(long long), cudaMemcpyDeviceToHost));
    if (reset) NGP_CUDA(cudaMemset(c->prof, 0, 8 * sizeof(long long)));
    return 0;
}

int ngp_ctx_set_scene(ngp_ctx *c, const uint8_t *bitfield, uint32_t cascade, uint32_t grid_size, const float *aabb_host, float min_near) {
    NGP_REQUIRE(c != nullptr, "ctx_set_scene: null ctx");
    NGP_DEVPTR(bitfield);
    NGP_REQUIRE(cascade >= 1 && cascade <= 8 && grid_size >= 2 && grid_size <= 1024, "ctx_set_scene: bad cascade/grid_size");
    NGP_REQUIRE(aabb_host != nullptr, "ctx_set_scene: aabb_host is null");
    c->bitfield = bitfield; c->cascade = cascade; c->grid_size = grid_size; c->min_near = min_near;
    for (int i = 0; i < 6; i++) c->aabb[i] = aabb_host[i];
    // coarse occupancy needs 64-cell blocks aligned in Morton order (power-of-two grid) and an 8-byte aligned bitfield
    const bool pow2 = (grid_size & (grid_size - 1)) == 0 && grid_size >= 32;
    const uint32_t words = pow2 && (reinterpret_cast<uintptr_t>(bitfield) % 8 == 0) ? cascade * (grid_size * grid_size * grid_size / 2048) : 0;
    if (words != c->coarse_words) {
        cudaFree(c->coarse); c->coarse = nullptr; c->coarse_words = 0;
        if (words) {
            NGP_CUDA(cudaMalloc(&c->coarse, sizeof(uint32_t) * words));
            c->coarse_words = words;
        }
    }
    c->has_scene = true;
    return 0;
}

int ngp_network_forward(ngp_ctx *c, const float *xyzs, const float *dirs, uint32_t M, int apply_density_scale, float *sigmas,
                        float *rgbs, void *stream) {
    NGP_REQUIRE(c != nullptr && c->has_net, "network_forward: ngp_ctx_set_network has not been called");
    if (M == 0) return 0;
    NGP_DEVPTR(xyzs); NGP_DEVPTR(dirs); NGP_DEVPTR(sigmas); NGP_DEVPTR(rgbs);
    const uint32_t n_tiles = div_up(M, kTile);
    const uint32_t grid = n_tiles < (uint32_t)c->fwd_blocks ? n_tiles : (uint32_t)c->fwd_blocks;
    NetConst nc = c->nc;
    if (!apply_density_scale) nc.density_scale = 1.0f;
    k_network_forward<<<grid, kTile, c->fwd_smem, as_stream(stream)>>>(nc, xyzs, dirs, M, sigmas, rgbs, n_tiles);
    return after_launch("network_forward");
}

int ngp_network_forward_tc(ngp_ctx *c, const float *xyzs, const float *dirs, uint32_t M, int apply_density_scale, float *sigmas,
                           float *rgbs, void *stream) {
    NGP_REQUIRE(c != nullptr && c->has_net, "network_forward_tc: ngp_ctx_set_network has not been called");
    if (M == 0) return 0;
    NGP_DEVPTR(xyzs); NGP_DEVPTR(dirs); NGP_DEVPTR(sigmas); NGP_DEVPTR(rgbs);
    const uint32_t n_tiles = div_up(M, (uint32_t)tc::kRows);
    const uint32_t cap = (uint32_t)c->sm_count * 4;
    const uint32_t grid = n_tiles < cap ? n_tiles : cap;
    NetConst nc = c->nc;
    if (!apply_density_scale) nc.density_scale = 1.0f;
    k_network_forward_tc<false><<<grid, tc::kRows, c->fwd_tc_smem, as_stream(stream)>>>(nc, xyzs, dirs, M, sigmas, rgbs, n_tiles);
    return after_launch("network_forward_tc");
}

int ngp_density_forward(ngp_ctx *c, const float *xyzs, uint32_t M, int apply_density_scale, float *sigmas, void *geo_feat_f16, void *stream) {
    NGP_REQUIRE(c != nullptr && c->has_net, "density_forward: ngp_ctx_set_network has not been called");
    if (M == 0) return 0;
    NGP_DEVPTR(xyzs); NGP_DEVPTR(sigmas);
    if (geo_feat_f16) NGP_DEVPTR(geo_feat_f16);
    const uint32_t n_tiles = div_up(M, (uint32_t)tc::kRows);
    const uint32_t cap = (uint32_t)c->sm_count * 4;
    const uint32_t grid = n_tiles < cap ? n_tiles : cap;
    NetConst nc = c->nc;
    if (!apply_density_scale) nc.density_scale = 1.0f;
    k_network_forward_tc<true><<<grid, tc::kRows, c->fwd_tc_smem, as_stream(stream)>>>(nc, xyzs, nullptr, M, sigmas, geo_feat_f16, n_tiles);
    return after_launch("density_forward");
}

// ---- differentiable small-batch render (diff_render.cuh)
static int diff_args(ngp_ctx *c, const char *what, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma, uint32_t max_steps,
                     float bg_color, diff::DiffArgs &da) {
    NGP_REQUIRE(c != nullptr && c->has_net && c->has_scene, "%s: ngp_ctx_set_network / ngp_ctx_set_scene have not been called", what);
    NGP_REQUIRE(max_steps >= 1, "%s: max_steps must be >= 1", what);
    NGP_DEVPTR(rays_o); NGP_DEVPTR(rays_d);
    da = diff::DiffArgs{};
    da.rays_o = rays_o; da.rays_d = rays_d; da.N = N; da.max_steps = max_steps;
    da.mc = make_march_const(c->nc.bound, dt_gamma, max_steps, c->cascade, c->grid_size, c->bitfield);
    for (int i = 0; i < 6; i++) da.aabb[i] = c->aabb[i];
    da.min_near = c->min_near; da.bg = bg_color;
    da.wpack_t = c->wpack_t;
    return 0;
}
static uint32_t diff_grid(const ngp_ctx *c, uint32_t units) {   // units = rays or 32-point tiles, one per warp
    const uint32_t want = div_up(units, (uint32_t)diff::kWarps), cap = (uint32_t)c->sm_count * 4u;
    return want < cap ? want : cap;
}

int ngp_render_diff_forward(ngp_ctx *c, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma, uint32_t max_steps,
                            float bg_color, float *image, float *depth, float *weights_sum, float *finals, int32_t *samples_per_ray,
                            void *stream) {
    if (N == 0) return 0;
    diff::DiffArgs da;
    if (int rc = diff_args(c, "render_diff_forward", rays_o, rays_d, N, dt_gamma, max_steps, bg_color, da)) return rc;
    NGP_DEVPTR(image); NGP_DEVPTR(depth); NGP_DEVPTR(weights_sum); NGP_DEVPTR(finals);
    NGP_REQUIRE((reinterpret_cast<uintptr_t>(finals) & 15) == 0, "render_diff_forward: finals must be 16-byte aligned");
    if (samples_per_ray) NGP_DEVPTR(samples_per_ray);
    da.image = image; da.depth = depth; da.weights_sum = weights_sum; da.finals = reinterpret_cast<float4 *>(finals);
    da.samples_per_ray = samples_per_ray;
    // kGeneral = false: every level is a dense grid that fits its table or a power-of-two hash table (every NGP configuration); the
    // general instantiation carries the integer-division index paths as well
    if (c->fast_levels) diff::k_diff_render<false, false><<<diff_grid(c, N), diff::kThreads, c->diff_smem[0], as_stream(stream)>>>(c->nc, da);
    else diff::k_diff_render<false, true><<<diff_grid(c, N), diff::kThreads, c->diff_smem[0], as_stream(stream)>>>(c->nc, da);
    return after_launch("render_diff_forward");
}

int ngp_render_diff_backward(ngp_ctx *c, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma, uint32_t max_steps,
                             float bg_color, const float *finals, const float *grad_image, const float *grad_weights_sum,
                             float *grad_rays_o, float *grad_rays_d, void *stream) {
    if (N == 0) return 0;
    diff::DiffArgs da;
    if (int rc = diff_args(c, "render_diff_backward", rays_o, rays_d, N, dt_gamma, max_steps, bg_color, da)) return rc;
    NGP_DEVPTR(finals); NGP_DEVPTR(grad_image); NGP_DEVPTR(grad_rays_o); NGP_DEVPTR(grad_rays_d);
    NGP_REQUIRE((reinterpret_cast<uintptr_t>(finals) & 15) == 0, "render_diff_backward: finals must be 16-byte aligned");
    if (grad_weights_sum) NGP_DEVPTR(grad_weights_sum);
    da.finals = const_cast<float4 *>(reinterpret_cast<const float4 *>(finals));
    da.g_image = grad_image; da.g_ws = grad_weights_sum; da.g_o = grad_rays_o; da.g_d = grad_rays_d;
    if (c->fast_levels) diff::k_diff_render<true, false><<<diff_grid(c, N), diff::kThreads, c->diff_smem[1], as_stream(stream)>>>(c->nc, da);
    else diff::k_diff_render<true, true><<<diff_grid(c, N), diff::kThreads, c->diff_smem[1], as_stream(stream)>>>(c->nc, da);
    return after_launch("render_diff_backward");
}

int ngp_density_backward(ngp_ctx *c, const float *xyzs, const float *grad_sigmas, uint32_t M, int apply_density_scale, float *grad_xyzs,
                         void *stream) {
    NGP_REQUIRE(c != nullptr && c->has_net, "density_backward: ngp_ctx_set_network has not been called");
    if (M == 0) return 0;
    NGP_DEVPTR(xyzs); NGP_DEVPTR(grad_sigmas); NGP_DEVPTR(grad_xyzs);
    NetConst nc = c->nc;
    if (!apply_density_scale) nc.density_scale = 1.0f;
    const uint32_t blocks = diff_grid(c, div_up(M, 32u));
    if (c->fast_levels) diff::k_density_backward<false><<<blocks, diff::kThreads, c->diff_smem[1], as_stream(stream)>>>(nc, c->wpack_t, xyzs, grad_sigmas, M, grad_xyzs);
    else diff::k_density_backward<true><<<blocks, diff::kThreads, c->diff_smem[1], as_stream(stream)>>>(nc, c->wpack_t, xyzs, grad_sigmas, M, grad_xyzs);
    return after_launch("density_backward");
}

static int ensure_workspace(ngp_ctx *c, uint32_t N) {
    // sample buffer of one epoch: rows_per_ray * N rows of 8 B (streamed, evict-first); >= N + 256 rows always fits one iteration.
    // 64 rows per ray save an epoch on cfg2 A (-2 %) but lengthen the epochs that get replayed on terminating scenes.
    unsigned long long rpr = c->opt.rows_per_ray < 3u ? 3u : c->opt.rows_per_ray;
    const uint32_t cap = N <= c->capN ? c->capN : N + N / 8 + 1024;
    unsigned long long rows = rpr * cap + 1024;
    if (rows > 0x7fffffffull) rows = 0x7fffffffull;
    if (N <= c->capN && rows <= c->capRows) return 0;
    NGP_REQUIRE(rows >= 2ull * N + 512, "render: %u rays do not fit the sample buffer", N);
    // all or nothing: the old buffers go first (they may be most of the memory), every pointer is nulled, and a failure part-way
    // frees what it got and leaves an empty workspace behind
    free_all(c->nears, c->fars, c->state[0], c->state[1], c->n_emit, c->alive[0], c->alive[1], c->smp);
    c->capN = 0; c->capRows = 0;
    cudaError_t e = cudaSuccess;
    auto get = [&](auto *&ptr, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(&ptr, bytes); };
    get(c->nears, sizeof(float) * cap);
    get(c->fars, sizeof(float) * cap);
    get(c->state[0], sizeof(float4) * 2 * (size_t)cap);
    get(c->state[1], sizeof(float4) * 2 * (size_t)cap);
    get(c->n_emit, sizeof(uint32_t) * cap);
    get(c->alive[0], sizeof(int) * cap);
    get(c->alive[1], sizeof(int) * cap);
    get(c->smp, sizeof(float2) * (size_t)rows);
    if (e != cudaSuccess) {
        free_all(c->nears, c->fars, c->state[0], c->state[1], c->n_emit, c->alive[0], c->alive[1], c->smp);
        cudaGetLastError();   // the failed cudaMalloc must not surface later as a bogus launch error
        set_error("render: workspace for %u rays (%llu sample rows) could not be allocated: %s", N, rows, cudaGetErrorString(e));
        return 2;
    }
    c->capN = cap;
    c->capRows = (uint32_t)rows;
    return 0;
}

struct RenderExtras {   // ngp_render_uq's optional outputs
    float *aggregated_density = nullptr;
    float *sample_dump = nullptr;
    uint32_t dump_rows = 0;
    int32_t *ray_info = nullptr;
    int64_t *dump_header = nullptr;
    // rays from poses: rays_o / rays_d passed to render_core are then OUTPUT buffers the prologue fills
    const float *poses = nullptr;
    uint32_t img_h = 0, img_w = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0;
};
static int render_core(ngp_ctx *c, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma, uint32_t max_steps, float bg_color,
                       float *image, float *depth, float *weights_sum, int32_t *samples_per_ray, int64_t *stats, const RenderExtras &ex, void *stream);

int ngp_render(ngp_ctx *c, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma, uint32_t max_steps, float bg_color,
               float *image, float *depth, float *weights_sum, int32_t *samples_per_ray, int64_t *stats, void *stream) {
    return render_core(c, rays_o, rays_d, N, dt_gamma, max_steps, bg_color, image, depth, weights_sum, samples_per_ray, stats, RenderExtras(), stream);
}

int ngp_render_sample_rows(ngp_ctx *c, uint32_t N, uint64_t *rows) {
    NGP_REQUIRE(c != nullptr && rows != nullptr, "render_sample_rows: null argument");
    unsigned long long rpr = c->opt.rows_per_ray < 3u ? 3u : c->opt.rows_per_ray;
    const uint32_t cap = N <= c->capN ? c->capN : N + N / 8 + 1024;
    unsigned long long r = rpr * cap + 1024;
    *rows = r > 0x7fffffffull ? 0x7fffffffull : r;
    return 0;
}

int ngp_render_uq(ngp_ctx *c, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma, uint32_t max_steps, float bg_color,
                  float *image, float *depth, float *weights_sum, int32_t *samples_per_ray, int64_t *stats, float *aggregated_density,
                  float *sample_dump, uint64_t dump_rows, int32_t *ray_info, int64_t *dump_header, void *stream) {
    NGP_REQUIRE(c != nullptr, "render_uq: null ctx");
    RenderExtras ex;
    if (aggregated_density) { NGP_DEVPTR(aggregated_density); ex.aggregated_density = aggregated_density; }
    if (sample_dump) {
        NGP_DEVPTR(sample_dump); NGP_DEVPTR(ray_info); NGP_DEVPTR(dump_header);
        ex.sample_dump = sample_dump; ex.dump_rows = dump_rows > 0x7fffffffull ? 0x7fffffffu : (uint32_t)dump_rows;
        ex.ray_info = ray_info; ex.dump_header = dump_header;
    }
    if (N && ex.aggregated_density && N > c->capAgg) {
        free_all(c->agg[0], c->agg[1]);
        c->capAgg = 0;
        NGP_CUDA(cudaMalloc(&c->agg[0], sizeof(float) * (size_t)N));
        NGP_CUDA(cudaMalloc(&c->agg[1], sizeof(float) * (size_t)N));
        c->capAgg = N;
    }
    return render_core(c, rays_o, rays_d, N, dt_gamma, max_steps, bg_color, image, depth, weights_sum, samples_per_ray, stats, ex, stream);
}

static int render_core(ngp_ctx *c, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma, uint32_t max_steps, float bg_color,
                       float *image, float *depth, float *weights_sum, int32_t *samples_per_ray, int64_t *stats, const RenderExtras &ex, void *stream) {
    NGP_REQUIRE(c != nullptr && c->has_net && c->has_scene, "render: ngp_ctx_set_network / ngp_ctx_set_scene have not been called");
    NGP_REQUIRE(max_steps >= 1, "render: max_steps must be >= 1");
    NGP_REQUIRE(c->fast_levels, "render: a hash-grid level is neither a dense grid that fits its table nor a power-of-two hash table; "
                                "the persistent renderer is not built for that (use the stand-alone ops: NeRFRenderer.render(fused=False))");
    if (N == 0) return 0;
    NGP_DEVPTR(rays_o); NGP_DEVPTR(rays_d); NGP_DEVPTR(image); NGP_DEVPTR(depth); NGP_DEVPTR(weights_sum);
    if (int rc = ensure_workspace(c, N)) return rc;
    cudaStream_t st = as_stream(stream);
#ifdef NGP_WATCHDOG
    {
        const int zero = 0;
        NGP_CUDA(cudaMemcpyToSymbolAsync(tc::g_wd_abort, &zero, sizeof(int), 0, cudaMemcpyHostToDevice, st));
    }
#endif
    NGP_CUDA(cudaMemsetAsync(c->ctr, 0, sizeof(RenderCtr), st));
    RenderArgs ra{};
    ra.rays_o = rays_o; ra.rays_d = rays_d; ra.N = N;
    ra.poses = ex.poses; ra.gen_o = ex.poses ? const_cast<float *>(rays_o) : nullptr; ra.gen_d = ex.poses ? const_cast<float *>(rays_d) : nullptr;
    ra.img_w = ex.img_w; ra.img_hw = ex.img_h * ex.img_w; ra.fx = ex.fx; ra.fy = ex.fy; ra.cx = ex.cx; ra.cy = ex.cy; ra.min_near = c->min_near; ra.bg_color = bg_color; ra.max_steps = max_steps;
    for (int i = 0; i < 6; i++) ra.aabb[i] = c->aabb[i];
    ra.mc = make_march_const(c->nc.bound, dt_gamma, max_steps, c->cascade, c->grid_size, c->bitfield);
    ra.nears = c->nears; ra.fars = c->fars; ra.state[0] = c->state[0]; ra.state[1] = c->state[1]; ra.n_emit = c->n_emit;
    ra.alive[0] = c->alive[0]; ra.alive[1] = c->alive[1]; ra.smp = c->smp; ra.cap_rows = c->capRows;
    const RenderOptions &op = c->opt;
    ra.max_epoch = op.max_epoch;
    ra.force_kshift = op.kshift;
    ra.tile_balance = op.tile_balance;
    ra.t_div = op.t_div < 1u ? 1u : op.t_div;
    for (uint32_t k = 1; k <= (uint32_t)kMaxStep; k++) {
        const uint32_t n = (uint32_t)kMaxSamplesPerEpoch / k, t = (op.t_cap ? op.t_cap : (uint32_t)kMaxSamplesPerEpoch) / k;
        ra.cap_n[k - 1] = n < 1u ? 1u : n;
        ra.cap_t[k - 1] = t < 1u ? 1u : t;
    }
    ra.e0_cap = op.e0_cap;
    ra.rate0 = N / (op.rate0_div ? op.rate0_div : 1u) + 1u;
    ra.early_cap = op.early_cap < 1u ? 1u : op.early_cap; ra.early_steps = op.early_steps;
    ra.horizon_div = (float)(op.horizon_div < 1u ? 1u : op.horizon_div);
    ra.ctr = c->ctr; ra.image = image; ra.depth = depth; ra.weights_sum = weights_sum;
    ra.samples_per_ray = samples_per_ray; ra.stats = reinterpret_cast<long long *>(stats); ra.prof = c->prof;
    ra.agg[0] = c->agg[0]; ra.agg[1] = c->agg[1]; ra.agg_out = ex.aggregated_density;
    ra.dump = reinterpret_cast<float4 *>(ex.sample_dump); ra.dump_rows = ex.dump_rows;
    ra.ray_info = reinterpret_cast<int4 *>(ex.ray_info); ra.dump_hdr = reinterpret_cast<long long *>(ex.dump_header);
    ra.coarse = nullptr; ra.coarse_words = 0;
    size_t smem = c->render_smem;
    if (!op.no_coarse && c->coarse_words && smem + (NGP_COARSE_SMEM ? c->coarse_words * 4 : 0) <= c->render_smem_max) {
        const uint32_t n_blocks = c->coarse_words * 32;
        k_coarse_occupancy<<<div_up(n_blocks, 256), 256, 0, st>>>(reinterpret_cast<const unsigned long long *>(c->bitfield), n_blocks, c->coarse);
        if (int rc = after_launch("coarse_occupancy")) return rc;
        ra.coarse = c->coarse; ra.coarse_words = c->coarse_words;
        if (NGP_COARSE_SMEM) smem += c->coarse_words * 4;
    }
    NetConst nc = c->nc;
    void *args[] = {&nc, &ra};
    const void *fn = c->prof ? (const void *)k_render<true> : (const void *)k_render<false>;
    NGP_CUDA(cudaLaunchCooperativeKernel(fn, dim3(c->render_blocks), dim3(kRenderThreads), args, smem, st));
    return after_launch("render");
}

static int ensure_host_staging(ngp_ctx *c, uint32_t N, uint32_t F) {
    if (!c->h_error) NGP_CUDA(cudaMallocHost(&c->h_error, sizeof(uint32_t)));
    if (N > c->capHost) {   // all or nothing, as ensure_workspace
        free_all(c->d_o, c->d_d, c->d_img, c->d_depth, c->d_ws);
        c->capHost = 0;
        cudaError_t e = cudaSuccess;
        auto get = [&](float *&ptr, size_t n) { if (e == cudaSuccess) e = cudaMalloc(&ptr, sizeof(float) * n); };
        get(c->d_o, 3 * (size_t)N); get(c->d_d, 3 * (size_t)N); get(c->d_img, 3 * (size_t)N); get(c->d_depth, N); get(c->d_ws, N);
        if (e != cudaSuccess) {
            free_all(c->d_o, c->d_d, c->d_img, c->d_depth, c->d_ws);
            cudaGetLastError();
            set_error("render_host: staging for %u rays could not be allocated: %s", N, cudaGetErrorString(e));
            return 2;
        }
        c->capHost = N;
    }
    if (F > c->capPoses) {
        free_all(c->d_poses);
        c->capPoses = 0;
        NGP_CUDA(cudaMalloc(&c->d_poses, sizeof(float) * 16 * (size_t)F));
        c->capPoses = F;
    }
    return 0;
}

// image + depth back to the host, then the kernel's own consistency word: a render whose death histogram and live list disagreed
// (RenderCtr::error) is reported, not returned as an image
static int finish_host_render(ngp_ctx *c, uint32_t N, float *image_host, float *depth_host, cudaStream_t st, const char *what) {
    NGP_CUDA(cudaMemcpyAsync(image_host, c->d_img, sizeof(float) * 3 * (size_t)N, cudaMemcpyDeviceToHost, st));
    NGP_CUDA(cudaMemcpyAsync(depth_host, c->d_depth, sizeof(float) * (size_t)N, cudaMemcpyDeviceToHost, st));
    NGP_CUDA(cudaMemcpyAsync(c->h_error, &c->ctr->error, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    NGP_CUDA(cudaStreamSynchronize(st));
    NGP_REQUIRE(*c->h_error == 0u, "%s: the renderer's schedule check failed (death histogram and live list disagree); the image is poisoned with NaN", what);
    return 0;
}

int ngp_render_host(ngp_ctx *c, const float *rays_o_host, const float *rays_d_host, uint32_t N, float dt_gamma, uint32_t max_steps,
                    float bg_color, float *image_host, float *depth_host, void *stream) {
    NGP_REQUIRE(c != nullptr, "render_host: null ctx");
    NGP_REQUIRE(rays_o_host && rays_d_host && image_host && depth_host, "render_host: null host pointer");
    if (N == 0) return 0;
    if (int rc = ensure_host_staging(c, N, 0)) return rc;
    cudaStream_t st = as_stream(stream);
    NGP_CUDA(cudaMemcpyAsync(c->d_o, rays_o_host, sizeof(float) * 3 * (size_t)N, cudaMemcpyHostToDevice, st));
    NGP_CUDA(cudaMemcpyAsync(c->d_d, rays_d_host, sizeof(float) * 3 * (size_t)N, cudaMemcpyHostToDevice, st));
    if (int rc = ngp_render(c, c->d_o, c->d_d, N, dt_gamma, max_steps, bg_color, c->d_img, c->d_depth, c->d_ws, nullptr, nullptr, stream)) return rc;
    return finish_host_render(c, N, image_host, depth_host, st, "render_host");
}

int ngp_render_poses(ngp_ctx *c, const float *poses, uint32_t F, uint32_t H, uint32_t W, float fx, float fy, float cx, float cy, float dt_gamma,
                     uint32_t max_steps, float bg_color, float *image, float *depth, float *weights_sum, int32_t *samples_per_ray, int64_t *stats,
                     void *stream) {
    NGP_REQUIRE(c != nullptr, "render_poses: null ctx");
    const unsigned long long n = (unsigned long long)F * H * W;
    if (n == 0) return 0;
    NGP_REQUIRE(n <= 0xffffffffull, "render_poses: %llu rays exceed the 32-bit ray index of the renderer", n);
    NGP_REQUIRE(fx != 0.f && fy != 0.f, "render_poses: zero focal length");
    NGP_DEVPTR(poses);
    const uint32_t N = (uint32_t)n;
    if (int rc = ensure_host_staging(c, N, 0)) return rc;   // d_o / d_d: where the prologue leaves the rays for the later phases
    RenderExtras ex;
    ex.poses = poses; ex.img_h = H; ex.img_w = W; ex.fx = fx; ex.fy = fy; ex.cx = cx; ex.cy = cy;
    return render_core(c, c->d_o, c->d_d, N, dt_gamma, max_steps, bg_color, image, depth, weights_sum, samples_per_ray, stats, ex, stream);
}

int ngp_render_poses_host(ngp_ctx *c, const float *poses_host, uint32_t F, uint32_t H, uint32_t W, float fx, float fy, float cx, float cy,
                          float dt_gamma, uint32_t max_steps, float bg_color, float *image_host, float *depth_host, void *stream) {
    NGP_REQUIRE(c != nullptr, "render_poses_host: null ctx");
    NGP_REQUIRE(poses_host && image_host && depth_host, "render_poses_host: null host pointer");
    const unsigned long long n = (unsigned long long)F * H * W;
    if (n == 0) return 0;
    NGP_REQUIRE(n <= 0xffffffffull, "render_poses_host: %llu rays exceed the 32-bit ray index of the renderer", n);
    const uint32_t N = (uint32_t)n;
    if (int rc = ensure_host_staging(c, N, F)) return rc;
    cudaStream_t st = as_stream(stream);
    NGP_CUDA(cudaMemcpyAsync(c->d_poses, poses_host, sizeof(float) * 16 * (size_t)F, cudaMemcpyHostToDevice, st));   // 64 B per frame
    NGP_REQUIRE(fx != 0.f && fy != 0.f, "render_poses_host: zero focal length");
    RenderExtras ex;   // the renderer's prologue generates the rays (nerf/utils.py:52-116) into the staging buffers itself
    ex.poses = c->d_poses; ex.img_h = H; ex.img_w = W; ex.fx = fx; ex.fy = fy; ex.cx = cx; ex.cy = cy;
    if (int rc = render_core(c, c->d_o, c->d_d, N, dt_gamma, max_steps, bg_color, c->d_img, c->d_depth, c->d_ws, nullptr, nullptr, ex, stream)) return rc;
    return finish_host_render(c, N, image_host, depth_host, st, "render_poses_host");
}

// Tunables of the renderer's epoch policy, per context (no process-wide state).  Unknown names are an error.
int ngp_ctx_set_option(ngp_ctx *c, const char *name, uint64_t value) {
    NGP_REQUIRE(c != nullptr && name != nullptr, "ctx_set_option: null argument");
    const struct { const char *name; uint32_t RenderOptions::*field; } table[] = {
        {"rows_per_ray", &RenderOptions::rows_per_ray}, {"max_epoch", &RenderOptions::max_epoch}, {"kshift", &RenderOptions::kshift}, {"tile_balance", &RenderOptions::tile_balance},
        {"t_div", &RenderOptions::t_div}, {"t_cap", &RenderOptions::t_cap}, {"e0_cap", &RenderOptions::e0_cap},
        {"rate0_div", &RenderOptions::rate0_div}, {"no_coarse", &RenderOptions::no_coarse}, {"early_cap", &RenderOptions::early_cap},
        {"early_steps", &RenderOptions::early_steps}, {"horizon_div", &RenderOptions::horizon_div},
    };
    for (const auto &t : table)
        if (strcmp(t.name, name) == 0) {
            NGP_REQUIRE(value <= 0xffffffffull, "ctx_set_option: %s = %llu does not fit 32 bits", name, (unsigned long long)value);
            c->opt.*(t.field) = (uint32_t)value;
            return 0;
        }
    NGP_REQUIRE(false, "ctx_set_option: unknown option '%s' (rows_per_ray, max_epoch, kshift, tile_balance, t_div, t_cap, e0_cap, rate0_div, no_coarse, early_cap, early_steps, horizon_div)", name);
    return 1;
}

}  // extern "C"
