// fused.cu -- the B200 fast path (no upstream counterpart):
//   ngp_network_forward : hash-grid encode -> sigma MLP -> SH -> colour MLP -> exp/sigmoid, one kernel, nothing but
//                         xyz/dir in and sigma/rgb out touches HBM        (nerf/network_ff.py:51-74)
//   ngp_render          : NeRFRenderer.run_cuda's whole inference loop (nerf/renderer.py:261-284,329-386) as ONE
//                         persistent cooperative launch, one 16-warp CTA per SM.  Warps are autonomous: a warp pulls a
//                         group of 32 live rays and, per sample, runs {march_one -> hash-grid gather -> both MLPs on
//                         tensor cores -> composite} with lane r owning ray r / row r; nothing but o/d in and the ray
//                         accumulators out touches HBM.  The grid meets once per iteration because the reference's
//                         `n_step = clamp(N // n_alive, 1, 8)` schedule is global and the per-ray sample counts (an
//                         integer output we must match) depend on it.
#include <cooperative_groups.h>
#include <float.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>

#include "common.cuh"
#include "device_math.cuh"
#include "scan.cuh"

namespace cg = cooperative_groups;

namespace ngp {

constexpr int kTile = 128;        // sample rows per tile == threads per CTA
constexpr int kEncStride = 40;    // halves per staged encoding row (32 + 8 pad: conflict-free fragment loads)
constexpr int kShStride = 24;     // halves per staged SH row (16 + 8 pad)

struct NetConst {
    const __half2 *table;
    const uint2 *wpack;            // global, fragment order: sigma net then colour net
    uint32_t wpack_count;          // uint2 elements
    uint32_t nh_sigma, nh_color;   // hidden (64x64) matmuls per net = num_layers - 1
    float bound, two_bound, density_scale;
    LevelConst lv[kMaxLevels];
};

struct RenderCtr {
    uint32_t n_alive[3];    // live rays of iteration k at [k % 3]
    uint32_t cursor[3];     // MARCH phase: next unclaimed entry of the live list
    uint32_t ticket[3];     // EVALUATE phase: next unclaimed 32-ray tile
    uint32_t pad;
    unsigned long long samples;
};

// ---------------------------------------------------------------------------------------------- device pieces

__device__ __forceinline__ void copy_wpack_to_smem(const NetConst &nc, uint2 *s_w) {
    const uint4 *src = reinterpret_cast<const uint4 *>(nc.wpack);
    uint4 *dst = reinterpret_cast<uint4 *>(s_w);
    for (uint32_t i = threadIdx.x; i < nc.wpack_count / 2; i += blockDim.x) dst[i] = __ldg(src + i);
}

// 16-level hash-grid encoding of one point into 16 half2 (= 4 x uint4) of shared memory.
__device__ __forceinline__ void encode_point(const NetConst &nc, float x, float y, float z, uint4 *dst) {
    // grid.py:144  inputs = (inputs + bound) / (2 * bound)
    const float u0 = __fdiv_rn(__fadd_rn(x, nc.bound), nc.two_bound);
    const float u1 = __fdiv_rn(__fadd_rn(y, nc.bound), nc.two_bound);
    const float u2 = __fdiv_rn(__fadd_rn(z, nc.bound), nc.two_bound);
    if ((u0 < 0 || u0 > 1) || (u1 < 0 || u1 > 1) || (u2 < 0 || u2 > 1)) {  // gridencoder.cu:99-123
        dst[0] = dst[1] = dst[2] = dst[3] = make_uint4(0, 0, 0, 0);
        return;
    }
    // 4 levels (32 independent 4-byte gathers) per trip; the trip loop is deliberately NOT unrolled: the renderer's 16 warps
    // sit in different phases, and a 4x smaller gather body keeps them inside the instruction cache
#pragma unroll 1
    for (int grp = 0; grp < 4; grp++) {
        float f[4][3];
        __half2 v[4][8];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const LevelConst &lc = nc.lv[grp * 4 + j];
            const float scale = lc.scale;
            float p0 = __fmaf_rn(u0, scale, 0.5f), p1 = __fmaf_rn(u1, scale, 0.5f), p2 = __fmaf_rn(u2, scale, 0.5f);
            const uint32_t g0 = (uint32_t)floorf(p0), g1 = (uint32_t)floorf(p1), g2 = (uint32_t)floorf(p2);
            f[j][0] = __fsub_rn(p0, (float)g0); f[j][1] = __fsub_rn(p1, (float)g1); f[j][2] = __fsub_rn(p2, (float)g2);
            const __half2 *tl = nc.table + lc.offset;
            uint32_t row[8];
            cell_rows(lc, g0, g1, g2, row);
#pragma unroll
            for (int k = 0; k < 8; k++) v[j][k] = __ldg(tl + row[k]);
        }
        uint32_t packed[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float p0 = f[j][0], p1 = f[j][1], p2 = f[j][2];
            const float q0 = 1.0f - p0, q1 = 1.0f - p1, q2 = 1.0f - p2;
            const float w00 = q0 * q1, w10 = p0 * q1, w01 = q0 * p1, w11 = p0 * p1;   // (w_x * w_y) * w_z, gridencoder.cu:148-160
            const float w[8] = {w00 * q2, w10 * q2, w01 * q2, w11 * q2, w00 * p2, w10 * p2, w01 * p2, w11 * p2};
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const float2 g = __half22float2(v[j][k]);
                a0 = fmaf(w[k], g.x, a0);
                a1 = fmaf(w[k], g.y, a1);
            }
            packed[j] = pack_half2(a0, a1);
        }
        dst[grp] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
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

// ---------------------------------------------------------------------------------------------- persistent renderer
// NeRFRenderer.run_cuda's inference loop (nerf/renderer.py:329-376) as ONE persistent cooperative launch, one 16-warp CTA per
// SM.  The reference's global schedule `n_step = clamp(N // n_alive, 1, 8)` (renderer.py:357) is kept exactly -- per-ray sample
// counts and the rays_t recurrence depend on it -- so the grid meets twice per iteration:
//
//   MARCH     lane-level dynamic scheduling: every LANE owns one live ray, probes the occupancy bitfield one cell per trip in
//             lockstep with the other 31 lanes of its warp (so the warp stays converged whatever mix of occupied / empty cells
//             its rays cross), and the moment its ray has its n_step samples (or leaves the box) the lane takes the next live
//             ray from the list (warp ballot + prefix, one atomic per warp per refill).  Samples go to an L2-resident buffer in
//             the reference's row order (row = n * n_step + s, raymarching.cu:732-734): 20 B per sample, no dirs.
//   EVALUATE  warp w pulls tiles of 32 live rays = 32 * n_step consecutive sample rows (consecutive rows are consecutive samples
//   + COMPOSITE  of one ray and share hash-grid cells): 16-level gather -> sigma MLP -> SH + colour MLP, 32 rows at a time on
//             tensor cores with the weights resident in shared memory; then lane r composites ray r front to back in registers
//             with early termination (raymarching.cu:828-913) and the survivors are appended to the next live list with one
//             ballot + one atomicAdd per warp (list order is not the reference's, which no per-ray output depends on; the
//             stand-alone compact_alive op keeps the stable order).
// xyz / deltas make one trip through L2; dirs / sigmas / rgbs never exist in memory.
constexpr int kRenderThreads = 512;
constexpr int kRenderWarps = kRenderThreads / 32;
constexpr int kMaxStep = 8;                    // renderer.py:357
constexpr int kTileRows = 32 * kMaxStep;       // sample rows of one 32-ray tile
constexpr int kRefillLanes = 8;                // a warp refills its idle lanes from the live list once this many are idle

struct WarpSmem {
    float sigma[kTileRows], rg[kTileRows], b[kTileRows];   // network outputs per row (rg = half2 bits: rgb are half-rounded)
    float dt[kTileRows], dl[kTileRows];                    // deltas[:, 0], deltas[:, 1]
    __half enc[32 * kEncStride];               // A-operand staging of the current 32 rows; first 16 B of a row = (sigma, r, g, b) after the MLPs
    __half sh[32 * kShStride];                 // degree-4 SH of lane r's ray direction
};
static_assert(sizeof(WarpSmem) % 16 == 0 && offsetof(WarpSmem, enc) % 16 == 0 && offsetof(WarpSmem, sh) % 16 == 0, "WarpSmem alignment");

struct RenderArgs {
    const float *rays_o, *rays_d;
    uint32_t N;
    float min_near, bg_color;
    uint32_t max_steps;
    float aabb[6];
    MarchConst mc;
    float *nears, *fars, *rays_t;
    int *alive[2];
    const uint32_t *coarse;   // [coarse_words] 4x4x4-block occupancy bits (Morton order) or NULL
    uint32_t coarse_words;
    float4 *smp_a;     // [N] (x, y, z, dt) per sample row of the current iteration
    float *smp_b;      // [N] t_new - last_t
    RenderCtr *ctr;
    float *image, *depth, *weights_sum;
    int *samples_per_ray;
    long long *stats;
    long long *prof;   // optional [8]: warp-cycles per phase (diagnostics, ngp_ctx_phase_cycles)
};

// MARCH phase of one iteration for one warp (see the header comment).  Returns the samples this lane emitted.
__device__ __forceinline__ uint32_t march_phase(const RenderArgs &ra, const uint32_t *s_coarse, int lane, uint32_t lt_mask,
                                                const int *__restrict__ alive_in, uint32_t n_alive, uint32_t n_step, uint32_t *cursor) {
    bool has = false, list_done = false;
    uint32_t n = 0, emitted = 0, total = 0;
    int index = 0;
    float ox = 0.f, oy = 0.f, oz = 0.f, tm = 0.f, last_t = 0.f, far = 0.f;
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
                    index = __ldcg(alive_in + n);
                    const float *o = ra.rays_o + (size_t)index * 3, *d = ra.rays_d + (size_t)index * 3;
                    ox = __ldg(o); oy = __ldg(o + 1); oz = __ldg(o + 2);
                    rd = make_ray_dir(__ldg(d), __ldg(d + 1), __ldg(d + 2));
                    tm = __ldcg(ra.rays_t + index);
                    last_t = tm;
                    far = __ldcg(ra.fars + index);
                    emitted = 0;
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
            if (tm < far && march_probe(ra.mc, ox, oy, oz, rd, tm, last_t, x, y, z, dt, dl, s_coarse)) {
                const size_t row = (size_t)n * n_step + emitted;
                __stcg(ra.smp_a + row, make_float4(x, y, z, dt));
                __stcg(ra.smp_b + row, dl);
                emitted++;
            }
            if (emitted == n_step || !(tm < far)) {      // chunk complete (raymarching.cu:752) or ray exhausted
                for (uint32_t s = emitted; s < n_step; s++) {   // un-emitted rows read as delta == 0 (raymarching.py:329-331)
                    const size_t row = (size_t)n * n_step + s;
                    __stcg(ra.smp_a + row, make_float4(0.f, 0.f, 0.f, 0.f));
                    __stcg(ra.smp_b + row, 0.f);
                }
                if (ra.samples_per_ray && emitted) ra.samples_per_ray[index] += (int)emitted;
                total += emitted;
                has = false;
            }
        }
        __syncwarp();
    }
    return total;
}

// EVALUATE + COMPOSITE phase of one iteration for one warp (see the header comment).
template <bool kProf>
__device__ __forceinline__ void eval_phase(const NetConst &nc, const RenderArgs &ra, const uint2 *__restrict__ s_w, WarpSmem &ws, int lane,
                                           uint32_t lt_mask, const int *__restrict__ alive_in, int *__restrict__ alive_out, uint32_t n_alive,
                                           uint32_t n_step, uint32_t *ticket, uint32_t *next_count, long long (&pc)[8]) {
    const uint32_t n_tiles = (n_alive + 31) / 32;
    long long c0 = kProf ? clock64() : 0, c1;
#define NGP_PROF(k) if (kProf) { c1 = clock64(); pc[k] += c1 - c0; c0 = c1; }
    uint32_t next_tile = 0;
    if (lane == 0) next_tile = atomicAdd(ticket, 1u);
    for (;;) {
        const uint32_t tile = __shfl_sync(0xffffffffu, next_tile, 0);
        if (tile >= n_tiles) break;
        if (lane == 0) next_tile = atomicAdd(ticket, 1u);   // claimed now, needed a whole tile later: its latency is hidden
        const uint32_t n = tile * 32 + lane;
        const bool has_ray = n < n_alive;
        int index = 0;
        RayAcc a = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (has_ray) {
            index = __ldcg(alive_in + n);
            const float *d = ra.rays_d + (size_t)index * 3;
            uint4 *srow = reinterpret_cast<uint4 *>(ws.sh + lane * kShStride);
            sh_to_half(__ldg(d), __ldg(d + 1), __ldg(d + 2), srow[0], srow[1]);   // the direction is constant along the ray
            a.t = __ldcg(ra.rays_t + index); a.ws = __ldcg(ra.weights_sum + index); a.d = __ldcg(ra.depth + index);
            a.r = __ldcg(ra.image + (size_t)index * 3); a.g = __ldcg(ra.image + (size_t)index * 3 + 1); a.b = __ldcg(ra.image + (size_t)index * 3 + 2);
        }
        const uint32_t tile_rays = min(32u, n_alive - tile * 32);
        const uint32_t rows = tile_rays * n_step;
        const size_t row_base = (size_t)tile * 32 * n_step;
        NGP_PROF(0)
        for (uint32_t row0 = 0; row0 < rows; row0 += 32) {
            const uint32_t r = row0 + lane;
            float4 sa = make_float4(0.f, 0.f, 0.f, 0.f);
            float sb = 0.f;
            if (r < rows) {
                sa = __ldcg(ra.smp_a + row_base + r);
                sb = __ldcg(ra.smp_b + row_base + r);
            }
            ws.dt[r] = sa.w; ws.dl[r] = sb;
            const bool live = sa.w != 0.f;                       // delta == 0 is the reference's end-of-ray sentinel
            if (live) encode_point(nc, sa.x, sa.y, sa.z, reinterpret_cast<uint4 *>(ws.enc + lane * kEncStride));
            __syncwarp();
            NGP_PROF(3)
            if (__any_sync(0xffffffffu, live)) {
                mlp_rows32(nc, s_w, ws.enc,
                           [&](int rr) {
                               const uint32_t ray = min((row0 + (uint32_t)rr) / n_step, 31u);
                               return ws.sh + ray * kShStride;
                           },
                           0, lane, [&](int rr, float sg) { reinterpret_cast<float *>(ws.enc + rr * kEncStride)[0] = sg; },
                           [&](int rr, int t, float v0, float v1) {
                               float *o = reinterpret_cast<float *>(ws.enc + rr * kEncStride);
                               if (t == 0) { o[1] = v0; o[2] = v1; } else { o[3] = v0; }
                           });
                __syncwarp();
                if (live) {
                    const float4 o = *reinterpret_cast<const float4 *>(ws.enc + lane * kEncStride);
                    ws.sigma[r] = o.x; ws.rg[r] = __uint_as_float(pack_half2(o.y, o.z)); ws.b[r] = o.w;   // rgb are half-rounded: exact
                }
            }
            __syncwarp();
            NGP_PROF(4)
        }
        // ---- composite (raymarching.cu:828-913): lane r <-> ray r, rows r * n_step .. + n_step - 1
        bool survives = false;
        if (has_ray) {
            uint32_t s = 0;
            while (s < n_step) {
                const uint32_t r = lane * n_step + s;
                const float dt = ws.dt[r];
                if (dt == 0.f) break;
                const uint32_t rg = __float_as_uint(ws.rg[r]);
                const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&rg));
                if (composite_sample(a, ws.sigma[r], c.x, c.y, ws.b[r], dt, ws.dl[r])) break;
                s++;
            }
            survives = s == n_step;
            if (survives) __stcg(ra.rays_t + index, a.t);                       // raymarching.cu:905-911
            __stcg(ra.weights_sum + index, a.ws); __stcg(ra.depth + index, a.d);
            __stcg(ra.image + (size_t)index * 3, a.r); __stcg(ra.image + (size_t)index * 3 + 1, a.g); __stcg(ra.image + (size_t)index * 3 + 2, a.b);
        }
        const uint32_t ballot = __ballot_sync(0xffffffffu, survives);   // renderer.py:369
        if (ballot) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(next_count, (uint32_t)__popc(ballot));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (survives) __stcg(alive_out + base + __popc(ballot & lt_mask), index);
        }
        __syncwarp();   // ws is rewritten by the next tile
        NGP_PROF(5)
    }
#undef NGP_PROF
}

template <bool kProf>
__global__ void __launch_bounds__(kRenderThreads, 1) k_render(NetConst nc, RenderArgs ra) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint2 *s_w = reinterpret_cast<uint2 *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    WarpSmem &ws = reinterpret_cast<WarpSmem *>(s_w + nc.wpack_count)[warp];
    uint32_t *s_coarse = ra.coarse ? reinterpret_cast<uint32_t *>(reinterpret_cast<WarpSmem *>(s_w + nc.wpack_count) + kRenderWarps) : nullptr;
    const uint32_t N = ra.N;
    const uint32_t lt_mask = (1u << lane) - 1u;
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long k_start = kProf ? clock64() : 0;
    copy_wpack_to_smem(nc, s_w);
    if (s_coarse)
        for (uint32_t i = tid; i < ra.coarse_words; i += blockDim.x) s_coarse[i] = __ldg(ra.coarse + i);

    // ---- prologue: near/far, ray state, first live list (renderer.py:273-274, 335-343)
    for (uint32_t n = blockIdx.x * blockDim.x + tid; n < N; n += gridDim.x * blockDim.x) {
        const float *o = ra.rays_o + (size_t)n * 3, *d = ra.rays_d + (size_t)n * 3;
        float near, far;
        if (!slab_near_far(__ldg(o), __ldg(o + 1), __ldg(o + 2), __ldg(d), __ldg(d + 1), __ldg(d + 2), ra.aabb, ra.min_near, near, far))
            near = far = FLT_MAX;
        ra.nears[n] = near; ra.fars[n] = far; ra.rays_t[n] = near;
        ra.alive[0][n] = (int)n;
        ra.weights_sum[n] = 0.f; ra.depth[n] = 0.f;
        ra.image[(size_t)n * 3] = 0.f; ra.image[(size_t)n * 3 + 1] = 0.f; ra.image[(size_t)n * 3 + 2] = 0.f;
        if (ra.samples_per_ray) ra.samples_per_ray[n] = 0;
    }
    if (blockIdx.x == 0 && tid == 0) ra.ctr->n_alive[0] = N;   // all counters were zeroed by the host before the launch
    grid.sync();

    uint32_t step = 0, iter = 0, tot_emitted = 0;
    while (step < ra.max_steps) {
        const uint32_t n_alive = __ldcg(&ra.ctr->n_alive[iter % 3]);
        if (n_alive == 0) break;
        uint32_t n_step = N / n_alive;                                  // renderer.py:357
        n_step = n_step > kMaxStep ? kMaxStep : (n_step < 1 ? 1 : n_step);
        const int *alive_in = ra.alive[iter & 1];
        int *alive_out = ra.alive[(iter & 1) ^ 1];
        const uint32_t slot = iter % 3, nslot = (iter + 1) % 3;
        if (blockIdx.x == 0 && tid == 0) {   // none of these words is touched by anyone else during this iteration
            ra.ctr->cursor[nslot] = 0;
            ra.ctr->ticket[nslot] = 0;
            ra.ctr->n_alive[(iter + 2) % 3] = 0;
        }
        long long g0 = kProf ? clock64() : 0;
        tot_emitted += march_phase(ra, s_coarse, lane, lt_mask, alive_in, n_alive, n_step, &ra.ctr->cursor[slot]);
        if (kProf) { const long long g1 = clock64(); pc[1] += g1 - g0; g0 = g1; }
        grid.sync();
        if (kProf) pc[2] += clock64() - g0;
        eval_phase<kProf>(nc, ra, s_w, ws, lane, lt_mask, alive_in, alive_out, n_alive, n_step, &ra.ctr->ticket[slot], &ra.ctr->n_alive[nslot], pc);
        if (kProf) g0 = clock64();
        grid.sync();
        if (kProf) pc[6] += clock64() - g0;
        step += n_step;
        iter++;
    }

    // ---- epilogue: background + depth normalisation (renderer.py:375-376); all accumulators are final after the last barrier
    for (uint32_t n = blockIdx.x * blockDim.x + tid; n < N; n += gridDim.x * blockDim.x) {
        const float ws_n = __ldcg(ra.weights_sum + n);
        const float k = __fmul_rn(__fsub_rn(1.0f, ws_n), ra.bg_color);
        float *im = ra.image + (size_t)n * 3;
        im[0] = __fadd_rn(__ldcg(im), k); im[1] = __fadd_rn(__ldcg(im + 1), k); im[2] = __fadd_rn(__ldcg(im + 2), k);
        const float near = ra.nears[n], far = ra.fars[n];
        const float dd = __fsub_rn(__ldcg(ra.depth + n), near);
        ra.depth[n] = __fdiv_rn(dd < 0.f ? 0.f : dd, __fsub_rn(far, near));   // 0/0 = NaN for misses, like the reference
    }
    if (kProf && lane == 0) {
        pc[7] = clock64() - k_start;
        pc[0] += pc[7] - (pc[0] + pc[1] + pc[2] + pc[3] + pc[4] + pc[5] + pc[6]);   // + prologue, epilogue, loop control
        for (int k = 0; k < 8; k++) atomicAdd(reinterpret_cast<unsigned long long *>(ra.prof + k), (unsigned long long)pc[k]);
    }
    if (ra.stats) {
        const uint32_t wemit = warp_sum(tot_emitted);
        if (lane == 0) atomicAdd(&ra.ctr->samples, (unsigned long long)wemit);
        grid.sync();   // every warp's counters are in
        if (blockIdx.x == 0 && tid == 0) {
            ra.stats[0] = (long long)iter;
            ra.stats[1] = (long long)__ldcg(&ra.ctr->samples);
            ra.stats[2] = 0;
            ra.stats[3] = 2 * (long long)iter;   // grid-wide barriers inside the loop
        }
    }
}

}  // namespace ngp

using namespace ngp;

// ---------------------------------------------------------------------------------------------- context
struct ngp_ctx {
    int device = 0, sm_count = 0;
    size_t max_smem_optin = 0;
    bool has_net = false, has_scene = false;
    NetConst nc{};
    uint2 *wpack = nullptr;
    // scene
    const uint8_t *bitfield = nullptr;
    uint32_t cascade = 1, grid_size = 128;
    float aabb[6] = {-1, -1, -1, 1, 1, 1};
    float min_near = 0.2f;
    // render workspace
    uint32_t capN = 0;
    float *nears = nullptr, *fars = nullptr, *rays_t = nullptr, *smp_b = nullptr;
    float4 *smp_a = nullptr;
    int *alive[2] = {nullptr, nullptr};
    RenderCtr *ctr = nullptr;
    // staging for ngp_render_host
    uint32_t capHost = 0;
    float *d_o = nullptr, *d_d = nullptr, *d_img = nullptr, *d_depth = nullptr, *d_ws = nullptr;
    uint32_t *coarse = nullptr;   // 4x4x4-block occupancy bits, rebuilt from the borrowed bitfield at every ngp_render
    uint32_t coarse_words = 0;
    long long *prof = nullptr;    // device [8], allocated by ngp_ctx_enable_profile
    int render_blocks = 0, fwd_blocks = 0;
    size_t render_smem = 0, fwd_smem = 0;
};

static size_t fwd_smem_bytes(const NetConst &nc) { return (size_t)nc.wpack_count * 8 + kTile * (kEncStride + kShStride) * 2; }
static size_t render_smem_bytes(const NetConst &nc) { return (size_t)nc.wpack_count * 8 + (size_t)kRenderWarps * sizeof(WarpSmem); }

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
    cudaFree(c->wpack); cudaFree(c->nears); cudaFree(c->fars); cudaFree(c->rays_t); cudaFree(c->alive[0]); cudaFree(c->alive[1]); cudaFree(c->smp_a); cudaFree(c->smp_b);
    cudaFree(c->ctr); cudaFree(c->prof); cudaFree(c->coarse); cudaFree(c->d_o); cudaFree(c->d_d); cudaFree(c->d_img); cudaFree(c->d_depth); cudaFree(c->d_ws);
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
    const uint32_t count = 512 + nc.nh_sigma * 1024 + 256 + 512 + nc.nh_color * 1024 + 256;
    if (c->wpack) { cudaFree(c->wpack); c->wpack = nullptr; }
    NGP_CUDA(cudaMalloc(&c->wpack, sizeof(uint2) * count));
    k_pack_network<<<1, 256, 0, st>>>(reinterpret_cast<const __half *>(w_sigma_f16), nc.nh_sigma, reinterpret_cast<const __half *>(w_color_f16),
                                      nc.nh_color, c->wpack);
    if (int rc = after_launch("pack_network")) return rc;
    nc.wpack = c->wpack;
    nc.wpack_count = count;
    c->fwd_smem = fwd_smem_bytes(nc);
    c->render_smem = render_smem_bytes(nc);
    NGP_CUDA(cudaFuncSetAttribute(k_network_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->fwd_smem));
    NGP_CUDA(cudaFuncSetAttribute(k_render<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->max_smem_optin));
    NGP_CUDA(cudaFuncSetAttribute(k_render<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->max_smem_optin));
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

static int ensure_workspace(ngp_ctx *c, uint32_t N) {
    if (N <= c->capN) return 0;
    cudaFree(c->nears); cudaFree(c->fars); cudaFree(c->rays_t); cudaFree(c->alive[0]); cudaFree(c->alive[1]); cudaFree(c->smp_a); cudaFree(c->smp_b);
    c->capN = 0;
    const uint32_t cap = N + N / 8 + 1024;
    NGP_CUDA(cudaMalloc(&c->nears, sizeof(float) * cap));
    NGP_CUDA(cudaMalloc(&c->fars, sizeof(float) * cap));
    NGP_CUDA(cudaMalloc(&c->rays_t, sizeof(float) * cap));
    NGP_CUDA(cudaMalloc(&c->alive[0], sizeof(int) * cap));
    NGP_CUDA(cudaMalloc(&c->alive[1], sizeof(int) * cap));
    NGP_CUDA(cudaMalloc(&c->smp_a, sizeof(float4) * cap));   // n_alive * n_step <= N rows per iteration
    NGP_CUDA(cudaMalloc(&c->smp_b, sizeof(float) * cap));
    c->capN = cap;
    return 0;
}

int ngp_render(ngp_ctx *c, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma, uint32_t max_steps, float bg_color,
               float *image, float *depth, float *weights_sum, int32_t *samples_per_ray, int64_t *stats, void *stream) {
    NGP_REQUIRE(c != nullptr && c->has_net && c->has_scene, "render: ngp_ctx_set_network / ngp_ctx_set_scene have not been called");
    NGP_REQUIRE(max_steps >= 1, "render: max_steps must be >= 1");
    if (N == 0) return 0;
    NGP_DEVPTR(rays_o); NGP_DEVPTR(rays_d); NGP_DEVPTR(image); NGP_DEVPTR(depth); NGP_DEVPTR(weights_sum);
    if (int rc = ensure_workspace(c, N)) return rc;
    cudaStream_t st = as_stream(stream);
    NGP_CUDA(cudaMemsetAsync(c->ctr, 0, sizeof(RenderCtr), st));
    RenderArgs ra{};
    ra.rays_o = rays_o; ra.rays_d = rays_d; ra.N = N; ra.min_near = c->min_near; ra.bg_color = bg_color; ra.max_steps = max_steps;
    for (int i = 0; i < 6; i++) ra.aabb[i] = c->aabb[i];
    ra.mc = make_march_const(c->nc.bound, dt_gamma, max_steps, c->cascade, c->grid_size, c->bitfield);
    ra.nears = c->nears; ra.fars = c->fars; ra.rays_t = c->rays_t; ra.alive[0] = c->alive[0]; ra.alive[1] = c->alive[1]; ra.smp_a = c->smp_a; ra.smp_b = c->smp_b;
    ra.ctr = c->ctr; ra.image = image; ra.depth = depth; ra.weights_sum = weights_sum;
    ra.samples_per_ray = samples_per_ray; ra.stats = reinterpret_cast<long long *>(stats); ra.prof = c->prof;
    ra.coarse = nullptr; ra.coarse_words = 0;
    size_t smem = c->render_smem;
    static const bool no_coarse = getenv("NGP_NO_COARSE") != nullptr;   // A/B switch for profiling only
    if (!no_coarse && c->coarse_words && smem + c->coarse_words * 4 <= c->max_smem_optin) {
        const uint32_t n_blocks = c->coarse_words * 32;
        k_coarse_occupancy<<<div_up(n_blocks, 256), 256, 0, st>>>(reinterpret_cast<const unsigned long long *>(c->bitfield), n_blocks, c->coarse);
        if (int rc = after_launch("coarse_occupancy")) return rc;
        ra.coarse = c->coarse; ra.coarse_words = c->coarse_words;
        smem += c->coarse_words * 4;
    }
    NetConst nc = c->nc;
    void *args[] = {&nc, &ra};
    NGP_CUDA(cudaLaunchCooperativeKernel(c->prof ? (const void *)k_render<true> : (const void *)k_render<false>, dim3(c->render_blocks), dim3(kRenderThreads), args, smem, st));
    return after_launch("render");
}

int ngp_render_host(ngp_ctx *c, const float *rays_o_host, const float *rays_d_host, uint32_t N, float dt_gamma, uint32_t max_steps,
                    float bg_color, float *image_host, float *depth_host, void *stream) {
    NGP_REQUIRE(c != nullptr, "render_host: null ctx");
    NGP_REQUIRE(rays_o_host && rays_d_host && image_host && depth_host, "render_host: null host pointer");
    if (N == 0) return 0;
    if (N > c->capHost) {
        cudaFree(c->d_o); cudaFree(c->d_d); cudaFree(c->d_img); cudaFree(c->d_depth); cudaFree(c->d_ws);
        c->capHost = 0;
        NGP_CUDA(cudaMalloc(&c->d_o, sizeof(float) * 3 * (size_t)N));
        NGP_CUDA(cudaMalloc(&c->d_d, sizeof(float) * 3 * (size_t)N));
        NGP_CUDA(cudaMalloc(&c->d_img, sizeof(float) * 3 * (size_t)N));
        NGP_CUDA(cudaMalloc(&c->d_depth, sizeof(float) * (size_t)N));
        NGP_CUDA(cudaMalloc(&c->d_ws, sizeof(float) * (size_t)N));
        c->capHost = N;
    }
    cudaStream_t st = as_stream(stream);
    NGP_CUDA(cudaMemcpyAsync(c->d_o, rays_o_host, sizeof(float) * 3 * (size_t)N, cudaMemcpyHostToDevice, st));
    NGP_CUDA(cudaMemcpyAsync(c->d_d, rays_d_host, sizeof(float) * 3 * (size_t)N, cudaMemcpyHostToDevice, st));
    if (int rc = ngp_render(c, c->d_o, c->d_d, N, dt_gamma, max_steps, bg_color, c->d_img, c->d_depth, c->d_ws, nullptr, nullptr, stream)) return rc;
    NGP_CUDA(cudaMemcpyAsync(image_host, c->d_img, sizeof(float) * 3 * (size_t)N, cudaMemcpyDeviceToHost, st));
    NGP_CUDA(cudaMemcpyAsync(depth_host, c->d_depth, sizeof(float) * (size_t)N, cudaMemcpyDeviceToHost, st));
    NGP_CUDA(cudaStreamSynchronize(st));
    return 0;
}

}  // extern "C"
