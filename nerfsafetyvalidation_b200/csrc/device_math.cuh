// device_math.cuh -- device functions shared by the stand-alone ops and the fused renderer.
//
// Bit-exactness contract: march_ray() and composite_sample() reproduce, operation for operation, the
// rounding sequence nvcc 12.9 emitted for the reference kernels (raymarching/src/raymarching.cu:706-814,
// :828-913; checked against `cuobjdump -sass` of the reference rebuilt for sm_100a).  Every contraction is
// spelled with __fmaf_rn/__fmul_rn/__fadd_rn so that this file does not depend on -fmad.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace ngp {

// ------------------------------------------------------------------------------------------------
// Morton helpers (raymarching.cu:58-83)
__host__ __device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__host__ __device__ __forceinline__ uint32_t morton3d(uint32_t x, uint32_t y, uint32_t z) {
    return expand_bits10(x) | (expand_bits10(y) << 1) | (expand_bits10(z) << 2);
}
__host__ __device__ __forceinline__ uint32_t compact_bits10(uint32_t x) {
    x &= 0x49249249u;
    x = (x | (x >> 2)) & 0xc30c30c3u;
    x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu;
    x = (x | (x >> 16)) & 0x0000ffffu;
    return x;
}

// ------------------------------------------------------------------------------------------------
// Ray / AABB slab test (raymarching.cu:93-147).  Returns false for a miss (caller writes FLT_MAX twice).
__device__ __forceinline__ bool slab_near_far(float ox, float oy, float oz, float dx, float dy, float dz,
                                              const float *__restrict__ aabb, float min_near, float &near,
                                              float &far) {
    const float rdx = __frcp_rn(dx), rdy = __frcp_rn(dy), rdz = __frcp_rn(dz);
    near = __fmul_rn(__fsub_rn(aabb[0], ox), rdx);
    far = __fmul_rn(__fsub_rn(aabb[3], ox), rdx);
    if (near > far) { const float c = near; near = far; far = c; }
    float near_y = __fmul_rn(__fsub_rn(aabb[1], oy), rdy);
    float far_y = __fmul_rn(__fsub_rn(aabb[4], oy), rdy);
    if (near_y > far_y) { const float c = near_y; near_y = far_y; far_y = c; }
    if (near > far_y || near_y > far) return false;
    if (near_y > near) near = near_y;
    if (far_y < far) far = far_y;
    float near_z = __fmul_rn(__fsub_rn(aabb[2], oz), rdz);
    float far_z = __fmul_rn(__fsub_rn(aabb[5], oz), rdz);
    if (near_z > far_z) { const float c = near_z; near_z = far_z; far_z = c; }
    if (near > far_z || near_z > far) return false;
    if (near_z > near) near = near_z;
    if (far_z < far) far = far_z;
    if (near < min_near) near = min_near;
    return true;
}

// ------------------------------------------------------------------------------------------------
// Occupancy-bitfield marcher (raymarching.cu:706-814)
struct MarchConst {
    float bound, dt_gamma, dt_min, dt_max;
    float Hf, Hm1f, rH, H3f, cm1f;  // (float)H, (float)(H-1), 1/(float)H, (float)(H*H*H), (float)(C-1)
    const uint8_t *grid;
};

__host__ __device__ __forceinline__ MarchConst make_march_const(float bound, float dt_gamma, uint32_t max_steps,
                                                                uint32_t C, uint32_t H, const uint8_t *grid) {
    MarchConst m;
    m.bound = bound;
    m.dt_gamma = dt_gamma;
    m.dt_min = 3.4641016f / (float)max_steps;                     // 2*SQRT3()/max_steps         (:745)
    m.dt_max = (3.4641016f * (float)(1 << (C - 1))) / (float)H;   // 2*SQRT3()*(1<<(C-1))/H      (:746)
    m.Hf = (float)H;
    m.Hm1f = (float)(H - 1);
    m.rH = 1.0f / (float)H;
    m.H3f = (float)(H * H * H);
    m.cm1f = (float)C - 1.0f;
    m.grid = grid;
    return m;
}

// frexpf exponent clamped to [0, C-1]   (mip_from_pos / mip_from_dt, raymarching.cu:44-56)
__device__ __forceinline__ int mip_level(float mx, float cm1f) {
    int e;
    frexpf(mx, &e);
    return (int)fminf(cm1f, fmaxf(0.0f, (float)e));
}

__device__ __forceinline__ float march_dt(float t, const MarchConst &m) {
    return fminf(m.dt_max, fmaxf(m.dt_min, __fmul_rn(t, m.dt_gamma)));
}

// cell coordinate along one axis: clamp(0.5 * (x * mip_rbound + 1) * H, 0, H-1) truncated.
// The reference evaluates the two products in double and narrows once; 0.5*a is exact and a*H is a
// 24-bit x <=11-bit product that double holds exactly, so one fp32 rounding of the same product is
// bit-identical -- and keeps the hot loop off the (slow) FP64 pipe.
__device__ __forceinline__ int cell_coord(float x, float mip_rbound, const MarchConst &m) {
    const float a = __fmaf_rn(x, mip_rbound, 1.0f);
    const float f = __fmul_rn(__fmul_rn(0.5f, a), m.Hf);
    return (int)fminf(m.Hm1f, fmaxf(0.0f, f));
}

__device__ __forceinline__ float cell_exit(int n, float sgn, float mip_bound, float x, float rd, const MarchConst &m) {
    float a = __fadd_rn((float)n, 0.5f);
    a = __fmaf_rn(sgn, 0.5f, a);
    a = __fmul_rn(a, m.rH);
    a = __fmaf_rn(a, 2.0f, -1.0f);
    a = __fmaf_rn(mip_bound, a, -x);
    return __fmul_rn(a, rd);
}

// Marches one ray from t for up to n_step occupied samples.  emit(step, x, y, z, dt, delta_t) is called per
// sample.  Returns the number of samples emitted; t_out is the marcher's own t afterwards (NOT what the
// reference stores in rays_t -- that is advanced by the compositor).
template <class Emit>
__device__ __forceinline__ uint32_t march_ray(const MarchConst &m, float ox, float oy, float oz, float dx, float dy,
                                              float dz, float t, float far, uint32_t n_step, Emit &&emit) {
    const float rdx = __frcp_rn(dx), rdy = __frcp_rn(dy), rdz = __frcp_rn(dz);
    const float sx = copysignf(1.0f, dx), sy = copysignf(1.0f, dy), sz = copysignf(1.0f, dz);
    uint32_t step = 0;
    float last_t = t;
    while (t < far && step < n_step) {
        const float x = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(dx, t, ox)));
        const float y = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(dy, t, oy)));
        const float z = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(dz, t, oz)));
        const float dt = march_dt(t, m);
        const int l_pos = mip_level(fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z))), m.cm1f);
        const int l_dt = mip_level(__fmul_rn(__fmul_rn(dt, m.Hf), 0.5f), m.cm1f);
        const int level = max(l_pos, l_dt);
        const float mip_bound = fminf((float)(1 << level), m.bound);
        const float mip_rbound = __frcp_rn(mip_bound);
        const int nx = cell_coord(x, mip_rbound, m);
        const int ny = cell_coord(y, mip_rbound, m);
        const int nz = cell_coord(z, mip_rbound, m);
        const uint32_t idx = (uint32_t)__fmaf_rn(m.H3f, (float)level, (float)morton3d(nx, ny, nz));
        const bool occ = (__ldg(m.grid + (idx >> 3)) >> (idx & 7)) & 1;
        if (occ) {
            t = __fadd_rn(t, dt);
            emit(step, x, y, z, dt, __fsub_rn(t, last_t));
            last_t = t;
            step++;
        } else {
            const float tx = cell_exit(nx, sx, mip_bound, x, rdx, m);
            const float ty = cell_exit(ny, sy, mip_bound, y, rdy, m);
            const float tz = cell_exit(nz, sz, mip_bound, z, rdz, m);
            const float tt = __fadd_rn(t, fmaxf(0.0f, fminf(tx, fminf(ty, tz))));
            do {
                t = __fadd_rn(t, march_dt(t, m));
            } while (t < tt);
        }
    }
    return step;
}

// ------------------------------------------------------------------------------------------------
// Front-to-back compositing of one sample (raymarching.cu:866-894).  Returns true when the ray terminates
// AFTER accumulating this sample (T < 1e-4 evaluated in double in the reference; (double)T < 1e-4 holds
// exactly when T <= 1e-4f because float(1e-4) < 1e-4 < nextafter(float(1e-4))).
struct RayAcc {
    float t, ws, d, r, g, b;
};
__device__ __forceinline__ bool composite_sample(RayAcc &a, float sigma, float cr, float cg, float cb, float dt,
                                                 float delta_t) {
    const float q = __fmul_rn(__fmul_rn(dt, sigma), -1.4426950216293334961f);
    float e;
    asm("ex2.approx.f32 %0, %1;" : "=f"(e) : "f"(q));  // __expf(-sigma*dt)
    const float alpha = __fsub_rn(1.0f, e);
    const float T = __fsub_rn(1.0f, a.ws);
    const float w = __fmul_rn(alpha, T);
    a.ws = __fadd_rn(a.ws, w);
    a.t = __fadd_rn(a.t, delta_t);
    a.d = __fmaf_rn(w, a.t, a.d);
    a.r = __fmaf_rn(w, cr, a.r);
    a.g = __fmaf_rn(w, cg, a.g);
    a.b = __fmaf_rn(w, cb, a.b);
    return T <= 1e-4f;
}

// ------------------------------------------------------------------------------------------------
// Hash grid (gridencoder/src/gridencoder.cu:35-72, :75-176), D = 3
struct LevelConst {
    uint32_t offset;   // rows before this level
    uint32_t size;     // rows in this level (hashmap_size)
    uint32_t stride;   // resolution + 1 (align_corners = false)
    uint32_t mode;     // 0 = dense x + y*s + z*s*s, 1 = hash & (size-1), 2 = hash % size, 3 = dense % size
    float scale;       // exp2f(level*S)*H - 1, as the device computes it
};
constexpr int kMaxLevels = 16;

__device__ __forceinline__ uint32_t hash3(uint32_t x, uint32_t y, uint32_t z) {
    return x ^ (y * 2654435761u) ^ (z * 805459861u);
}

// reference-faithful index (general gridtype / align_corners), rows not yet multiplied by C
__device__ __forceinline__ uint32_t grid_row_ref(uint32_t gridtype, bool align_corners, uint32_t hashmap_size,
                                                 uint32_t resolution, uint32_t x, uint32_t y, uint32_t z) {
    const uint32_t s = align_corners ? resolution : resolution + 1;
    const uint32_t p[3] = {x, y, z};
    uint32_t stride = 1, index = 0;
    bool go = true;
#pragma unroll
    for (int d = 0; d < 3; d++) {
        go = go && stride <= hashmap_size;  // the reference loop stops at the first level that overflows
        if (go) { index += p[d] * stride; stride *= s; }
    }
    if (gridtype == 0 && stride > hashmap_size) index = hash3(x, y, z);
    return index % hashmap_size;
}

__device__ __forceinline__ uint32_t grid_row_fast(const LevelConst &lc, uint32_t x, uint32_t y, uint32_t z) {
    if (lc.mode == 0) return x + (y + z * lc.stride) * lc.stride;
    if (lc.mode == 3) return (x + (y + z * lc.stride) * lc.stride) % lc.size;
    const uint32_t h = hash3(x, y, z);
    return lc.mode == 1 ? (h & (lc.size - 1)) : (h % lc.size);
}

// ------------------------------------------------------------------------------------------------
// Real spherical harmonics, degree 4 (16 values)  (shencoder/src/shencoder.cu:44-69)
__device__ __forceinline__ void sh_basis(float x, float y, float z, uint32_t degree, float *o) {
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    o[0] = 0.28209479177387814f;
    if (degree <= 1) return;
    o[1] = -0.48860251190291987f * y;
    o[2] = 0.48860251190291987f * z;
    o[3] = -0.48860251190291987f * x;
    if (degree <= 2) return;
    o[4] = 1.0925484305920792f * xy;
    o[5] = -1.0925484305920792f * yz;
    o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
    o[7] = -1.0925484305920792f * xz;
    o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
    if (degree <= 3) return;
    o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
    o[10] = 2.8906114426405538f * xy * z;
    o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
    o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
    o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
    o[14] = 1.4453057213202769f * z * (x2 - y2);
    o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
}

// ------------------------------------------------------------------------------------------------
// Tensor-core MLP building blocks: mma.sync.m16n8k16 f16 x f16 -> f32, activations chained in registers.
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t pack_half2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&h);
}

// One dense layer for MT 16-row tiles held by this warp.
//   A  [MT][KT][4] : fp16 A fragments (K = 16*KT)
//   wf            : this layer's weights in fragment order, uint2 per (ntile, ktile, lane)
//   C  [MT][NT][4] : fp32 accumulators out (N = 8*NT)
template <int MT, int KT, int NT>
__device__ __forceinline__ void mlp_layer(const uint32_t (&A)[MT][KT][4], const uint2 *__restrict__ wf, int lane,
                                          float (&C)[MT][NT][4]) {
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
#pragma unroll
        for (int mt = 0; mt < MT; mt++) {
            C[mt][nt][0] = 0.f; C[mt][nt][1] = 0.f; C[mt][nt][2] = 0.f; C[mt][nt][3] = 0.f;
        }
#pragma unroll
        for (int kt = 0; kt < KT; kt++) {
            const uint2 b = wf[(nt * KT + kt) * 32 + lane];
#pragma unroll
            for (int mt = 0; mt < MT; mt++) mma16816(C[mt][nt], A[mt][kt], b.x, b.y);
        }
    }
}

// ReLU + fp16 rounding; accumulators of a 64-wide layer become the next layer's A fragments.
template <int MT, int NT>
__device__ __forceinline__ void relu_to_frag(const float (&C)[MT][NT][4], uint32_t (&A)[MT][NT / 2][4]) {
#pragma unroll
    for (int mt = 0; mt < MT; mt++)
#pragma unroll
        for (int kt = 0; kt < NT / 2; kt++) {
            A[mt][kt][0] = pack_half2(fmaxf(C[mt][2 * kt][0], 0.f), fmaxf(C[mt][2 * kt][1], 0.f));
            A[mt][kt][1] = pack_half2(fmaxf(C[mt][2 * kt][2], 0.f), fmaxf(C[mt][2 * kt][3], 0.f));
            A[mt][kt][2] = pack_half2(fmaxf(C[mt][2 * kt + 1][0], 0.f), fmaxf(C[mt][2 * kt + 1][1], 0.f));
            A[mt][kt][3] = pack_half2(fmaxf(C[mt][2 * kt + 1][2], 0.f), fmaxf(C[mt][2 * kt + 1][3], 0.f));
        }
}

// Packs one row-major [N][K] half weight matrix into mma B-fragment order (uint2 per (ntile,ktile,lane)).
// `remap`: optional source-column remap for the fused colour layer (see fused.cu); -1 => zero.
__device__ __forceinline__ void pack_weight_frags(const __half *__restrict__ W, int N, int K, uint2 *__restrict__ dst,
                                                  int tid, int nthreads, bool shift_cols_from16) {
    const int KT = K / 16, NT = N / 8;
    for (int i = tid; i < NT * KT * 32; i += nthreads) {
        const int lane = i & 31, tile = i >> 5, kt = tile % KT, nt = tile / KT;
        const int g = lane >> 2, t = lane & 3;
        const int n = nt * 8 + g;
        uint32_t v[2];
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int k = kt * 16 + 2 * t + 8 * r;
            __half h0, h1;
            if (!shift_cols_from16) {
                h0 = W[n * K + k];
                h1 = W[n * K + k + 1];
            } else {
                // column k of the packed matrix reads source column k (k < 16), nothing (k == 16), k-1 (k > 16)
                h0 = (k < 16) ? W[n * K + k] : (k == 16 ? __float2half(0.f) : W[n * K + k - 1]);
                const int k1 = k + 1;
                h1 = (k1 < 16) ? W[n * K + k1] : (k1 == 16 ? __float2half(0.f) : W[n * K + k1 - 1]);
            }
            const __half2 p = __halves2half2(h0, h1);
            v[r] = *reinterpret_cast<const uint32_t *>(&p);
        }
        dst[i] = make_uint2(v[0], v[1]);
    }
}

}  // namespace ngp
