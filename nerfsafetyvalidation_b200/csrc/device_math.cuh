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
// Camera ray of pixel p (row-major, H x W) of frame pose P (cam2world [4,4] row-major): nerf/utils.py:52-116 with N = -1 (every
// pixel, centre at +0.5): direction ((i - cx) / fx, (j - cy) / fy, 1) normalised and rotated, origin = translation.  Operation order
// is fixed (IEEE mul / add / div / sqrt, no contraction) so that ngp_get_rays, the renderer's prologue and the oracle agree bit for bit.
__device__ __forceinline__ void pose_ray(const float *__restrict__ P, uint32_t p, uint32_t W, float fx, float fy, float cx, float cy, float (&o)[3],
                                         float (&d)[3]) {
    const float i = __fadd_rn((float)(p % W), 0.5f), j = __fadd_rn((float)(p / W), 0.5f);
    const float x = __fdiv_rn(__fsub_rn(i, cx), fx), y = __fdiv_rn(__fsub_rn(j, cy), fy);
    const float nrm = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), 1.0f));
    const float d0 = __fdiv_rn(x, nrm), d1 = __fdiv_rn(y, nrm), d2 = __fdiv_rn(1.0f, nrm);
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float r0 = __ldg(P + 4 * k), r1 = __ldg(P + 4 * k + 1), r2 = __ldg(P + 4 * k + 2);
        d[k] = __fadd_rn(__fadd_rn(__fmul_rn(d0, r0), __fmul_rn(d1, r1)), __fmul_rn(d2, r2));
        o[k] = __ldg(P + 4 * k + 3);
    }
}

// ------------------------------------------------------------------------------------------------
// Occupancy-bitfield marcher (raymarching.cu:706-814)
struct MarchConst {
    float bound, dt_gamma, dt_min, dt_max;
    float Hf, Hm1f, rH, H3f, cm1f;  // (float)H, (float)(H-1), 1/(float)H, (float)(H*H*H), (float)(C-1)
    float mip_b[8], mip_rb[8];      // fminf(1 << level, bound) and its correctly rounded reciprocal (:768-769)
    const uint8_t *grid;
};

__host__ __device__ __forceinline__ float __frcp_rn_host_or_device(float v) {
#ifdef __CUDA_ARCH__
    return __frcp_rn(v);
#else
    return 1.0f / v;   // IEEE division on the host == rcp.rn
#endif
}
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
    for (int l = 0; l < 8; l++) {
        m.mip_b[l] = fminf((float)(1 << l), bound);
        m.mip_rb[l] = __frcp_rn_host_or_device(m.mip_b[l]);
    }
    m.grid = grid;
    return m;
}

// frexpf exponent clamped to [0, C-1]   (mip_from_pos / mip_from_dt, raymarching.cu:44-56).
// frexpf's exponent of a normal float is (biased exponent - 126); zero and denormals give <= -126 (clamped to 0 anyway)
// and CUDA's frexpf returns exponent 0 for inf / nan, so the whole call reduces to integer work on the bit pattern.
__device__ __forceinline__ int mip_level(float mx, float cm1f) {
    const int field = (int)((__float_as_uint(mx) >> 23) & 0xffu);
    const int e = field == 255 ? 0 : field - 126;
    return min((int)cm1f, max(0, e));
}

__device__ __forceinline__ float march_dt(float t, const MarchConst &m) {
    return fminf(m.dt_max, fmaxf(m.dt_min, __fmul_rn(t, m.dt_gamma)));
}
// The (x, y, z, dt) the marcher attaches to a sample taken at ray parameter t (raymarching.cu:756-761): both are pure functions of
// t and the ray, so the fused renderer's sample buffer stores t alone and every consumer recomputes them bit for bit.
__device__ __forceinline__ float sample_dt(const MarchConst &m, float t) {
    return m.dt_gamma == 0.0f ? fminf(m.dt_max, fmaxf(m.dt_min, 0.0f)) : march_dt(t, m);   // NaN-safe: fmaxf drops the NaN of inf * 0
}
__device__ __forceinline__ float sample_pos(const MarchConst &m, float o, float d, float t) {
    return fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(d, t, o)));
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

struct RayDir {   // per-ray constants of the marcher (raymarching.cu:727-731,804-806)
    float dx, dy, dz, rdx, rdy, rdz, sx, sy, sz;
};
__device__ __forceinline__ RayDir make_ray_dir(float dx, float dy, float dz) {
    RayDir r;
    r.dx = dx; r.dy = dy; r.dz = dz;
    r.rdx = __frcp_rn(dx); r.rdy = __frcp_rn(dy); r.rdz = __frcp_rn(dz);
    r.sx = copysignf(1.0f, dx); r.sy = copysignf(1.0f, dy); r.sz = copysignf(1.0f, dz);
    return r;
}

// ONE trip of the reference's marching loop body (raymarching.cu:754-812), precondition t < far: probes the cell under
// the ray at t.  Occupied: returns true with the sample (x, y, z, dt, delta_t = t_new - last_t) and t advanced by dt.
// Empty: returns false with t advanced, in dt increments, past the cell's exit.  The loop state is just (t, last_t), so a
// caller may interleave the probes of 32 rays in lockstep (the persistent renderer) or run them to completion (march_ray).
// `coarse` (optional, shared memory): one bit per aligned 4x4x4 block of cells = 64 consecutive Morton indices, set iff any of
// them is occupied.  A clear bit answers the probe without the dependent global load; the result is the reference's either way.
__device__ __forceinline__ bool march_probe(const MarchConst &m, float ox, float oy, float oz, const RayDir &rd, float &t, float &last_t,
                                            float &x, float &y, float &z, float &dt, float &delta_t, const uint32_t *coarse = nullptr) {
    const bool one_level = m.cm1f == 0.0f;     // cascade 1 (bound <= 1): the mip level is always 0
    const bool const_dt = m.dt_gamma == 0.0f;  // dt = clamp(t * 0, dt_min, dt_max) for every t (NaN-safe: fmaxf drops the NaN)
    const float dt0 = fminf(m.dt_max, fmaxf(m.dt_min, 0.0f));
    x = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(rd.dx, t, ox)));
    y = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(rd.dy, t, oy)));
    z = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(rd.dz, t, oz)));
    dt = const_dt ? dt0 : march_dt(t, m);
    int level = 0;
    if (!one_level) {
        const int l_pos = mip_level(fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z))), m.cm1f);
        const int l_dt = mip_level(__fmul_rn(__fmul_rn(dt, m.Hf), 0.5f), m.cm1f);
        level = max(l_pos, l_dt);
    }
    const float mip_bound = m.mip_b[level];
    const float mip_rbound = m.mip_rb[level];
    const int nx = cell_coord(x, mip_rbound, m);
    const int ny = cell_coord(y, mip_rbound, m);
    const int nz = cell_coord(z, mip_rbound, m);
    const uint32_t idx = (uint32_t)__fmaf_rn(m.H3f, (float)level, (float)morton3d(nx, ny, nz));
    bool occ = true;
    if (coarse) occ = (coarse[idx >> 11] >> ((idx >> 6) & 31u)) & 1u;
    if (occ) occ = (__ldg(m.grid + (idx >> 3)) >> (idx & 7)) & 1;
    if (occ) {
        t = __fadd_rn(t, dt);
        delta_t = __fsub_rn(t, last_t);
        last_t = t;
        return true;
    }
    const float tx = cell_exit(nx, rd.sx, mip_bound, x, rd.rdx, m);
    const float ty = cell_exit(ny, rd.sy, mip_bound, y, rd.rdy, m);
    const float tz = cell_exit(nz, rd.sz, mip_bound, z, rd.rdz, m);
    const float tt = __fadd_rn(t, fmaxf(0.0f, fminf(tx, fminf(ty, tz))));
    if (const_dt) {
        do { t = __fadd_rn(t, dt0); } while (t < tt);
    } else {
        do { t = __fadd_rn(t, march_dt(t, m)); } while (t < tt);
    }
    return false;
}

// The cell test of march_probe WITHOUT advancing the ray (same expressions, operation for operation): is the cell under the ray at t
// occupied?  If not, tt <- the parameter at which the marcher's skip loop stops stepping (t + max(0, distance to the cell's exit),
// raymarching.cu:795-806).  For callers that test many candidate parameters of ONE ray in parallel (diff_render.cuh).
__device__ __forceinline__ bool march_peek(const MarchConst &m, float ox, float oy, float oz, const RayDir &rd, float t, float &tt) {
    const bool one_level = m.cm1f == 0.0f;
    const float x = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(rd.dx, t, ox)));
    const float y = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(rd.dy, t, oy)));
    const float z = fminf(m.bound, fmaxf(-m.bound, __fmaf_rn(rd.dz, t, oz)));
    int level = 0;
    if (!one_level) {
        const float dt = sample_dt(m, t);
        const int l_pos = mip_level(fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z))), m.cm1f);
        const int l_dt = mip_level(__fmul_rn(__fmul_rn(dt, m.Hf), 0.5f), m.cm1f);
        level = max(l_pos, l_dt);
    }
    const float mip_bound = m.mip_b[level];
    const float mip_rbound = m.mip_rb[level];
    const int nx = cell_coord(x, mip_rbound, m);
    const int ny = cell_coord(y, mip_rbound, m);
    const int nz = cell_coord(z, mip_rbound, m);
    const uint32_t idx = (uint32_t)__fmaf_rn(m.H3f, (float)level, (float)morton3d(nx, ny, nz));
    tt = t;
    if ((__ldg(m.grid + (idx >> 3)) >> (idx & 7)) & 1) return true;
    const float tx = cell_exit(nx, rd.sx, mip_bound, x, rd.rdx, m);
    const float ty = cell_exit(ny, rd.sy, mip_bound, y, rd.rdy, m);
    const float tz = cell_exit(nz, rd.sz, mip_bound, z, rd.rdz, m);
    tt = __fadd_rn(t, fmaxf(0.0f, fminf(tx, fminf(ty, tz))));
    return false;
}

// Marches one ray from t for up to n_step occupied samples (one reference chunk).  emit(step, x, y, z, dt, delta_t) is
// called per sample.  Returns the number of samples emitted.
template <class Emit>
__device__ __forceinline__ uint32_t march_ray(const MarchConst &m, float ox, float oy, float oz, float dx, float dy,
                                              float dz, float t, float far, uint32_t n_step, Emit &&emit) {
    const RayDir rd = make_ray_dir(dx, dy, dz);
    uint32_t step = 0;
    float last_t = t, x, y, z, dt, delta_t;
    while (t < far && step < n_step) {
        if (march_probe(m, ox, oy, oz, rd, t, last_t, x, y, z, dt, delta_t)) {
            emit(step, x, y, z, dt, delta_t);
            step++;
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
// w_out (optional): the sample's compositing weight alpha * T, for per-ray sums the reference's sampler path keeps
// (aggregated_density = sum of w * sigma, nerf/renderer.py:244).
__device__ __forceinline__ bool composite_sample(RayAcc &a, float sigma, float cr, float cg, float cb, float dt,
                                                 float delta_t, float *w_out = nullptr) {
    const float q = __fmul_rn(__fmul_rn(dt, sigma), -1.4426950216293334961f);
    float e;
    asm("ex2.approx.f32 %0, %1;" : "=f"(e) : "f"(q));  // __expf(-sigma*dt)
    const float alpha = __fsub_rn(1.0f, e);
    const float T = __fsub_rn(1.0f, a.ws);
    const float w = __fmul_rn(alpha, T);
    if (w_out) *w_out = w;
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

// The 8 table rows of the cell whose low corner is (g0, g1, g2); corner k = (k & 1, (k >> 1) & 1, (k >> 2) & 1), the order
// the reference accumulates in (gridencoder.cu:146-171).  lc.mode is uniform per level, so the branch never diverges, and the
// two common modes share the per-axis terms between corners instead of recomputing the hash / dense index 8 times.
template <bool kGeneral = true>
__device__ __forceinline__ void cell_rows(const LevelConst &lc, uint32_t g0, uint32_t g1, uint32_t g2, uint32_t (&row)[8]) {
    if (lc.mode == 1) {            // hashed level, power-of-two table: (x ^ y*P1 ^ z*P2) & (T - 1); (g+1)*P = g*P + P mod 2^32
        const uint32_t m = lc.size - 1u;
        const uint32_t x0 = g0, x1 = g0 + 1u;
        const uint32_t y0 = g1 * 2654435761u, y1 = y0 + 2654435761u;
        const uint32_t z0 = g2 * 805459861u, z1 = z0 + 805459861u;
        row[0] = (x0 ^ y0 ^ z0) & m; row[1] = (x1 ^ y0 ^ z0) & m; row[2] = (x0 ^ y1 ^ z0) & m; row[3] = (x1 ^ y1 ^ z0) & m;
        row[4] = (x0 ^ y0 ^ z1) & m; row[5] = (x1 ^ y0 ^ z1) & m; row[6] = (x0 ^ y1 ^ z1) & m; row[7] = (x1 ^ y1 ^ z1) & m;
    } else if (lc.mode == 0) {     // dense level that fits its table: x + y*s + z*s*s
        const uint32_t s1 = lc.stride, s2 = s1 * s1;
        const uint32_t b = g0 + (g1 + g2 * s1) * s1;
        row[0] = b; row[1] = b + 1u; row[2] = b + s1; row[3] = b + s1 + 1u;
        row[4] = b + s2; row[5] = b + s2 + 1u; row[6] = b + s2 + s1; row[7] = b + s2 + s1 + 1u;
    } else if (kGeneral) {         // non-power-of-two hash table or wrapped dense level: the general form (integer divisions; the
                                   // persistent renderer is built without it and refuses such tables, which no NGP configuration produces)
#pragma unroll
        for (int k = 0; k < 8; k++) row[k] = grid_row_fast(lc, g0 + (k & 1), g1 + ((k >> 1) & 1), g2 + ((k >> 2) & 1));
    } else {
#pragma unroll
        for (int k = 0; k < 8; k++) row[k] = 0u;
    }
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
    asm(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Packed fp32 pairs (Blackwell FFMA2 / FMUL2: one issue slot for two IEEE round-to-nearest operations, results bit-identical to
// two scalar instructions).
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long A = *reinterpret_cast<unsigned long long *>(&a), B = *reinterpret_cast<unsigned long long *>(&b),
                       C = *reinterpret_cast<unsigned long long *>(&c), D;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(D) : "l"(A), "l"(B), "l"(C));
    return *reinterpret_cast<float2 *>(&D);
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
    unsigned long long A = *reinterpret_cast<unsigned long long *>(&a), B = *reinterpret_cast<unsigned long long *>(&b), D;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(D) : "l"(A), "l"(B));
    return *reinterpret_cast<float2 *>(&D);
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
    for (int nt = 0; nt < NT; nt++)
#pragma unroll
        for (int mt = 0; mt < MT; mt++) {
            C[mt][nt][0] = 0.f; C[mt][nt][1] = 0.f; C[mt][nt][2] = 0.f; C[mt][nt][3] = 0.f;
        }
    // k-tile outermost: the MT * NT accumulators are independent, so back-to-back mma never wait on each other
#pragma unroll
    for (int kt = 0; kt < KT; kt++) {
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
            const uint2 b = wf[(nt * KT + kt) * 32 + lane];
#pragma unroll
            for (int mt = 0; mt < MT; mt++) mma16816(C[mt][nt], A[mt][kt], b.x, b.y);
        }
    }
}

// relu + round-to-nearest fp16 pair in one instruction (lo -> low half)
__device__ __forceinline__ uint32_t pack_half2_relu(float lo, float hi) {
    uint32_t d;
    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}

// ReLU + fp16 rounding; accumulators of a 64-wide layer become the next layer's A fragments.
template <int MT, int NT>
__device__ __forceinline__ void relu_to_frag(const float (&C)[MT][NT][4], uint32_t (&A)[MT][NT / 2][4]) {
#pragma unroll
    for (int mt = 0; mt < MT; mt++)
#pragma unroll
        for (int kt = 0; kt < NT / 2; kt++) {
            A[mt][kt][0] = pack_half2_relu(C[mt][2 * kt][0], C[mt][2 * kt][1]);
            A[mt][kt][1] = pack_half2_relu(C[mt][2 * kt][2], C[mt][2 * kt][3]);
            A[mt][kt][2] = pack_half2_relu(C[mt][2 * kt + 1][0], C[mt][2 * kt + 1][1]);
            A[mt][kt][3] = pack_half2_relu(C[mt][2 * kt + 1][2], C[mt][2 * kt + 1][3]);
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
