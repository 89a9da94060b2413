// ops_encode.cu -- stand-alone encoders behind the C ABI:
//   ngp_grid_encode_forward  (drop-in for _gridencoder.grid_encode_forward, D = 3, forward only)
//   ngp_grid_level_scales
//   ngp_sh_encode_forward    (drop-in for _shencoder.sh_encode_forward, degree <= 4)
#include "common.cuh"
#include "device_math.cuh"

namespace ngp {

constexpr int kEncThreads = 256;

template <typename T> struct Ld;
template <> struct Ld<float> { static __device__ __forceinline__ float get(const float *p) { return __ldg(p); } };
template <> struct Ld<__half> { static __device__ __forceinline__ float get(const __half *p) { return __half2float(__ldg(p)); } };

// gridencoder/src/gridencoder.cu:75-176.  One thread = one (point, level); blockIdx.y = level so a wave of
// CTAs walks one level's table slice at a time (same L2-friendly order as the reference), but with
// vectorised corner loads (C*sizeof(T) bytes per corner in one request when C == 2).
template <typename T, int C, bool REF_ROUND>
__global__ void __launch_bounds__(kEncThreads)
k_grid_encode(const float *__restrict__ inputs, const T *__restrict__ table, const int *__restrict__ offsets, T *__restrict__ outputs,
              uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners, int out_layout,
              int *__restrict__ idx_out) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint32_t level = blockIdx.y;
    const float u0 = inputs[(size_t)b * 3 + 0], u1 = inputs[(size_t)b * 3 + 1], u2 = inputs[(size_t)b * 3 + 2];
    T *out = outputs + (out_layout == 0 ? ((size_t)level * B + b) * C : ((size_t)b * L + level) * C);
    const bool oob = (u0 < 0 || u0 > 1) || (u1 < 0 || u1 > 1) || (u2 < 0 || u2 > 1);
    if (oob) {
#pragma unroll
        for (int ch = 0; ch < C; ch++) out[ch] = (T)0.0f;
        if (idx_out)
            for (int k = 0; k < 8; k++) idx_out[((size_t)level * B + b) * 8 + k] = -1;
        return;
    }
    const uint32_t off = (uint32_t)offsets[level];
    const uint32_t hashmap_size = (uint32_t)offsets[level + 1] - off;
    const float scale = __fmaf_rn(exp2f((float)level * S), (float)H, -1.0f);
    const uint32_t resolution = (uint32_t)ceilf(scale) + 1;
    const float half_off = align_corners ? 0.0f : 0.5f;
    float p0 = __fmaf_rn(u0, scale, half_off), p1 = __fmaf_rn(u1, scale, half_off), p2 = __fmaf_rn(u2, scale, half_off);
    const uint32_t g0 = (uint32_t)floorf(p0), g1 = (uint32_t)floorf(p1), g2 = (uint32_t)floorf(p2);
    p0 = __fsub_rn(p0, (float)g0); p1 = __fsub_rn(p1, (float)g1); p2 = __fsub_rn(p2, (float)g2);
    const float q0 = __fsub_rn(1.0f, p0), q1 = __fsub_rn(1.0f, p1), q2 = __fsub_rn(1.0f, p2);

    // all 8 rows first, then all loads (8 independent requests in flight), then the blend
    uint32_t rows[8];
#pragma unroll
    for (int k = 0; k < 8; k++)
        rows[k] = grid_row_ref(gridtype, align_corners, hashmap_size, resolution, g0 + (k & 1), g1 + ((k >> 1) & 1), g2 + ((k >> 2) & 1));
    float vals[8][C];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const T *e = table + ((size_t)off + rows[k]) * C;
        if constexpr (C == 2 && sizeof(T) == 2) {
            const __half2 h = __ldg(reinterpret_cast<const __half2 *>(e));
            vals[k][0] = __low2float(h); vals[k][1] = __high2float(h);
        } else if constexpr (C == 2 && sizeof(T) == 4) {
            const float2 f = __ldg(reinterpret_cast<const float2 *>(e));
            vals[k][0] = f.x; vals[k][1] = f.y;
        } else {
#pragma unroll
            for (int ch = 0; ch < C; ch++) vals[k][ch] = Ld<T>::get(e + ch);
        }
    }
    float acc[C];
#pragma unroll
    for (int ch = 0; ch < C; ch++) acc[ch] = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float w = __fmul_rn(__fmul_rn((k & 1) ? p0 : q0, (k & 2) ? p1 : q1), (k & 4) ? p2 : q2);
#pragma unroll
        for (int ch = 0; ch < C; ch++) {
            if constexpr (REF_ROUND) {  // scalar_t results[C] in the reference: half after the product, half after the add
                const float p = __half2float(__float2half_rn(__fmul_rn(w, vals[k][ch])));
                acc[ch] = __half2float(__float2half_rn(__fadd_rn(acc[ch], p)));
            } else {
                acc[ch] = __fmaf_rn(w, vals[k][ch], acc[ch]);
            }
        }
        if (idx_out) idx_out[((size_t)level * B + b) * 8 + k] = (int)rows[k];
    }
    if constexpr (C == 2 && sizeof(T) == 2) {
        *reinterpret_cast<__half2 *>(out) = __floats2half2_rn(acc[0], acc[1]);
    } else if constexpr (C == 2 && sizeof(T) == 4) {
        *reinterpret_cast<float2 *>(out) = make_float2(acc[0], acc[1]);
    } else {
#pragma unroll
        for (int ch = 0; ch < C; ch++) out[ch] = (T)acc[ch];
    }
}

__global__ void k_level_scales(uint32_t L, float S, uint32_t H, float *scales) {
    const uint32_t l = threadIdx.x;
    if (l < L) scales[l] = __fmaf_rn(exp2f((float)l * S), (float)H, -1.0f);
}

// shencoder/src/shencoder.cu:27-124 (forward, degree <= 4)
__global__ void __launch_bounds__(kEncThreads) k_sh_encode(const float *__restrict__ inputs, float *__restrict__ outputs, uint32_t B, uint32_t degree) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    float o[16];
    sh_basis(inputs[(size_t)b * 3], inputs[(size_t)b * 3 + 1], inputs[(size_t)b * 3 + 2], degree, o);
    const uint32_t n = degree * degree;
    float *out = outputs + (size_t)b * n;
    if (degree == 4) {
        float4 *o4 = reinterpret_cast<float4 *>(out);
        o4[0] = make_float4(o[0], o[1], o[2], o[3]);   o4[1] = make_float4(o[4], o[5], o[6], o[7]);
        o4[2] = make_float4(o[8], o[9], o[10], o[11]); o4[3] = make_float4(o[12], o[13], o[14], o[15]);
    } else {
        for (uint32_t i = 0; i < n; i++) out[i] = o[i];
    }
}

template <typename T, bool RR>
static int launch_grid(const float *inputs, const void *table, const int *offsets, void *outputs, uint32_t B, uint32_t C, uint32_t L,
                       float S, uint32_t H, uint32_t gridtype, bool ac, int out_layout, int *idx_out, cudaStream_t st) {
    const dim3 grid(div_up(B, kEncThreads), L);
    const T *t = reinterpret_cast<const T *>(table);
    T *o = reinterpret_cast<T *>(outputs);
    switch (C) {
        case 1: k_grid_encode<T, 1, RR><<<grid, kEncThreads, 0, st>>>(inputs, t, offsets, o, B, L, S, H, gridtype, ac, out_layout, idx_out); break;
        case 2: k_grid_encode<T, 2, RR><<<grid, kEncThreads, 0, st>>>(inputs, t, offsets, o, B, L, S, H, gridtype, ac, out_layout, idx_out); break;
        case 4: k_grid_encode<T, 4, RR><<<grid, kEncThreads, 0, st>>>(inputs, t, offsets, o, B, L, S, H, gridtype, ac, out_layout, idx_out); break;
        case 8: k_grid_encode<T, 8, RR><<<grid, kEncThreads, 0, st>>>(inputs, t, offsets, o, B, L, S, H, gridtype, ac, out_layout, idx_out); break;
        default: set_error("GridEncoding: C must be 1, 2, 4, or 8."); return 1;   // gridencoder.cu:355
    }
    return after_launch("grid_encode_forward");
}

}  // namespace ngp

using namespace ngp;

extern "C" {

int ngp_grid_encode_forward(const float *inputs01, const void *table, int table_dtype, const int32_t *offsets, void *outputs,
                            uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, uint32_t gridtype,
                            int align_corners, int out_layout, int ref_rounding, int32_t *idx_out, void *stream) {
    NGP_REQUIRE(D == 3, "grid_encode_forward: only D == 3 is built (the NeRF path); got D=%u", D);
    NGP_REQUIRE(L >= 1 && L <= 65535, "grid_encode_forward: bad number of levels %u", L);
    NGP_REQUIRE(table_dtype == NGP_DTYPE_F32 || table_dtype == NGP_DTYPE_F16, "grid_encode_forward: table dtype must be f32 or f16");
    NGP_REQUIRE(out_layout == 0 || out_layout == 1, "grid_encode_forward: out_layout must be 0 ([L,B,C]) or 1 ([B,L*C])");
    if (B == 0) return 0;
    NGP_DEVPTR(inputs01); NGP_DEVPTR(table); NGP_DEVPTR(offsets); NGP_DEVPTR(outputs);
    cudaStream_t st = as_stream(stream);
    if (table_dtype == NGP_DTYPE_F32)
        return launch_grid<float, false>(inputs01, table, offsets, outputs, B, C, L, S, H, gridtype, align_corners != 0, out_layout, idx_out, st);
    if (ref_rounding)
        return launch_grid<__half, true>(inputs01, table, offsets, outputs, B, C, L, S, H, gridtype, align_corners != 0, out_layout, idx_out, st);
    return launch_grid<__half, false>(inputs01, table, offsets, outputs, B, C, L, S, H, gridtype, align_corners != 0, out_layout, idx_out, st);
}

int ngp_grid_level_scales(uint32_t L, float S, uint32_t H, float *scales, void *stream) {
    NGP_REQUIRE(L >= 1 && L <= 1024, "grid_level_scales: bad L");
    NGP_DEVPTR(scales);
    k_level_scales<<<1, 1024, 0, as_stream(stream)>>>(L, S, H, scales);
    return after_launch("grid_level_scales");
}

int ngp_sh_encode_forward(const float *inputs, float *outputs, uint32_t B, uint32_t degree, void *stream) {
    NGP_REQUIRE(degree >= 1 && degree <= 4, "sh_encode_forward: degree must be in [1, 4] (the NeRF path uses 4); got %u", degree);
    if (B == 0) return 0;
    NGP_DEVPTR(inputs); NGP_DEVPTR(outputs);
    k_sh_encode<<<div_up(B, kEncThreads), kEncThreads, 0, as_stream(stream)>>>(inputs, outputs, B, degree);
    return after_launch("sh_encode_forward");
}

}  // extern "C"
