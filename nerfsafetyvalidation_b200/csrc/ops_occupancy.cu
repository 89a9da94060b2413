// ops_occupancy.cu -- occupancy-grid maintenance: what NeRFRenderer.update_extra_state and
// mark_untrained_grid (nerf/renderer.py:388-546) do with a Python loop nest over 64^3 / 128^3 blocks, meshgrid + cat + morton3D +
// advanced-index scatter, as three flat kernels that work in MORTON order, the order density_grid is stored in: the query points of a
// cascade are generated straight at their cell's Morton index (no coordinate tensors, no index tensors, no scatter), the EMA update
// is one pass, and the camera-coverage test of mark_untrained_grid walks the poses per cell in registers.
#include <float.h>
#include <stdint.h>

#include "common.cuh"
#include "device_math.cuh"

namespace ngp {

constexpr int kThreads = 256;

// xyz of a grid cell's query point exactly as the reference's torch expression evaluates it (renderer.py:481,488-490), one rounding
// per torch kernel:  ((2 * c) / (H - 1) - 1) * scale + (u * 2 - 1) * hgs      scale = float(bound - hgs), hgs = float(bound / H)
__device__ __forceinline__ float cell_point(int c, float Hm1, float scale, float hgs, float u) {
    const float base = __fmul_rn(__fsub_rn(__fdiv_rn(__fmul_rn(2.0f, (float)c), Hm1), 1.0f), scale);
    return __fadd_rn(base, __fmul_rn(__fsub_rn(__fmul_rn(u, 2.0f), 1.0f), hgs));
}

// coords == NULL: point m is the cell with Morton index m (m < H^3); its jitter is row (x * H + y) * H + z of `jitter` -- the row
//                 the reference's meshgrid(ij) enumeration gives that cell, so the same random stream lands on the same cells.
// coords != NULL: point i is cell coords[i] with jitter row i (the partial-update branch, renderer.py:497-523).
// jitter == NULL: cell centres.
__global__ void k_occupancy_points(const int32_t *__restrict__ coords, const float *__restrict__ jitter, uint32_t n, uint32_t H, float scale,
                                   float hgs, float *__restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int x, y, z;
    uint32_t row;
    if (coords) {
        x = coords[3 * (size_t)i]; y = coords[3 * (size_t)i + 1]; z = coords[3 * (size_t)i + 2];
        row = i;
    } else {
        x = (int)compact_bits10(i); y = (int)compact_bits10(i >> 1); z = (int)compact_bits10(i >> 2);
        row = ((uint32_t)x * H + (uint32_t)y) * H + (uint32_t)z;
    }
    const float u0 = jitter ? jitter[3 * (size_t)row] : 0.5f, u1 = jitter ? jitter[3 * (size_t)row + 1] : 0.5f, u2 = jitter ? jitter[3 * (size_t)row + 2] : 0.5f;
    const float Hm1 = (float)(H - 1);
    out[3 * (size_t)i] = cell_point(x, Hm1, scale, hgs, u0);
    out[3 * (size_t)i + 1] = cell_point(y, Hm1, scale, hgs, u1);
    out[3 * (size_t)i + 2] = cell_point(z, Hm1, scale, hgs, u2);
}

// renderer.py:531-532: where both are >= 0, grid = max(grid * decay, tmp); tmp < 0 marks cells this update did not query.
__global__ void k_density_grid_ema(float *__restrict__ grid, const float *__restrict__ tmp, uint32_t n, float decay) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float g = grid[i], t = tmp[i];
    if (g >= 0.0f && t >= 0.0f) grid[i] = fmaxf(__fmul_rn(g, decay), t);
}

// renderer.py:388-449: a cell no training camera sees gets density -1.  One thread per (cascade, Morton cell); the poses are
// cam2world [B,4,4] row-major.  cam = (world - t) @ R (world2cam with the rotation transposed, :430-431); covered iff
// z > 0, |x| < cx / fx * z + 2 hgs, |y| < cy / fy * z + 2 hgs for at least one pose.
__global__ void k_mark_untrained(const float *__restrict__ poses, uint32_t B, float kx, float ky, uint32_t H, uint32_t cascade, float bound,
                                 float *__restrict__ grid) {
    const uint32_t H3 = H * H * H;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= H3 * cascade) return;
    const uint32_t cas = i / H3, m = i - cas * H3;
    const float b = fminf((float)(1u << cas), bound);
    const float hgs = __fdiv_rn(b, (float)H), scale = __fsub_rn(b, hgs), Hm1 = (float)(H - 1);
    const float wx = cell_point((int)compact_bits10(m), Hm1, scale, 0.0f, 0.5f), wy = cell_point((int)compact_bits10(m >> 1), Hm1, scale, 0.0f, 0.5f),
                wz = cell_point((int)compact_bits10(m >> 2), Hm1, scale, 0.0f, 0.5f);
    const float margin = __fmul_rn(hgs, 2.0f);
    bool seen = false;
    for (uint32_t p = 0; p < B && !seen; p++) {
        const float *P = poses + 16 * (size_t)p;
        const float dx = __fsub_rn(wx, __ldg(P + 3)), dy = __fsub_rn(wy, __ldg(P + 7)), dz = __fsub_rn(wz, __ldg(P + 11));
        const float cx_ = dx * __ldg(P + 0) + dy * __ldg(P + 4) + dz * __ldg(P + 8);
        const float cy_ = dx * __ldg(P + 1) + dy * __ldg(P + 5) + dz * __ldg(P + 9);
        const float cz_ = dx * __ldg(P + 2) + dy * __ldg(P + 6) + dz * __ldg(P + 10);
        seen = cz_ > 0.0f && fabsf(cx_) < __fmaf_rn(kx, cz_, margin) && fabsf(cy_) < __fmaf_rn(ky, cz_, margin);
    }
    if (!seen) grid[i] = -1.0f;
}

}  // namespace ngp

using namespace ngp;

extern "C" {

int ngp_occupancy_points(const int32_t *coords, const float *jitter, uint32_t n, uint32_t H, float bound_c, float *xyzs, void *stream) {
    if (n == 0) return 0;
    NGP_DEVPTR(xyzs);
    if (coords) NGP_DEVPTR(coords);
    if (jitter) NGP_DEVPTR(jitter);
    NGP_REQUIRE(H >= 2 && H <= 1024, "occupancy_points: grid size must be in [2, 1024]");
    NGP_REQUIRE(coords != nullptr || n == H * H * H, "occupancy_points: without coords, n must be the whole grid (H^3 = %u)", H * H * H);
    // the reference computes `bound / H` and `bound - hgs` in Python doubles and hands them to torch as scalars (-> fp32)
    const double hgs = (double)bound_c / (double)H;
    k_occupancy_points<<<div_up(n, kThreads), kThreads, 0, as_stream(stream)>>>(coords, jitter, n, H, (float)((double)bound_c - hgs), (float)hgs, xyzs);
    return after_launch("occupancy_points");
}

int ngp_density_grid_ema(float *grid, const float *tmp, uint32_t n, float decay, void *stream) {
    if (n == 0) return 0;
    NGP_DEVPTR(grid); NGP_DEVPTR(tmp);
    k_density_grid_ema<<<div_up(n, kThreads), kThreads, 0, as_stream(stream)>>>(grid, tmp, n, decay);
    return after_launch("density_grid_ema");
}

int ngp_mark_untrained_grid(const float *poses, uint32_t B, float fx, float fy, float cx, float cy, uint32_t H, uint32_t cascade, float bound,
                            float *density_grid, void *stream) {
    NGP_DEVPTR(density_grid);
    if (B) NGP_DEVPTR(poses);
    NGP_REQUIRE(H >= 2 && H <= 1024 && cascade >= 1 && cascade <= 8, "mark_untrained_grid: bad grid size / cascade");
    const uint32_t n = H * H * H * cascade;
    k_mark_untrained<<<div_up(n, kThreads), kThreads, 0, as_stream(stream)>>>(poses, B, cx / fx, cy / fy, H, cascade, bound, density_grid);
    return after_launch("mark_untrained_grid");
}

}  // extern "C"
