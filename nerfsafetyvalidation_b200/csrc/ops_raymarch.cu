// ops_raymarch.cu -- stand-alone raymarching ops behind the C ABI (drop-in for the reference's
// _raymarching extension, inference + setup functions) and the sync-free stable compaction that replaces
// `rays_alive[rays_alive >= 0]` (nerf/renderer.py:369).
#include <float.h>
#include <mutex>

#include "common.cuh"
#include "device_math.cuh"
#include "scan.cuh"

namespace ngp {

constexpr int kThreads = 256;

// raymarching/src/raymarching.cu:93-147
__global__ void __launch_bounds__(kThreads) k_near_far(const float *__restrict__ rays_o, const float *__restrict__ rays_d,
                                                      const float *__restrict__ aabb, uint32_t N, float min_near,
                                                      float *__restrict__ nears, float *__restrict__ fars) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float *o = rays_o + (size_t)n * 3, *d = rays_d + (size_t)n * 3;
    float near, far;
    if (!slab_near_far(o[0], o[1], o[2], d[0], d[1], d[2], aabb, min_near, near, far)) near = far = FLT_MAX;
    nears[n] = near;
    fars[n] = far;
}

// nerf/utils.py:52-116 (N = -1: every pixel): pixel centre (+0.5) -> camera direction ((i - cx) / fx, (j - cy) / fy, 1), normalised,
// rotated by the cam2world pose; origin = pose translation.  One thread per ray; poses [F,4,4] row-major.  Operation order is fixed
// (IEEE mul / add / div / sqrt, no contraction) so the oracle reproduces it bit for bit; torch's norm + cuBLAS matmul agree to 1 ulp.
__global__ void __launch_bounds__(kThreads) k_get_rays(const float *__restrict__ poses, uint32_t F, uint32_t H, uint32_t W, float fx, float fy,
                                                       float cx, float cy, float *__restrict__ rays_o, float *__restrict__ rays_d) {
    const size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t HW = H * W;
    if (n >= (size_t)F * HW) return;
    const uint32_t f = (uint32_t)(n / HW), p = (uint32_t)(n % HW);
    float o[3], d[3];
    pose_ray(poses + (size_t)f * 16, p, W, fx, fy, cx, cy, o, d);
#pragma unroll
    for (int k = 0; k < 3; k++) { rays_o[n * 3 + k] = o[k]; rays_d[n * 3 + k] = d[k]; }
}

// raymarching/src/raymarching.cu:216-262
__global__ void __launch_bounds__(kThreads) k_morton3d(const int *__restrict__ coords, uint32_t N, int *__restrict__ indices) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    indices[n] = (int)morton3d(coords[n * 3], coords[n * 3 + 1], coords[n * 3 + 2]);
}
__global__ void __launch_bounds__(kThreads) k_morton3d_invert(const int *__restrict__ indices, uint32_t N, int *__restrict__ coords) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const int ind = indices[n];
    coords[n * 3 + 0] = (int)compact_bits10((uint32_t)(ind >> 0));
    coords[n * 3 + 1] = (int)compact_bits10((uint32_t)(ind >> 1));
    coords[n * 3 + 2] = (int)compact_bits10((uint32_t)(ind >> 2));
}

// raymarching/src/raymarching.cu:269-291 : one thread packs 8 densities (two 16-byte loads) into a byte
__global__ void __launch_bounds__(kThreads) k_packbits(const float4 *__restrict__ grid, uint32_t N, float thresh,
                                                      uint8_t *__restrict__ bitfield) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float4 a = __ldg(grid + 2 * (size_t)n), b = __ldg(grid + 2 * (size_t)n + 1);
    uint32_t bits = 0;
    bits |= (a.x > thresh) ? 1u : 0u;   bits |= (a.y > thresh) ? 2u : 0u;
    bits |= (a.z > thresh) ? 4u : 0u;   bits |= (a.w > thresh) ? 8u : 0u;
    bits |= (b.x > thresh) ? 16u : 0u;  bits |= (b.y > thresh) ? 32u : 0u;
    bits |= (b.z > thresh) ? 64u : 0u;  bits |= (b.w > thresh) ? 128u : 0u;
    bitfield[n] = (uint8_t)bits;
}

// raymarching/src/raymarching.cu:706-814
__global__ void __launch_bounds__(kThreads) k_march_rays(uint32_t n_alive, uint32_t n_step, const int *__restrict__ rays_alive,
                                                        const float *__restrict__ rays_t, const float *__restrict__ rays_o,
                                                        const float *__restrict__ rays_d, MarchConst mc,
                                                        const float *__restrict__ fars, float *__restrict__ xyzs,
                                                        float *__restrict__ dirs, float *__restrict__ deltas) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_alive) return;
    const int index = rays_alive[n];
    const float *o = rays_o + (size_t)index * 3, *d = rays_d + (size_t)index * 3;
    const float dx = d[0], dy = d[1], dz = d[2];
    float *px = xyzs + (size_t)n * n_step * 3, *pd = dirs + (size_t)n * n_step * 3, *pl = deltas + (size_t)n * n_step * 2;
    march_ray(mc, o[0], o[1], o[2], dx, dy, dz, rays_t[index], fars[index], n_step,
              [&](uint32_t s, float x, float y, float z, float dt, float delta_t) {
                  px[s * 3 + 0] = x; px[s * 3 + 1] = y; px[s * 3 + 2] = z;
                  pd[s * 3 + 0] = dx; pd[s * 3 + 1] = dy; pd[s * 3 + 2] = dz;
                  pl[s * 2 + 0] = dt; pl[s * 2 + 1] = delta_t;
              });
}

// raymarching/src/raymarching.cu:828-913
__global__ void __launch_bounds__(kThreads) k_composite_rays(uint32_t n_alive, uint32_t n_step, int *__restrict__ rays_alive,
                                                            float *__restrict__ rays_t, const float *__restrict__ sigmas,
                                                            const float *__restrict__ rgbs, const float *__restrict__ deltas,
                                                            float *__restrict__ weights_sum, float *__restrict__ depth,
                                                            float *__restrict__ image) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_alive) return;
    const int index = rays_alive[n];
    const float *sg = sigmas + (size_t)n * n_step, *c = rgbs + (size_t)n * n_step * 3, *dl = deltas + (size_t)n * n_step * 2;
    RayAcc a;
    a.t = rays_t[index]; a.ws = weights_sum[index]; a.d = depth[index];
    a.r = image[(size_t)index * 3 + 0]; a.g = image[(size_t)index * 3 + 1]; a.b = image[(size_t)index * 3 + 2];
    uint32_t step = 0;
    while (step < n_step) {
        const float dt = dl[step * 2];
        if (dt == 0.0f) break;
        if (composite_sample(a, sg[step], c[step * 3], c[step * 3 + 1], c[step * 3 + 2], dt, dl[step * 2 + 1])) break;
        step++;
    }
    if (step < n_step) rays_alive[n] = -1; else rays_t[index] = a.t;
    weights_sum[index] = a.ws; depth[index] = a.d;
    image[(size_t)index * 3 + 0] = a.r; image[(size_t)index * 3 + 1] = a.g; image[(size_t)index * 3 + 2] = a.b;
}

// ------------------------------------------------------------------------------------------------
// Stable stream compaction: keeps in[i] >= 0, order preserved, count left on the device.
// One pass: warp ballot -> block scan -> decoupled look-back across tiles (tiles take tickets so that every
// predecessor a tile may spin on is already resident).
constexpr int kCpThreads = 256, kCpItems = 8, kCpTile = kCpThreads * kCpItems;

__global__ void __launch_bounds__(kCpThreads) k_compact(const int *__restrict__ in, uint32_t n, int *__restrict__ out,
                                                       uint32_t *__restrict__ d_count, uint32_t *ticket,
                                                       unsigned long long *status, uint32_t n_tiles) {
    __shared__ uint32_t s_tile, s_base, s_warp[kCpThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t first = tile * kCpTile + tid * kCpItems;
    int v[kCpItems];
    uint32_t cnt = 0;
    if (first + kCpItems <= n) {
        const int4 a = __ldg(reinterpret_cast<const int4 *>(in + first)), b = __ldg(reinterpret_cast<const int4 *>(in + first) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < kCpItems; k++) v[k] = (first + k < n) ? in[first + k] : -1;
    }
#pragma unroll
    for (int k = 0; k < kCpItems; k++) cnt += v[k] >= 0;
    const uint32_t incl = warp_inclusive_sum(cnt, lane);
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kCpThreads / 32; w++) {
        const uint32_t c = s_warp[w];
        warp_off += (w < warp) ? c : 0;
        total += c;
    }
    if (warp == 0) {
        const uint32_t base = lookback_exclusive_prefix(status, tile, total, /*tag=*/1u, lane);
        if (lane == 0) {
            s_base = base;
            if (tile == n_tiles - 1) *d_count = base + total;
        }
    }
    __syncthreads();
    uint32_t pos = s_base + warp_off + incl - cnt;
#pragma unroll
    for (int k = 0; k < kCpItems; k++)
        if (v[k] >= 0) out[pos++] = v[k];
}

// Scratch of ngp_compact_alive_static: ticket word + look-back status words for at most kCompactStaticMaxN
// elements, resident in every device that loads the library.  In host code this name is only a handle for
// cudaGetSymbolAddress; the kernel gets the per-device address resolved below.
constexpr uint32_t kCompactStaticMaxN = 1u << 22;
constexpr size_t kCompactStaticBytes = 16 + 8 * (size_t)(kCompactStaticMaxN / kCpTile + 1);
static __device__ __align__(256) unsigned char g_compact_scratch[(kCompactStaticBytes + 255) / 256 * 256];

constexpr int kMaxDevices = 64;
static std::mutex g_compact_scratch_mu;
static void *g_compact_scratch_addr[kMaxDevices] = {};

// Device address of g_compact_scratch on the CURRENT device, resolved once per device.
static int compact_static_scratch(void **addr) {
    int dev = -1;
    NGP_CUDA(cudaGetDevice(&dev));
    NGP_REQUIRE(dev >= 0 && dev < kMaxDevices, "compact_alive_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_compact_scratch_mu);
    if (!g_compact_scratch_addr[dev]) {
        void *p = nullptr;
        NGP_CUDA(cudaGetSymbolAddress(&p, g_compact_scratch));
        NGP_REQUIRE(p != nullptr, "compact_alive_static: cudaGetSymbolAddress returned a null address on device %d", dev);
        g_compact_scratch_addr[dev] = p;
    }
    *addr = g_compact_scratch_addr[dev];
    return 0;
}

}  // namespace ngp

using namespace ngp;

extern "C" {

int ngp_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, uint32_t N, float min_near,
                           float *nears, float *fars, void *stream) {
    if (N == 0) return 0;
    NGP_DEVPTR(rays_o); NGP_DEVPTR(rays_d); NGP_DEVPTR(aabb); NGP_DEVPTR(nears); NGP_DEVPTR(fars);
    k_near_far<<<div_up(N, kThreads), kThreads, 0, as_stream(stream)>>>(rays_o, rays_d, aabb, N, min_near, nears, fars);
    return after_launch("near_far_from_aabb");
}

int ngp_get_rays(const float *poses, uint32_t F, uint32_t H, uint32_t W, float fx, float fy, float cx, float cy, float *rays_o,
                 float *rays_d, void *stream) {
    const unsigned long long n = (unsigned long long)F * H * W;
    if (n == 0) return 0;
    NGP_REQUIRE(n <= 0xffffffffull, "get_rays: %llu rays exceed the 32-bit ray index of the renderer", n);
    NGP_REQUIRE(fx != 0.f && fy != 0.f, "get_rays: zero focal length");
    NGP_DEVPTR(poses); NGP_DEVPTR(rays_o); NGP_DEVPTR(rays_d);
    k_get_rays<<<div_up((uint32_t)n, kThreads), kThreads, 0, as_stream(stream)>>>(poses, F, H, W, fx, fy, cx, cy, rays_o, rays_d);
    return after_launch("get_rays");
}

int ngp_morton3D(const int32_t *coords, uint32_t N, int32_t *indices, void *stream) {
    if (N == 0) return 0;
    NGP_DEVPTR(coords); NGP_DEVPTR(indices);
    k_morton3d<<<div_up(N, kThreads), kThreads, 0, as_stream(stream)>>>(coords, N, indices);
    return after_launch("morton3D");
}

int ngp_morton3D_invert(const int32_t *indices, uint32_t N, int32_t *coords, void *stream) {
    if (N == 0) return 0;
    NGP_DEVPTR(coords); NGP_DEVPTR(indices);
    k_morton3d_invert<<<div_up(N, kThreads), kThreads, 0, as_stream(stream)>>>(indices, N, coords);
    return after_launch("morton3D_invert");
}

int ngp_packbits(const float *grid, uint32_t N, float density_thresh, uint8_t *bitfield, void *stream) {
    if (N == 0) return 0;
    NGP_DEVPTR(grid); NGP_DEVPTR(bitfield);
    NGP_REQUIRE((reinterpret_cast<uintptr_t>(grid) & 15) == 0, "packbits: grid must be 16-byte aligned");
    k_packbits<<<div_up(N, kThreads), kThreads, 0, as_stream(stream)>>>(reinterpret_cast<const float4 *>(grid), N,
                                                                         density_thresh, bitfield);
    return after_launch("packbits");
}

int ngp_march_rays(uint32_t n_alive, uint32_t n_step, const int32_t *rays_alive, const float *rays_t, const float *rays_o,
                   const float *rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t H,
                   const uint8_t *grid, const float *nears, const float *fars, float *xyzs, float *dirs, float *deltas,
                   uint32_t perturb, void *stream) {
    (void)nears;
    NGP_REQUIRE(perturb == 0, "march_rays: perturb != 0 is not supported on the inference path (validate.py:290 passes perturb=False)");
    NGP_REQUIRE(n_step >= 1 && max_steps >= 1 && C >= 1 && C <= 8 && H >= 2, "march_rays: bad n_step/max_steps/C/H");
    if (n_alive == 0) return 0;
    NGP_DEVPTR(rays_alive); NGP_DEVPTR(rays_t); NGP_DEVPTR(rays_o); NGP_DEVPTR(rays_d); NGP_DEVPTR(grid);
    NGP_DEVPTR(fars); NGP_DEVPTR(xyzs); NGP_DEVPTR(dirs); NGP_DEVPTR(deltas);
    const MarchConst mc = make_march_const(bound, dt_gamma, max_steps, C, H, grid);
    k_march_rays<<<div_up(n_alive, kThreads), kThreads, 0, as_stream(stream)>>>(n_alive, n_step, rays_alive, rays_t, rays_o,
                                                                                 rays_d, mc, fars, xyzs, dirs, deltas);
    return after_launch("march_rays");
}

int ngp_composite_rays(uint32_t n_alive, uint32_t n_step, int32_t *rays_alive, float *rays_t, const float *sigmas,
                       const float *rgbs, const float *deltas, float *weights_sum, float *depth, float *image, void *stream) {
    if (n_alive == 0) return 0;
    NGP_REQUIRE(n_step >= 1, "composite_rays: n_step must be >= 1");
    NGP_DEVPTR(rays_alive); NGP_DEVPTR(rays_t); NGP_DEVPTR(sigmas); NGP_DEVPTR(rgbs); NGP_DEVPTR(deltas);
    NGP_DEVPTR(weights_sum); NGP_DEVPTR(depth); NGP_DEVPTR(image);
    k_composite_rays<<<div_up(n_alive, kThreads), kThreads, 0, as_stream(stream)>>>(n_alive, n_step, rays_alive, rays_t, sigmas,
                                                                                     rgbs, deltas, weights_sum, depth, image);
    return after_launch("composite_rays");
}

uint64_t ngp_compact_scratch_bytes(uint32_t n) { return 16 + 8ull * (div_up(n, kCpTile) + 1); }

int ngp_compact_alive(const int32_t *in, uint32_t n, int32_t *out, uint32_t *d_count, void *d_scratch, void *stream) {
    NGP_DEVPTR(d_count);
    if (n == 0) {
        NGP_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), as_stream(stream)));
        return 0;
    }
    NGP_DEVPTR(in); NGP_DEVPTR(out); NGP_DEVPTR(d_scratch);
    NGP_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(d_scratch) & 7) == 0,
                "compact_alive: `in` must be 16-byte aligned, scratch 8-byte aligned");
    const uint32_t n_tiles = div_up(n, kCpTile);
    NGP_CUDA(cudaMemsetAsync(d_scratch, 0, ngp_compact_scratch_bytes(n), as_stream(stream)));
    uint32_t *ticket = reinterpret_cast<uint32_t *>(d_scratch);
    unsigned long long *status = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(d_scratch) + 16);
    k_compact<<<n_tiles, kCpThreads, 0, as_stream(stream)>>>(in, n, out, d_count, ticket, status, n_tiles);
    return after_launch("compact_alive");
}

int ngp_compact_alive_static(const int32_t *in, uint32_t n, int32_t *out, uint32_t *d_count, void *stream) {
    NGP_DEVPTR(d_count);
    NGP_REQUIRE(n <= kCompactStaticMaxN,
                "compact_alive_static: n = %u exceeds the built-in scratch (at most %u elements); use ngp_compact_alive with caller scratch",
                n, kCompactStaticMaxN);
    if (n == 0) {
        NGP_CUDA(cudaMemsetAsync(d_count, 0, sizeof(uint32_t), as_stream(stream)));
        return 0;
    }
    NGP_DEVPTR(in); NGP_DEVPTR(out);
    NGP_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0, "compact_alive_static: `in` must be 16-byte aligned");
    void *scratch = nullptr;
    if (int rc = compact_static_scratch(&scratch)) return rc;
    const uint32_t n_tiles = div_up(n, kCpTile);
    // the array keeps the previous call's ticket and status words: clear the part this call reads
    NGP_CUDA(cudaMemsetAsync(scratch, 0, ngp_compact_scratch_bytes(n), as_stream(stream)));
    uint32_t *ticket = reinterpret_cast<uint32_t *>(scratch);
    unsigned long long *status = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(scratch) + 16);
    k_compact<<<n_tiles, kCpThreads, 0, as_stream(stream)>>>(in, n, out, d_count, ticket, status, n_tiles);
    return after_launch("compact_alive_static");
}

}  // extern "C"
