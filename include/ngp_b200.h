/*
 * ngp_b200.h -- C ABI of the B200-native Instant-NGP inference renderer.
 *
 * This is the drop-in boundary for the hot path of sisl/NeRFSafetyValidation
 * (NeRFRenderer.render(cuda_ray=True), inference branch).  Every entry point replaces one function the
 * reference exports from its pybind11 extensions (file:line below are relative to the reference root);
 * the signatures are the reference's, with at::Tensor replaced by a raw DEVICE pointer + sizes and an
 * explicit stream (the reference launches everything on the legacy default stream).
 *
 * Conventions
 *   - All pointers are device pointers unless the name ends in `_host`.
 *   - The caller owns every buffer; the library allocates nothing except inside an ngp_ctx
 *     (explicitly created / destroyed workspace for the fused path).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - Every function returns 0 on success, non-zero on error; ngp_last_error() then describes it
 *     (thread-local).  Nothing throws across the boundary.  There is NO CPU fallback.
 *   - Kernels are enqueued asynchronously; only *_host entry points and ngp_ctx_* synchronise.
 */
#ifndef NGP_B200_H
#define NGP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
/* the library is built with -fvisibility=hidden; exactly the entry points declared here are exported */
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define NGP_ABI_VERSION 1

#define NGP_DTYPE_F32 0
#define NGP_DTYPE_F16 1

int ngp_abi_version(void);
const char *ngp_last_error(void);
/* number of kernels this library has launched since load (bench.py's gpu_launches claim) */
uint64_t ngp_launch_count(void);

/* ---------------------------------------------------------------- raymarching (inference + setup)
 * replaces _raymarching.{near_far_from_aabb, morton3D, morton3D_invert, packbits, march_rays,
 * composite_rays}: raymarching/src/bindings.cpp:5-18, raymarching/src/raymarching.h:7-18 */

/* raymarching/src/raymarching.cu:93-158.  aabb: device float[6]. Misses write FLT_MAX to both. */
int ngp_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, uint32_t N,
                           float min_near, float *nears, float *fars, void *stream);
/* nerf/utils.py:52-116 (get_rays, N = -1: every pixel, centres at +0.5): poses = device float [F,4,4] row-major cam2world,
 * rays_o / rays_d = device float [F, H*W, 3].  Replaces the meshgrid + stack + norm + matmul ATen sequence the reference runs
 * before every render (validate.py:290 -> nav/estimator_helpers.py:236-243). */
int ngp_get_rays(const float *poses, uint32_t F, uint32_t H, uint32_t W, float fx, float fy, float cx, float cy,
                 float *rays_o, float *rays_d, void *stream);
/* raymarching/src/raymarching.cu:216-235 */
int ngp_morton3D(const int32_t *coords, uint32_t N, int32_t *indices, void *stream);
/* raymarching/src/raymarching.cu:240-262 */
int ngp_morton3D_invert(const int32_t *indices, uint32_t N, int32_t *coords, void *stream);
/* raymarching/src/raymarching.cu:269-302.  N = number of output BYTES (C*H^3/8). */
int ngp_packbits(const float *grid, uint32_t N, float density_thresh, uint8_t *bitfield, void *stream);
/* Occupancy-grid maintenance (nerf/renderer.py:388-546), in the Morton order density_grid is stored in.
 * ngp_occupancy_points: the query points of one cascade, xyz [n,3] fp32 =
 *     ((2 c) / (H - 1) - 1) * (bound_c - bound_c / H) + (2 u - 1) * bound_c / H          (renderer.py:481-490, same roundings)
 *   coords == NULL: n = H^3, point m = the cell with Morton index m, u = jitter row (x H + y) H + z (the row the reference's
 *                   meshgrid enumeration gives that cell: the same random stream lands on the same cells);
 *   coords int32 [n,3]: point i = cell coords[i], u = jitter row i (the partial-update branch, :497-523);
 *   jitter fp32 [rows,3] in [0,1) or NULL (cell centres).
 * ngp_density_grid_ema: grid = max(grid * decay, tmp) where both are >= 0 (:531-532), in place.
 * ngp_mark_untrained_grid: density_grid[c][m] = -1 for every cell no pose (cam2world [B,4,4] fp32, device) sees (:388-449). */
int ngp_occupancy_points(const int32_t *coords, const float *jitter, uint32_t n, uint32_t H, float bound_c,
                         float *xyzs, void *stream);
int ngp_density_grid_ema(float *grid, const float *tmp, uint32_t n, float decay, void *stream);
int ngp_mark_untrained_grid(const float *poses, uint32_t B, float fx, float fy, float cx, float cy, uint32_t H,
                            uint32_t cascade, float bound, float *density_grid, void *stream);
/* raymarching/src/raymarching.cu:706-825.  xyzs/dirs/deltas must be pre-zeroed by the caller
 * (raymarching.py:329-331).  perturb must be 0 (validate.py:290 never perturbs at inference). */
int ngp_march_rays(uint32_t n_alive, uint32_t n_step, const int32_t *rays_alive, const float *rays_t,
                   const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps,
                   uint32_t C, uint32_t H, const uint8_t *grid, const float *nears, const float *fars,
                   float *xyzs, float *dirs, float *deltas, uint32_t perturb, void *stream);
/* raymarching/src/raymarching.cu:828-922.  In place on rays_alive, rays_t, weights_sum, depth, image. */
int ngp_composite_rays(uint32_t n_alive, uint32_t n_step, int32_t *rays_alive, float *rays_t,
                       const float *sigmas, const float *rgbs, const float *deltas, float *weights_sum,
                       float *depth, float *image, void *stream);
/* nerf/renderer.py:369 `rays_alive = rays_alive[rays_alive >= 0]` as a stable, sync-free stream compaction
 * (warp ballot + block scan + decoupled look-back).  d_count: device uint32 receiving the survivor count.
 * d_scratch: device scratch of at least ngp_compact_scratch_bytes(n) bytes (contents irrelevant). */
uint64_t ngp_compact_scratch_bytes(uint32_t n);
int ngp_compact_alive(const int32_t *in, uint32_t n, int32_t *out, uint32_t *d_count, void *d_scratch,
                      void *stream);
/* No counterpart in the reference; added for callers that cannot
 * allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as
 * ngp_compact_alive, but the look-back scratch is a file-scope __device__ array of the library (about 16 KiB,
 * held by every GPU that loads the library until exit), so the call allocates nothing.  n above
 * NGP_COMPACT_STATIC_MAX_N is rejected with an error.  All calls on one device share the array: the call
 * clears what it reads on `stream` first, and two calls that overlap on the GPU (different streams without
 * an ordering between them) race -- serialise them or use ngp_compact_alive with separate scratch. */
#define NGP_COMPACT_STATIC_MAX_N (1u << 22)
int ngp_compact_alive_static(const int32_t *in, uint32_t n, int32_t *out, uint32_t *d_count, void *stream);

/* ---------------------------------------------------------------- encoders
 * replaces _gridencoder.grid_encode_forward (gridencoder/src/bindings.cpp:5-8, gridencoder.cu:423-446)
 * for D=3, calc_grad_inputs=False.  table_dtype selects float / half table AND output dtype, like the
 * reference's AT_DISPATCH.  out_layout 0 = [L,B,C] (the reference kernel's), 1 = [B,L*C] (what
 * grid.py:52 permutes to; saves the permute copy).  Half tables: ref_rounding=1 reproduces the reference's
 * round-to-half after every corner (gridencoder.cu:143,166) bit for bit; 0 accumulates in fp32.
 * idx_out: optional int32 [L,B,8] receiving the table rows gathered (integer parity hook), or NULL. */
int ngp_grid_encode_forward(const float *inputs01, const void *table, int table_dtype, const int32_t *offsets,
                            void *outputs, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H,
                            uint32_t gridtype, int align_corners, int out_layout, int ref_rounding,
                            int32_t *idx_out, void *stream);
/* per-level `scale = exp2f(level*S)*H - 1` exactly as the device computes it (gridencoder.cu:126) */
int ngp_grid_level_scales(uint32_t L, float S, uint32_t H, float *scales, void *stream);
/* replaces _shencoder.sh_encode_forward (shencoder/src/shencoder.cu:388-419), degree 1..4, fp32 */
int ngp_sh_encode_forward(const float *inputs, float *outputs, uint32_t B, uint32_t degree, void *stream);

/* ---------------------------------------------------------------- fully fused MLP
 * replaces _ffmlp.ffmlp_inference (ffmlp/src/ffmlp.cu:673-709) for hidden_dim=64, ReLU, output_dim=16:
 * inputs [B,in_dim] half (B % 128 == 0, in_dim in {32,64}), weights = contiguous row-major
 * [64][in_dim] + (num_layers-1) x [64][64] + [16][64] halves (ffmlp.py:121), outputs [B,16] half.
 * fp16 operands, fp32 accumulation on tensor cores. */
int ngp_ffmlp_inference(const void *inputs_f16, const void *weights_f16, uint32_t B, uint32_t input_dim,
                        uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                        uint32_t output_activation, void *outputs_f16, void *stream);

/* ---------------------------------------------------------------- fused path (no upstream counterpart)
 * One context = one model resident on one device: half hash table, offsets, both MLPs, occupancy
 * bitfield, plus a ray-state workspace that grows to the largest N seen. */
typedef struct ngp_ctx ngp_ctx;

int ngp_ctx_create(ngp_ctx **ctx);
int ngp_ctx_destroy(ngp_ctx *ctx);
/* Registers the network of nerf/network_ff.py:11-49 (hashgrid L levels x 2 ch -> sigma_net -> SH4 + colour net).
 * Pointers are borrowed (must stay valid); offsets_host is a HOST int32[L+1]. */
int ngp_ctx_set_network(ngp_ctx *ctx, const void *table_f16, const int32_t *offsets_dev,
                        const int32_t *offsets_host, uint32_t L, float S, uint32_t H,
                        const void *w_sigma_f16, uint32_t nl_sigma, const void *w_color_f16,
                        uint32_t nl_color, float bound, float density_scale, void *stream);
/* nerf/renderer.py:73-99 state: bitfield [cascade*grid^3/8] bytes (borrowed), aabb = HOST float[6]. */
int ngp_ctx_set_scene(ngp_ctx *ctx, const uint8_t *bitfield, uint32_t cascade, uint32_t grid_size,
                      const float *aabb_host, float min_near);
/* Tunables of ngp_render's epoch policy, per context (the library keeps no process-wide state and reads no environment
 * variables; no upstream counterpart).  name: "rows_per_ray" (sample-buffer rows per ray, default 64), "max_epoch" (cap on the
 * reference iterations per epoch, 0 = none; 1 = one barrier pair per reference iteration), "kshift" (log2 samples per ray per
 * pass, >= 4 = automatic), "tile_balance" (automatic tile geometry: warp tiles per pipeline below which tiles get shorter, default 4), "t_div", "t_cap", "e0_cap", "rate0_div", "early_cap", "early_steps", "horizon_div" (epoch-length policy),
 * "no_coarse" (1 = march without the coarse occupancy grid).  No option changes any output.  Unknown names are an error. */
int ngp_ctx_set_option(ngp_ctx *ctx, const char *name, uint64_t value);
/* Diagnostics for the persistent renderer (no upstream counterpart): when enabled, every warp of ngp_render accumulates
 * clock64() cycles per phase.  ngp_ctx_phase_cycles synchronises the device and copies the 8 sums to HOST memory:
 * {0 tile set-up / tile end + prologue / epilogue / loop control, 1 march, 2 grid barrier after march, 3 gather warps: sample
 *  load + hash-grid gather + operand store, 4 epilogue warps: activations + compositing, 5 gather warps waiting (operand slot,
 *  accumulators), 6 epilogue warps waiting (layer commits), 7 grid barrier after evaluate}, each summed over all warps of all
 * launches since the last reset. */
int ngp_ctx_enable_profile(ngp_ctx *ctx, int enable);
int ngp_ctx_phase_cycles(ngp_ctx *ctx, int64_t *out_host, int reset);
/* NeRFNetwork.forward (nerf/network_ff.py:51-74) in one kernel: xyzs,dirs [M,3] fp32 -> sigmas [M] fp32,
 * rgbs [M,3] fp32.  apply_density_scale != 0 also folds in `sigmas *= density_scale` (renderer.py:365).
 * Any M >= 0 (rows padded internally). */
int ngp_network_forward(ngp_ctx *ctx, const float *xyzs, const float *dirs, uint32_t M, int apply_density_scale,
                        float *sigmas, float *rgbs, void *stream);
/* Same contract as ngp_network_forward, both MLPs on tcgen05 (128-row tiles, accumulators in tensor memory). */
int ngp_network_forward_tc(ngp_ctx *ctx, const float *xyzs, const float *dirs, uint32_t M, int apply_density_scale,
                           float *sigmas, float *rgbs, void *stream);
/* NeRFNetwork.density (nerf/network_ff.py:76-89; `density_fn` of validate.py:288, the planner's and update_extra_state's query) in
 * one kernel: xyzs [M,3] fp32 -> sigmas [M] fp32 = exp(h0) (x density_scale if apply_density_scale != 0) and, if geo_feat_f16
 * is not NULL, geo_feat [M,15] half = h[1:16].  Hash-grid gather + sigma net on tcgen05; nothing else touches HBM. */
int ngp_density_forward(ngp_ctx *ctx, const float *xyzs, uint32_t M, int apply_density_scale, float *sigmas,
                        void *geo_feat_f16, void *stream);
/* NeRFRenderer.run_cuda, inference branch (nerf/renderer.py:261-284,329-386) as ONE persistent launch:
 * near/far -> loop{march -> encode -> MLPs -> composite -> stable compaction, global n_step schedule} ->
 * background mix + depth normalisation.  image [N,3], depth [N], weights_sum [N] fp32 (any may not be NULL).
 * samples_per_ray: optional int32 [N]; stats: optional device int64[4] = {iterations, samples, tiles, 0}. */
int ngp_render(ngp_ctx *ctx, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma,
               uint32_t max_steps, float bg_color, float *image, float *depth, float *weights_sum,
               int32_t *samples_per_ray, int64_t *stats, void *stream);
/* ngp_render with the per-ray / per-sample outputs the reference's uncertainty-quantification callers read (uncertain.py:80-91,
 * 185-189; nerf/renderer.py:244, 381-384), all optional (NULL = not wanted):
 *   aggregated_density [N] fp32     sum over the ray's samples of weight * sigma (renderer.py:244, which only the sampler path
 *                                   `run()` computes upstream)
 *   sample_dump [dump_rows][4] fp32 (sigma * density_scale, r, g, b) of EVERY sample the LAST epoch evaluated, at the sample's row of
 *                                   the epoch's sample buffer; dump_rows >= ngp_render_sample_rows(ctx, N).  With it:
 *   ray_info [N][4] int32           by ray id: (epoch tag, position in that epoch's live list, samples marched in it, iteration of
 *                                   the epoch in which the ray fell below T = 1e-4 or INT_MAX)
 *   dump_header [8] int64           of the last epoch: its tag, n_step, reference iterations it ran, log2 K, row slots per ray,
 *                                   live rays at its start, samples per ray, 0.  Sample i of live-list position n sits at row
 *                                   (n / R) * R * slots + (i / K) * 32 + (n % R) * K + i % K with R = 32 / K.
 * nerfsafetyvalidation_b200/nerf/renderer.py assembles the reference's `sigmas` / `rgbs` (its last iteration's tensors) from these. */
int ngp_render_uq(ngp_ctx *ctx, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma,
                  uint32_t max_steps, float bg_color, float *image, float *depth, float *weights_sum,
                  int32_t *samples_per_ray, int64_t *stats, float *aggregated_density, float *sample_dump,
                  uint64_t dump_rows, int32_t *ray_info, int64_t *dump_header, void *stream);
/* Rows of the epoch sample buffer a render of N rays uses (HOST out): the size sample_dump must have. */
int ngp_render_sample_rows(ngp_ctx *ctx, uint32_t N, uint64_t *rows);
/* Same, HOST buffers in and out: H2D of rays, render, D2H of image/depth; synchronous.  This is the call the
 * end-to-end number in bench.py is measured through. */
int ngp_render_host(ngp_ctx *ctx, const float *rays_o_host, const float *rays_d_host, uint32_t N,
                    float dt_gamma, uint32_t max_steps, float bg_color, float *image_host, float *depth_host,
                    void *stream);

/* ngp_render for F camera frames given as cam2world poses [F,4,4] fp32 (DEVICE): the rays of every pixel (nerf/utils.py:52-116, N = -1)
 * are generated in the render kernel's prologue -- no ray tensors cross the boundary (64 B per frame instead of 24 B per ray).
 * Outputs as ngp_render with N = F * H * W rays in frame-major, row-major pixel order. */
int ngp_render_poses(ngp_ctx *ctx, const float *poses, uint32_t F, uint32_t H, uint32_t W, float fx, float fy,
                     float cx, float cy, float dt_gamma, uint32_t max_steps, float bg_color, float *image,
                     float *depth, float *weights_sum, int32_t *samples_per_ray, int64_t *stats, void *stream);
/* Estimator.render_from_pose (nav/estimator_helpers.py:236-243) for a batch of F cameras: HOST poses [F,4,4] in (64 B per
 * frame instead of 24 B per ray), rays generated on the device (ngp_get_rays), one render over all F*H*W rays (one n_step
 * schedule, exactly like renderer.py:265-269 flattens a [F,N,3] batch), HOST image [F*H*W,3] / depth [F*H*W] out; synchronous. */
int ngp_render_poses_host(ngp_ctx *ctx, const float *poses_host, uint32_t F, uint32_t H, uint32_t W, float fx, float fy,
                          float cx, float cy, float dt_gamma, uint32_t max_steps, float bg_color, float *image_host,
                          float *depth_host, void *stream);

/* ---------------------------------------------------------------- differentiable small-batch path
 * What Estimator.measurement_fn (nav/estimator_helpers.py:191-225: <= 1024 rays, loss.backward() to the pose, 100 iterations)
 * needs from the CUDA-ray renderer, which upstream cannot give it: raymarching.py:161-230 (march_rays_train) has no backward, so
 * validate.py:117 falls back to the sampler `run()` whenever gradients w.r.t. rays are wanted.  Here the CUDA-ray render itself is
 * differentiable w.r.t. rays_o / rays_d.  One warp per ray, no workspace: the backward pass re-marches and re-evaluates.
 *
 * forward: the inference render of the same rays with the same march / network / compositor functions as ngp_render, each ray marched
 * as ONE chunk of at most max_steps samples (the reference restarts its marcher every n_step = clamp(N / n_alive, 1, 8) samples from
 * a re-summed t, raymarching.cu:723-725,870,905; after a large empty gap that t can be one ulp off, so a few rays in 10^5 take one
 * sample more or fewer there than here), plus `finals` [N,4] fp32 (16-byte aligned) = raw (r, g, b, weights_sum) per ray, which
 * the backward pass reads.  Also the faster renderer for small inference batches (no cooperative launch, no workspace): see
 * NeRFRenderer.small_batch_rays. */
int ngp_render_diff_forward(ngp_ctx *ctx, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma,
                            uint32_t max_steps, float bg_color, float *image, float *depth, float *weights_sum,
                            float *finals, int32_t *samples_per_ray, void *stream);
/* backward: grad_image [N,3], grad_weights_sum [N] or NULL -> grad_rays_o, grad_rays_d [N,3] (overwritten).  Chain: compositor
 * (raymarching.cu:606-699; depth carries no gradient there either) -> sigmoid / exp -> colour and density MLPs, input gradients only
 * (ffmlp.cu backward with calc_grad_inputs; fp16 operands, fp32 accumulation, per-row power-of-two rescaling before every fp16
 * conversion) -> SH Jacobian (shencoder.cu:126-355) and trilinear hash-grid input gradient (gridencoder.cu:180-223) -> x = o + t d,
 * dir = d with the sample parameters t held constant (near/far and the marcher are not differentiated upstream either).  The
 * network and scene registered with ctx must be the ones the forward call saw. */
int ngp_render_diff_backward(ngp_ctx *ctx, const float *rays_o, const float *rays_d, uint32_t N, float dt_gamma,
                             uint32_t max_steps, float bg_color, const float *finals, const float *grad_image,
                             const float *grad_weights_sum, float *grad_rays_o, float *grad_rays_d, void *stream);
/* Backward of ngp_density_forward w.r.t. the points (network_ff.py:76-89 under autograd: gridencoder/grid.py:62-90 with
 * calc_grad_inputs + ffmlp.py:51-96; the planner's density_fn queries, nav/quad_plot.py:223-300): grad_sigmas [M] -> grad_xyzs [M,3].
 * geo_feat carries no gradient. */
int ngp_density_backward(ngp_ctx *ctx, const float *xyzs, const float *grad_sigmas, uint32_t M, int apply_density_scale,
                         float *grad_xyzs, void *stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* NGP_B200_H */
