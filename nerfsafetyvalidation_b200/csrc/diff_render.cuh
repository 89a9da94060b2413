// diff_render.cuh -- the differentiable small-batch side of the fast path (included by fused.cu inside namespace ngp, after the
// mma.sync building blocks): what Estimator.measurement_fn (nav/estimator_helpers.py:191-225, <= 1024 rays, loss.backward() on the
// pose) and the planner's density queries (nav/quad_plot.py:223-300 through validate.py:288) need from the CUDA-ray path.
//
//   (second template argument of every kernel: kGeneral = the hash grid has levels that need the integer-division index paths;
//    false for every NGP table, which keeps those paths out of the instruction stream)
//   k_diff_render<false>   forward: one WARP per ray -- march (32 candidate parameters per step, see march_window) -> 32-sample chunks -> hash grid ->
//                          both MLPs (mma.sync, fp32 accumulate) -> front-to-back compositing with the T < 1e-4 stop.  Same functions
//                          as the inference renderer; a ray is marched as ONE chunk (the reference's batch-dependent chunk restarts can
//                          move one sample of a few rays in 10^5, see include/ngp_b200.h).
//   k_diff_render<true>    backward: re-runs the forward of every chunk (nothing but 16 B per ray is kept between the passes) and
//                          pushes d loss / d image through the compositor (raymarching.cu:606-699), the colour and density MLPs
//                          (input gradients only: ffmlp.cu's backward with calc_grad_inputs), the SH Jacobian (shencoder.cu:126-355)
//                          and the trilinear hash-grid input gradient (gridencoder.cu:180-223, dy_dx) to d loss / d rays_o, d rays_d.
//                          Sample positions t are constants of the derivative, as they are in the reference (near/far and the
//                          marcher have no backward there either): x = o + t d, dir = d.
//   k_density_backward     d loss / d xyz of NeRFNetwork.density (network_ff.py:76-89) for a [M,3] batch of points.
//
// Gradient arithmetic: the MLP backward runs on fp16 tensor-core operands with fp32 accumulation like the forward.  Upstream
// gradients of a pixel loss are far below fp16's range (2 (c - c*) / (3 N) * w ~ 1e-6), and the reference's half backward loses them;
// here every row carries a power-of-two scale that is re-derived from the row's largest entry before EVERY fp32 -> fp16 conversion
// (exact: the chain is linear per row once the ReLU masks are fixed), so the result does not depend on the loss's magnitude.
#pragma once

namespace diff {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;
constexpr int kMaxHidden = 4;        // hidden activations per net = num_layers <= 4
constexpr int kDencStride = 33;      // floats per staged d(encoding) row: conflict-free row reads

// ---------------------------------------------------------------------------------------------- transposed weight pack
// B fragments of W^T for the input-gradient matmuls dX = dY W: same order, sizes and offsets as k_pack_network's forward pack.
template <class Get>
__device__ __forceinline__ void pack_frags_fn(int N, int K, uint2 *__restrict__ dst, int tid, int nthreads, Get &&get) {
    const int KT = K / 16, NT = N / 8;
    for (int i = tid; i < NT * KT * 32; i += nthreads) {
        const int lane = i & 31, tile = i >> 5, kt = tile % KT, nt = tile / KT;
        const int g = lane >> 2, t = lane & 3;
        const int n = nt * 8 + g;
        uint32_t v[2];
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int k = kt * 16 + 2 * t + 8 * r;
            const __half2 p = __halves2half2(get(n, k), get(n, k + 1));
            v[r] = *reinterpret_cast<const uint32_t *>(&p);
        }
        dst[i] = make_uint2(v[0], v[1]);
    }
}

// W: row-major [out][in] (ffmlp.py:121).  Transposed operand T[n = in][k = out] = W[k][n].
__device__ __forceinline__ void pack_transposed(const __half *__restrict__ W, int out, int in, uint2 *dst, int tid, int nt) {
    pack_frags_fn(in, out, dst, tid, nt, [=](int n, int k) { return W[k * in + n]; });
}
// First colour layer: the forward pack shifts its input columns (packed column 16 <- zero, 17..31 <- source 16..30); T[n = packed in][k = out].
__device__ __forceinline__ void pack_transposed_shifted(const __half *__restrict__ W, uint2 *dst, int tid, int nt) {
    pack_frags_fn(32, 64, dst, tid, nt, [=](int n, int k) {
        return n < 16 ? W[k * 32 + n] : (n == 16 ? __float2half(0.f) : W[k * 32 + n - 1]);
    });
}

}  // namespace diff

__global__ void k_pack_network_transposed(const __half *__restrict__ w_sigma, uint32_t nh_sigma, const __half *__restrict__ w_color,
                                          uint32_t nh_color, uint2 *__restrict__ dst) {
    const int tid = threadIdx.x, nt = blockDim.x;
    uint2 *p = dst;
    diff::pack_transposed(w_sigma, 64, 32, p, tid, nt); p += 512;
    for (uint32_t k = 0; k < nh_sigma; k++) { diff::pack_transposed(w_sigma + 2048 + (size_t)k * 4096, 64, 64, p, tid, nt); p += 1024; }
    diff::pack_transposed(w_sigma + 2048 + (size_t)nh_sigma * 4096, 16, 64, p, tid, nt); p += 256;
    diff::pack_transposed_shifted(w_color, p, tid, nt); p += 512;
    for (uint32_t k = 0; k < nh_color; k++) { diff::pack_transposed(w_color + 2048 + (size_t)k * 4096, 64, 64, p, tid, nt); p += 1024; }
    diff::pack_transposed(w_color + 2048 + (size_t)nh_color * 4096, 16, 64, p, tid, nt);
}

namespace diff {

// ---------------------------------------------------------------------------------------------- forward with ReLU masks
// Bit (mt * 8 + nt) * 4 + e of a mask <- accumulator C[mt][nt][e] of a 64-wide hidden layer is positive (its ReLU passes).
__device__ __forceinline__ uint64_t relu_mask(const float (&C)[2][8][4]) {
    uint64_t m = 0;
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int nt = 0; nt < 8; nt++)
#pragma unroll
            for (int e = 0; e < 4; e++) m |= (uint64_t)(C[mt][nt][e] > 0.0f) << ((mt * 8 + nt) * 4 + e);
    return m;
}

struct NetOffsets {   // uint2 offsets into the forward pack and (identically) the transposed pack
    uint32_t s0, sh, sl, c0, ch, cl;
};
__device__ __forceinline__ NetOffsets net_offsets(const NetConst &nc) {
    NetOffsets o;
    o.s0 = 0; o.sh = 512; o.sl = o.sh + nc.nh_sigma * 1024; o.c0 = o.sl + 256; o.ch = o.c0 + 512; o.cl = o.ch + nc.nh_color * 1024;
    return o;
}

// A fragments of 32 staged fp16 rows of width 32 (stride in halves)
__device__ __forceinline__ void load_frags32(const __half *s_rows, int stride, int row0, int lane, uint32_t (&A0)[2][2][4]) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        const uint32_t *lo = reinterpret_cast<const uint32_t *>(s_rows + (row0 + mt * 16 + g) * stride);
        const uint32_t *hi = reinterpret_cast<const uint32_t *>(s_rows + (row0 + mt * 16 + g + 8) * stride);
#pragma unroll
        for (int kt = 0; kt < 2; kt++) {
            A0[mt][kt][0] = lo[kt * 8 + t]; A0[mt][kt][1] = hi[kt * 8 + t];
            A0[mt][kt][2] = lo[kt * 8 + 4 + t]; A0[mt][kt][3] = hi[kt * 8 + 4 + t];
        }
    }
}

// One net of network_ff.py:26-49 on this warp's 32 rows: [32] -> 64 -> (nh x 64) -> 16 output columns (O holds columns 0..15).
// masks[0] <- first layer's ReLU, masks[1 + k] <- hidden matmul k's.
// (inlined on purpose: a non-inlined copy shared by both nets halves the code but passes the fragments through local memory, which
// costs the warp-per-ray kernel 1.6x of its throughput at 16 000+ rays)
__device__ __forceinline__ void net_forward(const uint2 *__restrict__ w_first, const uint2 *__restrict__ w_hidden, const uint2 *__restrict__ w_last,
                                            uint32_t nh, const uint32_t (&A0)[2][2][4], int lane, float (&O)[2][2][4], uint64_t (&masks)[kMaxHidden]) {
    float C[2][8][4];
    uint32_t A[2][4][4];
    mlp_layer<2, 2, 8>(A0, w_first, lane, C);
    masks[0] = relu_mask(C);
    relu_to_frag<2, 8>(C, A);
    // rolled: one copy of the 64-MMA hidden layer in the instruction stream (these kernels run 4 warps per CTA; code size, not
    // issue rate, is what an unrolled chain of layers would cost them)
#pragma unroll 1
    for (uint32_t k = 0; k < nh; k++) {
        mlp_layer<2, 4, 8>(A, w_hidden + (size_t)k * 1024, lane, C);
        const uint64_t mk = relu_mask(C);
#pragma unroll
        for (int i = 1; i < kMaxHidden; i++) masks[i] = (uint32_t)i == k + 1 ? mk : masks[i];
        relu_to_frag<2, 8>(C, A);
    }
    mlp_layer<2, 4, 2>(A, w_last, lane, O);
}

// ---------------------------------------------------------------------------------------------- backward
// Power-of-two factor that brings m into [1, 2) (1 for zero / subnormal m).
__device__ __forceinline__ float pow2_normaliser(float m) {
    const uint32_t e = __float_as_uint(m) & 0x7f800000u;
    return e == 0u ? 1.0f : __uint_as_float(0x7f000000u - e);
}

// fp32 gradient accumulators of a layer with 8 NT columns -> fp16 A fragments of the next (earlier) layer's matmul.  Entries whose
// ReLU did not pass are dropped (kMask), then each ROW is rescaled by a power of two so that its largest entry lies in [1, 2);
// rs[mt][h] (row g + 8 h + 16 mt) accumulates the factor.
template <int NT, bool kMask>
__device__ __forceinline__ void grad_to_frag(float (&C)[2][NT][4], uint64_t mask, uint32_t (&A)[2][NT / 2][4], float (&rs)[2][2]) {
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        float m[2] = {0.f, 0.f};
#pragma unroll
        for (int nt = 0; nt < NT; nt++)
#pragma unroll
            for (int e = 0; e < 4; e++) {
                if (kMask && !((mask >> ((mt * 8 + nt) * 4 + e)) & 1ull)) C[mt][nt][e] = 0.0f;
                m[e >> 1] = fmaxf(m[e >> 1], fabsf(C[mt][nt][e]));
            }
        float f[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            m[h] = fmaxf(m[h], __shfl_xor_sync(0xffffffffu, m[h], 1));
            m[h] = fmaxf(m[h], __shfl_xor_sync(0xffffffffu, m[h], 2));
            f[h] = pow2_normaliser(m[h]);
            rs[mt][h] *= f[h];
        }
#pragma unroll
        for (int kt = 0; kt < NT / 2; kt++) {
            A[mt][kt][0] = pack_half2(C[mt][2 * kt][0] * f[0], C[mt][2 * kt][1] * f[0]);
            A[mt][kt][1] = pack_half2(C[mt][2 * kt][2] * f[1], C[mt][2 * kt][3] * f[1]);
            A[mt][kt][2] = pack_half2(C[mt][2 * kt + 1][0] * f[0], C[mt][2 * kt + 1][1] * f[0]);
            A[mt][kt][3] = pack_half2(C[mt][2 * kt + 1][2] * f[1], C[mt][2 * kt + 1][3] * f[1]);
        }
    }
}

// Input gradient of one net: G (fp32, [32 rows][16 output columns], row-scaled by rs) -> Cin ([32][32 input columns], scaled by the
// updated rs).  wt_* = the transposed pack.
__device__ __forceinline__ void net_backward(const uint2 *__restrict__ wt_first, const uint2 *__restrict__ wt_hidden, const uint2 *__restrict__ wt_last,
                                             uint32_t nh, float (&G)[2][2][4], const uint64_t (&masks)[kMaxHidden], int lane, float (&rs)[2][2],
                                             float (&Cin)[2][4][4]) {
    uint32_t Aout[2][1][4];
    grad_to_frag<2, false>(G, 0ull, Aout, rs);
    float C[2][8][4];
    uint32_t A[2][4][4];
    mlp_layer<2, 1, 8>(Aout, wt_last, lane, C);       // d(last hidden activation)
#pragma unroll 1
    for (int k = (int)nh - 1; k >= 0; k--) {
        uint64_t mk = masks[1];
#pragma unroll
        for (int i = 2; i < kMaxHidden; i++) mk = i == k + 1 ? masks[i] : mk;
        grad_to_frag<8, true>(C, mk, A, rs);
        mlp_layer<2, 4, 8>(A, wt_hidden + (size_t)k * 1024, lane, C);
    }
    grad_to_frag<8, true>(C, masks[0], A, rs);
    mlp_layer<2, 4, 4>(A, wt_first, lane, Cin);
}

// d(sum_k g[k] * SH_k(x, y, z)) / d(x, y, z) for the degree-4 basis of sh_basis() (device_math.cuh; shencoder.cu:126-355 keeps the
// same Jacobian as dy_dx).
__device__ __forceinline__ void sh_basis_grad(float x, float y, float z, const float *g, float &gx, float &gy, float &gz) {
    const float a = 0.48860251190291987f, b = 1.0925484305920792f, c = 0.94617469575755997f, e = 0.54627421529603959f;
    const float f = 0.59004358992664352f, h = 2.8906114426405538f, p = 0.45704579946446572f, q = 0.3731763325901154f, r = 1.4453057213202769f;
    const float x2 = x * x, y2 = y * y, z2 = z * z;
    gx = -a * g[3] + b * y * g[4] - b * z * g[7] + 2.0f * e * x * g[8] - 6.0f * f * x * y * g[9] + h * y * z * g[10] +
         p * (1.0f - 5.0f * z2) * g[13] + 2.0f * r * x * z * g[14] + f * (-3.0f * x2 + 3.0f * y2) * g[15];
    gy = -a * g[1] + b * x * g[4] - b * z * g[5] - 2.0f * e * y * g[8] + f * (-3.0f * x2 + 3.0f * y2) * g[9] + h * x * z * g[10] +
         p * (1.0f - 5.0f * z2) * g[11] - 2.0f * r * y * z * g[14] + 6.0f * f * x * y * g[15];
    gz = a * g[2] - b * y * g[5] + 2.0f * c * z * g[6] - b * x * g[7] + h * x * y * g[10] - 10.0f * p * y * z * g[11] +
         q * (15.0f * z2 - 3.0f) * g[12] - 10.0f * p * x * z * g[13] + r * (x2 - y2) * g[14];
}

// d(sum_j denc[j] * enc_j(x, y, z)) / d(x, y, z): the trilinear interpolation's input gradient (gridencoder.cu:180-223), fp32,
// corners re-gathered (they are L2-resident: the forward of the same chunk has just read them).  denc: 32 floats, column 2 l + c.
template <bool kGeneral>
__device__ __forceinline__ void grid_input_grad(const NetConst &nc, float x, float y, float z, const float *denc, float &gx, float &gy, float &gz) {
    gx = gy = gz = 0.0f;
    float u0, u1, u2;
    if (!encode_unit(nc, x, y, z, u0, u1, u2)) return;
#pragma unroll 4
    for (int l = 0; l < 16; l++) {
        const LevelConst &lc = nc.lv[l];
        const float scale = lc.scale;
        const float p0 = __fmaf_rn(u0, scale, 0.5f), p1 = __fmaf_rn(u1, scale, 0.5f), p2 = __fmaf_rn(u2, scale, 0.5f);
        const uint32_t g0 = (uint32_t)floorf(p0), g1 = (uint32_t)floorf(p1), g2 = (uint32_t)floorf(p2);
        const float f0 = p0 - (float)g0, f1 = p1 - (float)g1, f2 = p2 - (float)g2;
        uint32_t row[8];
        cell_rows<kGeneral>(lc, g0, g1, g2, row);
        float2 v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = __half22float2(__ldg(nc.lvbase[l] + row[k]));
        const float d0 = denc[2 * l], d1 = denc[2 * l + 1];
        float s[8];   // the level's contribution per corner: denc . v
#pragma unroll
        for (int k = 0; k < 8; k++) s[k] = d0 * v[k].x + d1 * v[k].y;
        const float e0 = 1.0f - f0, e1 = 1.0f - f1, e2 = 1.0f - f2;
        gx += scale * (e1 * e2 * (s[1] - s[0]) + f1 * e2 * (s[3] - s[2]) + e1 * f2 * (s[5] - s[4]) + f1 * f2 * (s[7] - s[6]));
        gy += scale * (e0 * e2 * (s[2] - s[0]) + f0 * e2 * (s[3] - s[1]) + e0 * f2 * (s[6] - s[4]) + f0 * f2 * (s[7] - s[5]));
        gz += scale * (e0 * e1 * (s[4] - s[0]) + f0 * e1 * (s[5] - s[1]) + e0 * f1 * (s[6] - s[2]) + f0 * f1 * (s[7] - s[3]));
    }
    const float r = 1.0f / nc.two_bound;   // grid.py:144: u = (x + bound) / (2 bound)
    gx *= r; gy *= r; gz *= r;
}

// ---------------------------------------------------------------------------------------------- one 32-row chunk
struct WarpSmem {
    __half *enc;    // [32][kEncStride] staged encodings
    __half *sh;     // [kShStride] the ray's SH row (one per warp: every sample of a ray shares its direction)
    float4 *out;    // [32] (sigma * density_scale, r, g, b)
    float2 *smp;    // [32] (t, delta_t)
    float *denc;    // [32][kDencStride] d loss / d encoding (backward only)
};
__host__ __device__ constexpr size_t warp_smem_bytes(bool backward) {
    return 32 * kEncStride * 2 + 64 + 32 * 16 + 32 * 8 + (backward ? 32 * kDencStride * 4 : 0);
}
__device__ __forceinline__ WarpSmem warp_smem(unsigned char *p, bool backward) {
    WarpSmem w;
    w.enc = reinterpret_cast<__half *>(p); p += 32 * kEncStride * 2;
    w.sh = reinterpret_cast<__half *>(p); p += 64;
    w.out = reinterpret_cast<float4 *>(p); p += 32 * 16;
    w.smp = reinterpret_cast<float2 *>(p); p += 32 * 8;
    w.denc = backward ? reinterpret_cast<float *>(p) : nullptr;
    return w;
}

__device__ __forceinline__ float sigmoid_half(float v) {   // network_ff.py:72 under autocast: half in, half out
    return round_half(__fdividef(1.0f, 1.0f + __expf(-round_half(v))));
}

// network_ff.py:51-74 on the 32 staged rows of this warp -> ws.out; the ReLU masks of both nets are kept for the backward.
// kColor = false: the density half only (network_ff.py:76-89): ws.out[row].x.
template <bool kColor>
__device__ __forceinline__ void eval_chunk(const NetConst &nc, const uint2 *__restrict__ s_w, const NetOffsets &of, const WarpSmem &ws, int lane,
                                           uint64_t (&ms)[kMaxHidden], uint64_t (&mc)[kMaxHidden]) {
    const int g = lane >> 2, t = lane & 3;
    uint32_t A0[2][2][4];
    load_frags32(ws.enc, kEncStride, 0, lane, A0);
    float O[2][2][4];
    net_forward(s_w + of.s0, s_w + of.sh, s_w + of.sl, nc.nh_sigma, A0, lane, O, ms);
    const uint32_t *shp = reinterpret_cast<const uint32_t *>(ws.sh);
    float *out = reinterpret_cast<float *>(ws.out);
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        uint32_t g0 = pack_half2(O[mt][0][0], O[mt][0][1]), g1 = pack_half2(O[mt][0][2], O[mt][0][3]);
        if (t == 0) {
            out[(mt * 16 + g) * 4] = expf(__half2float(__ushort_as_half((unsigned short)(g0 & 0xffffu)))) * nc.density_scale;
            out[(mt * 16 + g + 8) * 4] = expf(__half2float(__ushort_as_half((unsigned short)(g1 & 0xffffu)))) * nc.density_scale;
            g0 &= 0xffff0000u;
            g1 &= 0xffff0000u;
        }
        if (kColor) {
            A0[mt][0][0] = shp[t]; A0[mt][0][1] = shp[t]; A0[mt][0][2] = shp[4 + t]; A0[mt][0][3] = shp[4 + t];
            A0[mt][1][0] = g0; A0[mt][1][1] = g1;
            A0[mt][1][2] = pack_half2(O[mt][1][0], O[mt][1][1]); A0[mt][1][3] = pack_half2(O[mt][1][2], O[mt][1][3]);
        }
    }
    if (!kColor) return;
    float Oc[2][2][4];
    net_forward(s_w + of.c0, s_w + of.ch, s_w + of.cl, nc.nh_color, A0, lane, Oc, mc);
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        float *lo = out + (mt * 16 + g) * 4, *hi = out + (mt * 16 + g + 8) * 4;
        if (t == 0) {          // columns 0, 1 = r, g
            lo[1] = sigmoid_half(Oc[mt][0][0]); lo[2] = sigmoid_half(Oc[mt][0][1]);
            hi[1] = sigmoid_half(Oc[mt][0][2]); hi[2] = sigmoid_half(Oc[mt][0][3]);
        } else if (t == 1) {   // column 2 = b
            lo[3] = sigmoid_half(Oc[mt][0][0]);
            hi[3] = sigmoid_half(Oc[mt][0][2]);
        }
    }
}

// Backward of eval_chunk.  Lane i passes its row's d loss / d h0 (log-density, fp32) and, with kColor, d loss / d (pre-sigmoid r, g, b).
// ws.denc[row] <- d loss / d encoding (32 floats); sh_acc += this thread's share of sum over rows of d loss / d SH (columns
// 2t, 2t+1, 8+2t, 9+2t; reduce over the lanes with equal t afterwards).
template <bool kColor>
__device__ __forceinline__ void backward_chunk(const NetConst &nc, const uint2 *__restrict__ s_wt, const NetOffsets &of, const WarpSmem &ws, int lane,
                                               const uint64_t (&ms)[kMaxHidden], const uint64_t (&mc)[kMaxHidden], float dh0, float dp0, float dp1,
                                               float dp2, float (&sh_acc)[4]) {
    const int g = lane >> 2, t = lane & 3;
    float rs[2][2] = {{1.f, 1.f}, {1.f, 1.f}};
    float G[2][2][4];
    if (kColor) {
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int row = mt * 16 + g + 8 * h;
                const float a0 = __shfl_sync(0xffffffffu, dp0, row), a1 = __shfl_sync(0xffffffffu, dp1, row), a2 = __shfl_sync(0xffffffffu, dp2, row);
                G[mt][0][2 * h] = t == 0 ? a0 : (t == 1 ? a2 : 0.0f);
                G[mt][0][2 * h + 1] = t == 0 ? a1 : 0.0f;
                G[mt][1][2 * h] = 0.0f; G[mt][1][2 * h + 1] = 0.0f;
            }
        float Cin[2][4][4];
        net_backward(s_wt + of.c0, s_wt + of.ch, s_wt + of.cl, nc.nh_color, G, mc, lane, rs, Cin);
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float inv = 1.0f / rs[mt][e >> 1];
                sh_acc[e & 1] += Cin[mt][0][e] * inv;
                sh_acc[2 + (e & 1)] += Cin[mt][1][e] * inv;
                G[mt][0][e] = Cin[mt][2][e];     // colour-input columns 16..31 = sigma-net output columns 0..15 (column 0 forced to zero there)
                G[mt][1][e] = Cin[mt][3][e];
            }
    } else {
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
            for (int e = 0; e < 4; e++) G[mt][0][e] = G[mt][1][e] = 0.0f;
    }
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const float d = __shfl_sync(0xffffffffu, dh0, mt * 16 + g + 8 * h);
            if (t == 0) G[mt][0][2 * h] = d * rs[mt][h];   // output column 0 = log-density
        }
    float Cenc[2][4][4];
    net_backward(s_wt + of.s0, s_wt + of.sh, s_wt + of.sl, nc.nh_sigma, G, ms, lane, rs, Cenc);
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int nt = 0; nt < 4; nt++)
#pragma unroll
            for (int e = 0; e < 4; e++)
                ws.denc[(mt * 16 + g + 8 * (e >> 1)) * kDencStride + nt * 8 + 2 * t + (e & 1)] = Cenc[mt][nt][e] / rs[mt][e >> 1];
}

// ---------------------------------------------------------------------------------------------- march, one warp per ray
// The reference's marcher (raymarching.cu:754-812) visits a SUBSET of one fixed sequence of parameters per ray, whatever the
// occupancy: t_0 = near, t_{k+1} = t_k + dt(t_k) -- an occupied cell advances by dt(t), an empty one repeats the same step until the
// cell's exit is passed.  So a warp tests 32 consecutive candidates of its ray at once (one occupancy load per lane, all in flight)
// and then replays the marcher's control flow on the 32 answers: an occupied candidate is a sample and hands over to the next
// candidate; an empty one hands over to the first later candidate at or beyond its cell's exit.  Same samples, same (t, delta_t), as
// the serial loop -- without 32 dependent loads per chunk.
struct MarchState {
    float t_base;    // first candidate of the next window
    float last_t;    // the marcher's last_t: the parameter after the previous sample
    float skip_to;   // an empty cell's exit that lies beyond the previous window (-FLT_MAX: none)
    bool done;       // the ray has left [near, far) or used up its sample budget
};

// Tests the next 32 candidates and appends the samples found, in order, to ws.smp[have ..); returns the new fill <= 32.  `budget`
// = samples the ray may still take (max_steps).  When the window holds more samples than the chunk has room for, the march is
// rewound to the candidate after the last sample taken (the marcher visits it next in any case).
__device__ __forceinline__ uint32_t march_window(const MarchConst &m, float ox, float oy, float oz, const RayDir &rd, float far,
                                                 uint32_t budget, uint32_t have, MarchState &st, const WarpSmem &ws, int lane) {
    constexpr uint32_t kFull = 0xffffffffu;
    float cur = st.t_base, tk = 0.f, tn = 0.f;
#pragma unroll 4
    for (int i = 0; i < 32; i++) {                     // every lane walks the sequence; lane i keeps candidate i and its successor
        const float nx = __fadd_rn(cur, sample_dt(m, cur));
        if (lane == i) { tk = cur; tn = nx; }
        cur = nx;
    }
    const bool valid = tk < far;
    float tt = tk;
    bool occ = false;
    if (valid) occ = march_peek(m, ox, oy, oz, rd, tk, tt);
    const uint32_t occ_mask = __ballot_sync(kFull, occ), valid_mask = __ballot_sync(kFull, valid);
    const uint32_t reach = __ballot_sync(kFull, tk >= st.skip_to);
    int j = reach ? __ffs(reach) - 1 : 32;
    float carry = j < 32 ? -FLT_MAX : st.skip_to;
    uint32_t visited = 0;
    bool done = false;
    while (j < 32) {                                   // warp-uniform replay of the marcher's hand-overs
        if (!((valid_mask >> j) & 1u)) { done = true; break; }
        if ((occ_mask >> j) & 1u) {                    // a run of occupied candidates: each is visited and hands over to the next
            const uint32_t rest = ~occ_mask & (kFull << j);
            const int e = rest ? __ffs(rest) - 1 : 32;
            visited |= (e == 32 ? kFull : ((1u << e) - 1u)) & (kFull << j);
            j = e;
        } else {                                       // empty cell: first later candidate at or beyond the cell's exit (at least one step)
            const float t_exit = __shfl_sync(kFull, tt, j);
            const uint32_t later = __ballot_sync(kFull, tk >= t_exit) & (j == 31 ? 0u : (kFull << (j + 1)));
            if (later) j = __ffs(later) - 1;
            else { j = 32; carry = t_exit; }
        }
    }
    uint32_t samples = visited & occ_mask;
    uint32_t c = (uint32_t)__popc(samples);
    const uint32_t room = min(32u - have, budget);
    bool rewind = false;
    if (c > room) {                                    // take the first `room` samples only
        samples &= (1u << __fns(samples, 0, room + 1)) - 1u;
        c = room;
        rewind = true;
    }
    if (c >= budget) done = true;                      // the ray's max_steps-th sample ends it
    else if (rewind) done = false;
    const uint32_t below = samples & ((1u << lane) - 1u);
    const float prev_next = __shfl_sync(kFull, tn, below ? 31 - __clz(below) : 0);
    if ((samples >> lane) & 1u) ws.smp[have + __popc(below)] = make_float2(tk, __fsub_rn(tn, below ? prev_next : st.last_t));
    float next_base = cur;
    if (samples) {
        st.last_t = __shfl_sync(kFull, tn, 31 - __clz(samples));
        if (rewind) { next_base = st.last_t; carry = -FLT_MAX; }   // the candidate after an occupied one is visited next
    }
    st.t_base = next_base;
    st.skip_to = carry;
    st.done = done;
    return have + c;
}

// ---------------------------------------------------------------------------------------------- kernels
struct DiffArgs {
    const float *rays_o, *rays_d;
    uint32_t N, max_steps;
    MarchConst mc;
    float aabb[6];
    float min_near, bg;
    const uint2 *wpack_t;
    // forward outputs
    float *image, *depth, *weights_sum;
    float4 *finals;            // [N] raw (r, g, b, weights_sum) before the background mix: what the backward pass needs of the forward
    int32_t *samples_per_ray;  // optional
    // backward
    const float *g_image;      // [N,3] d loss / d image
    const float *g_ws;         // [N] d loss / d weights_sum or nullptr
    float *g_o, *g_d;          // [N,3]
};

template <bool kBackward, bool kGeneral>
__global__ void __launch_bounds__(kThreads) k_diff_render(NetConst nc, DiffArgs ra) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint2 *s_w = reinterpret_cast<uint2 *>(smem_raw);
    uint2 *s_wt = s_w + nc.wpack_count;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    copy_wpack_to_smem(nc, s_w);
    if (kBackward) {
        const uint4 *src = reinterpret_cast<const uint4 *>(ra.wpack_t);
        uint4 *dst = reinterpret_cast<uint4 *>(s_wt);
        for (uint32_t i = threadIdx.x; i < nc.wpack_count / 2; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const WarpSmem ws = warp_smem(smem_raw + (size_t)nc.wpack_count * 8 * (kBackward ? 2 : 1) + warp * warp_smem_bytes(kBackward), kBackward);
    const NetOffsets of = net_offsets(nc);
    const MarchConst &m = ra.mc;

    for (uint32_t ray = blockIdx.x * kWarps + warp; ray < ra.N; ray += gridDim.x * kWarps) {
        const float ox = __ldg(ra.rays_o + 3 * (size_t)ray), oy = __ldg(ra.rays_o + 3 * (size_t)ray + 1), oz = __ldg(ra.rays_o + 3 * (size_t)ray + 2);
        const float dx = __ldg(ra.rays_d + 3 * (size_t)ray), dy = __ldg(ra.rays_d + 3 * (size_t)ray + 1), dz = __ldg(ra.rays_d + 3 * (size_t)ray + 2);
        float near, far;
        if (!slab_near_far(ox, oy, oz, dx, dy, dz, ra.aabb, ra.min_near, near, far)) near = far = FLT_MAX;
        if (lane == 0) {
            uint4 lo, hi;
            sh_to_half(dx, dy, dz, lo, hi);
            reinterpret_cast<uint4 *>(ws.sh)[0] = lo;
            reinterpret_cast<uint4 *>(ws.sh)[1] = hi;
        }
        const RayDir rd = make_ray_dir(dx, dy, dz);
        RayAcc acc = {near, 0.f, 0.f, 0.f, 0.f, 0.f};
        MarchState st = {near, near, -FLT_MAX, false};
        uint32_t total = 0;
        bool done = false;
        float4 fin = make_float4(0.f, 0.f, 0.f, 0.f);
        float gI0 = 0.f, gI1 = 0.f, gI2 = 0.f, gws = 0.f;
        if (kBackward) {
            fin = __ldg(ra.finals + ray);
            gI0 = __ldg(ra.g_image + 3 * (size_t)ray); gI1 = __ldg(ra.g_image + 3 * (size_t)ray + 1); gI2 = __ldg(ra.g_image + 3 * (size_t)ray + 2);
            // image = acc + (1 - weights_sum) * bg (renderer.py:374): the background term reaches sigma through weights_sum
            gws = (ra.g_ws ? __ldg(ra.g_ws + ray) : 0.0f) - ra.bg * (gI0 + gI1 + gI2);
        }
        float go0 = 0.f, go1 = 0.f, go2 = 0.f, gd0 = 0.f, gd1 = 0.f, gd2 = 0.f;
        float sh_acc[4] = {0.f, 0.f, 0.f, 0.f};

        while (!done) {
            // march: windows of 32 candidate parameters until the chunk holds 32 samples (or the ray ends)
            uint32_t c = 0;
            while (c < 32u && !st.done) c = march_window(m, ox, oy, oz, rd, far, ra.max_steps - total - c, c, st, ws, lane);
            if (c == 0) break;
            if (st.done) done = true;       // this chunk is the ray's last
            total += c;
            __syncwarp();
            const bool valid = (uint32_t)lane < c;
            float ts = 0.f, x = 0.f, y = 0.f, z = 0.f;
            uint4 *erow = reinterpret_cast<uint4 *>(ws.enc + lane * kEncStride);
            if (valid) {
                ts = ws.smp[lane].x;
                x = sample_pos(m, ox, dx, ts); y = sample_pos(m, oy, dy, ts); z = sample_pos(m, oz, dz, ts);
                encode_point<4, kGeneral>(nc, x, y, z, erow);   // 32 loads in flight per trip: latency, not registers, is what this kernel is short of
            } else {
                erow[0] = erow[1] = erow[2] = erow[3] = make_uint4(0, 0, 0, 0);
            }
            __syncwarp();
            uint64_t ms[kMaxHidden] = {0, 0, 0, 0}, mc[kMaxHidden] = {0, 0, 0, 0};
            eval_chunk<true>(nc, s_w, of, ws, lane, ms, mc);
            __syncwarp();
            // composite in sample order (raymarching.cu:866-894): every lane carries the ray's accumulators, lane j keeps sample j's terms
            float w_i = 0.f, Ta_i = 0.f, R0 = 0.f, R1 = 0.f, R2 = 0.f;
            uint32_t n_valid = c;
            for (uint32_t j = 0; j < c; j++) {
                const float4 o4 = ws.out[j];
                const float2 s = ws.smp[j];
                float w;
                const bool stop = composite_sample(acc, o4.x, o4.y, o4.z, o4.w, sample_dt(m, s.x), s.y, &w);
                if (kBackward && j == (uint32_t)lane) { w_i = w; Ta_i = 1.0f - acc.ws; R0 = acc.r; R1 = acc.g; R2 = acc.b; }
                if (stop) { done = true; n_valid = j + 1; break; }
            }
            if (kBackward) {
                float dh0 = 0.f, dp0 = 0.f, dp1 = 0.f, dp2 = 0.f;
                const bool used = (uint32_t)lane < n_valid;
                if (used) {
                    const float4 o4 = ws.out[lane];
                    // raymarching.cu:606-699 (composite_rays_train_backward), depth left out like there
                    dp0 = gI0 * w_i * o4.y * (1.0f - o4.y);
                    dp1 = gI1 * w_i * o4.z * (1.0f - o4.z);
                    dp2 = gI2 * w_i * o4.w * (1.0f - o4.w);
                    const float dsig = sample_dt(m, ts) * (gI0 * (Ta_i * o4.y - (fin.x - R0)) + gI1 * (Ta_i * o4.z - (fin.y - R1)) +
                                                           gI2 * (Ta_i * o4.w - (fin.z - R2)) + gws * (1.0f - fin.w));
                    dh0 = dsig * o4.x;   // sigma = exp(h0) * density_scale
                }
                backward_chunk<true>(nc, s_wt, of, ws, lane, ms, mc, dh0, dp0, dp1, dp2, sh_acc);
                __syncwarp();
                if (used) {
                    float gx, gy, gz;
                    grid_input_grad<kGeneral>(nc, x, y, z, ws.denc + lane * kDencStride, gx, gy, gz);
                    // x = clamp(o + t d, -bound, bound) (raymarching.cu:756-758): no gradient through a clamped coordinate
                    if (fabsf(__fmaf_rn(dx, ts, ox)) > m.bound) gx = 0.f;
                    if (fabsf(__fmaf_rn(dy, ts, oy)) > m.bound) gy = 0.f;
                    if (fabsf(__fmaf_rn(dz, ts, oz)) > m.bound) gz = 0.f;
                    go0 += gx; go1 += gy; go2 += gz;
                    gd0 += ts * gx; gd1 += ts * gy; gd2 += ts * gz;
                }
            }
            __syncwarp();
        }

        if (!kBackward) {
            if (lane == 0) {
                const float bgw = (1.0f - acc.ws) * ra.bg;          // renderer.py:374-376
                ra.image[3 * (size_t)ray] = acc.r + bgw; ra.image[3 * (size_t)ray + 1] = acc.g + bgw; ra.image[3 * (size_t)ray + 2] = acc.b + bgw;
                ra.depth[ray] = fmaxf(acc.d - near, 0.0f) / (far - near);
                ra.weights_sum[ray] = acc.ws;
                ra.finals[ray] = make_float4(acc.r, acc.g, acc.b, acc.ws);
                if (ra.samples_per_ray) ra.samples_per_ray[ray] = (int32_t)total;
            }
        } else {
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                go0 += __shfl_xor_sync(0xffffffffu, go0, o); go1 += __shfl_xor_sync(0xffffffffu, go1, o); go2 += __shfl_xor_sync(0xffffffffu, go2, o);
                gd0 += __shfl_xor_sync(0xffffffffu, gd0, o); gd1 += __shfl_xor_sync(0xffffffffu, gd1, o); gd2 += __shfl_xor_sync(0xffffffffu, gd2, o);
            }
#pragma unroll
            for (int k = 0; k < 4; k++)
#pragma unroll
                for (int o = 16; o >= 4; o >>= 1) sh_acc[k] += __shfl_xor_sync(0xffffffffu, sh_acc[k], o);
            float gsh[16];
#pragma unroll
            for (int col = 0; col < 16; col++) gsh[col] = __shfl_sync(0xffffffffu, sh_acc[(col >> 3) * 2 + (col & 1)], (col & 7) >> 1);
            float sx, sy, sz;
            sh_basis_grad(dx, dy, dz, gsh, sx, sy, sz);
            if (lane == 0) {
                ra.g_o[3 * (size_t)ray] = go0; ra.g_o[3 * (size_t)ray + 1] = go1; ra.g_o[3 * (size_t)ray + 2] = go2;
                ra.g_d[3 * (size_t)ray] = gd0 + sx; ra.g_d[3 * (size_t)ray + 1] = gd1 + sy; ra.g_d[3 * (size_t)ray + 2] = gd2 + sz;
            }
        }
        __syncwarp();
    }
}

// d loss / d xyz of sigma = exp(h0) (* density_scale) for M points (network_ff.py:76-89): one warp per 32 points.
template <bool kGeneral>
__global__ void __launch_bounds__(kThreads) k_density_backward(NetConst nc, const uint2 *__restrict__ wpack_t, const float *__restrict__ xyzs,
                                                              const float *__restrict__ g_sigma, uint32_t M, float *__restrict__ g_xyz) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint2 *s_w = reinterpret_cast<uint2 *>(smem_raw);
    uint2 *s_wt = s_w + nc.wpack_count;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    copy_wpack_to_smem(nc, s_w);
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(wpack_t);
        uint4 *dst = reinterpret_cast<uint4 *>(s_wt);
        for (uint32_t i = threadIdx.x; i < nc.wpack_count / 2; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const WarpSmem ws = warp_smem(smem_raw + (size_t)nc.wpack_count * 16 + warp * warp_smem_bytes(true), true);
    const NetOffsets of = net_offsets(nc);
    const uint32_t n_tiles = (M + 31u) / 32u;
    for (uint32_t tile = blockIdx.x * kWarps + warp; tile < n_tiles; tile += gridDim.x * kWarps) {
        const uint32_t i = tile * 32u + lane;
        const bool valid = i < M;
        float x = 0.f, y = 0.f, z = 0.f;
        uint4 *erow = reinterpret_cast<uint4 *>(ws.enc + lane * kEncStride);
        if (valid) {
            x = __ldg(xyzs + 3 * (size_t)i); y = __ldg(xyzs + 3 * (size_t)i + 1); z = __ldg(xyzs + 3 * (size_t)i + 2);
            encode_point<4, kGeneral>(nc, x, y, z, erow);   // 32 loads in flight per trip: latency, not registers, is what this kernel is short of
        } else {
            erow[0] = erow[1] = erow[2] = erow[3] = make_uint4(0, 0, 0, 0);
        }
        __syncwarp();
        uint64_t ms[kMaxHidden] = {0, 0, 0, 0}, mc[kMaxHidden] = {0, 0, 0, 0};
        eval_chunk<false>(nc, s_w, of, ws, lane, ms, mc);
        __syncwarp();
        const float dh0 = valid ? __ldg(g_sigma + i) * ws.out[lane].x : 0.0f;
        float sh_acc[4] = {0.f, 0.f, 0.f, 0.f};
        backward_chunk<false>(nc, s_wt, of, ws, lane, ms, mc, dh0, 0.f, 0.f, 0.f, sh_acc);
        __syncwarp();
        if (valid) {
            float gx, gy, gz;
            grid_input_grad<kGeneral>(nc, x, y, z, ws.denc + lane * kDencStride, gx, gy, gz);
            g_xyz[3 * (size_t)i] = gx; g_xyz[3 * (size_t)i + 1] = gy; g_xyz[3 * (size_t)i + 2] = gz;
        }
        __syncwarp();
    }
}

}  // namespace diff
