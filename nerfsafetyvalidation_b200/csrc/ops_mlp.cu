// ops_mlp.cu -- ngp_ffmlp_inference: drop-in for _ffmlp.ffmlp_inference (ffmlp/src/ffmlp.cu:673-709) for the
// 64-wide ReLU networks of nerf/network_ff.py.  fp16 operands, fp32 accumulation (mma.sync.m16n8k16),
// weights packed once per CTA into B-fragment order in shared memory, activations chained in registers:
// a warp owns 32 rows and carries them through every layer without touching shared or global memory.
#include "common.cuh"
#include "device_math.cuh"

namespace ngp {

constexpr int kMlpThreads = 128;          // 4 warps
constexpr int kMlpRowsPerWarp = 32;       // MT = 2
constexpr int kMlpRowsPerCta = 128;

template <int KT_IN>
__global__ void __launch_bounds__(kMlpThreads)
k_ffmlp(const __half *__restrict__ inputs, const __half *__restrict__ weights, uint32_t B, uint32_t n_hidden, __half *__restrict__ outputs,
        uint32_t n_tiles) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint2 *wf = reinterpret_cast<uint2 *>(smem_raw);
    constexpr int IN = KT_IN * 16;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;

    // pack: layer 0 [64][IN], hidden k [64][64], last [16][64]; one uint2 = 4 halves
    uint2 *wf0 = wf;
    uint2 *wfh = wf0 + 64 * IN / 4;
    uint2 *wfl = wfh + (size_t)n_hidden * 64 * 64 / 4;
    pack_weight_frags(weights, 64, IN, wf0, tid, kMlpThreads, false);
    for (uint32_t k = 0; k < n_hidden; k++)
        pack_weight_frags(weights + 64 * IN + (size_t)k * 4096, 64, 64, wfh + (size_t)k * 1024, tid, kMlpThreads, false);
    pack_weight_frags(weights + 64 * IN + (size_t)n_hidden * 4096, 16, 64, wfl, tid, kMlpThreads, false);
    __syncthreads();

    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint32_t row0 = tile * kMlpRowsPerCta + warp * kMlpRowsPerWarp;
        uint32_t A0[2][KT_IN][4];
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
            const uint32_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
            const uint32_t *p_lo = reinterpret_cast<const uint32_t *>(inputs + (size_t)r_lo * IN);
            const uint32_t *p_hi = reinterpret_cast<const uint32_t *>(inputs + (size_t)r_hi * IN);
#pragma unroll
            for (int kt = 0; kt < KT_IN; kt++) {
                A0[mt][kt][0] = r_lo < B ? __ldg(p_lo + kt * 8 + t) : 0u;
                A0[mt][kt][1] = r_hi < B ? __ldg(p_hi + kt * 8 + t) : 0u;
                A0[mt][kt][2] = r_lo < B ? __ldg(p_lo + kt * 8 + 4 + t) : 0u;
                A0[mt][kt][3] = r_hi < B ? __ldg(p_hi + kt * 8 + 4 + t) : 0u;
            }
        }
        float C[2][8][4];
        uint32_t A[2][4][4];
        mlp_layer<2, KT_IN, 8>(A0, wf0, lane, C);
        relu_to_frag<2, 8>(C, A);
        for (uint32_t k = 0; k < n_hidden; k++) {
            mlp_layer<2, 4, 8>(A, wfh + (size_t)k * 1024, lane, C);
            relu_to_frag<2, 8>(C, A);
        }
        float O[2][2][4];
        mlp_layer<2, 4, 2>(A, wfl, lane, O);
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
            const uint32_t r_lo = row0 + mt * 16 + g, r_hi = r_lo + 8;
#pragma unroll
            for (int nt = 0; nt < 2; nt++) {
                if (r_lo < B) *reinterpret_cast<uint32_t *>(outputs + (size_t)r_lo * 16 + nt * 8 + 2 * t) = pack_half2(O[mt][nt][0], O[mt][nt][1]);
                if (r_hi < B) *reinterpret_cast<uint32_t *>(outputs + (size_t)r_hi * 16 + nt * 8 + 2 * t) = pack_half2(O[mt][nt][2], O[mt][nt][3]);
            }
        }
    }
}

}  // namespace ngp

using namespace ngp;

extern "C" int ngp_ffmlp_inference(const void *inputs_f16, const void *weights_f16, uint32_t B, uint32_t input_dim, uint32_t output_dim,
                                   uint32_t hidden_dim, uint32_t num_layers, uint32_t activation, uint32_t output_activation,
                                   void *outputs_f16, void *stream) {
    NGP_REQUIRE(hidden_dim == 64, "ffmlp_inference: only hidden_dim == 64 is built (network_ff.py uses 64); got %u", hidden_dim);
    NGP_REQUIRE(input_dim == 16 || input_dim == 32 || input_dim == 48 || input_dim == 64,
                "ffmlp_inference: input_dim must be 16*m, m in 1..4; got %u", input_dim);
    NGP_REQUIRE(output_dim == 16, "ffmlp_inference: padded output_dim must be 16 (ffmlp.py:114); got %u", output_dim);
    NGP_REQUIRE(num_layers >= 2 && num_layers <= 8, "ffmlp_inference: num_layers must be in [2, 8]; got %u", num_layers);
    NGP_REQUIRE(activation == 0, "ffmlp_inference: only ReLU hidden activation (0) is built; got %u", activation);
    NGP_REQUIRE(output_activation == 6, "ffmlp_inference: only output activation 'none' (6) is built; got %u", output_activation);
    if (B == 0) return 0;
    NGP_DEVPTR(inputs_f16); NGP_DEVPTR(weights_f16); NGP_DEVPTR(outputs_f16);
    const uint32_t n_hidden = num_layers - 1;
    const size_t smem = sizeof(__half) * (64 * input_dim + (size_t)n_hidden * 4096 + 16 * 64);
    const uint32_t n_tiles = div_up(B, kMlpRowsPerCta);
    int dev = 0, sms = 0;
    NGP_CUDA(cudaGetDevice(&dev));
    NGP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const uint32_t grid = n_tiles < (uint32_t)sms * 4 ? n_tiles : (uint32_t)sms * 4;
    const __half *in = reinterpret_cast<const __half *>(inputs_f16), *w = reinterpret_cast<const __half *>(weights_f16);
    __half *out = reinterpret_cast<__half *>(outputs_f16);
    cudaStream_t st = as_stream(stream);
#define NGP_LAUNCH_MLP(KT)                                                                                         \
    do {                                                                                                           \
        NGP_CUDA(cudaFuncSetAttribute(k_ffmlp<KT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
        k_ffmlp<KT><<<grid, kMlpThreads, smem, st>>>(in, w, B, n_hidden, out, n_tiles);                            \
    } while (0)
    switch (input_dim / 16) {
        case 1: NGP_LAUNCH_MLP(1); break;
        case 2: NGP_LAUNCH_MLP(2); break;
        case 3: NGP_LAUNCH_MLP(3); break;
        default: NGP_LAUNCH_MLP(4); break;
    }
#undef NGP_LAUNCH_MLP
    return after_launch("ffmlp_inference");
}
