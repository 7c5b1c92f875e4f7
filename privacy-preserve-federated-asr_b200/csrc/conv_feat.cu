// Waveform feature extractor, layer 0: Conv1d(1 -> C, k, stride) + LayerNorm(C) + GELU fused (HF:51-75).
// HBM-bound on its output ([B, Tout, C]); one warp per output frame, filter bank in shared memory,
// the k input samples of a frame are loaded once and broadcast with shuffles.
#include "common.cuh"

namespace b200asr {

template <typename T, int NP>
__global__ void __launch_bounds__(256) conv0_ln_gelu_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                            const float* __restrict__ bias,
                                                            const float* __restrict__ ln_w,
                                                            const float* __restrict__ ln_b, T* __restrict__ y,
                                                            long long frames, int L, int Tout, int C, int k,
                                                            int stride, float eps) {
    extern __shared__ float smem[];
    float* w_s = smem;              // [k][C]  (transposed so that a warp reads consecutive channels)
    float* bias_s = w_s + k * C;    // [C]
    float* lnw_s = bias_s + C;      // [C]
    float* lnb_s = lnw_s + C;       // [C]
    for (int i = threadIdx.x; i < k * C; i += blockDim.x) {
        const int c = i / k, j = i % k;
        w_s[j * C + c] = w[i];
    }
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        bias_s[c] = bias ? bias[c] : 0.f;
        lnw_s[c] = ln_w[c];
        lnb_s[c] = ln_b[c];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const float inv_c = 1.0f / (float)C;
    for (long long f = warp; f < frames; f += nwarps) {
        const int b = (int)(f / Tout), t = (int)(f % Tout);
        const float* xp = x + (long long)b * L + (long long)t * stride;
        const float xv = (lane < k) ? __ldg(xp + lane) : 0.f;
        float acc[NP][2];
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const float2 bb = *reinterpret_cast<const float2*>(bias_s + 2 * lane + 64 * i);
            acc[i][0] = bb.x;
            acc[i][1] = bb.y;
        }
        for (int j = 0; j < k; ++j) {
            const float xj = __shfl_sync(0xffffffffu, xv, j);
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const float2 ww = *reinterpret_cast<const float2*>(w_s + j * C + 2 * lane + 64 * i);
                acc[i][0] = fmaf(ww.x, xj, acc[i][0]);
                acc[i][1] = fmaf(ww.y, xj, acc[i][1]);
            }
        }
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < NP; ++i) sum += acc[i][0] + acc[i][1];
        const float mean = warp_sum(sum) * inv_c;
        float sq = 0.f;
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const float d0 = acc[i][0] - mean, d1 = acc[i][1] - mean;
            sq += d0 * d0 + d1 * d1;
        }
        const float rstd = rsqrtf(warp_sum(sq) * inv_c + eps);
        T* yp = y + f * C;
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int c = 2 * lane + 64 * i;
            const float o0 = gelu_for<T>((acc[i][0] - mean) * rstd * lnw_s[c] + lnb_s[c]);
            const float o1 = gelu_for<T>((acc[i][1] - mean) * rstd * lnw_s[c + 1] + lnb_s[c + 1]);
            if constexpr (sizeof(T) == 2) {
                *reinterpret_cast<uint32_t*>(yp + c) = bf2_pack(o0, o1);
            } else {
                *reinterpret_cast<float2*>(yp + c) = make_float2(o0, o1);
            }
        }
    }
}

}  // namespace b200asr

using namespace b200asr;

extern "C" int b200asr_conv0_ln_gelu_fwd(const float* x, const float* w, const float* bias, const float* ln_w,
                                         const float* ln_b, void* y, int B, int L, int C, int k, int stride,
                                         float eps, int dtype, void* stream) {
    B200_REQUIRE(x && w && ln_w && ln_b && y, "conv0: null pointer");
    B200_REQUIRE(B > 0 && L >= k && k >= 1 && k <= 32 && stride >= 1, "conv0: bad shape B=%d L=%d k=%d", B, L, k);
    B200_REQUIRE(C % 64 == 0 && C <= 1024, "conv0: C=%d must be a multiple of 64 and <= 1024", C);
    const int Tout = (L - k) / stride + 1;
    const long long frames = (long long)B * Tout;
    const size_t smem = (size_t)(k * C + 3 * C) * sizeof(float);
    const int np = C / 64;
    long long want = (frames + 7) / 8;
    int grid = (int)(want < (long long)num_sms() * 8 ? want : (long long)num_sms() * 8);
    if (grid < 1) grid = 1;
    cudaStream_t st = as_stream(stream);
#define CONV0_CASE(TT, NPV)                                                                                  \
    do {                                                                                                     \
        if (smem > 48 * 1024)                                                                                \
            cudaFuncSetAttribute(conv0_ln_gelu_kernel<TT, NPV>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                 (int)smem);                                                                 \
        conv0_ln_gelu_kernel<TT, NPV><<<grid, 256, smem, st>>>(x, w, bias, ln_w, ln_b, (TT*)y, frames, L, Tout, C, \
                                                                k, stride, eps);                             \
    } while (0)
    B200_DISPATCH_DTYPE(dtype, {
        switch (np) {
            case 1: CONV0_CASE(T, 1); break;
            case 2: CONV0_CASE(T, 2); break;
            case 4: CONV0_CASE(T, 4); break;
            case 8: CONV0_CASE(T, 8); break;
            case 16: CONV0_CASE(T, 16); break;
            default:
                set_error("conv0: unsupported channel count %d", C);
                return B200ASR_ERR_UNSUPPORTED;
        }
    });
#undef CONV0_CASE
    return check_launch("conv0_ln_gelu_fwd");
}
