// Waveform feature extractor, layer 0: Conv1d(1 -> C, k, stride) + LayerNorm(C) + GELU fused (HF:51-75).
// HBM-bound on its output ([B, Tout, C]); one warp per output frame, filter bank in shared memory,
// the k input samples of a frame are loaded once and broadcast with shuffles.
#include "common.cuh"

namespace b200asr {

// FR consecutive frames per warp iteration: every filter tap pair read from shared memory feeds FR accumulators (the
// one-frame version re-read the whole 20 KB filter bank per frame and was shared-memory-bandwidth bound at 14% of
// the HBM peak); the arithmetic runs on the packed fp32 pipe (FFMA2), two channels per instruction.
constexpr int CONV0_FR = 4;

template <typename T, int NP>
__global__ void __launch_bounds__(256) conv0_ln_gelu_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                            const float* __restrict__ bias,
                                                            const float* __restrict__ ln_w,
                                                            const float* __restrict__ ln_b, T* __restrict__ y,
                                                            long long ngroups, int L, int Tout, int C, int k,
                                                            int stride, float eps) {
    constexpr int FR = CONV0_FR;
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
    const int gpu_ = (Tout + FR - 1) / FR;  // frame groups per utterance
    const int span = (FR - 1) * stride + k; // waveform samples one group touches (<= 32, checked by the host)
    for (long long gi = warp; gi < ngroups; gi += nwarps) {
        const int b = (int)(gi / gpu_), t0 = (int)(gi % gpu_) * FR;
        const long long xo = (long long)b * L + (long long)t0 * stride;
        const float xv = (lane < span && xo + lane < (long long)(b + 1) * L) ? __ldg(x + xo + lane) : 0.f;
        float2 acc[FR][NP];
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const float2 bb = *reinterpret_cast<const float2*>(bias_s + 2 * lane + 64 * i);
#pragma unroll
            for (int f = 0; f < FR; ++f) acc[f][i] = bb;
        }
        for (int j = 0; j < k; ++j) {
            float2 xf[FR];
#pragma unroll
            for (int f = 0; f < FR; ++f) {
                const float xj = __shfl_sync(0xffffffffu, xv, f * stride + j);
                xf[f] = make_float2(xj, xj);
            }
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const float2 ww = *reinterpret_cast<const float2*>(w_s + j * C + 2 * lane + 64 * i);
#pragma unroll
                for (int f = 0; f < FR; ++f) acc[f][i] = __ffma2_rn(ww, xf[f], acc[f][i]);
            }
        }
#pragma unroll
        for (int f = 0; f < FR; ++f) {
            const int t = t0 + f;
            float2 s2 = make_float2(0.f, 0.f);
#pragma unroll
            for (int i = 0; i < NP; ++i) s2 = __fadd2_rn(s2, acc[f][i]);
            const float mean = warp_sum(s2.x + s2.y) * inv_c;
            const float2 nm = make_float2(-mean, -mean);
            float2 q2 = make_float2(0.f, 0.f);
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                acc[f][i] = __fadd2_rn(acc[f][i], nm);
                q2 = __ffma2_rn(acc[f][i], acc[f][i], q2);
            }
            const float rstd = rsqrtf(warp_sum(q2.x + q2.y) * inv_c + eps);
            if (t >= Tout) continue;  // warp-uniform
            const float2 r2 = make_float2(rstd, rstd);
            T* yp = y + ((long long)b * Tout + t) * C;
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int c = 2 * lane + 64 * i;
                const float2 gw = __fmul2_rn(*reinterpret_cast<const float2*>(lnw_s + c), r2);
                float2 o = __ffma2_rn(acc[f][i], gw, *reinterpret_cast<const float2*>(lnb_s + c));
                if constexpr (sizeof(T) == 2) {
                    o = gelu_fast2(o);
                    *reinterpret_cast<uint32_t*>(yp + c) = bf2_pack(o.x, o.y);
                } else {
                    *reinterpret_cast<float2*>(yp + c) = make_float2(gelu_erf(o.x), gelu_erf(o.y));
                }
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
    B200_REQUIRE((CONV0_FR - 1) * stride + k <= 32, "conv0: (FR-1)*stride + k = %d exceeds one warp of samples",
                 (CONV0_FR - 1) * stride + k);
    const long long frames = (long long)B * ((Tout + CONV0_FR - 1) / CONV0_FR);  // frame groups
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
