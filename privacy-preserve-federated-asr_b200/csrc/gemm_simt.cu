// FP32-FMA GEMM (parity mode and skinny / unaligned products that are HBM-bound anyway: lm_head N=32,
// dementia_head N=2).  No TF32, no tensor cores: fp32 results are reproducible against the CPU oracle to
// rounding.  64x64 tile, BK=16, 256 threads, 4x4 micro-tile, either operand K-major or MN-major.
#include "common.cuh"
#include "gemm.h"

namespace b200asr {

namespace {
constexpr int BM = 64, BN = 64, BK = 16, PAD = 4;
}

template <typename T>
__device__ __forceinline__ float ld_elem(const T* p, long long idx) { return to_f<T>(p[idx]); }

template <typename TA, typename TC>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const b200asr_gemm_args a) {
    __shared__ float As[BK][BM + PAD];
    __shared__ float Bs[BK][BN + PAD];
    const int tid = threadIdx.x;
    const int z = blockIdx.z;
    const int b0 = z % a.batch0, b1 = z / a.batch0;
    const TA* A = (const TA*)a.A + (long long)b0 * a.sa0 + (long long)b1 * a.sa1;
    const TA* B = (const TA*)a.B + (long long)b0 * a.sb0 + (long long)b1 * a.sb1;
    const long long coff = (long long)b0 * a.sc0 + (long long)b1 * a.sc1;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int ty = tid / 16, tx = tid % 16;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < a.K; k0 += BK) {
        // ---- A tile
        if (a.a_kmajor) {
            const int kk = tid % BK, mm = tid / BK;  // 16 rows per pass
#pragma unroll
            for (int p = 0; p < BM / 16; ++p) {
                const int m = m0 + mm + p * 16, k = k0 + kk;
                As[kk][mm + p * 16] = (m < a.M && k < a.K) ? ld_elem(A, (long long)m * a.lda + k) : 0.f;
            }
        } else {
            const int mm = tid % BM, kk = tid / BM;  // 4 k per pass
#pragma unroll
            for (int p = 0; p < BK / 4; ++p) {
                const int m = m0 + mm, k = k0 + kk + p * 4;
                As[kk + p * 4][mm] = (m < a.M && k < a.K) ? ld_elem(A, (long long)k * a.lda + m) : 0.f;
            }
        }
        // ---- B tile
        if (a.b_kmajor) {
            const int kk = tid % BK, nn = tid / BK;
#pragma unroll
            for (int p = 0; p < BN / 16; ++p) {
                const int n = n0 + nn + p * 16, k = k0 + kk;
                Bs[kk][nn + p * 16] = (n < a.N && k < a.K) ? ld_elem(B, (long long)n * a.ldb + k) : 0.f;
            }
        } else {
            const int nn = tid % BN, kk = tid / BN;
#pragma unroll
            for (int p = 0; p < BK / 4; ++p) {
                const int n = n0 + nn, k = k0 + kk + p * 4;
                Bs[kk + p * 4][nn] = (n < a.N && k < a.K) ? ld_elem(B, (long long)k * a.ldb + n) : 0.f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 av = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            const float ar[4] = {av.x, av.y, av.z, av.w};
            const float br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
        }
        __syncthreads();
    }
    // ---- epilogue
    TC* C = (TC*)a.C + coff;
    const TC* R = a.residual ? (const TC*)a.residual + coff : nullptr;
    TC* AO = a.aux_out ? (TC*)a.aux_out + coff : nullptr;
    const TC* AI = a.aux_in ? (const TC*)a.aux_in + coff : nullptr;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= a.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n >= a.N) continue;
            const long long idx = (long long)m * a.ldc + n;
            float v = acc[i][j] * a.alpha;
            if (a.bias) v += a.bias[n];
            if (AO) AO[idx] = from_f<TC>(a.act == B200ASR_ACT_GELU_SAVE_GRAD ? gelu_erf_grad(v) : v);
            if (a.act != B200ASR_ACT_NONE) v = gelu_erf(v);
            if (a.mul_gelu_grad) v *= (a.mul_gelu_grad == 2) ? to_f<TC>(AI[idx]) : gelu_erf_grad(to_f<TC>(AI[idx]));
            if (R) v += to_f<TC>(R[idx]);
            C[idx] = from_f<TC>(v);
        }
    }
}

int gemm_simt_launch(const b200asr_gemm_args& a, cudaStream_t st) {
    dim3 grid(ceil_div(a.N, BN), ceil_div(a.M, BM), a.batch0 * a.batch1);
    if (grid.y > 65535 || grid.z > 65535) {
        set_error("gemm(simt): grid too large (M=%d, batch=%d)", a.M, a.batch0 * a.batch1);
        return B200ASR_ERR_UNSUPPORTED;
    }
    if (a.dtype == B200ASR_F32 && a.c_dtype == B200ASR_F32) gemm_simt_kernel<float, float><<<grid, 256, 0, st>>>(a);
    else if (a.dtype == B200ASR_BF16 && a.c_dtype == B200ASR_BF16)
        gemm_simt_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, st>>>(a);
    else if (a.dtype == B200ASR_BF16 && a.c_dtype == B200ASR_F32)
        gemm_simt_kernel<__nv_bfloat16, float><<<grid, 256, 0, st>>>(a);
    else
        gemm_simt_kernel<float, __nv_bfloat16><<<grid, 256, 0, st>>>(a);
    return check_launch("gemm_simt");
}

}  // namespace b200asr
