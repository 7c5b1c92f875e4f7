// Grouped positional convolution (HF:89-128), channels-last, round-1 implementation on the FP32 FMA pipes:
// per (utterance, group, 64-frame tile) the input window [(64 + k - 1) x Cg] is staged once in shared memory
// and every tap's [Cg x Cg] filter slice streams through a second buffer.  The same kernel computes the data
// gradient (taps reversed, filter slice transposed); the weight gradient reduces over time per (group, tap).
#include "common.cuh"
#include "gemm.h"

namespace b200asr {

namespace {
constexpr int PC_TT = 64;      // frames per tile
constexpr int PC_THREADS = 256;
}

// w_pack layout: [groups][k][Cg (row)][Cg (col)]  -- for forward rows = in-channel i, cols = out-channel o;
// for the data gradient rows = o, cols = i.
template <typename T>
__global__ void posconv_pack_kernel(const T* __restrict__ w, T* __restrict__ wf, T* __restrict__ wb, int C, int Cg, int k) {
    const long long total = (long long)C * Cg * k;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int j = (int)(idx % k);
        const int i = (int)((idx / k) % Cg);
        const int co = (int)(idx / ((long long)k * Cg));
        const int g = co / Cg, o = co % Cg;
        const T v = w[idx];
        wf[(((long long)g * k + j) * Cg + i) * Cg + o] = v;
        wb[(((long long)g * k + j) * Cg + o) * Cg + i] = v;
    }
}

// tensor-core layouts (64 channels per group, bf16): K-major rows of k*64 per output column of the implicit GEMM
//   wf[g][o][j][i] = W[g*64+o][i][j]          (forward:  n = o, taps ascending)
//   wb[g][i][j][o] = W[g*64+o][i][k-1-j]      (data gradient: n = i, taps reversed)
__global__ void posconv_pack_tc_kernel(const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ wf,
                                       __nv_bfloat16* __restrict__ wb, int C, int k) {
    const long long total = (long long)C * 64 * k;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int j = (int)(idx % k);
        const int i = (int)((idx / k) % 64);
        const int co = (int)(idx / ((long long)k * 64));
        const int g = co / 64, o = co % 64;
        const __nv_bfloat16 v = w[idx];
        wf[(((long long)g * 64 + o) * k + j) * 64 + i] = v;
        wb[(((long long)g * 64 + i) * k + (k - 1 - j)) * 64 + o] = v;
    }
}

// dw[c][i][j] = dwt[c][j][i]   (tap-major fp32 result of the tensor-core weight gradient -> HF layout [C, Cg, k])
__global__ void posconv_dw_permute_kernel(const float* __restrict__ dwt, float* __restrict__ dw, long long total, int Cg, int k) {
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int j = (int)(idx % k);
        const int i = (int)((idx / k) % Cg);
        const long long c = idx / ((long long)k * Cg);
        dw[idx] = dwt[(c * k + j) * Cg + i];
    }
}

static bool posconv_tc_ok(int dtype, int C, int groups, const void* a, const void* b) {
    return dtype == B200ASR_BF16 && C == groups * 64 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0;
}

// y[b, t, g*Cg + c] = bias[c] + sum_j sum_r W[g][jj][r][c] * x[b, t + j - pad, g*Cg + r],  jj = flip ? k-1-j : j
template <typename T, int CG>
__global__ void __launch_bounds__(PC_THREADS) posconv_kernel(const T* __restrict__ x, const T* __restrict__ wp,
                                                             const float* __restrict__ bias, T* __restrict__ y, int T_,
                                                             int C, int k, int pad, int flip) {
    extern __shared__ float smem[];
    float* xs = smem;                             // [(PC_TT + k - 1)][CG + 1]
    float* ws = xs + (PC_TT + k - 1) * (CG + 1);  // [CG][CG]
    const int b = blockIdx.z, g = blockIdx.y, t0 = blockIdx.x * PC_TT;
    const int tid = threadIdx.x;
    constexpr int TN = CG / 16;  // output channels per thread
    const int ty = tid / 16, tx = tid % 16;  // 16 x 16 threads: 4 frames x TN channels each
    const int rows = PC_TT + k - 1;
    for (int idx = tid; idx < rows * CG; idx += PC_THREADS) {
        const int r = idx / CG, c = idx % CG;
        const int t = t0 + r - pad;
        xs[r * (CG + 1) + c] = (t >= 0 && t < T_) ? to_f<T>(x[((long long)b * T_ + t) * C + g * CG + c]) : 0.f;
    }
    float acc[4][TN];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int n = 0; n < TN; ++n) acc[i][n] = 0.f;
    for (int j = 0; j < k; ++j) {
        __syncthreads();
        const int jj = flip ? k - 1 - j : j;
        const T* wsrc = wp + ((long long)g * k + jj) * CG * CG;
        for (int idx = tid; idx < CG * CG; idx += PC_THREADS) ws[idx] = to_f<T>(wsrc[idx]);
        __syncthreads();
#pragma unroll 4
        for (int r = 0; r < CG; ++r) {
            float xv[4], wv[TN];
#pragma unroll
            for (int i = 0; i < 4; ++i) xv[i] = xs[(ty * 4 + i + j) * (CG + 1) + r];
#pragma unroll
            for (int n = 0; n < TN; ++n) wv[n] = ws[r * CG + tx * TN + n];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int n = 0; n < TN; ++n) acc[i][n] = fmaf(xv[i], wv[n], acc[i][n]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int t = t0 + ty * 4 + i;
        if (t >= T_) continue;
#pragma unroll
        for (int n = 0; n < TN; ++n) {
            const int c = g * CG + tx * TN + n;
            y[((long long)b * T_ + t) * C + c] = from_f<T>(acc[i][n] + (bias ? bias[c] : 0.f));
        }
    }
}

// dW[g*Cg + o, i, j] += sum_t dy[b, t, g*Cg + o] * x[b, t + j - pad, g*Cg + i]     (grid: k, groups, B)
template <typename T, int CG>
__global__ void __launch_bounds__(PC_THREADS) posconv_wgrad_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                                                                   float* __restrict__ dw, int T_, int C, int k, int pad) {
    constexpr int TC = 32;  // frames per chunk
    __shared__ float dys[TC][CG + 1];
    __shared__ float xss[TC][CG + 1];
    const int j = blockIdx.x, g = blockIdx.y, b = blockIdx.z;
    const int tid = threadIdx.x;
    constexpr int TN = CG / 16;
    const int ty = tid / 16, tx = tid % 16;  // o = ty*TN + .., i = tx*TN + ..
    float acc[TN][TN];
#pragma unroll
    for (int a = 0; a < TN; ++a)
#pragma unroll
        for (int c = 0; c < TN; ++c) acc[a][c] = 0.f;
    for (int t0 = 0; t0 < T_; t0 += TC) {
        __syncthreads();
        for (int idx = tid; idx < TC * CG; idx += PC_THREADS) {
            const int r = idx / CG, c = idx % CG;
            const int t = t0 + r;
            const int tx_ = t + j - pad;
            dys[r][c] = (t < T_) ? to_f<T>(dy[((long long)b * T_ + t) * C + g * CG + c]) : 0.f;
            xss[r][c] = (t < T_ && tx_ >= 0 && tx_ < T_) ? to_f<T>(x[((long long)b * T_ + tx_) * C + g * CG + c]) : 0.f;
        }
        __syncthreads();
#pragma unroll 4
        for (int r = 0; r < TC; ++r) {
            float dv[TN], xv[TN];
#pragma unroll
            for (int a = 0; a < TN; ++a) dv[a] = dys[r][ty * TN + a];
#pragma unroll
            for (int c = 0; c < TN; ++c) xv[c] = xss[r][tx * TN + c];
#pragma unroll
            for (int a = 0; a < TN; ++a)
#pragma unroll
                for (int c = 0; c < TN; ++c) acc[a][c] = fmaf(dv[a], xv[c], acc[a][c]);
        }
    }
#pragma unroll
    for (int a = 0; a < TN; ++a)
#pragma unroll
        for (int c = 0; c < TN; ++c) {
            const int o = ty * TN + a, i = tx * TN + c;
            atomicAdd(&dw[((long long)(g * CG + o) * CG + i) * k + j], acc[a][c]);
        }
}

template <typename T>
static int launch_posconv(const void* x, const void* wp, const float* bias, void* y, int B, int T_, int C, int groups, int k,
                          int pad, int flip, cudaStream_t st) {
    const int Cg = C / groups;
    dim3 grid(ceil_div(T_, PC_TT), groups, B);
    const size_t smem = ((size_t)(PC_TT + k - 1) * (Cg + 1) + (size_t)Cg * Cg) * sizeof(float);
#define PC_CASE(CGV)                                                                                                 \
    do {                                                                                                             \
        if (smem > 48 * 1024) cudaFuncSetAttribute(posconv_kernel<T, CGV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        posconv_kernel<T, CGV><<<grid, PC_THREADS, smem, st>>>((const T*)x, (const T*)wp, bias, (T*)y, T_, C, k, pad, flip); \
    } while (0)
    if (Cg == 64) PC_CASE(64);
    else if (Cg == 32) PC_CASE(32);
    else if (Cg == 16) PC_CASE(16);
    else if (Cg == 80) PC_CASE(80);
    else { set_error("posconv: channels per group %d unsupported (16/32/64/80)", Cg); return B200ASR_ERR_UNSUPPORTED; }
#undef PC_CASE
    return check_launch("posconv");
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

int b200asr_posconv_pack_weight(const void* w, void* w_fwd, void* w_bwd, int C, int groups, int k, int dtype, void* stream) {
    B200_REQUIRE(w && w_fwd && w_bwd, "posconv_pack_weight: null pointer");
    B200_REQUIRE(C > 0 && groups > 0 && C % groups == 0 && k > 0, "posconv_pack_weight: bad shape");
    const int Cg = C / groups;
    const long long total = (long long)C * Cg * k;
    long long g = (total + 255) / 256;
    if (g > num_sms() * 8) g = num_sms() * 8;
    if (posconv_tc_ok(dtype, C, groups, w_fwd, w_bwd)) {
        posconv_pack_tc_kernel<<<(int)g, 256, 0, as_stream(stream)>>>((const __nv_bfloat16*)w, (__nv_bfloat16*)w_fwd,
                                                                       (__nv_bfloat16*)w_bwd, C, k);
        return check_launch("posconv_pack_weight");
    }
    B200_DISPATCH_DTYPE(dtype, {
        posconv_pack_kernel<T><<<(int)g, 256, 0, as_stream(stream)>>>((const T*)w, (T*)w_fwd, (T*)w_bwd, C, Cg, k);
    });
    return check_launch("posconv_pack_weight");
}

int b200asr_posconv_fwd(const void* x, const void* w_fwd, const float* bias, void* y, int B, int T_, int C, int groups, int k,
                        int dtype, void* stream) {
    B200_REQUIRE(x && w_fwd && y, "posconv_fwd: null pointer");
    B200_REQUIRE(B > 0 && T_ > 0 && C > 0 && groups > 0 && C % groups == 0 && k > 0, "posconv_fwd: bad shape");
    if (posconv_tc_ok(dtype, C, groups, x, w_fwd) && (reinterpret_cast<uintptr_t>(y) & 15) == 0)
        return posconv_tc_taps(x, w_fwd, bias, y, B, T_, C, groups, k, k / 2, as_stream(stream));
    B200_DISPATCH_DTYPE(dtype, { return launch_posconv<T>(x, w_fwd, bias, y, B, T_, C, groups, k, k / 2, 0, as_stream(stream)); });
    return B200ASR_OK;
}

size_t b200asr_posconv_bwd_workspace_bytes(int C, int groups, int k) {
    return (size_t)C * (C / (groups > 0 ? groups : 1)) * k * sizeof(float);
}

int b200asr_posconv_bwd(const void* x, const void* w_bwd, const void* dy, void* dx, float* dw, void* workspace,
                        size_t workspace_bytes, int B, int T_, int C, int groups, int k, int dtype, void* stream) {
    B200_REQUIRE(x && w_bwd && dy, "posconv_bwd: null pointer");
    B200_REQUIRE(B > 0 && T_ > 0 && C > 0 && groups > 0 && C % groups == 0 && k > 0, "posconv_bwd: bad shape");
    cudaStream_t st = as_stream(stream);
    const int pad = k / 2;
    const int Cg = C / groups;
    if (posconv_tc_ok(dtype, C, groups, x, dy) && (!dx || (reinterpret_cast<uintptr_t>(dx) & 15) == 0) &&
        (reinterpret_cast<uintptr_t>(w_bwd) & 15) == 0) {
        if (dx) {
            int rc = posconv_tc_taps(dy, w_bwd, nullptr, dx, B, T_, C, groups, k, k - 1 - pad, st);
            if (rc) return rc;
        }
        if (dw) {
            B200_REQUIRE(workspace && workspace_bytes >= b200asr_posconv_bwd_workspace_bytes(C, groups, k),
                         "posconv_bwd: workspace too small");
            int rc = posconv_tc_wgrad(x, dy, (float*)workspace, B, T_, C, groups, k, pad, st);
            if (rc) return rc;
            const long long total = (long long)C * Cg * k;
            long long g = (total + 255) / 256;
            if (g > num_sms() * 8) g = num_sms() * 8;
            posconv_dw_permute_kernel<<<(int)g, 256, 0, st>>>((const float*)workspace, dw, total, Cg, k);
            return check_launch("posconv_dw_permute");
        }
        return B200ASR_OK;
    }
    if (dx) {
        // dx[t] = sum_j dy[t - j + pad] W_j^T = sum_j' dy[t + j' - (k-1-pad)] W_{k-1-j'}^T
        int rc;
        B200_DISPATCH_DTYPE(dtype, { rc = launch_posconv<T>(dy, w_bwd, nullptr, dx, B, T_, C, groups, k, k - 1 - pad, 1, st); });
        if (rc) return rc;
    }
    if (dw) {
        cudaError_t e = cudaMemsetAsync(dw, 0, (size_t)C * Cg * k * sizeof(float), st);
        if (e != cudaSuccess) { set_error("posconv_bwd: memset: %s", cudaGetErrorString(e)); return B200ASR_ERR_CUDA; }
        dim3 grid(k, groups, B);
        B200_DISPATCH_DTYPE(dtype, {
            if (Cg == 64) posconv_wgrad_kernel<T, 64><<<grid, PC_THREADS, 0, st>>>((const T*)x, (const T*)dy, dw, T_, C, k, pad);
            else if (Cg == 32) posconv_wgrad_kernel<T, 32><<<grid, PC_THREADS, 0, st>>>((const T*)x, (const T*)dy, dw, T_, C, k, pad);
            else if (Cg == 16) posconv_wgrad_kernel<T, 16><<<grid, PC_THREADS, 0, st>>>((const T*)x, (const T*)dy, dw, T_, C, k, pad);
            else if (Cg == 80) posconv_wgrad_kernel<T, 80><<<grid, PC_THREADS, 0, st>>>((const T*)x, (const T*)dy, dw, T_, C, k, pad);
            else { set_error("posconv_bwd: channels per group %d unsupported", Cg); return B200ASR_ERR_UNSUPPORTED; }
        });
        return check_launch("posconv_wgrad");
    }
    return B200ASR_OK;
}

}  // extern "C"
