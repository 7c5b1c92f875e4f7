// Kernels only the Wav2Vec2 / Hubert members of the model family need (BASELINE configs[4], reached from
// centralized/finetune_ASRs.py:154-172):
//   * conv layer 0 with GroupNorm(num_groups = C) + GELU      (HF wav2vec2 Wav2Vec2GroupNormConvLayer)
//   * weight_norm(dim=2) of the positional conv filter        (HF wav2vec2 Wav2Vec2PositionalConvEmbedding)
//   * y = res + gelu(c) and its backward                      (pos-conv activation + residual, no LayerNorm in between)
// All HBM-bound; reductions use fixed chunking (deterministic, no floating-point atomics).
#include "common.cuh"

namespace b200asr {

// ------------------------------------------------------------------------------------------------ conv0 + GroupNorm
// Statistics pass: per (utterance, channel) sum and sum of squares of the conv output over time, one block per
// (time chunk, utterance); thread = channel, the chunk's waveform samples staged in shared memory (broadcast reads).
constexpr int GN_CHUNK = 256;  // frames per block

__global__ void __launch_bounds__(256) conv0_gn_stats_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                             const float* __restrict__ bias, float* __restrict__ partial,
                                                             int L, int Tout, int C, int k, int stride, int nchunk) {
    extern __shared__ float smem[];
    float* w_s = smem;           // [k][C]
    float* x_s = w_s + k * C;    // [GN_CHUNK * stride + k]
    const int b = blockIdx.y, chunk = blockIdx.x;
    const int t0 = chunk * GN_CHUNK;
    const int nt = min(GN_CHUNK, Tout - t0);
    for (int i = threadIdx.x; i < k * C; i += blockDim.x) {
        const int c = i / k, j = i % k;
        w_s[j * C + c] = w[i];
    }
    const int nx = (nt - 1) * stride + k;
    const float* xp = x + (long long)b * L + (long long)t0 * stride;
    for (int i = threadIdx.x; i < nx; i += blockDim.x) x_s[i] = xp[i];
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float bc = bias ? bias[c] : 0.f;
        float s = 0.f, q = 0.f;
        for (int t = 0; t < nt; ++t) {
            float acc = bc;
            for (int j = 0; j < k; ++j) acc = fmaf(w_s[j * C + c], x_s[t * stride + j], acc);
            s += acc;
            q = fmaf(acc, acc, q);
        }
        float* p = partial + (((long long)b * nchunk + chunk) * C + c) * 2;
        p[0] = s;
        p[1] = q;
    }
}

// scale[b][c] = rstd * gamma, shift[b][c] = beta - mean * rstd * gamma (chunks combined in double)
__global__ void conv0_gn_finalize_kernel(const float* __restrict__ partial, const float* __restrict__ gamma,
                                         const float* __restrict__ beta, float* __restrict__ scale,
                                         float* __restrict__ shift, int B, int C, int nchunk, int Tout, float eps) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * C) return;
    const int b = i / C, c = i % C;
    double s = 0.0, q = 0.0;
    for (int ch = 0; ch < nchunk; ++ch) {
        const float* p = partial + (((long long)b * nchunk + ch) * C + c) * 2;
        s += (double)p[0];
        q += (double)p[1];
    }
    const double mean = s / Tout;
    double var = q / Tout - mean * mean;  // biased variance, as torch group_norm
    if (var < 0.0) var = 0.0;
    const float rstd = (float)(1.0 / sqrt(var + (double)eps));
    const float g = gamma[c] * rstd;
    scale[i] = g;
    shift[i] = beta[c] - (float)mean * g;
}

// Apply pass: recompute the conv (10 MACs per output, cheaper than a round trip of the fp32 conv output through
// HBM), normalise with the per-(utterance, channel) scale/shift, GELU, write channels-last.  One warp per frame.
template <typename T, int NP>
__global__ void __launch_bounds__(256) conv0_gn_apply_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                             const float* __restrict__ bias,
                                                             const float* __restrict__ scale,
                                                             const float* __restrict__ shift, T* __restrict__ y,
                                                             long long frames, int L, int Tout, int C, int k, int stride) {
    extern __shared__ float smem[];
    float* w_s = smem;            // [k][C]
    float* bias_s = w_s + k * C;  // [C]
    for (int i = threadIdx.x; i < k * C; i += blockDim.x) {
        const int c = i / k, j = i % k;
        w_s[j * C + c] = w[i];
    }
    for (int c = threadIdx.x; c < C; c += blockDim.x) bias_s[c] = bias ? bias[c] : 0.f;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
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
        const float* sc = scale + (long long)b * C;
        const float* sh = shift + (long long)b * C;
        T* yp = y + f * C;
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int c = 2 * lane + 64 * i;
            const float2 s2 = __ldg(reinterpret_cast<const float2*>(sc + c));
            const float2 h2 = __ldg(reinterpret_cast<const float2*>(sh + c));
            const float o0 = gelu_for<T>(fmaf(acc[i][0], s2.x, h2.x));
            const float o1 = gelu_for<T>(fmaf(acc[i][1], s2.y, h2.y));
            if constexpr (sizeof(T) == 2) {
                *reinterpret_cast<uint32_t*>(yp + c) = bf2_pack(o0, o1);
            } else {
                *reinterpret_cast<float2*>(yp + c) = make_float2(o0, o1);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ weight_norm(dim=2)
// v [rows, k] fp32 (rows = Cout * Cin/groups, tap index fastest), g [k].  w[r][j] = g[j] * v[r][j] / ||v[:, j]||.
// Column reductions in two deterministic stages: WN_CHUNKS row chunks -> partial[chunk][k] -> one block finalises.
constexpr int WN_CHUNKS = 128;

// partial[chunk][j] = sum over the chunk's rows of a[r][j] * b[r][j]   (b == a gives the squared norm)
__global__ void __launch_bounds__(256) wn_coldot_partial_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                                float* __restrict__ partial, int rows, int k) {
    const int chunk = blockIdx.x;
    const int rows_per = (rows + WN_CHUNKS - 1) / WN_CHUNKS;
    const int r0 = chunk * rows_per, r1 = min(rows, r0 + rows_per);
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        float acc = 0.f;
        for (int r = r0; r < r1; ++r) acc = fmaf(a[(long long)r * k + j], b[(long long)r * k + j], acc);
        partial[chunk * k + j] = acc;
    }
}

// mode 0: out0[j] = sqrt(sum)           (the per-tap norm)
// mode 1: out0[j] = sum / norm[j]       (dg = <dW, v> / ||v||);  out1[j] = sum (kept for dv)
__global__ void wn_finalize_kernel(const float* __restrict__ partial, const float* __restrict__ norm,
                                   float* __restrict__ out0, float* __restrict__ out1, int k, int mode) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    double s = 0.0;
    for (int c = 0; c < WN_CHUNKS; ++c) s += (double)partial[c * k + j];
    if (mode == 0) {
        out0[j] = (float)sqrt(s);
    } else {
        out0[j] = (float)(s / (double)norm[j]);
        out1[j] = (float)s;
    }
}

__global__ void wn_apply_kernel(const float* __restrict__ v, const float* __restrict__ g, const float* __restrict__ norm,
                                float* __restrict__ w, long long n, int k) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int j = (int)(i % k);
        w[i] = v[i] * (g[j] / norm[j]);
    }
}

// dv = g/||v|| * dW - g * <dW, v> / ||v||^3 * v
__global__ void wn_bwd_apply_kernel(const float* __restrict__ dw, const float* __restrict__ v, const float* __restrict__ g,
                                    const float* __restrict__ norm, const float* __restrict__ dot,
                                    float* __restrict__ dv, long long n, int k) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int j = (int)(i % k);
        const float nj = norm[j], gj = g[j];
        const float a = gj / nj;
        const float bcoef = gj * dot[j] / (nj * nj * nj);
        dv[i] = fmaf(a, dw[i], -bcoef * v[i]);
    }
}

// ------------------------------------------------------------------------------------------------ res + gelu(c)
template <typename T>
__global__ void __launch_bounds__(256) gelu_add_kernel(const T* __restrict__ c, const T* __restrict__ res,
                                                       T* __restrict__ y, long long nvec) {
    constexpr int N = Vec16<T>::N;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < nvec; i += stride) {
        Vec16<T> vc, vr, vy;
        float fc[N], fr[N];
        vc.load(c + i * N);
        vc.unpack(fc);
        if (res) {
            vr.load(res + i * N);
            vr.unpack(fr);
        }
#pragma unroll
        for (int e = 0; e < N; ++e) fc[e] = gelu_for<T>(fc[e]) + (res ? fr[e] : 0.f);
        vy.pack(fc);
        vy.store(y + i * N);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) gelu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ c,
                                                       T* __restrict__ dc, long long nvec) {
    constexpr int N = Vec16<T>::N;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < nvec; i += stride) {
        Vec16<T> vd, vc, vo;
        float fd[N], fc[N];
        vd.load(dy + i * N);
        vd.unpack(fd);
        vc.load(c + i * N);
        vc.unpack(fc);
#pragma unroll
        for (int e = 0; e < N; ++e) fd[e] *= gelu_grad_for<T>(fc[e]);
        vo.pack(fd);
        vo.store(dc + i * N);
    }
}

static int ew_grid(long long n, int threads) {
    long long want = (n + threads - 1) / threads;
    const long long cap = (long long)num_sms() * 16;
    if (want > cap) want = cap;
    return (int)(want < 1 ? 1 : want);
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

size_t b200asr_conv0_gn_gelu_workspace_bytes(int B, int L, int C, int k, int stride) {
    if (B <= 0 || L < k || stride < 1) return 0;
    const int Tout = (L - k) / stride + 1;
    const int nchunk = (Tout + GN_CHUNK - 1) / GN_CHUNK;
    return ((size_t)B * nchunk * C * 2 + (size_t)B * C * 2) * sizeof(float);
}

int b200asr_conv0_gn_gelu_fwd(const float* x, const float* w, const float* bias, const float* gn_w, const float* gn_b,
                              void* y, void* workspace, size_t workspace_bytes, int B, int L, int C, int k, int stride,
                              float eps, int dtype, void* stream) {
    B200_REQUIRE(x && w && gn_w && gn_b && y && workspace, "conv0_gn: null pointer");
    B200_REQUIRE(B > 0 && L >= k && k >= 1 && k <= 32 && stride >= 1 && stride <= 32, "conv0_gn: bad shape B=%d L=%d k=%d", B, L, k);
    B200_REQUIRE(C % 64 == 0 && C <= 1024, "conv0_gn: C=%d must be a multiple of 64 and <= 1024", C);
    B200_REQUIRE(workspace_bytes >= b200asr_conv0_gn_gelu_workspace_bytes(B, L, C, k, stride), "conv0_gn: workspace too small");
    const int Tout = (L - k) / stride + 1;
    const int nchunk = (Tout + GN_CHUNK - 1) / GN_CHUNK;
    float* partial = (float*)workspace;
    float* scale = partial + (size_t)B * nchunk * C * 2;
    float* shift = scale + (size_t)B * C;
    cudaStream_t st = as_stream(stream);
    const size_t smem_stats = (size_t)(k * C + GN_CHUNK * stride + k) * sizeof(float);
    if (smem_stats > 48 * 1024)
        cudaFuncSetAttribute(conv0_gn_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_stats);
    conv0_gn_stats_kernel<<<dim3(nchunk, B), 256, smem_stats, st>>>(x, w, bias, partial, L, Tout, C, k, stride, nchunk);
    int rc = check_launch("conv0_gn_stats");
    if (rc) return rc;
    conv0_gn_finalize_kernel<<<ceil_div(B * C, 256), 256, 0, st>>>(partial, gn_w, gn_b, scale, shift, B, C, nchunk, Tout, eps);
    rc = check_launch("conv0_gn_finalize");
    if (rc) return rc;
    const long long frames = (long long)B * Tout;
    const size_t smem = (size_t)(k * C + C) * sizeof(float);
    const int np = C / 64;
    long long want = (frames + 7) / 8;
    int grid = (int)(want < (long long)num_sms() * 8 ? want : (long long)num_sms() * 8);
    if (grid < 1) grid = 1;
#define GN_CASE(TT, NPV)                                                                                              \
    do {                                                                                                              \
        if (smem > 48 * 1024)                                                                                         \
            cudaFuncSetAttribute(conv0_gn_apply_kernel<TT, NPV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        conv0_gn_apply_kernel<TT, NPV><<<grid, 256, smem, st>>>(x, w, bias, scale, shift, (TT*)y, frames, L, Tout, C, k, stride); \
    } while (0)
    B200_DISPATCH_DTYPE(dtype, {
        switch (np) {
            case 1: GN_CASE(T, 1); break;
            case 2: GN_CASE(T, 2); break;
            case 4: GN_CASE(T, 4); break;
            case 8: GN_CASE(T, 8); break;
            case 16: GN_CASE(T, 16); break;
            default:
                set_error("conv0_gn: unsupported channel count %d", C);
                return B200ASR_ERR_UNSUPPORTED;
        }
    });
#undef GN_CASE
    return check_launch("conv0_gn_apply");
}

size_t b200asr_weight_norm_workspace_bytes(int k) { return (size_t)(WN_CHUNKS + 1) * (size_t)(k > 0 ? k : 0) * sizeof(float); }

int b200asr_weight_norm_fwd(const float* v, const float* g, float* w, float* norm, void* workspace, size_t workspace_bytes,
                            int rows, int k, void* stream) {
    B200_REQUIRE(v && g && w && norm && workspace, "weight_norm_fwd: null pointer");
    B200_REQUIRE(rows > 0 && k > 0, "weight_norm_fwd: bad shape rows=%d k=%d", rows, k);
    B200_REQUIRE(workspace_bytes >= b200asr_weight_norm_workspace_bytes(k), "weight_norm_fwd: workspace too small");
    cudaStream_t st = as_stream(stream);
    float* partial = (float*)workspace;
    wn_coldot_partial_kernel<<<WN_CHUNKS, 256, 0, st>>>(v, v, partial, rows, k);
    int rc = check_launch("weight_norm_sumsq");
    if (rc) return rc;
    wn_finalize_kernel<<<ceil_div(k, 128), 128, 0, st>>>(partial, nullptr, norm, nullptr, k, 0);
    rc = check_launch("weight_norm_finalize");
    if (rc) return rc;
    const long long n = (long long)rows * k;
    wn_apply_kernel<<<ew_grid(n, 256), 256, 0, st>>>(v, g, norm, w, n, k);
    return check_launch("weight_norm_apply");
}

int b200asr_weight_norm_bwd(const float* dw, const float* v, const float* g, const float* norm, float* dv, float* dg,
                            void* workspace, size_t workspace_bytes, int rows, int k, void* stream) {
    B200_REQUIRE(dw && v && g && norm && dv && dg && workspace, "weight_norm_bwd: null pointer");
    B200_REQUIRE(rows > 0 && k > 0, "weight_norm_bwd: bad shape rows=%d k=%d", rows, k);
    B200_REQUIRE(workspace_bytes >= b200asr_weight_norm_workspace_bytes(k), "weight_norm_bwd: workspace too small");
    cudaStream_t st = as_stream(stream);
    float* partial = (float*)workspace;
    float* dot = partial + (size_t)WN_CHUNKS * k;
    wn_coldot_partial_kernel<<<WN_CHUNKS, 256, 0, st>>>(dw, v, partial, rows, k);
    int rc = check_launch("weight_norm_dot");
    if (rc) return rc;
    wn_finalize_kernel<<<ceil_div(k, 128), 128, 0, st>>>(partial, norm, dg, dot, k, 1);
    rc = check_launch("weight_norm_bwd_finalize");
    if (rc) return rc;
    const long long n = (long long)rows * k;
    wn_bwd_apply_kernel<<<ew_grid(n, 256), 256, 0, st>>>(dw, v, g, norm, dot, dv, n, k);
    return check_launch("weight_norm_bwd_apply");
}

int b200asr_gelu_add_fwd(const void* c, const void* res, void* y, long long n, int dtype, void* stream) {
    B200_REQUIRE(c && y && n > 0, "gelu_add_fwd: bad argument");
    cudaStream_t st = as_stream(stream);
    B200_DISPATCH_DTYPE(dtype, {
        constexpr int N = Vec16<T>::N;
        B200_REQUIRE(n % N == 0, "gelu_add_fwd: n=%lld must be a multiple of %d", n, N);
        gelu_add_kernel<T><<<ew_grid(n / N, 256), 256, 0, st>>>((const T*)c, (const T*)res, (T*)y, n / N);
    });
    return check_launch("gelu_add_fwd");
}

int b200asr_gelu_bwd(const void* dy, const void* c, void* dc, long long n, int dtype, void* stream) {
    B200_REQUIRE(dy && c && dc && n > 0, "gelu_bwd: bad argument");
    cudaStream_t st = as_stream(stream);
    B200_DISPATCH_DTYPE(dtype, {
        constexpr int N = Vec16<T>::N;
        B200_REQUIRE(n % N == 0, "gelu_bwd: n=%lld must be a multiple of %d", n, N);
        gelu_bwd_kernel<T><<<ew_grid(n / N, 256), 256, 0, st>>>((const T*)dy, (const T*)c, (T*)dc, n / N);
    });
    return check_launch("gelu_bwd");
}

}  // extern "C"
