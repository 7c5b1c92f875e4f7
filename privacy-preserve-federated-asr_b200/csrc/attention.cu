// Multi-head self-attention (HF:189-197).  Round-1 implementation: composed from the batched GEMM back ends
// (tcgen05 for bf16, fp32 FMA for parity mode) + row-softmax kernels, the scores living in a caller-provided
// workspace; backward recomputes the probabilities from the saved log-sum-exp (no [B,H,T,T] tensor is kept
// between forward and backward).
#include "attn.h"
#include "common.cuh"
#include "gemm.h"

namespace b200asr {

// one warp per score row; in place.  lse_out (optional) = log sum exp of the (already scaled) logits.
template <typename T>
__global__ void __launch_bounds__(256) softmax_rows_kernel(T* __restrict__ s, const int* __restrict__ klen,
                                                           float* __restrict__ lse_out, int H, int Tq, int Tk,
                                                           int ldk, long long rows) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long row = warp; row < rows; row += nwarps) {
        const int b = (int)(row / ((long long)H * Tq));
        const int kl = klen ? min(klen[b], Tk) : Tk;
        T* p = s + row * ldk;
        float mx = -INFINITY;
        for (int j = lane; j < kl; j += 32) mx = fmaxf(mx, to_f<T>(p[j]));
        mx = warp_max(mx);
        float sum = 0.f;
        for (int j = lane; j < kl; j += 32) sum += expf(to_f<T>(p[j]) - mx);
        sum = warp_sum(sum);
        const float inv = sum > 0.f ? 1.0f / sum : 0.f;
        for (int j = lane; j < ldk; j += 32) {
            const float v = (j < kl) ? expf(to_f<T>(p[j]) - mx) * inv : 0.f;
            p[j] = from_f<T>(v);
        }
        if (lse_out && lane == 0) lse_out[row] = (sum > 0.f) ? mx + logf(sum) : -INFINITY;
    }
}

// if lse != nullptr: `p` holds scaled logits S on entry and is overwritten with P = exp(S - lse) (masked);
// dp is overwritten with dS = P * (dP - rowsum(dP * P)).
template <typename T>
__global__ void __launch_bounds__(256) softmax_rows_bwd_kernel(T* __restrict__ p, T* __restrict__ dp,
                                                               const float* __restrict__ lse,
                                                               const int* __restrict__ klen, int H, int Tq, int Tk,
                                                               int ldk, long long rows) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long row = warp; row < rows; row += nwarps) {
        const int b = (int)(row / ((long long)H * Tq));
        const int kl = klen ? min(klen[b], Tk) : Tk;
        T* pr = p + row * ldk;
        T* dr = dp + row * ldk;
        float dot = 0.f;
        if (lse) {
            const float l = lse[row];
            for (int j = lane; j < ldk; j += 32) {
                float pv = 0.f;
                if (j < kl) pv = to_f<T>(from_f<T>(expf(to_f<T>(pr[j]) - l)));
                pr[j] = from_f<T>(pv);
                if (j < kl) dot += pv * to_f<T>(dr[j]);
            }
        } else {
            for (int j = lane; j < kl; j += 32) dot += to_f<T>(pr[j]) * to_f<T>(dr[j]);
        }
        dot = warp_sum(dot);
        __syncwarp();
        for (int j = lane; j < ldk; j += 32) {
            const float v = (j < kl) ? to_f<T>(pr[j]) * (to_f<T>(dr[j]) - dot) : 0.f;
            dr[j] = from_f<T>(v);
        }
    }
}

// Composed path: x[b,h,q,k] = keep(b,h,q,k) ? x * inv_keep : 0 on a [B,H,T,ldk] plane (probabilities in the forward,
// dP and then P in the backward), keep bits in the layout of the fused kernels (attn.h).
template <typename T>
__global__ void __launch_bounds__(256) apply_keep_kernel(T* __restrict__ x, const unsigned* __restrict__ keep, int Tq, int Tk,
                                                         int ldk, int KW, int TQ, float inv_keep, long long rows) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long row = warp; row < rows; row += nwarps) {
        const long long bh = row / Tq;
        const int q = (int)(row - bh * Tq);
        T* xr = x + row * ldk;
        for (int k0 = 0; k0 < Tk; k0 += 32) {
            const unsigned w = keep[(bh * KW + (k0 >> 5)) * (long long)TQ + q];
            const int k = k0 + lane;
            if (k < Tk) xr[k] = from_f<T>((w >> lane) & 1u ? to_f<T>(xr[k]) * inv_keep : 0.f);
        }
    }
}

static int apply_keep(void* x, const unsigned* keep, int B, int H, int Tn, int ldk, float p_drop, int dtype, cudaStream_t st);

static int softmax_grid(long long rows) {
    long long g = (rows + 7) / 8;
    const long long cap = (long long)num_sms() * 8;
    return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

static int g_mha_fused = 1;  // test hook (b200asr_mha_set_fused): 0 = composed GEMM + softmax path

static int round_up8(int x) { return (x + 7) & ~7; }

static size_t elem_size(int dtype) { return dtype == B200ASR_BF16 ? 2 : 4; }

// S[b,h] = alpha * Q[b,h] K[b,h]^T   (scores [B,H,T,ldk])
static int gemm_qk(const void* qkv, void* S, int B, int T, int H, int dh, int ldk, float alpha, int dtype,
                   cudaStream_t st) {
    const int D = H * dh;
    const size_t es = elem_size(dtype);
    b200asr_gemm_args g = {};
    g.A = qkv;
    g.B = (const char*)qkv + (size_t)D * es;
    g.C = S;
    g.M = T; g.N = T; g.K = dh;
    g.lda = 3 * D; g.ldb = 3 * D; g.ldc = ldk;
    g.a_kmajor = 1; g.b_kmajor = 1;
    g.alpha = alpha;
    g.batch0 = H; g.batch1 = B;
    g.sa0 = dh; g.sb0 = dh; g.sc0 = (long long)T * ldk;
    g.sa1 = (long long)T * 3 * D; g.sb1 = (long long)T * 3 * D; g.sc1 = (long long)H * T * ldk;
    g.dtype = dtype; g.c_dtype = dtype;
    return gemm_dispatch(g, st);
}

static int apply_keep(void* x, const unsigned* keep, int B, int H, int Tn, int ldk, float p_drop, int dtype, cudaStream_t st) {
    const long long rows = (long long)B * H * Tn;
    const int kw = attn_keep_kw(Tn), tq = attn_keep_tq(Tn);
    B200_DISPATCH_DTYPE(dtype, {
        apply_keep_kernel<T><<<softmax_grid(rows), 256, 0, st>>>((T*)x, keep, Tn, Tn, ldk, kw, tq, 1.0f / (1.0f - p_drop), rows);
    });
    return check_launch("attn_apply_keep");
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

int b200asr_softmax_rows(void* s, const int* klen, float* lse, int B, int H, int Tq, int Tk, int ldk, int dtype,
                         void* stream) {
    B200_REQUIRE(s, "softmax_rows: null pointer");
    B200_REQUIRE(ldk >= Tk && Tk > 0, "softmax_rows: bad Tk/ldk");
    const long long rows = (long long)B * H * Tq;
    if (rows <= 0) return B200ASR_OK;
    B200_DISPATCH_DTYPE(dtype, {
        softmax_rows_kernel<T><<<softmax_grid(rows), 256, 0, as_stream(stream)>>>((T*)s, klen, lse, H, Tq, Tk, ldk, rows);
    });
    return check_launch("softmax_rows");
}

int b200asr_softmax_rows_bwd(void* p, void* dp, const float* lse, const int* klen, int B, int H, int Tq, int Tk,
                             int ldk, int dtype, void* stream) {
    B200_REQUIRE(p && dp, "softmax_rows_bwd: null pointer");
    const long long rows = (long long)B * H * Tq;
    if (rows <= 0) return B200ASR_OK;
    B200_DISPATCH_DTYPE(dtype, {
        softmax_rows_bwd_kernel<T><<<softmax_grid(rows), 256, 0, as_stream(stream)>>>((T*)p, (T*)dp, lse, klen, H, Tq, Tk,
                                                                                   ldk, rows);
    });
    return check_launch("softmax_rows_bwd");
}

int b200asr_mha_set_fused(int on) {
    const int old = g_mha_fused;
    g_mha_fused = on ? 1 : 0;
    return old;
}

size_t b200asr_mha_fwd_workspace_bytes(int B, int T, int H, int dh, int dtype) {
    (void)dh;
    return (size_t)B * H * T * round_up8(T) * elem_size(dtype);
}

size_t b200asr_attn_keep_words(int B, int T, int H) {
    return (size_t)B * H * attn_keep_kw(T) * attn_keep_tq(T);
}

int b200asr_attn_dropout_mask(unsigned* keep, int n_sites, int B, int T, int H, float p, unsigned long long seed,
                              unsigned long long site0, unsigned long long site_stride, unsigned long long offset,
                              const unsigned long long* dev_offset, void* stream) {
    B200_REQUIRE(keep && n_sites > 0 && B > 0 && T > 0 && H > 0, "attn_dropout_mask: bad argument");
    B200_REQUIRE(p >= 0.f && p < 1.f, "attn_dropout_mask: p=%f must be in [0, 1)", (double)p);
    return attn_keep_mask_launch(keep, n_sites, B, T, H, p, seed, site0, site_stride, offset, dev_offset, as_stream(stream));
}

int b200asr_mha_fwd(const void* qkv, void* out, float* lse, const int* klen, void* workspace, size_t workspace_bytes,
                    int B, int T, int H, int dh, float scale, int dtype, void* stream) {
    return b200asr_mha_fwd_dropout(qkv, out, lse, klen, nullptr, 0.f, workspace, workspace_bytes, B, T, H, dh, scale, dtype,
                                   stream);
}

int b200asr_mha_fwd_dropout(const void* qkv, void* out, float* lse, const int* klen, const unsigned* keep, float p_drop,
                            void* workspace, size_t workspace_bytes, int B, int T, int H, int dh, float scale, int dtype,
                            void* stream) {
    B200_REQUIRE(qkv && out && lse && workspace, "mha_fwd: null pointer");
    B200_REQUIRE(B > 0 && T > 0 && H > 0 && dh > 0 && dh % 8 == 0, "mha_fwd: bad shape");
    B200_REQUIRE(workspace_bytes >= b200asr_mha_fwd_workspace_bytes(B, T, H, dh, dtype), "mha_fwd: workspace too small");
    B200_REQUIRE(!keep || (p_drop >= 0.f && p_drop < 1.f), "mha_fwd: p_drop=%f must be in [0, 1)", (double)p_drop);
    cudaStream_t st = as_stream(stream);
    if (g_mha_fused && mha_tc_eligible(dtype, dh, qkv))
        return mha_fwd_tc_launch(qkv, out, lse, klen, keep, p_drop, B, T, H, scale, st);
    const int D = H * dh, ldk = round_up8(T);
    const size_t es = elem_size(dtype);
    int rc = gemm_qk(qkv, workspace, B, T, H, dh, ldk, scale, dtype, st);
    if (rc) return rc;
    rc = b200asr_softmax_rows(workspace, klen, lse, B, H, T, T, ldk, dtype, stream);
    if (rc) return rc;
    if (keep) {   // dropout on the probabilities (HF:189-197); the log-sum-exp stays that of the undropped softmax
        rc = apply_keep(workspace, keep, B, H, T, ldk, p_drop, dtype, st);
        if (rc) return rc;
    }
    // out[b, t, h*dh + n] = sum_k P[b,h,t,k] V[b,k,h,n]
    b200asr_gemm_args g = {};
    g.A = workspace;
    g.B = (const char*)qkv + (size_t)2 * D * es;
    g.C = out;
    g.M = T; g.N = dh; g.K = T;
    g.lda = ldk; g.ldb = 3 * D; g.ldc = D;
    g.a_kmajor = 1; g.b_kmajor = 0;
    g.alpha = 1.f;
    g.batch0 = H; g.batch1 = B;
    g.sa0 = (long long)T * ldk; g.sb0 = dh; g.sc0 = dh;
    g.sa1 = (long long)H * T * ldk; g.sb1 = (long long)T * 3 * D; g.sc1 = (long long)T * D;
    g.dtype = dtype; g.c_dtype = dtype;
    return gemm_dispatch(g, st);
}

size_t b200asr_mha_bwd_workspace_bytes(int B, int T, int H, int dh, int dtype) {
    const size_t composed = 2 * (size_t)B * H * T * round_up8(T) * elem_size(dtype) + 512;
    const size_t fused = (dtype == B200ASR_BF16 && dh == 64) ? mha_bwd_tc_workspace_bytes(B, T, H) : 0;
    return composed > fused ? composed : fused;
}

int b200asr_mha_bwd(const void* qkv, const void* out, const void* dout, const float* lse, const int* klen, void* dqkv,
                    void* workspace, size_t workspace_bytes, int B, int T, int H, int dh, float scale, int dtype,
                    void* stream) {
    return b200asr_mha_bwd_dropout(qkv, out, dout, lse, klen, nullptr, 0.f, dqkv, workspace, workspace_bytes, B, T, H, dh,
                                   scale, dtype, stream);
}

int b200asr_mha_bwd_dropout(const void* qkv, const void* out, const void* dout, const float* lse, const int* klen,
                            const unsigned* keep, float p_drop, void* dqkv, void* workspace, size_t workspace_bytes, int B,
                            int T, int H, int dh, float scale, int dtype, void* stream) {
    (void)out;
    B200_REQUIRE(qkv && dout && lse && dqkv && workspace, "mha_bwd: null pointer");
    B200_REQUIRE(B > 0 && T > 0 && H > 0 && dh > 0 && dh % 8 == 0, "mha_bwd: bad shape");
    B200_REQUIRE(workspace_bytes >= b200asr_mha_bwd_workspace_bytes(B, T, H, dh, dtype), "mha_bwd: workspace too small");
    cudaStream_t st = as_stream(stream);
    if (g_mha_fused && mha_tc_eligible(dtype, dh, qkv) && ((reinterpret_cast<uintptr_t>(dout) | reinterpret_cast<uintptr_t>(dqkv)) & 15) == 0) {
        B200_REQUIRE(out, "mha_bwd: the fused path needs the forward output");
        return mha_bwd_tc_launch(qkv, out, dout, lse, klen, keep, p_drop, dqkv, workspace, B, T, H, scale, st);
    }
    const int D = H * dh, ldk = round_up8(T);
    const size_t es = elem_size(dtype);
    const size_t plane = (size_t)B * H * T * ldk * es;
    void* P = workspace;
    void* dP = (char*)workspace + plane;
    const char* q = (const char*)qkv;
    const char* k = q + (size_t)D * es;
    const char* v = q + (size_t)2 * D * es;
    char* dq = (char*)dqkv;
    char* dk = dq + (size_t)D * es;
    char* dv = dq + (size_t)2 * D * es;
    const long long sS0 = (long long)T * ldk, sS1 = (long long)H * T * ldk;
    const long long sQ1 = (long long)T * 3 * D, sO1 = (long long)T * D;

    int rc = gemm_qk(qkv, P, B, T, H, dh, ldk, scale, dtype, st);  // S (scaled logits)
    if (rc) return rc;
    b200asr_gemm_args g = {};
    g.dtype = dtype; g.c_dtype = dtype; g.batch0 = H; g.batch1 = B; g.alpha = 1.f;
    // dP[t, key] = sum_n dO[t, n] V[key, n]
    g.A = dout; g.B = v; g.C = dP;
    g.M = T; g.N = T; g.K = dh; g.lda = D; g.ldb = 3 * D; g.ldc = ldk; g.a_kmajor = 1; g.b_kmajor = 1;
    g.sa0 = dh; g.sa1 = sO1; g.sb0 = dh; g.sb1 = sQ1; g.sc0 = sS0; g.sc1 = sS1;
    rc = gemm_dispatch(g, st);
    if (rc) return rc;
    if (keep) {   // dP = dP_drop o M / (1 - p): the softmax backward below then sees the gradient w.r.t. the undropped P
        rc = apply_keep(dP, keep, B, H, T, ldk, p_drop, dtype, st);
        if (rc) return rc;
    }
    rc = b200asr_softmax_rows_bwd(P, dP, lse, klen, B, H, T, T, ldk, dtype, stream);  // P, dS
    if (rc) return rc;
    // dQ[t, n] = scale * sum_key dS[t, key] K[key, n]
    g.A = dP; g.B = k; g.C = dq; g.alpha = scale;
    g.M = T; g.N = dh; g.K = T; g.lda = ldk; g.ldb = 3 * D; g.ldc = 3 * D; g.a_kmajor = 1; g.b_kmajor = 0;
    g.sa0 = sS0; g.sa1 = sS1; g.sb0 = dh; g.sb1 = sQ1; g.sc0 = dh; g.sc1 = sQ1;
    rc = gemm_dispatch(g, st);
    if (rc) return rc;
    // dK[key, n] = scale * sum_t dS[t, key] Q[t, n]
    g.A = dP; g.B = q; g.C = dk; g.alpha = scale;
    g.a_kmajor = 0; g.b_kmajor = 0;
    rc = gemm_dispatch(g, st);
    if (rc) return rc;
    // dV[key, n] = sum_t P_drop[t, key] dO[t, n]
    if (keep) {
        rc = apply_keep(P, keep, B, H, T, ldk, p_drop, dtype, st);
        if (rc) return rc;
    }
    g.A = P; g.B = dout; g.C = dv; g.alpha = 1.f;
    g.ldb = D; g.sb0 = dh; g.sb1 = sO1;
    return gemm_dispatch(g, st);
}

}  // extern "C"
