// DACS toggling network kernels: fused Gumbel-softmax hard mask (+ masked hidden states) forward/backward,
// mean over time of AD logits, RecallLoss / CE on [B, C] logits, CosFace diversity term.  All HBM-bound
// elementwise / reduction work: 128-bit loads, one pass over the data.
#include "common.cuh"

namespace b200asr {

// ------------------------------------------------------------------------------------------------
// gumbel hard mask, fused.  One thread = one 16-byte vector of channels of one token.
struct MaskOut { float masked, soft, mask; };

__device__ __forceinline__ MaskOut gumbel_hard(float s0, float s1, float g0, float g1, float inv_tau, float ratio,
                                               float h) {
    if (ratio != 0.f) s0 = (s1 - s0) * ratio + s0;          // models.py:435-441 TOGGLE_RATIO lerp of channel 0
    const float z0 = (s0 + g0) * inv_tau, z1 = (s1 + g1) * inv_tau;
    const float mx = fmaxf(z0, z1);
    const float e0 = expf(z0 - mx), e1 = expf(z1 - mx);
    const float den = e0 + e1;
    const float p0 = e0 / den, p1 = e1 / den;
    const float hard = (p0 >= p1) ? 1.f : 0.f;              // argmax of y_soft, first index on ties
    const float m = (hard - p0) + p0;                        // y_hard - y_soft.detach() + y_soft  (models.py:126)
    MaskOut o;
    o.masked = m * h;
    o.soft = p0;
    o.mask = m;
    return o;
}

__device__ __forceinline__ float gumbel_from_u32(uint32_t r) {
    // U in (0, 1]; E = -log(U) ~ Exp(1); g = -log(E)   (models.py:111-116)
    const float u = ((float)(r >> 8) + 1.0f) * (1.0f / 16777216.0f);
    const float e = fmaxf(-__logf(u), 1e-20f);   // (MUFU lg2: the noise is random, its last bits do not matter)
    return -__logf(e);
}

template <typename T>
__global__ void __launch_bounds__(256) dacs_mask_fwd_kernel(const T* __restrict__ scores, const T* __restrict__ h,
                                                            const float* __restrict__ noise_lm,
                                                            const float* __restrict__ noise_ad, T* __restrict__ lm_masked,
                                                            T* __restrict__ ad_masked, T* __restrict__ soft_lm,
                                                            T* __restrict__ soft_ad, T* __restrict__ lm_mask,
                                                            T* __restrict__ ad_mask, long long M, int d, float inv_tau,
                                                            float ratio, unsigned long long seed,
                                                            unsigned long long offset,
                                                            const unsigned long long* __restrict__ dev_offset) {
    constexpr int N = Vec16<T>::N;
    if (dev_offset) offset += *dev_offset;  // per-call counter kept on the device: a captured launch draws fresh noise
    const int vec_per_row = d / N;
    const long long total = M * vec_per_row;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long m = i / vec_per_row;
        const int c = (int)(i % vec_per_row) * N;
        const T* srow = scores + m * 4 * d;
        Vec16<T> v0, v1, v2, v3, vh;
        v0.load(srow + c);
        v1.load(srow + d + c);
        v2.load(srow + 2 * d + c);
        v3.load(srow + 3 * d + c);
        vh.load(h + m * d + c);
        float s0[N], s1[N], a0[N], a1[N], hh[N];
        v0.unpack(s0); v1.unpack(s1); v2.unpack(a0); v3.unpack(a1); vh.unpack(hh);
        float o_lm[N], o_ad[N], p_lm[N], p_ad[N], m_lm[N], m_ad[N];
#pragma unroll
        for (int j = 0; j < N; ++j) {
            float g0, g1, g2, g3;
            const long long e = m * d + c + j;
            if (noise_lm) {
                g0 = noise_lm[2 * e]; g1 = noise_lm[2 * e + 1];
                g2 = noise_ad[2 * e]; g3 = noise_ad[2 * e + 1];
            } else {
                // one Philox4x32-10 call per element = its four uniforms (counter: element index and call offset, key: seed).
                // (Round 1 went through curand_init per element: its 64-byte state and skip-ahead made the kernel ALU bound
                // at 21 % of the HBM rate.)
                const uint4 r = philox4x32_10(make_uint4((uint32_t)e, (uint32_t)((unsigned long long)e >> 32), (uint32_t)offset,
                                                         (uint32_t)(offset >> 32)),
                                              make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
                g0 = gumbel_from_u32(r.x); g1 = gumbel_from_u32(r.y);
                g2 = gumbel_from_u32(r.z); g3 = gumbel_from_u32(r.w);
            }
            const MaskOut l = gumbel_hard(s0[j], s1[j], g0, g1, inv_tau, ratio, hh[j]);
            const MaskOut a = gumbel_hard(a0[j], a1[j], g2, g3, inv_tau, ratio, hh[j]);
            o_lm[j] = l.masked; p_lm[j] = l.soft; m_lm[j] = l.mask;
            o_ad[j] = a.masked; p_ad[j] = a.soft; m_ad[j] = a.mask;
        }
        Vec16<T> w;
        const long long off = m * d + c;
        w.pack(o_lm); w.store(lm_masked + off);
        w.pack(o_ad); w.store(ad_masked + off);
        w.pack(p_lm); w.store(soft_lm + off);
        w.pack(p_ad); w.store(soft_ad + off);
        if (lm_mask) { w.pack(m_lm); w.store(lm_mask + off); }
        if (ad_mask) { w.pack(m_ad); w.store(ad_mask + off); }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) dacs_mask_bwd_kernel(const T* __restrict__ d_lm, const T* __restrict__ d_ad,
                                                            const T* __restrict__ h, const T* __restrict__ soft_lm,
                                                            const T* __restrict__ soft_ad, const T* __restrict__ lm_mask,
                                                            const T* __restrict__ ad_mask, T* __restrict__ d_scores,
                                                            T* __restrict__ dh, long long M, int d, float inv_tau,
                                                            float ratio) {
    constexpr int N = Vec16<T>::N;
    const int vec_per_row = d / N;
    const long long total = M * vec_per_row;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long m = i / vec_per_row;
        const int c = (int)(i % vec_per_row) * N;
        const long long off = m * d + c;
        Vec16<T> a, b, vh, pl, pa;
        a.load(d_lm + off); b.load(d_ad + off); vh.load(h + off); pl.load(soft_lm + off); pa.load(soft_ad + off);
        float gl[N], ga[N], hh[N], p_l[N], p_a[N];
        a.unpack(gl); b.unpack(ga); vh.unpack(hh); pl.unpack(p_l); pa.unpack(p_a);
        float ds0[N], ds1[N], da0[N], da1[N];
#pragma unroll
        for (int j = 0; j < N; ++j) {
            // mask = hard - sg(p0) + p0  =>  d p0 = d mask = d masked * h;  p0 = softmax(z)[0]
            const float dzl = p_l[j] * (1.f - p_l[j]) * gl[j] * hh[j] * inv_tau;  // d z0 (= -d z1), already / tau
            const float dza = p_a[j] * (1.f - p_a[j]) * ga[j] * hh[j] * inv_tau;
            ds0[j] = dzl * (1.f - ratio);
            ds1[j] = dzl * ratio - dzl;
            da0[j] = dza * (1.f - ratio);
            da1[j] = dza * ratio - dza;
        }
        T* drow = d_scores + m * 4 * d;
        Vec16<T> w;
        w.pack(ds0); w.store(drow + c);
        w.pack(ds1); w.store(drow + d + c);
        w.pack(da0); w.store(drow + 2 * d + c);
        w.pack(da1); w.store(drow + 3 * d + c);
        if (dh) {
            float ml[N], ma[N], o[N];
            if (lm_mask) {
                Vec16<T> x, y;
                x.load(lm_mask + off); y.load(ad_mask + off);
                x.unpack(ml); y.unpack(ma);
            } else {
#pragma unroll
                for (int j = 0; j < N; ++j) { ml[j] = p_l[j] >= 0.5f ? 1.f : 0.f; ma[j] = p_a[j] >= 0.5f ? 1.f : 0.f; }
            }
#pragma unroll
            for (int j = 0; j < N; ++j) o[j] = ml[j] * gl[j] + ma[j] * ga[j];
            w.pack(o); w.store(dh + off);
        }
    }
}

// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void mean_time_fwd_kernel(const T* __restrict__ x, float* __restrict__ out, int T_, int C) {
    __shared__ float red[32];
    const int b = blockIdx.x;
    for (int c = 0; c < C; ++c) {
        float s = 0.f;
        for (int t = threadIdx.x; t < T_; t += blockDim.x) s += to_f<T>(x[((long long)b * T_ + t) * C + c]);
        s = block_sum(s, red);
        if (threadIdx.x == 0) out[b * C + c] = s / (float)T_;
        __syncthreads();
    }
}

template <typename T>
__global__ void mean_time_bwd_kernel(const float* __restrict__ dout, T* __restrict__ dx, int T_, int C, float scale,
                                     long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const int b = (int)(i / ((long long)T_ * C));
        dx[i] = from_f<T>(dout[b * C + c] * scale / (float)T_);
    }
}

// ------------------------------------------------------------------------------------------------
// RecallLoss (models.py:187-260) / CE on [B, C] logits, C <= 8.  Single block; thread per row.
struct AdLossW { float w[8]; };

__global__ void ad_loss_kernel(const float* __restrict__ logits, const long long* __restrict__ target,
                               float* __restrict__ loss, float* __restrict__ dlogits, int B, int C, int kind, AdLossW ww) {
    __shared__ float red[32];
    const float eps = 1e-5f;
    float local = 0.f;
    for (int n = threadIdx.x; n < B; n += blockDim.x) {
        float x[8], pt[8], g[8];
        float mx = -INFINITY;
        for (int c = 0; c < C; ++c) { x[c] = logits[n * C + c]; mx = fmaxf(mx, x[c]); }
        float sum = 0.f;
        for (int c = 0; c < C; ++c) sum += expf(x[c] - mx);
        const float lz = mx + logf(sum);
        for (int c = 0; c < C; ++c) pt[c] = expf(x[c] - lz);
        const int y = (int)target[n];
        const float invN = 1.0f / (float)B;
        if (kind == 5) {  // cross entropy (mean)
            local += -(x[y] - lz) * invN;
            for (int c = 0; c < C; ++c) g[c] = 0.f;
            if (dlogits)
                for (int c = 0; c < C; ++c) dlogits[n * C + c] = (pt[c] - (c == y ? 1.f : 0.f)) * invN;
            continue;
        }
        for (int c = 0; c < C; ++c) {
            const float o = (c == y) ? 1.f : 0.f;
            const float tp = pt[c] * o;
            const float r = (tp + eps) / (o + eps);
            const float p = (tp + eps) / (pt[c] + eps);
            const float dr = o / (o + eps);                                              // d recall / d pt
            const float dp = (o * (pt[c] + eps) - (tp + eps)) / ((pt[c] + eps) * (pt[c] + eps));  // d precision / d pt
            const float w = ww.w[c];
            float val, dval;  // per-element contribution (already weighted) and its derivative w.r.t. pt
            switch (kind) {
                case 0: val = (1.f - r) * w; dval = -dr * w; break;                      // recall
                case 1: val = (1.f - p) * w; dval = -dp * w; break;                      // prec
                case 2: {                                                                // f1: 1 - mean(f1*w*C)
                    const float f = 2.f * r * p / (r + p);
                    const float df = 2.f * ((dr * p + r * dp) * (r + p) - r * p * (dr + dp)) / ((r + p) * (r + p));
                    val = -f * w; dval = -df * w; break;
                }
                case 3: val = -r * w; dval = -dr * w; break;                             // recall_ori: 1 - mean(..)
                default: val = -p * w; dval = -dp * w; break;                            // prec_ori
            }
            local += val * invN;
            g[c] = dval * invN;
        }
        if (dlogits) {
            float dot = 0.f;
            for (int c = 0; c < C; ++c) dot += g[c] * pt[c];
            for (int c = 0; c < C; ++c) dlogits[n * C + c] = pt[c] * (g[c] - dot);
        }
    }
    local = block_sum(local, red);
    if (threadIdx.x == 0) loss[0] = local + ((kind >= 2 && kind <= 4) ? 1.f : 0.f);
}

// ------------------------------------------------------------------------------------------------
// CosFace: one warp per row; rows [0,n) from x1 (label 0), [n,2n) from x2 (label 1).
template <typename T>
__global__ void __launch_bounds__(256) cosface_kernel(const T* __restrict__ x1, const T* __restrict__ x2,
                                                      const float* __restrict__ fcw, float* __restrict__ row_loss,
                                                      T* __restrict__ dx1, T* __restrict__ dx2, long long n, int d, float s,
                                                      float mrg) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const float gcoef = 1.0f / (float)(2 * n);
    for (long long r = warp; r < 2 * n; r += nwarps) {
        const int y = r >= n ? 1 : 0;
        const T* x = y ? x2 + (r - n) * d : x1 + r * d;
        float nn = 0.f, d0 = 0.f, d1 = 0.f;
        for (int c = lane; c < d; c += 32) {
            const float v = to_f<T>(x[c]);
            nn += v * v;
            d0 += v * fcw[c];
            d1 += v * fcw[d + c];
        }
        nn = warp_sum(nn); d0 = warp_sum(d0); d1 = warp_sum(d1);
        const float norm = fmaxf(sqrtf(nn), 1e-12f);  // F.normalize eps
        const float wf0 = d0 / norm, wf1 = d1 / norm;
        const float wy = y ? wf1 : wf0, wo = y ? wf0 : wf1;
        const float num = s * (wy - mrg);
        const float ea = expf(num), eb = expf(s * wo);
        const float L = num - logf(ea + eb);
        if (lane == 0) row_loss[r] = L;
        T* dx = y ? (dx2 ? dx2 + (r - n) * d : nullptr) : (dx1 ? dx1 + r * d : nullptr);
        if (dx) {
            const float q = eb / (ea + eb);
            const float gy = -s * q * gcoef, go = s * q * gcoef;  // d(-mean L) / d wf_y, d wf_other
            const float* Wy = fcw + (y ? d : 0);
            const float* Wo = fcw + (y ? 0 : d);
            const float inv = 1.0f / norm;
            // F.normalize: x / max(||x||, eps).  Below eps the denominator is a constant, so the reference's
            // autograd has no projection term there (d xn / dx = 1 / eps).
            const bool clamped = sqrtf(nn) < 1e-12f;
            for (int c = lane; c < d; c += 32) {
                const float xn = to_f<T>(x[c]) * inv;
                const float g = clamped ? (gy * Wy[c] + go * Wo[c]) * inv
                                        : (gy * (Wy[c] - xn * wy) + go * (Wo[c] - xn * wo)) * inv;
                dx[c] = from_f<T>(g);
            }
        }
    }
}

__global__ void neg_mean_kernel(const float* __restrict__ v, long long n, float* __restrict__ out) {
    __shared__ float red[32];
    float s = 0.f;
    for (long long i = threadIdx.x; i < n; i += blockDim.x) s += v[i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) out[0] = -s / (float)n;
}

static int ew_grid(long long total, int threads) {
    long long g = (total + threads - 1) / threads;
    const long long cap = (long long)num_sms() * 8;
    return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

int b200asr_dacs_mask_fwd_ctr(const void* scores, const void* h, const float* noise_lm, const float* noise_ad,
                              void* lm_masked, void* ad_masked, void* soft_lm, void* soft_ad, void* lm_mask, void* ad_mask,
                              long long M, int d, float tau, float toggle_ratio, unsigned long long seed,
                              unsigned long long offset, const unsigned long long* dev_offset, int dtype, void* stream);

int b200asr_dacs_mask_fwd(const void* scores, const void* h, const float* noise_lm, const float* noise_ad, void* lm_masked,
                          void* ad_masked, void* soft_lm, void* soft_ad, void* lm_mask, void* ad_mask, long long M, int d,
                          float tau, float toggle_ratio, unsigned long long seed, unsigned long long offset, int dtype,
                          void* stream) {
    return b200asr_dacs_mask_fwd_ctr(scores, h, noise_lm, noise_ad, lm_masked, ad_masked, soft_lm, soft_ad, lm_mask, ad_mask, M,
                                     d, tau, toggle_ratio, seed, offset, nullptr, dtype, stream);
}

int b200asr_dacs_mask_fwd_ctr(const void* scores, const void* h, const float* noise_lm, const float* noise_ad,
                              void* lm_masked, void* ad_masked, void* soft_lm, void* soft_ad, void* lm_mask, void* ad_mask,
                              long long M, int d, float tau, float toggle_ratio, unsigned long long seed,
                              unsigned long long offset, const unsigned long long* dev_offset, int dtype, void* stream) {
    B200_REQUIRE(scores && h && lm_masked && ad_masked && soft_lm && soft_ad, "dacs_mask_fwd: null pointer");
    B200_REQUIRE((noise_lm == nullptr) == (noise_ad == nullptr), "dacs_mask_fwd: give both noise tensors or neither");
    B200_REQUIRE(M > 0 && d > 0 && d % (dtype == B200ASR_BF16 ? 8 : 4) == 0, "dacs_mask_fwd: bad shape M=%lld d=%d", M, d);
    B200_REQUIRE(tau > 0.f, "dacs_mask_fwd: tau must be > 0");
    B200_DISPATCH_DTYPE(dtype, {
        const long long total = M * (d / Vec16<T>::N);
        dacs_mask_fwd_kernel<T><<<ew_grid(total, 256), 256, 0, as_stream(stream)>>>(
            (const T*)scores, (const T*)h, noise_lm, noise_ad, (T*)lm_masked, (T*)ad_masked, (T*)soft_lm, (T*)soft_ad,
            (T*)lm_mask, (T*)ad_mask, M, d, 1.0f / tau, toggle_ratio, seed, offset, dev_offset);
    });
    return check_launch("dacs_mask_fwd");
}

int b200asr_dacs_mask_bwd(const void* d_lm_masked, const void* d_ad_masked, const void* h, const void* soft_lm,
                          const void* soft_ad, const void* lm_mask, const void* ad_mask, void* d_scores, void* dh, long long M,
                          int d, float tau, float toggle_ratio, int dtype, void* stream) {
    B200_REQUIRE(d_lm_masked && d_ad_masked && h && soft_lm && soft_ad && d_scores, "dacs_mask_bwd: null pointer");
    B200_REQUIRE((lm_mask == nullptr) == (ad_mask == nullptr), "dacs_mask_bwd: give both masks or neither");
    B200_REQUIRE(M > 0 && d > 0 && d % (dtype == B200ASR_BF16 ? 8 : 4) == 0, "dacs_mask_bwd: bad shape");
    B200_DISPATCH_DTYPE(dtype, {
        const long long total = M * (d / Vec16<T>::N);
        dacs_mask_bwd_kernel<T><<<ew_grid(total, 256), 256, 0, as_stream(stream)>>>(
            (const T*)d_lm_masked, (const T*)d_ad_masked, (const T*)h, (const T*)soft_lm, (const T*)soft_ad,
            (const T*)lm_mask, (const T*)ad_mask, (T*)d_scores, (T*)dh, M, d, 1.0f / tau, toggle_ratio);
    });
    return check_launch("dacs_mask_bwd");
}

int b200asr_mean_time_fwd(const void* x, float* out, int B, int T_, int C, int dtype, void* stream) {
    B200_REQUIRE(x && out && B > 0 && T_ > 0 && C > 0, "mean_time_fwd: bad args");
    B200_DISPATCH_DTYPE(dtype, { mean_time_fwd_kernel<T><<<B, 256, 0, as_stream(stream)>>>((const T*)x, out, T_, C); });
    return check_launch("mean_time_fwd");
}

int b200asr_mean_time_bwd(const float* dout, void* dx, int B, int T_, int C, float scale, int dtype, void* stream) {
    B200_REQUIRE(dout && dx && B > 0 && T_ > 0 && C > 0, "mean_time_bwd: bad args");
    const long long total = (long long)B * T_ * C;
    B200_DISPATCH_DTYPE(dtype, {
        mean_time_bwd_kernel<T><<<ew_grid(total, 256), 256, 0, as_stream(stream)>>>(dout, (T*)dx, T_, C, scale, total);
    });
    return check_launch("mean_time_bwd");
}

int b200asr_ad_loss(const float* logits, const long long* target, float* loss, float* dlogits, int B, int C, int kind,
                    const float* w_host, void* stream) {
    B200_REQUIRE(logits && target && loss, "ad_loss: null pointer");
    B200_REQUIRE(B > 0 && C >= 2 && C <= 8, "ad_loss: C=%d unsupported", C);
    B200_REQUIRE(kind >= 0 && kind <= 5, "ad_loss: unknown kind %d", kind);
    AdLossW ww;
    float sum = 0.f;
    for (int c = 0; c < C; ++c) sum += w_host ? w_host[c] : 1.f;
    for (int c = 0; c < 8; ++c) ww.w[c] = (c < C) ? (w_host ? w_host[c] : 1.f) / sum : 0.f;
    ad_loss_kernel<<<1, 128, 0, as_stream(stream)>>>(logits, target, loss, dlogits, B, C, kind, ww);
    return check_launch("ad_loss");
}

int b200asr_cosface_fwd_bwd(const void* x1, const void* x2, const float* fc_w, float* loss, void* dx1, void* dx2,
                            void* workspace, size_t workspace_bytes, long long n, int d, float s, float m, int dtype,
                            void* stream) {
    B200_REQUIRE(x1 && x2 && fc_w && loss && workspace, "cosface: null pointer");
    B200_REQUIRE(n > 0 && d > 0, "cosface: bad shape");
    B200_REQUIRE(workspace_bytes >= (size_t)(2 * n) * sizeof(float), "cosface: workspace too small");
    cudaStream_t st = as_stream(stream);
    B200_DISPATCH_DTYPE(dtype, {
        cosface_kernel<T><<<ew_grid(2 * n * 32, 256), 256, 0, st>>>((const T*)x1, (const T*)x2, fc_w, (float*)workspace,
                                                                  (T*)dx1, (T*)dx2, n, d, s, m);
    });
    int rc = check_launch("cosface");
    if (rc) return rc;
    neg_mean_kernel<<<1, 1024, 0, st>>>((const float*)workspace, 2 * n, loss);
    return check_launch("cosface_reduce");
}

}  // extern "C"
