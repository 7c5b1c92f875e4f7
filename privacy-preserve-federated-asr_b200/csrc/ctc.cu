// CTC loss: fp32 log-softmax + log-space alpha/beta recursions + gradient w.r.t. the logits, and greedy decode.
// One CTA per (head, utterance).  The log-probabilities of the utterance live in shared memory (T*V*4 bytes);
// warp 0 runs the alpha recursion while warp 1 runs beta concurrently (one warp per recursion: each lane owns
// SPL consecutive lattice states, the two neighbour states come from the previous lane by shuffle); then all
// warps combine alpha+beta into the gradient.  Latency-bound (T sequential steps), not bandwidth-bound.
#include "common.cuh"

namespace b200asr {

namespace {
constexpr int CTC_THREADS = 256;
constexpr float NEG_INF = -INFINITY;

__device__ __forceinline__ float lse2(float a, float b) {
    const float m = fmaxf(a, b);
    if (m == NEG_INF) return NEG_INF;
    return m + logf(expf(a - m) + expf(b - m));
}
__device__ __forceinline__ float lse3(float a, float b, float c) {
    const float m = fmaxf(fmaxf(a, b), c);
    if (m == NEG_INF) return NEG_INF;
    return m + logf(expf(a - m) + expf(b - m) + expf(c - m));
}

// forward (DIR=+1) or backward (DIR=-1) lattice recursion by one warp.
// state s of lane: s = lane*SPL + i.  ext[s] = blank (s even) or target[(s-1)/2] (s odd).
template <int SPL, int DIR>
__device__ void lattice_pass(const float* __restrict__ lp, int V, const int* __restrict__ tgt, int L, int Tb, int blank,
                             float* __restrict__ out /* [T][Sld] */, int Sld) {
    const int lane = threadIdx.x & 31;
    const int S = 2 * L + 1;
    int lab[SPL];
    bool skip[SPL];  // may take the 2-step transition (from s-2 forward / from s+2 backward)
#pragma unroll
    for (int i = 0; i < SPL; ++i) {
        const int s = lane * SPL + i;
        lab[i] = blank;
        skip[i] = false;
        if (s < S && (s & 1)) {
            const int l = tgt[(s - 1) >> 1];
            lab[i] = l;
            if (DIR > 0) skip[i] = (s >= 3) && (tgt[(s - 3) >> 1] != l);
            else skip[i] = (s + 2 < S) && (tgt[(s + 1) >> 1] != l);
        }
    }
    float a[SPL];
    const int t0 = DIR > 0 ? 0 : Tb - 1;
#pragma unroll
    for (int i = 0; i < SPL; ++i) {
        const int s = lane * SPL + i;
        bool init;
        if (DIR > 0) init = (s == 0) || (s == 1 && S > 1);
        else init = (s == S - 1) || (s == S - 2 && S > 1);
        a[i] = (init && s < S) ? lp[t0 * V + lab[i]] : NEG_INF;
        if (s < S) out[(long long)t0 * Sld + s] = a[i];
    }
    for (int step = 1; step < Tb; ++step) {
        const int t = DIR > 0 ? step : Tb - 1 - step;
        float n1, n2;  // neighbour lane's two boundary states
        if (DIR > 0) {
            n1 = __shfl_up_sync(0xffffffffu, a[SPL - 1], 1);
            n2 = __shfl_up_sync(0xffffffffu, a[SPL - 2], 1);
            if (lane == 0) n1 = n2 = NEG_INF;
        } else {
            n1 = __shfl_down_sync(0xffffffffu, a[0], 1);
            n2 = __shfl_down_sync(0xffffffffu, a[1], 1);
            if (lane == 31) n1 = n2 = NEG_INF;
        }
        float nxt[SPL];
#pragma unroll
        for (int i = 0; i < SPL; ++i) {
            float p1, p2;
            if (DIR > 0) {
                p1 = (i >= 1) ? a[i - 1] : n1;
                p2 = (i >= 2) ? a[i - 2] : (i == 1 ? n1 : n2);
                if (i == 0) { p1 = n1; p2 = n2; }
            } else {
                p1 = (i + 1 < SPL) ? a[i + 1] : n1;
                p2 = (i + 2 < SPL) ? a[i + 2] : (i + 2 == SPL ? n1 : n2);
            }
            const float v = skip[i] ? lse3(a[i], p1, p2) : lse2(a[i], p1);
            nxt[i] = v + lp[t * V + lab[i]];
        }
#pragma unroll
        for (int i = 0; i < SPL; ++i) {
            const int s = lane * SPL + i;
            a[i] = (s < S) ? nxt[i] : NEG_INF;
            if (s < S) out[(long long)t * Sld + s] = a[i];
        }
    }
}

template <int DIR>
__device__ void lattice_dispatch(int spl, const float* lp, int V, const int* tgt, int L, int Tb, int blank, float* out,
                                 int Sld) {
    switch (spl) {
        case 2: lattice_pass<2, DIR>(lp, V, tgt, L, Tb, blank, out, Sld); break;
        case 4: lattice_pass<4, DIR>(lp, V, tgt, L, Tb, blank, out, Sld); break;
        case 8: lattice_pass<8, DIR>(lp, V, tgt, L, Tb, blank, out, Sld); break;
        default: lattice_pass<16, DIR>(lp, V, tgt, L, Tb, blank, out, Sld); break;
    }
}
}  // namespace

template <typename T>
__global__ void __launch_bounds__(CTC_THREADS) ctc_kernel(const T* __restrict__ logits, const int* __restrict__ targets,
                                                          const int* __restrict__ tgt_off, const int* __restrict__ tgt_len,
                                                          const int* __restrict__ in_len, const float* __restrict__ grad_scale,
                                                          float* __restrict__ loss, T* __restrict__ grad,
                                                          float* __restrict__ ws, int B, int T_, int V, int max_tgt,
                                                          int blank, int zero_infinity, int spl) {
    extern __shared__ float sm[];
    const int b = blockIdx.x, head = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int Sld = 2 * max_tgt + 1;
    const int L = min(tgt_len[b], max_tgt);
    const int S = 2 * L + 1;
    const int Tb = max(0, min(in_len[b], T_));
    float* lp = sm;                          // [T][V]
    int* tgt_s = (int*)(lp + (size_t)T_ * V);  // [max_tgt]
    float* acc_s = (float*)(tgt_s + max_tgt + 1);  // [nwarps][V]
    const long long hb = (long long)head * B + b;
    const T* x = logits + hb * T_ * V;
    float* alpha = ws + hb * 2 * T_ * Sld;
    float* beta = alpha + (long long)T_ * Sld;

    for (int i = threadIdx.x; i < L; i += blockDim.x) tgt_s[i] = targets[tgt_off[b] + i];
    // ---- phase 1: fp32 log-softmax of the valid frames
    for (int t = warp; t < Tb; t += nwarps) {
        float mx = NEG_INF;
        for (int v = lane; v < V; v += 32) mx = fmaxf(mx, to_f<T>(x[(long long)t * V + v]));
        mx = warp_max(mx);
        float sum = 0.f;
        for (int v = lane; v < V; v += 32) sum += expf(to_f<T>(x[(long long)t * V + v]) - mx);
        sum = warp_sum(sum);
        const float lz = mx + logf(sum);
        for (int v = lane; v < V; v += 32) lp[t * V + v] = to_f<T>(x[(long long)t * V + v]) - lz;
    }
    __syncthreads();
    // ---- phase 2: alpha (warp 0) and beta (warp 1) concurrently
    if (Tb > 0) {
        if (warp == 0) lattice_dispatch<+1>(spl, lp, V, tgt_s, L, Tb, blank, alpha, Sld);
        else if (warp == 1) lattice_dispatch<-1>(spl, lp, V, tgt_s, L, Tb, blank, beta, Sld);
    }
    __threadfence_block();
    __syncthreads();
    // ---- phase 3: loss and gradient
    float nll = INFINITY;
    if (Tb > 0) {
        const float aT = alpha[(long long)(Tb - 1) * Sld + S - 1];
        const float aT2 = S > 1 ? alpha[(long long)(Tb - 1) * Sld + S - 2] : NEG_INF;
        nll = -lse2(aT, aT2);
    }
    if (Tb == 0 && L == 0) nll = 0.f;  // empty input, empty target (torch convention)
    const bool infeasible = !(nll < INFINITY);
    if (threadIdx.x == 0) loss[hb] = (infeasible && zero_infinity) ? 0.f : nll;
    if (!grad) return;
    const float gs = grad_scale ? grad_scale[hb] : 1.f;
    T* g = grad + hb * T_ * V;
    float* acc = acc_s + warp * V;
    for (int t = warp; t < T_; t += nwarps) {
        if (t >= Tb || infeasible) {
            // torch: gradient of padded frames is 0; zero_infinity zeroes infeasible utterances (NaN otherwise)
            for (int v = lane; v < V; v += 32)
                g[(long long)t * V + v] = from_f<T>((t < Tb && !zero_infinity) ? NAN : 0.f);
            continue;
        }
        const float* ar = alpha + (long long)t * Sld;
        const float* br = beta + (long long)t * Sld;
        float mx = NEG_INF;
        for (int s = lane; s < S; s += 32) mx = fmaxf(mx, ar[s] + br[s]);
        mx = warp_max(mx);
        for (int v = lane; v < V; v += 32) acc[v] = 0.f;
        __syncwarp();
        for (int s = lane; s < S; s += 32) {
            const float ab = ar[s] + br[s];
            if (ab > NEG_INF) {
                const int l = (s & 1) ? tgt_s[(s - 1) >> 1] : blank;
                atomicAdd(&acc[l], expf(ab - mx));
            }
        }
        __syncwarp();
        for (int v = lane; v < V; v += 32) {
            const float l = lp[t * V + v];
            const float a = acc[v];
            // d nll / d logit = softmax - exp(log(sum alpha*beta) + nll - lp)      (ATen ctc_loss backward form)
            const float occ = a > 0.f ? expf(mx + logf(a) + nll - l) : 0.f;
            g[(long long)t * V + v] = from_f<T>((expf(l) - occ) * gs);
        }
        __syncwarp();
    }
}

template <typename T>
__global__ void __launch_bounds__(256) greedy_kernel(const T* __restrict__ logits, const int* __restrict__ in_len,
                                                     int* __restrict__ ids, int* __restrict__ collapsed,
                                                     int* __restrict__ clen, int T_, int V, int blank) {
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int Tb = in_len ? min(in_len[b], T_) : T_;
    const T* x = logits + (long long)b * T_ * V;
    for (int t = warp; t < T_; t += nwarps) {
        float best = NEG_INF;
        int bi = 0x7fffffff;
        for (int v = lane; v < V; v += 32) {
            const float f = to_f<T>(x[(long long)t * V + v]);
            if (f > best || (f == best && v < bi)) { best = f; bi = v; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) ids[(long long)b * T_ + t] = bi;
    }
    __threadfence_block();
    __syncthreads();
    if (threadIdx.x == 0) {
        int n = 0, prev = -1;
        for (int t = 0; t < Tb; ++t) {
            const int id = ids[(long long)b * T_ + t];
            if (id != prev && id != blank) collapsed[(long long)b * T_ + n++] = id;
            prev = id;
        }
        clen[b] = n;
        for (int t = n; t < T_; ++t) collapsed[(long long)b * T_ + t] = -1;
    }
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

size_t b200asr_ctc_workspace_bytes(int n_heads, int B, int T, int max_tgt_len) {
    return (size_t)n_heads * B * 2 * T * (2 * max_tgt_len + 1) * sizeof(float);
}

int b200asr_ctc_fwd_bwd(const void* logits, const int* targets, const int* tgt_off, const int* tgt_len,
                        const int* in_len, const float* grad_scale, float* loss, void* grad, void* workspace,
                        size_t workspace_bytes, int n_heads, int B, int T_, int V, int max_tgt_len, int blank,
                        int zero_infinity, int dtype, void* stream) {
    B200_REQUIRE(logits && targets && tgt_off && tgt_len && in_len && loss && workspace, "ctc: null pointer");
    B200_REQUIRE(n_heads >= 1 && B >= 1 && T_ >= 1 && V >= 2, "ctc: bad shape");
    B200_REQUIRE(blank >= 0 && blank < V, "ctc: blank out of range");
    if (max_tgt_len < 1) max_tgt_len = 1;
    B200_REQUIRE(workspace_bytes >= b200asr_ctc_workspace_bytes(n_heads, B, T_, max_tgt_len), "ctc: workspace too small");
    const int S = 2 * max_tgt_len + 1;
    int spl = 2;
    while (spl * 32 < S) spl *= 2;
    if (spl > 16) {
        set_error("ctc: target length %d exceeds the supported 255", max_tgt_len);
        return B200ASR_ERR_UNSUPPORTED;
    }
    const size_t smem = ((size_t)T_ * V + (max_tgt_len + 1) + (CTC_THREADS / 32) * V) * sizeof(float);
    if (smem > 227 * 1024) {
        set_error("ctc: T*V=%d*%d log-probabilities do not fit in shared memory", T_, V);
        return B200ASR_ERR_UNSUPPORTED;
    }
    dim3 grid(B, n_heads);
    cudaStream_t st = as_stream(stream);
    B200_DISPATCH_DTYPE(dtype, {
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(ctc_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) {
                set_error("ctc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
                return B200ASR_ERR_CUDA;
            }
        }
        ctc_kernel<T><<<grid, CTC_THREADS, smem, st>>>((const T*)logits, targets, tgt_off, tgt_len, in_len, grad_scale, loss,
                                                        (T*)grad, (float*)workspace, B, T_, V, max_tgt_len, blank,
                                                        zero_infinity, spl);
    });
    return check_launch("ctc_fwd_bwd");
}

int b200asr_greedy_decode(const void* logits, const int* in_len, int* out_ids, int* collapsed, int* collapsed_len, int B,
                          int T_, int V, int blank, int dtype, void* stream) {
    B200_REQUIRE(logits && out_ids && collapsed && collapsed_len, "greedy_decode: null pointer");
    B200_REQUIRE(B >= 1 && T_ >= 1 && V >= 1, "greedy_decode: bad shape");
    B200_DISPATCH_DTYPE(dtype, {
        greedy_kernel<T><<<B, 256, 0, as_stream(stream)>>>((const T*)logits, in_len, out_ids, collapsed, collapsed_len, T_, V,
                                                          blank);
    });
    return check_launch("greedy_decode");
}

}  // extern "C"
