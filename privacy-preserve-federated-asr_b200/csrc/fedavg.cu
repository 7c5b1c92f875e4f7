// FedAvg on flat fp32 parameter buffers (federated/src/utils.py:240-249) and the fused AdamW step that shares
// the flat layout.  HBM-bound streaming kernels: 128-bit loads, grid = multiple of the SM count, one pass.
#include <string.h>

#include "common.cuh"

namespace b200asr {

constexpr int FEDAVG_MAX_K = 16;
struct FedPtrs {
    const float* in[FEDAVG_MAX_K];
    float w[FEDAVG_MAX_K];
};
struct PeerPtrs {
    float* buf[FEDAVG_MAX_K];
    float w[FEDAVG_MAX_K];
};

__device__ __forceinline__ float4 ld_stream(const float* p) {
    float4 r;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// uniform = 1: ((in0 + in1) + in2 ...) / K  (bit-exact with the reference's sequential sum then true division)
template <int UNIFORM>
__global__ void __launch_bounds__(256) fedavg_kernel(const FedPtrs p, float* __restrict__ out, int K, long long n) {
    const long long nvec = n / 4;
    const float kf = (float)K;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        float4 acc = ld_stream(p.in[0] + 4 * i);
        if (!UNIFORM) { acc.x *= p.w[0]; acc.y *= p.w[0]; acc.z *= p.w[0]; acc.w *= p.w[0]; }
        for (int j = 1; j < K; ++j) {
            const float4 v = ld_stream(p.in[j] + 4 * i);
            if (UNIFORM) { acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
            else {
                acc.x = fmaf(p.w[j], v.x, acc.x); acc.y = fmaf(p.w[j], v.y, acc.y);
                acc.z = fmaf(p.w[j], v.z, acc.z); acc.w = fmaf(p.w[j], v.w, acc.w);
            }
        }
        if (UNIFORM) { acc.x = __fdiv_rn(acc.x, kf); acc.y = __fdiv_rn(acc.y, kf); acc.z = __fdiv_rn(acc.z, kf); acc.w = __fdiv_rn(acc.w, kf); }
        *reinterpret_cast<float4*>(out + 4 * i) = acc;
    }
    // tail (n % 4 elements)
    const long long t0 = nvec * 4;
    for (long long i = t0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float acc = UNIFORM ? p.in[0][i] : p.w[0] * p.in[0][i];
        for (int j = 1; j < K; ++j) acc = UNIFORM ? acc + p.in[j][i] : fmaf(p.w[j], p.in[j][i], acc);
        out[i] = UNIFORM ? __fdiv_rn(acc, kf) : acc;
    }
}

__global__ void __launch_bounds__(256) prescale_kernel(float* __restrict__ x, float w, long long n) {
    const long long nvec = n / 4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        float4 v = *reinterpret_cast<float4*>(x + 4 * i);
        v.x *= w; v.y *= w; v.z *= w; v.w *= w;
        *reinterpret_cast<float4*>(x + 4 * i) = v;
    }
    for (long long i = nvec * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        x[i] *= w;
}

// Rank `rank` owns slice [lo, hi): reads that slice from every peer (NVLink loads), reduces, writes the result
// into every peer's slice (NVLink stores).  One kernel = reduce-scatter + all-gather with the scaling fused in.
template <int UNIFORM>
__global__ void __launch_bounds__(256) fedavg_peer_kernel(const PeerPtrs p, int world, long long lo, long long hi,
                                                          const int* __restrict__ abort_flag) {
    // the barrier in front of this kernel gave up on a peer (straggler / dead client): leave every buffer untouched --
    // a partial average written into the peers would be silently wrong on all ranks (PeerFedAvg.check() raises)
    if (abort_flag && *abort_flag) return;
    const float kf = (float)world;
    const long long nvec = (hi - lo) / 4;  // lo, hi are multiples of 4 except the global tail handled by the last rank
    constexpr int U = 4;                   // independent 16-byte loads per peer in flight per thread (NVLink latency)
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += U * stride) {
        float4 acc[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * stride;
            acc[u] = (i < nvec) ? __ldcs(reinterpret_cast<const float4*>(p.buf[0] + lo + 4 * i)) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (!UNIFORM) { acc[u].x *= p.w[0]; acc[u].y *= p.w[0]; acc[u].z *= p.w[0]; acc[u].w *= p.w[0]; }
        }
        for (int j = 1; j < world; ++j) {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const long long i = i0 + u * stride;
                v[u] = (i < nvec) ? __ldcs(reinterpret_cast<const float4*>(p.buf[j] + lo + 4 * i)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (UNIFORM) { acc[u].x += v[u].x; acc[u].y += v[u].y; acc[u].z += v[u].z; acc[u].w += v[u].w; }
                else {
                    acc[u].x = fmaf(p.w[j], v[u].x, acc[u].x); acc[u].y = fmaf(p.w[j], v[u].y, acc[u].y);
                    acc[u].z = fmaf(p.w[j], v[u].z, acc[u].z); acc[u].w = fmaf(p.w[j], v[u].w, acc[u].w);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * stride;
            if (i >= nvec) continue;
            if (UNIFORM) {
                acc[u].x = __fdiv_rn(acc[u].x, kf); acc[u].y = __fdiv_rn(acc[u].y, kf);
                acc[u].z = __fdiv_rn(acc[u].z, kf); acc[u].w = __fdiv_rn(acc[u].w, kf);
            }
            for (int j = 0; j < world; ++j) __stcs(reinterpret_cast<float4*>(p.buf[j] + lo + 4 * i), acc[u]);
        }
    }
    for (long long e = lo + nvec * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; e < hi; e += (long long)gridDim.x * blockDim.x) {
        float acc = UNIFORM ? p.buf[0][e] : p.w[0] * p.buf[0][e];
        for (int j = 1; j < world; ++j) acc = UNIFORM ? acc + p.buf[j][e] : fmaf(p.w[j], p.buf[j][e], acc);
        if (UNIFORM) acc = __fdiv_rn(acc, kf);
        for (int j = 0; j < world; ++j) p.buf[j][e] = acc;
    }
}

// NVLS variant: the flat buffers of all ranks are bound to ONE multicast object (NVSwitch); rank r owns slice [lo, hi):
//   multimem.ld_reduce  -- the switch fetches the slice from every GPU and returns the SUM (in-switch reduction)
//   multimem.st         -- the switch writes the averaged value into every GPU's buffer
// so each rank moves 4P/N bytes in and 4P/N bytes out over its own links, whatever N is (the peer kernel above moves
// 4P (N-1)/N each way).  The reduction order inside the switch is fixed by the fabric, not left-to-right: equal to the
// reference's sequential sum to fp32 rounding (<= 1e-6 relative), not bit for bit -- the peer kernel stays the bit-exact path.
__global__ void __launch_bounds__(256) fedavg_nvls_kernel(float* __restrict__ mc, int world, long long lo, long long hi,
                                                          const int* __restrict__ abort_flag) {
    if (abort_flag && *abort_flag) return;
    const float kf = (float)world;
    const long long nvec = (hi - lo) / 4;
    constexpr int U = 4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += U * stride) {
        float4 acc[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * stride;
            acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i < nvec)
                asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(acc[u].x), "=f"(acc[u].y), "=f"(acc[u].z), "=f"(acc[u].w)
                             : "l"(mc + lo + 4 * i) : "memory");
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * stride;
            if (i >= nvec) continue;
            acc[u].x = __fdiv_rn(acc[u].x, kf); acc[u].y = __fdiv_rn(acc[u].y, kf);
            acc[u].z = __fdiv_rn(acc[u].z, kf); acc[u].w = __fdiv_rn(acc[u].w, kf);
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                         ::"l"(mc + lo + 4 * i), "f"(acc[u].x), "f"(acc[u].y), "f"(acc[u].z), "f"(acc[u].w) : "memory");
        }
    }
    __threadfence_system();
}

// Cross-rank barrier over peer memory (one tiny block): thread t publishes `epoch` into peer t's flag slot [rank]
// (NVLink store, release at system scope) and spins on the local slot [t] until peer t has published the same epoch.
// Stream order on each rank + this barrier = "every rank's preceding kernels are complete and visible".  The spin is
// bounded (~2 s of SM clock): on expiry *timeout is set and the kernel returns instead of hanging the GPU.
struct PeerFlags {
    unsigned* f[FEDAVG_MAX_K];
};

static long long g_barrier_spin_clocks = 4000000000LL;  // ~2 s of SM clock (b200asr_fedavg_peer_barrier_set_timeout)

__global__ void fedavg_peer_barrier_kernel(const PeerFlags s, int world, int rank, unsigned epoch, int* timeout,
                                           long long spin_clocks) {
    const int t = threadIdx.x;
    if (t >= world) return;
    __threadfence_system();
    // a rank whose earlier barrier gave up publishes POISON instead of the epoch: every peer that meets it marks its own
    // timeout flag as well, so a torn round is reported by PeerFedAvg.check() on ALL ranks, not only on the straggler's
    constexpr unsigned POISON = 0xffffffffu;
    const bool poisoned = timeout && *reinterpret_cast<volatile int*>(timeout) != 0;
    unsigned* remote = s.f[t] + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(remote), "r"(poisoned ? POISON : epoch) : "memory");
    const unsigned* local = s.f[rank] + t;
    const long long t0 = clock64();
    for (;;) {
        unsigned v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(local) : "memory");
        if (v == POISON) {
            if (timeout) *timeout = 1;
            break;
        }
        if ((int)(v - epoch) >= 0) break;
        if (clock64() - t0 > spin_clocks) {
            if (timeout) *timeout = 1;
            break;
        }
        __nanosleep(200);
    }
    __threadfence_system();
}

// ---------------------------------------------------------------- AdamW (transformers.AdamW, correct_bias=True)
__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                    float* __restrict__ v, __nv_bfloat16* __restrict__ shadow, long long n,
                                                    float lr, float b1, float b2, float eps, float wd, float step_size,
                                                    const float* __restrict__ gnorm_sq, float max_norm,
                                                    const float* __restrict__ hyper) {
    if (hyper) {  // per-step scalars from device memory: the launch can be replayed from a CUDA graph
        lr = hyper[0];
        step_size = hyper[1];
    }
    float clip = 1.f;
    if (gnorm_sq) {
        const float nrm = sqrtf(*gnorm_sq);
        clip = fminf(1.f, max_norm / (nrm + 1e-6f));  // torch.nn.utils.clip_grad_norm_
    }
    const long long nvec = n / 4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        float4 pv = *reinterpret_cast<float4*>(p + 4 * i);
        const float4 gv = *reinterpret_cast<const float4*>(g + 4 * i);
        float4 mv = *reinterpret_cast<float4*>(m + 4 * i);
        float4 vv = *reinterpret_cast<float4*>(v + 4 * i);
        float pa[4] = {pv.x, pv.y, pv.z, pv.w};
        const float ga[4] = {gv.x, gv.y, gv.z, gv.w};
        float ma[4] = {mv.x, mv.y, mv.z, mv.w};
        float va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float gg = ga[j] * clip;
            ma[j] = b1 * ma[j] + (1.f - b1) * gg;
            va[j] = b2 * va[j] + (1.f - b2) * gg * gg;
            pa[j] -= step_size * ma[j] / (sqrtf(va[j]) + eps);
            pa[j] -= lr * wd * pa[j];
        }
        *reinterpret_cast<float4*>(p + 4 * i) = make_float4(pa[0], pa[1], pa[2], pa[3]);
        *reinterpret_cast<float4*>(m + 4 * i) = make_float4(ma[0], ma[1], ma[2], ma[3]);
        *reinterpret_cast<float4*>(v + 4 * i) = make_float4(va[0], va[1], va[2], va[3]);
        if (shadow) {
            uint2 o;
            o.x = bf2_pack(pa[0], pa[1]);
            o.y = bf2_pack(pa[2], pa[3]);
            *reinterpret_cast<uint2*>(shadow + 4 * i) = o;
        }
    }
    for (long long i = nvec * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float gg = g[i] * clip;
        const float mm = b1 * m[i] + (1.f - b1) * gg;
        const float vv = b2 * v[i] + (1.f - b2) * gg * gg;
        float pp = p[i] - step_size * mm / (sqrtf(vv) + eps);
        pp -= lr * wd * pp;
        p[i] = pp; m[i] = mm; v[i] = vv;
        if (shadow) shadow[i] = __float2bfloat16_rn(pp);
    }
}

// Per-parameter hyper-parameters (transformers.AdamW keeps state["step"] per parameter and skips parameters without a
// gradient): the flat buffer is cut into chunks that never straddle a parameter, `chunks[c] = {first vec4, vec4 count,
// parameter index, 0}`; `hyper[2 * param] = {lr, lr * sqrt(1 - b2^t) / (1 - b1^t)}` with a NEGATIVE step size meaning
// "no gradient this step: leave p / m / v / shadow alone" (a block dropped by LayerDrop, an unused head).  One block per
// chunk (grid-stride), so a launch captured in a CUDA graph serves every pattern of idle parameters.
__global__ void __launch_bounds__(256) adamw_chunks_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                           float* __restrict__ m, float* __restrict__ v,
                                                           __nv_bfloat16* __restrict__ shadow,
                                                           const int4* __restrict__ chunks, int nchunks,
                                                           const float* __restrict__ hyper, float b1, float b2, float eps,
                                                           float wd, const float* __restrict__ gnorm_sq, float max_norm) {
    float clip = 1.f;
    if (gnorm_sq) {
        const float nrm = sqrtf(*gnorm_sq);
        clip = fminf(1.f, max_norm / (nrm + 1e-6f));
    }
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const int4 ch = __ldg(chunks + c);
        const float lr = __ldg(hyper + 2 * ch.z), step_size = __ldg(hyper + 2 * ch.z + 1);
        if (step_size < 0.f) continue;
        for (int j = threadIdx.x; j < ch.y; j += blockDim.x) {
            const long long i = (long long)ch.x + j;
            float4 pv = *reinterpret_cast<float4*>(p + 4 * i);
            const float4 gv = *reinterpret_cast<const float4*>(g + 4 * i);
            float4 mv = *reinterpret_cast<float4*>(m + 4 * i);
            float4 vv = *reinterpret_cast<float4*>(v + 4 * i);
            float pa[4] = {pv.x, pv.y, pv.z, pv.w};
            const float ga[4] = {gv.x, gv.y, gv.z, gv.w};
            float ma[4] = {mv.x, mv.y, mv.z, mv.w};
            float va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {   // same arithmetic, in the same order, as adamw_kernel
                const float gg = ga[k] * clip;
                ma[k] = b1 * ma[k] + (1.f - b1) * gg;
                va[k] = b2 * va[k] + (1.f - b2) * gg * gg;
                pa[k] -= step_size * ma[k] / (sqrtf(va[k]) + eps);
                pa[k] -= lr * wd * pa[k];
            }
            *reinterpret_cast<float4*>(p + 4 * i) = make_float4(pa[0], pa[1], pa[2], pa[3]);
            *reinterpret_cast<float4*>(m + 4 * i) = make_float4(ma[0], ma[1], ma[2], ma[3]);
            *reinterpret_cast<float4*>(v + 4 * i) = make_float4(va[0], va[1], va[2], va[3]);
            if (shadow) {
                uint2 o;
                o.x = bf2_pack(pa[0], pa[1]);
                o.y = bf2_pack(pa[2], pa[3]);
                *reinterpret_cast<uint2*>(shadow + 4 * i) = o;
            }
        }
    }
}

// Deterministic: every block parks its partial sum, the last block to finish adds them in block order (no
// floating-point atomics, so the clipping coefficient -- and with it the whole update -- is reproducible bit for bit).
constexpr int SUMSQ_MAX_BLOCKS = 1024;
__device__ float g_sumsq_partial[SUMSQ_MAX_BLOCKS];
__device__ unsigned int g_sumsq_arrived = 0;

__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ x, long long n, float* __restrict__ out,
                                                    float* __restrict__ ws_partial, unsigned int* __restrict__ ws_arrived) {
    // ws_*: caller-owned partials + arrival counter (b200asr_sumsq_ws); the module-global pair serves b200asr_sumsq,
    // which is therefore limited to one launch in flight per device
    float* partial = ws_partial ? ws_partial : g_sumsq_partial;
    unsigned int* arrived = ws_arrived ? ws_arrived : &g_sumsq_arrived;
    __shared__ float red[32];
    __shared__ bool is_last;
    float s = 0.f;
    const long long nvec = n / 4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        const float4 v = *reinterpret_cast<const float4*>(x + 4 * i);
        s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    for (long long i = nvec * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        s += x[i] * x[i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = s;
        __threadfence();
        is_last = (atomicAdd(arrived, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float t = 0.f;
        for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) t += __ldcg(&partial[i]);
        t = block_sum(t, red);
        if (threadIdx.x == 0) {
            out[0] += t;
            *arrived = 0;
        }
    }
}

static int stream_grid(long long n) {
    long long g = (n / 4 + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

// Grid cap of the multi-GPU exchange kernels (b200asr_fedavg_set_max_blocks; 0 = none).  An exchange that runs UNDER other
// work (the layer-wise pipelined FedAvg overlaps the next round's forward) is link bound with a fraction of the SMs; the
// cap leaves the rest to the compute kernels.
static int g_fedavg_max_blocks = 0;
static int exchange_grid(long long n) {
    int g = stream_grid(n);
    if (g_fedavg_max_blocks > 0 && g > g_fedavg_max_blocks) g = g_fedavg_max_blocks;
    return g;
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

int b200asr_fedavg_scale_accum(const float* const* inputs, const float* weights, float* out, int K, long long n,
                               void* stream) {
    B200_REQUIRE(inputs && out, "fedavg: null pointer");
    B200_REQUIRE(K >= 1 && K <= FEDAVG_MAX_K, "fedavg: K=%d out of range [1,%d]", K, FEDAVG_MAX_K);
    if (n <= 0) return B200ASR_OK;
    FedPtrs p;
    for (int j = 0; j < FEDAVG_MAX_K; ++j) {
        p.in[j] = j < K ? inputs[j] : nullptr;
        p.w[j] = (j < K && weights) ? weights[j] : 0.f;
        if (j < K) {
            B200_REQUIRE(inputs[j], "fedavg: null client buffer %d", j);
            B200_REQUIRE(((uintptr_t)inputs[j] & 15) == 0, "fedavg: client buffer %d not 16-byte aligned", j);
        }
    }
    B200_REQUIRE(((uintptr_t)out & 15) == 0, "fedavg: out not 16-byte aligned");
    if (weights) fedavg_kernel<0><<<stream_grid(n), 256, 0, as_stream(stream)>>>(p, out, K, n);
    else fedavg_kernel<1><<<stream_grid(n), 256, 0, as_stream(stream)>>>(p, out, K, n);
    return check_launch("fedavg_scale_accum");
}

int b200asr_fedavg_prescale(float* buf, float w, long long n, void* stream) {
    B200_REQUIRE(buf && ((uintptr_t)buf & 15) == 0, "fedavg_prescale: bad pointer");
    if (n <= 0) return B200ASR_OK;
    prescale_kernel<<<stream_grid(n), 256, 0, as_stream(stream)>>>(buf, w, n);
    return check_launch("fedavg_prescale");
}

static int fedavg_peer_reduce_impl(float* const* peer_bufs, const float* weights, int world, int rank, long long n,
                                   const int* abort_flag, void* stream) {
    B200_REQUIRE(peer_bufs, "fedavg_peer_reduce: null pointer");
    B200_REQUIRE(world >= 1 && world <= FEDAVG_MAX_K && rank >= 0 && rank < world, "fedavg_peer_reduce: bad world/rank");
    if (n <= 0) return B200ASR_OK;
    PeerPtrs p;
    for (int j = 0; j < FEDAVG_MAX_K; ++j) {
        p.buf[j] = j < world ? peer_bufs[j] : nullptr;
        p.w[j] = (j < world && weights) ? weights[j] : 0.f;
        if (j < world) B200_REQUIRE(peer_bufs[j] && ((uintptr_t)peer_bufs[j] & 15) == 0, "fedavg_peer_reduce: bad peer %d", j);
    }
    // slice boundaries on multiples of 4 elements (16 B); the last rank takes the remainder
    const long long per = ((n / world) / 4) * 4;
    const long long lo = per * rank;
    const long long hi = (rank == world - 1) ? n : per * (rank + 1);
    if (hi <= lo) return B200ASR_OK;
    if (weights) fedavg_peer_kernel<0><<<exchange_grid(hi - lo), 256, 0, as_stream(stream)>>>(p, world, lo, hi, abort_flag);
    else fedavg_peer_kernel<1><<<exchange_grid(hi - lo), 256, 0, as_stream(stream)>>>(p, world, lo, hi, abort_flag);
    return check_launch("fedavg_peer_reduce");
}

int b200asr_fedavg_peer_reduce(float* const* peer_bufs, const float* weights, int world, int rank, long long n,
                               void* stream) {
    return fedavg_peer_reduce_impl(peer_bufs, weights, world, rank, n, nullptr, stream);
}

int b200asr_fedavg_peer_reduce_guarded(float* const* peer_bufs, const float* weights, int world, int rank, long long n,
                                       const int* abort_flag, void* stream) {
    return fedavg_peer_reduce_impl(peer_bufs, weights, world, rank, n, abort_flag, stream);
}

int b200asr_fedavg_nvls_reduce(float* multicast_buf, int world, int rank, long long n, const int* abort_flag, int max_blocks,
                               void* stream) {
    B200_REQUIRE(multicast_buf && ((uintptr_t)multicast_buf & 15) == 0, "fedavg_nvls_reduce: bad multicast pointer");
    B200_REQUIRE(world >= 1 && world <= FEDAVG_MAX_K && rank >= 0 && rank < world, "fedavg_nvls_reduce: bad world/rank");
    B200_REQUIRE(n % 4 == 0, "fedavg_nvls_reduce: n must be a multiple of 4 (FlatParams pads every segment)");
    if (n <= 0) return B200ASR_OK;
    const long long per = ((n / world) / 4) * 4;
    const long long lo = per * rank;
    const long long hi = (rank == world - 1) ? n : per * (rank + 1);
    if (hi <= lo) return B200ASR_OK;
    int grid = exchange_grid(hi - lo);
    if (max_blocks > 0 && grid > max_blocks) grid = max_blocks;   // leave SMs to work overlapped with the exchange
    fedavg_nvls_kernel<<<grid, 256, 0, as_stream(stream)>>>(multicast_buf, world, lo, hi, abort_flag);
    return check_launch("fedavg_nvls_reduce");
}

int b200asr_fedavg_peer_barrier(unsigned* const* peer_flags, int world, int rank, unsigned epoch, int* timeout_flag,
                                void* stream) {
    B200_REQUIRE(peer_flags, "fedavg_peer_barrier: null pointer");
    B200_REQUIRE(world >= 1 && world <= FEDAVG_MAX_K && rank >= 0 && rank < world, "fedavg_peer_barrier: bad world/rank");
    PeerFlags s;
    for (int j = 0; j < FEDAVG_MAX_K; ++j) {
        s.f[j] = j < world ? peer_flags[j] : nullptr;
        if (j < world) B200_REQUIRE(peer_flags[j], "fedavg_peer_barrier: null flag pointer for rank %d", j);
    }
    fedavg_peer_barrier_kernel<<<1, 32, 0, as_stream(stream)>>>(s, world, rank, epoch, timeout_flag, g_barrier_spin_clocks);
    return check_launch("fedavg_peer_barrier");
}

int b200asr_fedavg_set_max_blocks(int max_blocks) {
    const int old = g_fedavg_max_blocks;
    g_fedavg_max_blocks = max_blocks > 0 ? max_blocks : 0;
    return old;
}

int b200asr_fedavg_peer_barrier_set_timeout(double seconds) {
    B200_REQUIRE(seconds > 0.0 && seconds < 600.0, "fedavg_peer_barrier_set_timeout: seconds out of range");
    g_barrier_spin_clocks = (long long)(seconds * 1.9e9);
    return B200ASR_OK;
}

// ---- peer-visible allocations + CUDA IPC plumbing (one federated client per process / GPU)
int b200asr_peer_alloc(void** ptr, size_t bytes) {
    B200_REQUIRE(ptr && bytes > 0, "peer_alloc: bad argument");
    cudaError_t e = cudaMalloc(ptr, bytes);
    if (e == cudaSuccess) e = cudaMemset(*ptr, 0, bytes);
    if (e != cudaSuccess) {
        set_error("peer_alloc(%zu): %s", bytes, cudaGetErrorString(e));
        (void)cudaGetLastError();
        return B200ASR_ERR_CUDA;
    }
    return B200ASR_OK;
}

int b200asr_peer_free(void* ptr) {
    if (!ptr) return B200ASR_OK;
    cudaError_t e = cudaFree(ptr);
    if (e != cudaSuccess) {
        set_error("peer_free: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return B200ASR_ERR_CUDA;
    }
    return B200ASR_OK;
}

int b200asr_ipc_export(const void* ptr, unsigned char handle[64]) {
    B200_REQUIRE(ptr && handle, "ipc_export: null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void*>(ptr));
    if (e != cudaSuccess) {
        set_error("ipc_export: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return B200ASR_ERR_CUDA;
    }
    memcpy(handle, &h, 64);
    return B200ASR_OK;
}

int b200asr_ipc_open(const unsigned char handle[64], void** ptr) {
    B200_REQUIRE(ptr && handle, "ipc_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    cudaError_t e = cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        set_error("ipc_open: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return B200ASR_ERR_CUDA;
    }
    return B200ASR_OK;
}

int b200asr_ipc_close(void* ptr) {
    if (!ptr) return B200ASR_OK;
    cudaError_t e = cudaIpcCloseMemHandle(ptr);
    if (e != cudaSuccess) {
        set_error("ipc_close: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return B200ASR_ERR_CUDA;
    }
    return B200ASR_OK;
}

int b200asr_adamw_step(float* p, const float* grad, float* m, float* v, void* shadow_bf16, long long n, float lr, float beta1,
                       float beta2, float eps, float weight_decay, int step, const float* gnorm_sq, float max_norm,
                       void* stream) {
    B200_REQUIRE(p && grad && m && v, "adamw: null pointer");
    B200_REQUIRE(step >= 1, "adamw: step must be >= 1");
    B200_REQUIRE((((uintptr_t)p | (uintptr_t)grad | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "adamw: buffers must be 16-byte aligned");
    B200_REQUIRE(!shadow_bf16 || ((uintptr_t)shadow_bf16 & 7) == 0, "adamw: shadow must be 8-byte aligned");
    if (n <= 0) return B200ASR_OK;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    const float step_size = (float)((double)lr * sqrt(bc2) / bc1);
    adamw_kernel<<<stream_grid(n), 256, 0, as_stream(stream)>>>(p, grad, m, v, (__nv_bfloat16*)shadow_bf16, n, lr, beta1, beta2,
                                                                eps, weight_decay, step_size, gnorm_sq, max_norm, nullptr);
    return check_launch("adamw_step");
}

int b200asr_adamw_step_dev(float* p, const float* grad, float* m, float* v, void* shadow_bf16, long long n, const float* hyper,
                           float beta1, float beta2, float eps, float weight_decay, const float* gnorm_sq, float max_norm,
                           void* stream) {
    B200_REQUIRE(p && grad && m && v && hyper, "adamw_dev: null pointer");
    B200_REQUIRE((((uintptr_t)p | (uintptr_t)grad | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "adamw_dev: buffers must be 16-byte aligned");
    B200_REQUIRE(!shadow_bf16 || ((uintptr_t)shadow_bf16 & 7) == 0, "adamw_dev: shadow must be 8-byte aligned");
    if (n <= 0) return B200ASR_OK;
    adamw_kernel<<<stream_grid(n), 256, 0, as_stream(stream)>>>(p, grad, m, v, (__nv_bfloat16*)shadow_bf16, n, 0.f, beta1, beta2,
                                                                eps, weight_decay, 0.f, gnorm_sq, max_norm, hyper);
    return check_launch("adamw_step_dev");
}

static int sumsq_grid(long long n) {
    long long g = (n / 4 + 255) / 256;
    long long cap = (long long)num_sms() * 4;
    if (cap > SUMSQ_MAX_BLOCKS) cap = SUMSQ_MAX_BLOCKS;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

int b200asr_sumsq(const float* x, long long n, float* out, void* stream) {
    B200_REQUIRE(x && out && ((uintptr_t)x & 15) == 0, "sumsq: bad pointer");
    if (n <= 0) return B200ASR_OK;
    sumsq_kernel<<<sumsq_grid(n), 256, 0, as_stream(stream)>>>(x, n, out, nullptr, nullptr);
    return check_launch("sumsq");
}

size_t b200asr_sumsq_workspace_bytes(void) { return (size_t)(SUMSQ_MAX_BLOCKS + 4) * sizeof(float); }

int b200asr_sumsq_ws(const float* x, long long n, float* out, void* workspace, size_t workspace_bytes, void* stream) {
    B200_REQUIRE(x && out && workspace && ((uintptr_t)x & 15) == 0, "sumsq_ws: bad pointer");
    B200_REQUIRE(workspace_bytes >= b200asr_sumsq_workspace_bytes(), "sumsq_ws: workspace too small");
    if (n <= 0) return B200ASR_OK;
    float* partial = (float*)workspace;
    unsigned int* arrived = (unsigned int*)(partial + SUMSQ_MAX_BLOCKS);   // must be zero before the first use
    sumsq_kernel<<<sumsq_grid(n), 256, 0, as_stream(stream)>>>(x, n, out, partial, arrived);
    return check_launch("sumsq_ws");
}

int b200asr_adamw_step_chunks(float* p, const float* grad, float* m, float* v, void* shadow_bf16, const int* chunks,
                              int nchunks, const float* hyper, float beta1, float beta2, float eps, float weight_decay,
                              const float* gnorm_sq, float max_norm, void* stream) {
    B200_REQUIRE(p && grad && m && v && chunks && hyper, "adamw_chunks: null pointer");
    B200_REQUIRE((((uintptr_t)p | (uintptr_t)grad | (uintptr_t)m | (uintptr_t)v | (uintptr_t)chunks) & 15) == 0,
                 "adamw_chunks: buffers must be 16-byte aligned");
    B200_REQUIRE(!shadow_bf16 || ((uintptr_t)shadow_bf16 & 7) == 0, "adamw_chunks: shadow must be 8-byte aligned");
    if (nchunks <= 0) return B200ASR_OK;
    int grid = nchunks;
    const int cap = num_sms() * 8;
    if (grid > cap) grid = cap;
    adamw_chunks_kernel<<<grid, 256, 0, as_stream(stream)>>>(p, grad, m, v, (__nv_bfloat16*)shadow_bf16, (const int4*)chunks,
                                                             nchunks, hyper, beta1, beta2, eps, weight_decay, gnorm_sq, max_norm);
    return check_launch("adamw_step_chunks");
}

}  // extern "C"
