// LayerNorm (+residual, +GELU) forward/backward, column sums.  HBM-bound: one warp per row, 128-bit
// loads, values held in registers between the statistics pass and the normalise pass (one read, one write).
#include <mutex>

#include "common.cuh"

namespace b200asr {

constexpr int LN_BWD_BLOCKS = 148;  // one 512-thread block per SM; fixed so that the workspace size is device independent

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// (was 296)   // 2 x 148 SMs; fixed so that the workspace size is device independent
constexpr int COLSUM_CHUNKS = 64;

// per-lane helpers: N consecutive fp32 affine parameters starting at column c (c is a multiple of N >= 4)
template <int N> __device__ __forceinline__ void load_params(const float* __restrict__ p, int c, float* out) {
#pragma unroll
    for (int j = 0; j < N; j += 4) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(p + c + j));
        out[j] = v.x; out[j + 1] = v.y; out[j + 2] = v.z; out[j + 3] = v.w;
    }
}

// DROP: x is first passed through an inverted dropout (mask from DropKeys, see common.cuh) -- y = LN(res + drop(x)),
// sum_out = res + drop(x): the hidden_dropout + residual add + LayerNorm of a post-LN encoder block in one pass.
template <typename T, int NV, bool DROP>
__global__ void __launch_bounds__(128) ln_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res,
                                                     const float* __restrict__ w, const float* __restrict__ b,
                                                     T* __restrict__ y, T* __restrict__ sum_out,
                                                     float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                     int rows, int cols, float eps, int act, DropKeys dk) {
    constexpr int N = Vec16<T>::N;
    pdl_enter();
    unsigned long long drop_offset = 0;
    if (DROP) drop_offset = dk.offset + (dk.dev_offset ? *dk.dev_offset : 0ull);
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const int nvec = cols / N;
    const float inv_cols = 1.0f / (float)cols;
    for (int row = warp; row < rows; row += nwarps) {
        const long long base = (long long)row * cols;
        float v[NV][N];
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vi = lane + 32 * i;
            if (vi < nvec) {
                Vec16<T> a;
                a.load(x + base + (long long)vi * N);
                a.unpack(v[i]);
                if (DROP) {
                    uint32_t u[N];
                    dropout_uniforms<N>((unsigned long long)(base + (long long)vi * N), dk.key, drop_offset, u);
#pragma unroll
                    for (int j = 0; j < N; ++j) v[i][j] = (u[j] >= dk.thresh16) ? v[i][j] * dk.scale : 0.f;
                }
                if (res) {
                    Vec16<T> r;
                    float rf[N];
                    r.load(res + base + (long long)vi * N);
                    r.unpack(rf);
#pragma unroll
                    for (int j = 0; j < N; ++j) v[i][j] += rf[j];
                }
                if (sum_out) {
                    Vec16<T> s;
                    s.pack(v[i]);
                    s.store(sum_out + base + (long long)vi * N);
                    s.unpack(v[i]);  // statistics from the stored (rounded) value: backward sees the same numbers
                }
#pragma unroll
                for (int j = 0; j < N; ++j) sum += v[i][j];
            } else {
#pragma unroll
                for (int j = 0; j < N; ++j) v[i][j] = 0.f;
            }
        }
        const float mean = warp_sum(sum) * inv_cols;
        float sq = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vi = lane + 32 * i;
            if (vi < nvec) {
#pragma unroll
                for (int j = 0; j < N; ++j) {
                    const float d = v[i][j] - mean;
                    sq += d * d;
                }
            }
        }
        const float rstd = rsqrtf(warp_sum(sq) * inv_cols + eps);
        if (lane == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vi = lane + 32 * i;
            if (vi < nvec) {
                float o[N], wv[N], bv[N];
                if (w) {
                    load_params<N>(w, vi * N, wv);
                    load_params<N>(b, vi * N, bv);
                }
#pragma unroll
                for (int j = 0; j < N; ++j) {
                    float t = (v[i][j] - mean) * rstd;
                    if (w) t = fmaf(t, wv[j], bv[j]);
                    if (act == B200ASR_ACT_GELU) t = gelu_for<T>(t);
                    o[j] = t;
                }
                Vec16<T> ov;
                ov.pack(o);
                ov.store(y + base + (long long)vi * N);
            }
        }
    }
}

// LayerNorm backward.  A GROUP of GS threads owns one row at a time (one 16-byte vector per thread, so the three
// column accumulators -- dw += dz*xhat, db += dz and (CS) the column sums of dx, i.e. the bias gradient of the
// nn.Linear that produced the LayerNorm input -- cost 3 x N registers); a block runs G groups on different rows and the
// next row of a group is prefetched while the current one is reduced.  The row statistics cross the warps of a group
// through shared memory + a named barrier.  Groups add their accumulators into the block's shared sums in a fixed
// order, blocks write partials, reduce_partials_kernel finishes: deterministic, no floating-point atomics.
static __device__ unsigned long long* g_ln_trace_dev = nullptr;   // diagnostics (b200asr_ln_set_trace): 8 stamps per block
#define LTRACE(slot) do { if (g_ln_trace_dev && threadIdx.x == 0) g_ln_trace_dev[blockIdx.x * 8 + (slot)] = (unsigned long long)clock64(); } while (0)

// DR: a second output dx_drop = dropout(dx) with the keys `dk` (the hidden dropout in front of this LayerNorm's residual
// sum, HF:310-312 / 344-345: dx continues along the residual path, dx_drop is the gradient of the dropped Linear output),
// bit-identical to a b200asr_dropout pass over the stored dx; the CS column sums are then those of dx_drop (= that
// Linear's bias gradient).
template <typename T, bool CS, bool DR = false>
__global__ void __launch_bounds__(512, 1) ln_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ s,
                                                        const float* __restrict__ w, const float* __restrict__ b,
                                                        const float* __restrict__ mean_in,
                                                        const float* __restrict__ rstd_in,
                                                        const T* __restrict__ dres_add, T* __restrict__ dx,
                                                        float* __restrict__ partial, int rows, int cols, int act, int GS,
                                                        int G, T* __restrict__ dx_drop = nullptr, DropKeys dk = DropKeys{}) {
    constexpr int N = Vec16<T>::N;
    [[maybe_unused]] unsigned long long drop_off = 0;
    LTRACE(0);
    pdl_enter();
    LTRACE(1);
    if constexpr (DR) drop_off = dk.offset + (dk.dev_offset ? *dk.dev_offset : 0ull);
    extern __shared__ float sm[];
    float* xch = sm;                    // [2][G][16][2] per-warp row statistics, double buffered by row parity
    float* sacc = sm + 2 * 8 * 16 * 2;  // [G][3][N][GS] accumulators of the groups (end phase)
    const int g = threadIdx.x / GS, tg = threadIdx.x - g * GS;
    const int lane = threadIdx.x & 31, wg = tg >> 5, nwg = GS >> 5;
    const int c0 = tg * N;
    const bool active = c0 < cols;
    const float inv_cols = 1.0f / (float)cols;
    float aw[N], ab[N], ac[N], wv[N], bv[N];
#pragma unroll
    for (int j = 0; j < N; ++j) { aw[j] = ab[j] = ac[j] = 0.f; wv[j] = 1.f; bv[j] = 0.f; }
    if (w && active) {
        load_params<N>(w, c0, wv);
        load_params<N>(b, c0, bv);
    }
    const int ngroups = gridDim.x * G;
    int row = blockIdx.x * G + g;
    // PF rows of this group in flight (operands + row statistics): with one block per SM the HBM pipe needs ~45 KB
    // outstanding per SM; one row ahead (16 KB per block) left the kernel latency-bound at ~20% of the HBM peak.
    constexpr int PF = 3;
    Vec16<T> a_q[PF], g_q[PF];
    float m_q[PF], r_q[PF];
#pragma unroll
    for (int d = 0; d < PF; ++d) {
        a_q[d].raw = make_uint4(0, 0, 0, 0);
        g_q[d].raw = make_uint4(0, 0, 0, 0);
        m_q[d] = 0.f; r_q[d] = 0.f;
        const int r = row + d * ngroups;
        if (r < rows) {
            m_q[d] = __ldg(mean_in + r);
            r_q[d] = __ldg(rstd_in + r);
            if (active) {
                a_q[d].load(s + (long long)r * cols + c0);
                g_q[d].load(dy + (long long)r * cols + c0);
            }
        }
    }
    LTRACE(2);
    int parity = 0, nrows_done = 0;
    for (; row < rows; row += ngroups, parity ^= 1) {
        const long long base = (long long)row * cols + c0;
        Vec16<T> a = a_q[0], gg = g_q[0];
        const float mean = m_q[0], rstd = r_q[0];
#pragma unroll
        for (int d = 0; d + 1 < PF; ++d) { a_q[d] = a_q[d + 1]; g_q[d] = g_q[d + 1]; m_q[d] = m_q[d + 1]; r_q[d] = r_q[d + 1]; }
        const int nrow = row + PF * ngroups;
        if (nrow < rows) {
            m_q[PF - 1] = __ldg(mean_in + nrow);
            r_q[PF - 1] = __ldg(rstd_in + nrow);
            if (active) {
                a_q[PF - 1].load(s + (long long)nrow * cols + c0);
                g_q[PF - 1].load(dy + (long long)nrow * cols + c0);
            }
        }
        float xh[N], dz[N];
        a.unpack(xh);
        gg.unpack(dz);
        float s1 = 0.f, s2 = 0.f;
        if (active) {
#pragma unroll
            for (int j = 0; j < N; ++j) {
                const float h = (xh[j] - mean) * rstd;
                float d = dz[j];
                if (act == B200ASR_ACT_GELU) d *= gelu_grad_for<T>(w ? fmaf(h, wv[j], bv[j]) : h);
                aw[j] = fmaf(d, h, aw[j]);
                ab[j] += d;
                d *= wv[j];
                xh[j] = h;
                dz[j] = d;
                s1 += d;
                s2 = fmaf(d, h, s2);
            }
        }
        if (nrows_done++ == 0) LTRACE(3);
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        if (nwg > 1) {
            float* x2 = xch + ((parity * G + g) * 16) * 2;
            if (lane == 0) { x2[2 * wg] = s1; x2[2 * wg + 1] = s2; }
            named_bar_sync(1 + g, GS);
            s1 = 0.f; s2 = 0.f;
            for (int i = 0; i < nwg; ++i) { s1 += x2[2 * i]; s2 += x2[2 * i + 1]; }
        }
        s1 *= inv_cols;
        s2 *= inv_cols;
        if (active) {
            float o[N];
#pragma unroll
            for (int j = 0; j < N; ++j) o[j] = rstd * (dz[j] - s1 - xh[j] * s2);
            if (dres_add) {
                Vec16<T> r;
                float rf[N];
                r.load(dres_add + base);
                r.unpack(rf);
#pragma unroll
                for (int j = 0; j < N; ++j) o[j] += rf[j];
            }
            Vec16<T> ov;
            ov.pack(o);
            ov.store(dx + base);
            if constexpr (DR) {
                uint32_t u[N];
                dropout_uniforms<N>((unsigned long long)base, dk.key, drop_off, u);
                ov.unpack(o);   // the stored (rounded) values, as a separate dropout pass would read them
#pragma unroll
                for (int j = 0; j < N; ++j) o[j] = (u[j] >= dk.thresh16) ? o[j] * dk.scale : 0.f;
                ov.pack(o);
                ov.store(dx_drop + base);
            }
            if (CS) {
                ov.unpack(o);  // sum what was stored (the rounded values the consumer GEMMs read)
#pragma unroll
                for (int j = 0; j < N; ++j) ac[j] += o[j];
            }
        }
    }
    LTRACE(4);
    if (partial) {
        // every group parks its accumulators in its own region, element (k, column c0 + j) at [((g * 3 + k) * N + j) * GS + tg]:
        // consecutive threads hit consecutive banks (the first version added group after group into [3][cols] with an
        // 8-way bank conflict on every access and G block barriers: 7.2K of the block's 23K clk, tools/ln_trace.py)
        float* mine = sacc + (long long)g * 3 * N * GS;
#pragma unroll
        for (int j = 0; j < N; ++j) {
            mine[(0 * N + j) * GS + tg] = aw[j];
            mine[(1 * N + j) * GS + tg] = ab[j];
            mine[(2 * N + j) * GS + tg] = CS ? ac[j] : 0.f;
        }
        __syncthreads();
        LTRACE(5);
        float* out = partial + (long long)blockIdx.x * 3 * cols;
        // walk the staging layout itself (consecutive lanes = consecutive banks); the 32-byte-strided global stores of the
        // 12 KB partial are cheaper than an 8-way conflict on every shared read
        for (int kj = g; kj < 3 * N; kj += G) {   // thread (g, tg) takes rows g, g + G, ... of the [3 N][GS] staging layout
            const int i = kj * GS + tg;
            const int k = kj / N, j = kj - k * N;
            const int c = tg * N + j;
            if (c >= cols) continue;
            float v = 0.f;
            for (int gg = 0; gg < G; ++gg) v += sacc[gg * 3 * N * GS + i];   // fixed order: deterministic
            out[k * cols + c] = v;
        }
    }
    LTRACE(6);
}

// out[c] = sum_{p < nparts} partial[p * stride + c] for c < ncols_total; columns [0, cols) go to out0, [cols, 2 cols)
// to out1, [2 cols, 3 cols) to out2.  One block = 32 columns x 32 warps; warp w sums partials w, w+32, ... (coalesced 128-byte
// rows, all loads of a thread in flight together), then a fixed-order shared-memory combine: deterministic.
__global__ void __launch_bounds__(1024) reduce_partials_kernel(const float* __restrict__ partial, float* __restrict__ out0,
                                                               float* __restrict__ out1, float* __restrict__ out2,
                                                               int nparts, int cols, int ncols_total, int stride) {
    __shared__ float red[32][33];
    pdl_enter();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    // 32 warps: warp w owns partials w, w+32, ...; at the 148 partials of this library that is one batch of <= 5 independent
    // loads (one L2 round trip) where 8 warps needed five dependent batches -- this kernel is pure latency
    float a[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (c < ncols_total) {
        for (int p0 = w; p0 < nparts; p0 += 160) {
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const int p = p0 + 32 * k;
                if (p < nparts) a[k] += partial[(long long)p * stride + c];
            }
        }
    }
    red[w][lane] = ((a[0] + a[1]) + (a[2] + a[3])) + a[4];
    __syncthreads();
    if (w == 0 && c < ncols_total) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) t += red[i][lane];
        if (c < cols) { if (out0) out0[c] = t; }
        else if (c < 2 * cols) { if (out1) out1[c - cols] = t; }
        else if (out2) out2[c - 2 * cols] = t;
    }
}

template <typename T>
__global__ void colsum_partial_kernel(const T* __restrict__ x, float* __restrict__ partial, int rows, int cols,
                                      int ld, int rows_per_chunk) {
    // thread = one column; blockIdx.y = row chunk.  Warps read 32 consecutive columns: coalesced.
    pdl_enter();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cols) return;
    const int r0 = blockIdx.y * rows_per_chunk;
    const int r1 = min(rows, r0 + rows_per_chunk);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    int r = r0;
    for (; r + 3 < r1; r += 4) {
        a0 += to_f<T>(x[(long long)r * ld + c]);
        a1 += to_f<T>(x[(long long)(r + 1) * ld + c]);
        a2 += to_f<T>(x[(long long)(r + 2) * ld + c]);
        a3 += to_f<T>(x[(long long)(r + 3) * ld + c]);
    }
    for (; r < r1; ++r) a0 += to_f<T>(x[(long long)r * ld + c]);
    partial[(long long)blockIdx.y * cols + c] = (a0 + a1) + (a2 + a3);
}

// vectorised variant: a block = 32 column vectors (16 B each) x 8 row lanes; needs cols % N == 0, ld % N == 0 and a
// 16-byte aligned base.  partial[chunk][cols].
// DROP: x first goes through an inverted dropout (DropKeys, common.cuh; needs ld == cols so that the element index is
// r * cols + c); the dropped values are written to y and it is their (rounded) column sums that are accumulated --
// `do = dropout(ds); db = colsum(do)` of a block's backward in one pass over ds.
template <typename T, bool DROP>
__device__ __forceinline__ void colsum_load(const T* __restrict__ x, T* __restrict__ y, long long e0, float* f,
                                            const DropKeys& dk, unsigned long long drop_offset) {
    constexpr int N = Vec16<T>::N;
    Vec16<T> a;
    a.load(x + e0);
    a.unpack(f);
    if (DROP) {
        uint32_t u[N];
        dropout_uniforms<N>((unsigned long long)e0, dk.key, drop_offset, u);
#pragma unroll
        for (int j = 0; j < N; ++j) f[j] = (u[j] >= dk.thresh16) ? f[j] * dk.scale : 0.f;
        a.pack(f);
        a.store(y + e0);
        a.unpack(f);  // sum what was stored
    }
}

template <typename T, bool DROP>
__global__ void __launch_bounds__(256) colsum_partial_vec_kernel(const T* __restrict__ x, float* __restrict__ partial,
                                                                 int rows, int cols, int ld, int rows_per_chunk,
                                                                 T* __restrict__ y, DropKeys dk) {
    constexpr int N = Vec16<T>::N;
    __shared__ float red[8][32][N + 1];
    pdl_enter();
    unsigned long long drop_offset = 0;
    if (DROP) drop_offset = dk.offset + (dk.dev_offset ? *dk.dev_offset : 0ull);
    const int cv = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int c = (blockIdx.x * 32 + cv) * N;
    const int r0 = blockIdx.y * rows_per_chunk;
    const int r1 = min(rows, r0 + rows_per_chunk);
    float acc[N], acc2[N];
#pragma unroll
    for (int j = 0; j < N; ++j) acc[j] = acc2[j] = 0.f;
    if (c < cols) {
        int r = r0 + rl;
        for (; r + 8 < r1; r += 16) {
            float af[N], bf[N];
            colsum_load<T, DROP>(x, y, (long long)r * ld + c, af, dk, drop_offset);
            colsum_load<T, DROP>(x, y, (long long)(r + 8) * ld + c, bf, dk, drop_offset);
#pragma unroll
            for (int j = 0; j < N; ++j) { acc[j] += af[j]; acc2[j] += bf[j]; }
        }
        for (; r < r1; r += 8) {
            float af[N];
            colsum_load<T, DROP>(x, y, (long long)r * ld + c, af, dk, drop_offset);
#pragma unroll
            for (int j = 0; j < N; ++j) acc[j] += af[j];
        }
    }
#pragma unroll
    for (int j = 0; j < N; ++j) red[rl][cv][j] = acc[j] + acc2[j];
    __syncthreads();
    if (rl == 0 && c < cols) {
#pragma unroll
        for (int j = 0; j < N; ++j) {
            float t = 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) t += red[i][cv][j];
            partial[(long long)blockIdx.y * cols + c + j] = t;
        }
    }
}

template <typename T>
static int launch_ln_fwd(const void* x, const void* res, const float* w, const float* b, void* y, void* sum_out,
                         float* mean, float* rstd, int rows, int cols, float eps, int act, cudaStream_t st,
                         const DropKeys* drop = nullptr) {
    constexpr int N = Vec16<T>::N;
    const int nvec = cols / N;
    const int nv = ceil_div(nvec, 32);
    const int warps_per_block = 4;
    int grid = ceil_div(rows, warps_per_block);
    const int cap = num_sms() * 16;
    if (grid > cap) grid = cap;
    const DropKeys dk = drop ? *drop : DropKeys{};
#define LN_FWD_CASE(NVV)                                                                                              \
    do {                                                                                                              \
        if (drop)                                                                                                     \
            launch_pdl(ln_fwd_kernel<T, NVV, true>, dim3(grid), dim3(128), 0, st, (const T*)x, (const T*)res, w, b,   \
                       (T*)y, (T*)sum_out, mean, rstd, rows, cols, eps, act, dk);                                     \
        else                                                                                                          \
            launch_pdl(ln_fwd_kernel<T, NVV, false>, dim3(grid), dim3(128), 0, st, (const T*)x, (const T*)res, w, b,  \
                       (T*)y, (T*)sum_out, mean, rstd, rows, cols, eps, act, dk);                                     \
    } while (0)
    if (nv <= 1) LN_FWD_CASE(1);
    else if (nv <= 2) LN_FWD_CASE(2);
    else if (nv <= 4) LN_FWD_CASE(4);
    else if (nv <= 8) LN_FWD_CASE(8);
    else if (nv <= 16) LN_FWD_CASE(16);
    else { set_error("layernorm: cols=%d too large", cols); return B200ASR_ERR_UNSUPPORTED; }
#undef LN_FWD_CASE
    return check_launch("layernorm_fwd");
}

template <typename T>
static int launch_ln_bwd(const void* dy, const void* s, const float* w, const float* b, const float* mean,
                         const float* rstd, const void* dres, void* dx, float* partial, bool colsum, int rows, int cols,
                         int act, cudaStream_t st, void* dx_drop = nullptr, const DropKeys* drop = nullptr) {
    constexpr int N = Vec16<T>::N;
    const int GS = ceil_div(ceil_div(cols, N), 32) * 32;  // threads per row
    if (GS > 512) { set_error("layernorm_bwd: cols=%d too large", cols); return B200ASR_ERR_UNSUPPORTED; }
    int G = 512 / GS;
    if (G > 8) G = 8;
    int grid = ceil_div(rows, G);
    if (grid > LN_BWD_BLOCKS) grid = LN_BWD_BLOCKS;
    if (partial) grid = LN_BWD_BLOCKS;  // finalize sums exactly LN_BWD_BLOCKS partials
    const size_t smem = ((size_t)G * 3 * N * GS + 2 * 8 * 16 * 2) * sizeof(float);   // 50 KB at 512 threads (bf16)
    static bool attr_set = false;   // per instantiation
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(ln_bwd_kernel<T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(ln_bwd_kernel<T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(ln_bwd_kernel<T, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(ln_bwd_kernel<T, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        if (e != cudaSuccess) {
            set_error("layernorm_bwd: cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
            return B200ASR_ERR_CUDA;
        }
        attr_set = true;
    }
    if (drop && dx_drop) {
        if (colsum)
            launch_pdl(ln_bwd_kernel<T, true, true>, dim3(grid), dim3(G * GS), smem, st, (const T*)dy, (const T*)s, w, b, mean,
                       rstd, (const T*)dres, (T*)dx, partial, rows, cols, act, GS, G, (T*)dx_drop, *drop);
        else
            launch_pdl(ln_bwd_kernel<T, false, true>, dim3(grid), dim3(G * GS), smem, st, (const T*)dy, (const T*)s, w, b, mean,
                       rstd, (const T*)dres, (T*)dx, partial, rows, cols, act, GS, G, (T*)dx_drop, *drop);
    } else if (colsum)
        launch_pdl(ln_bwd_kernel<T, true>, dim3(grid), dim3(G * GS), smem, st, (const T*)dy, (const T*)s, w, b, mean, rstd,
                   (const T*)dres, (T*)dx, partial, rows, cols, act, GS, G, (T*)nullptr, DropKeys{});
    else
        launch_pdl(ln_bwd_kernel<T, false>, dim3(grid), dim3(G * GS), smem, st, (const T*)dy, (const T*)s, w, b, mean, rstd,
                   (const T*)dres, (T*)dx, partial, rows, cols, act, GS, G, (T*)nullptr, DropKeys{});
    return check_launch("layernorm_bwd");
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

int b200asr_layernorm_fwd(const void* x, const void* res, const float* w, const float* b, void* y, void* sum_out,
                          float* mean, float* rstd, int rows, int cols, float eps, int act, int dtype,
                          void* stream) {
    if (rows <= 0) return B200ASR_OK;
    B200_REQUIRE(x && y, "layernorm_fwd: null pointer");
    B200_REQUIRE((w == nullptr) == (b == nullptr), "layernorm_fwd: w and b must both be given or both NULL");
    B200_REQUIRE(cols > 0 && cols % (dtype == B200ASR_BF16 ? 8 : 4) == 0,
                 "layernorm_fwd: cols=%d must be a multiple of the 16-byte vector width", cols);
    B200_REQUIRE(!w || (((uintptr_t)w | (uintptr_t)b) & 15) == 0, "layernorm_fwd: w/b must be 16-byte aligned");
    B200_DISPATCH_DTYPE(dtype, {
        return launch_ln_fwd<T>(x, res, w, b, y, sum_out, mean, rstd, rows, cols, eps, act, as_stream(stream));
    });
    return B200ASR_OK;
}

int b200asr_layernorm_dropout_fwd(const void* x, const void* res, const float* w, const float* b, void* y, void* sum_out,
                                  float* mean, float* rstd, int rows, int cols, float eps, int act, float p,
                                  unsigned long long seed, unsigned long long site, unsigned long long offset,
                                  const unsigned long long* dev_offset, int dtype, void* stream) {
    if (rows <= 0) return B200ASR_OK;
    B200_REQUIRE(x && y, "layernorm_dropout_fwd: null pointer");
    B200_REQUIRE((w == nullptr) == (b == nullptr), "layernorm_dropout_fwd: w and b must both be given or both NULL");
    B200_REQUIRE(cols > 0 && cols % 8 == 0, "layernorm_dropout_fwd: cols=%d must be a multiple of 8", cols);
    B200_REQUIRE(!w || (((uintptr_t)w | (uintptr_t)b) & 15) == 0, "layernorm_dropout_fwd: w/b must be 16-byte aligned");
    B200_REQUIRE(p >= 0.f && p < 1.f, "layernorm_dropout_fwd: p=%f must be in [0, 1)", (double)p);
    const DropKeys dk = make_drop_keys(p, seed, site, offset, dev_offset);
    B200_DISPATCH_DTYPE(dtype, {
        return launch_ln_fwd<T>(x, res, w, b, y, sum_out, mean, rstd, rows, cols, eps, act, as_stream(stream), &dk);
    });
    return B200ASR_OK;
}

size_t b200asr_layernorm_bwd_workspace_bytes(int cols) { return (size_t)LN_BWD_BLOCKS * 3 * cols * sizeof(float); }

static int layernorm_bwd_impl(const void* dy, const void* s, const float* w, const float* b, const float* mean,
                              const float* rstd, const void* dres_add, void* dx, float* dw, float* db, float* dx_colsum,
                              void* workspace, size_t workspace_bytes, int rows, int cols, int act, int dtype,
                              void* stream, void* dx_drop, const DropKeys* drop);

int b200asr_layernorm_bwd(const void* dy, const void* s, const float* w, const float* b, const float* mean,
                          const float* rstd, const void* dres_add, void* dx, float* dw, float* db, float* dx_colsum,
                          void* workspace, size_t workspace_bytes, int rows, int cols, int act, int dtype,
                          void* stream) {
    return layernorm_bwd_impl(dy, s, w, b, mean, rstd, dres_add, dx, dw, db, dx_colsum, workspace, workspace_bytes, rows, cols,
                              act, dtype, stream, nullptr, nullptr);
}

int b200asr_layernorm_bwd_dropout(const void* dy, const void* s, const float* w, const float* b, const float* mean,
                                  const float* rstd, const void* dres_add, void* dx, void* dx_drop, float* dw, float* db,
                                  float* dx_drop_colsum, void* workspace, size_t workspace_bytes, int rows, int cols, int act,
                                  float p, unsigned long long seed, unsigned long long site, unsigned long long offset,
                                  const unsigned long long* dev_offset, int dtype, void* stream) {
    B200_REQUIRE(dx_drop, "layernorm_bwd_dropout: null dx_drop");
    B200_REQUIRE(p >= 0.f && p < 1.f, "layernorm_bwd_dropout: p=%f must be in [0, 1)", (double)p);
    const DropKeys k = make_drop_keys(p, seed, site, offset, dev_offset);
    return layernorm_bwd_impl(dy, s, w, b, mean, rstd, dres_add, dx, dw, db, dx_drop_colsum, workspace, workspace_bytes, rows,
                              cols, act, dtype, stream, dx_drop, &k);
}

static int layernorm_bwd_impl(const void* dy, const void* s, const float* w, const float* b, const float* mean,
                              const float* rstd, const void* dres_add, void* dx, float* dw, float* db, float* dx_colsum,
                              void* workspace, size_t workspace_bytes, int rows, int cols, int act, int dtype,
                              void* stream, void* dx_drop, const DropKeys* drop) {
    if (rows <= 0) return B200ASR_OK;
    B200_REQUIRE(dy && s && mean && rstd && dx, "layernorm_bwd: null pointer");
    B200_REQUIRE(cols > 0 && cols % (dtype == B200ASR_BF16 ? 8 : 4) == 0, "layernorm_bwd: bad cols=%d", cols);
    B200_REQUIRE(!w || (((uintptr_t)w | (uintptr_t)b) & 15) == 0, "layernorm_bwd: w/b must be 16-byte aligned");
    float* partial = nullptr;
    if (dw || db || dx_colsum) {
        B200_REQUIRE(workspace && workspace_bytes >= b200asr_layernorm_bwd_workspace_bytes(cols),
                     "layernorm_bwd: workspace too small");
        B200_REQUIRE((size_t)3 * cols * sizeof(float) <= 40 * 1024, "layernorm_bwd: cols=%d too large", cols);
        partial = (float*)workspace;
    }
    int rc;
    B200_DISPATCH_DTYPE(dtype, {
        rc = launch_ln_bwd<T>(dy, s, w, b, mean, rstd, dres_add, dx, partial, dx_colsum != nullptr, rows, cols, act,
                              as_stream(stream), dx_drop, drop);
    });
    if (rc != B200ASR_OK) return rc;
    if (partial) {
        const int total = dx_colsum ? 3 * cols : 2 * cols;
        launch_pdl(reduce_partials_kernel, dim3(ceil_div(total, 32)), dim3(1024), 0, as_stream(stream), (const float*)partial, dw,
                   db, dx_colsum, (int)LN_BWD_BLOCKS, cols, total, 3 * cols);
        return check_launch("layernorm_bwd_reduce");
    }
    return B200ASR_OK;
}

size_t b200asr_colsum_workspace_bytes(int cols) { return (size_t)COLSUM_CHUNKS * cols * sizeof(float); }

int b200asr_colsum(const void* x, float* out, void* workspace, size_t workspace_bytes, int rows, int cols, int ld,
                   int dtype, void* stream) {
    B200_REQUIRE(x && out && workspace, "colsum: null pointer");
    B200_REQUIRE(rows > 0 && cols > 0 && ld >= cols, "colsum: bad shape");
    B200_REQUIRE(workspace_bytes >= b200asr_colsum_workspace_bytes(cols), "colsum: workspace too small");
    const int nvec = dtype == B200ASR_BF16 ? 8 : 4;
    int nchunks;
    if (cols % nvec == 0 && ld % nvec == 0 && ((uintptr_t)x & 15) == 0) {
        const int gx = ceil_div(cols, 32 * nvec);
        int chunks = (4 * num_sms()) / gx;
        if (chunks > COLSUM_CHUNKS) chunks = COLSUM_CHUNKS;
        if (chunks > ceil_div(rows, 8)) chunks = ceil_div(rows, 8);
        if (chunks < 1) chunks = 1;
        const int rpc = ceil_div(rows, chunks);
        nchunks = ceil_div(rows, rpc);
        dim3 grid(gx, nchunks);
        B200_DISPATCH_DTYPE(dtype, {
            launch_pdl(colsum_partial_vec_kernel<T, false>, grid, dim3(256), 0, as_stream(stream), (const T*)x, (float*)workspace, rows, cols, ld, rpc,
                       (T*)nullptr, DropKeys{});
        });
    } else {
        const int chunks = rows < COLSUM_CHUNKS ? rows : COLSUM_CHUNKS;
        const int rpc = ceil_div(rows, chunks);
        nchunks = ceil_div(rows, rpc);
        dim3 grid(ceil_div(cols, 128), nchunks);
        B200_DISPATCH_DTYPE(dtype, {
            launch_pdl(colsum_partial_kernel<T>, grid, dim3(128), 0, as_stream(stream), (const T*)x, (float*)workspace, rows, cols, ld, rpc);
        });
    }
    int rc = check_launch("colsum_partial");
    if (rc != B200ASR_OK) return rc;
    launch_pdl(reduce_partials_kernel, dim3(ceil_div(cols, 32)), dim3(1024), 0, as_stream(stream), (const float*)workspace, out,
               (float*)nullptr, (float*)nullptr, nchunks, cols, cols, cols);
    return check_launch("colsum_reduce");
}

// ---- b200asr_colsum without a caller-provided workspace and without any allocation --------------------------------------
// The partial sums live in a file-scope device array (COLSUM_CHUNKS x B200ASR_COLSUM_STATIC_MAX_COLS floats = 2 MiB, held
// by every GPU that loads the library until the process exits).  In host code `g_colsum_static_scratch` is only the host
// side's handle of the symbol, not a device address: the address is looked up once per device, with that device current,
// by cudaGetSymbolAddress, and that pointer is what the kernels get.
static __device__ __align__(256) float g_colsum_static_scratch[COLSUM_CHUNKS * B200ASR_COLSUM_STATIC_MAX_COLS];

static int colsum_static_scratch(float** out_ptr) {
    constexpr int MAX_DEV = 64;
    static float* ptrs[MAX_DEV] = {};
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
        set_error("colsum_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return B200ASR_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        float* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_colsum_static_scratch);
        if (e != cudaSuccess || !ptr) {
            set_error("colsum_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return B200ASR_ERR_CUDA;
        }
        ptrs[dev] = ptr;
    }
    *out_ptr = ptrs[dev];
    return B200ASR_OK;
}

int b200asr_colsum_static_max_cols(void) { return B200ASR_COLSUM_STATIC_MAX_COLS; }

int b200asr_colsum_static(const void* x, float* out, int rows, int cols, int ld, int dtype, void* stream) {
    B200_REQUIRE(x && out, "colsum_static: null pointer");
    B200_REQUIRE(rows > 0 && cols > 0 && ld >= cols, "colsum_static: bad shape");
    B200_REQUIRE(cols <= B200ASR_COLSUM_STATIC_MAX_COLS,
                 "colsum_static: cols=%d exceeds the built-in scratch (max %d columns); use b200asr_colsum with a workspace",
                 cols, B200ASR_COLSUM_STATIC_MAX_COLS);
    float* scratch = nullptr;
    const int rc = colsum_static_scratch(&scratch);
    if (rc != B200ASR_OK) return rc;
    // nothing has to start at zero: every partial the reduction reads is written by the partial kernel of this same call
    return b200asr_colsum(x, out, scratch, sizeof(float) * COLSUM_CHUNKS * B200ASR_COLSUM_STATIC_MAX_COLS, rows, cols, ld, dtype,
                          stream);
}

int b200asr_dropout_colsum(const void* x, void* y, float* out, void* workspace, size_t workspace_bytes, int rows, int cols,
                           float p, unsigned long long seed, unsigned long long site, unsigned long long offset,
                           const unsigned long long* dev_offset, int dtype, void* stream) {
    B200_REQUIRE(x && y && out && workspace, "dropout_colsum: null pointer");
    B200_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0, "dropout_colsum: bad shape rows=%d cols=%d (cols %% 8 == 0)", rows, cols);
    B200_REQUIRE((((uintptr_t)x | (uintptr_t)y) & 15) == 0, "dropout_colsum: buffers must be 16-byte aligned");
    B200_REQUIRE(workspace_bytes >= b200asr_colsum_workspace_bytes(cols), "dropout_colsum: workspace too small");
    B200_REQUIRE(p >= 0.f && p < 1.f, "dropout_colsum: p=%f must be in [0, 1)", (double)p);
    const DropKeys dk = make_drop_keys(p, seed, site, offset, dev_offset);
    const int nvec = dtype == B200ASR_BF16 ? 8 : 4;
    const int gx = ceil_div(cols, 32 * nvec);
    int chunks = (4 * num_sms()) / gx;
    if (chunks > COLSUM_CHUNKS) chunks = COLSUM_CHUNKS;
    if (chunks > ceil_div(rows, 8)) chunks = ceil_div(rows, 8);
    if (chunks < 1) chunks = 1;
    const int rpc = ceil_div(rows, chunks);
    const int nchunks = ceil_div(rows, rpc);
    dim3 grid(gx, nchunks);
    B200_DISPATCH_DTYPE(dtype, {
        launch_pdl(colsum_partial_vec_kernel<T, true>, grid, dim3(256), 0, as_stream(stream), (const T*)x, (float*)workspace, rows, cols,
                   cols, rpc, (T*)y, dk);
    });
    int rc = check_launch("dropout_colsum_partial");
    if (rc != B200ASR_OK) return rc;
    launch_pdl(reduce_partials_kernel, dim3(ceil_div(cols, 32)), dim3(1024), 0, as_stream(stream), (const float*)workspace, out,
               (float*)nullptr, (float*)nullptr, nchunks, cols, cols, cols);
    return check_launch("dropout_colsum_reduce");
}

}  // extern "C"

// diagnostics: while `buf` (device memory, 8 u64 per block of the LayerNorm backward grid = 148 * 8 * 8 bytes) is set,
// ln_bwd_kernel records SM-clock stamps of thread 0: 0 entry, 1 dependency wait done, 2 first rows requested, 3 first row
// reduced, 4 row loop done, 5 block barrier passed, 6 partials written; NULL stops
extern "C" int b200asr_ln_set_trace(void* buf) {
    unsigned long long* p = (unsigned long long*)buf;
    cudaError_t e = cudaMemcpyToSymbol(b200asr::g_ln_trace_dev, &p, sizeof(p));
    if (e != cudaSuccess) {
        b200asr::set_error("ln_set_trace: %s", cudaGetErrorString(e));
        return B200ASR_ERR_CUDA;
    }
    return 0;
}
