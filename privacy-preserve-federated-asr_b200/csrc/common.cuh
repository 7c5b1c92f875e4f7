// Shared device/host helpers for the b200asr sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#include "../../include/b200asr.h"

namespace b200asr {

// ---- error plumbing (no exceptions cross the C boundary) -------------------
void set_error(const char* fmt, ...);
int check_launch(const char* what);  // cudaGetLastError -> status

#define B200_REQUIRE(cond, ...)                       \
    do {                                              \
        if (!(cond)) {                                \
            ::b200asr::set_error(__VA_ARGS__);        \
            return B200ASR_ERR_BAD_ARG;               \
        }                                             \
    } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

int num_sms();

// ---- dtype helpers -----------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// 16-byte vector of T (4 floats or 8 bf16) loaded/stored as one transaction
template <typename T> struct Vec16 {
    static constexpr int N = 16 / sizeof(T);
    uint4 raw;
    __device__ __forceinline__ void load(const T* p) { raw = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void store(T* p) const { *reinterpret_cast<uint4*>(p) = raw; }
    __device__ __forceinline__ void unpack(float* f) const;
    __device__ __forceinline__ void pack(const float* f);
};
template <> __device__ __forceinline__ void Vec16<float>::unpack(float* f) const {
    f[0] = __uint_as_float(raw.x); f[1] = __uint_as_float(raw.y);
    f[2] = __uint_as_float(raw.z); f[3] = __uint_as_float(raw.w);
}
template <> __device__ __forceinline__ void Vec16<float>::pack(const float* f) {
    raw.x = __float_as_uint(f[0]); raw.y = __float_as_uint(f[1]);
    raw.z = __float_as_uint(f[2]); raw.w = __float_as_uint(f[3]);
}
__device__ __forceinline__ void bf2_unpack(uint32_t u, float& a, float& b) {
    a = __uint_as_float(u << 16);
    b = __uint_as_float(u & 0xffff0000u);
}
__device__ __forceinline__ uint32_t bf2_pack(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
template <> __device__ __forceinline__ void Vec16<__nv_bfloat16>::unpack(float* f) const {
    bf2_unpack(raw.x, f[0], f[1]); bf2_unpack(raw.y, f[2], f[3]);
    bf2_unpack(raw.z, f[4], f[5]); bf2_unpack(raw.w, f[6], f[7]);
}
template <> __device__ __forceinline__ void Vec16<__nv_bfloat16>::pack(const float* f) {
    raw.x = bf2_pack(f[0], f[1]); raw.y = bf2_pack(f[2], f[3]);
    raw.z = bf2_pack(f[4], f[5]); raw.w = bf2_pack(f[6], f[7]);
}

// ---- math --------------------------------------------------------------------
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_erf_grad(float x) {
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
    const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
    return cdf + x * pdf;
}

// erf-GELU for the bf16 tensor-core epilogues: Abramowitz-Stegun 7.1.26 (|erf error| <= 1.5e-7, far below bf16
// rounding) with one MUFU exp and one MUFU rcp shared between the value and the derivative.
__device__ __forceinline__ void erf_fast_parts(float x, float& erfv, float& e) {
    const float z = fabsf(x) * 0.70710678118654752440f;
    const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
    e = __expf(-z * z);  // = exp(-x^2 / 2)
    const float poly = fmaf(fmaf(fmaf(fmaf(1.061405429f, t, -1.453152027f), t, 1.421413741f), t, -0.284496736f), t, 0.254829592f) * t;
    erfv = copysignf(fmaf(-poly, e, 1.0f), x);
}
__device__ __forceinline__ float gelu_fast(float x) {
    float er, e;
    erf_fast_parts(x, er, e);
    return 0.5f * x * (1.0f + er);
}
__device__ __forceinline__ float gelu_grad_fast(float x) {
    float er, e;
    erf_fast_parts(x, er, e);
    return fmaf(x * 0.39894228040143267794f, e, 0.5f * (1.0f + er));
}

// Two-at-a-time forms on the packed fp32 pipe of sm_100 (FFMA2 / FMUL2 / FADD2: one issue slot for two lanes' worth of
// work) -- same arithmetic as erf_fast_parts, evaluated for an (x.x, x.y) pair.  The GEMM epilogues are issue-slot bound
// on their K = 1024 shapes (profiles/r01_gemm_epilogue_ablation.txt), which is what this halves.
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void erf_fast_parts2(float2 x, float2& erfv, float2& e) {
    const float2 ax = make_float2(fabsf(x.x), fabsf(x.y));
    const float2 z = __fmul2_rn(ax, make_float2(0.70710678118654752440f, 0.70710678118654752440f));
    const float2 den = __ffma2_rn(make_float2(0.3275911f, 0.3275911f), z, make_float2(1.0f, 1.0f));
    const float2 t = make_float2(rcp_approx(den.x), rcp_approx(den.y));
    const float2 arg = __fmul2_rn(__fmul2_rn(z, make_float2(-1.4426950408889634f, -1.4426950408889634f)), z);
    e = make_float2(ex2_approx(arg.x), ex2_approx(arg.y));  // = exp(-x^2 / 2)
    float2 poly = __ffma2_rn(make_float2(1.061405429f, 1.061405429f), t, make_float2(-1.453152027f, -1.453152027f));
    poly = __ffma2_rn(poly, t, make_float2(1.421413741f, 1.421413741f));
    poly = __ffma2_rn(poly, t, make_float2(-0.284496736f, -0.284496736f));
    poly = __ffma2_rn(poly, t, make_float2(0.254829592f, 0.254829592f));
    poly = __fmul2_rn(poly, t);
    const float2 er = __ffma2_rn(make_float2(-poly.x, -poly.y), e, make_float2(1.0f, 1.0f));
    erfv = make_float2(copysignf(er.x, x.x), copysignf(er.y, x.y));
}
__device__ __forceinline__ float2 gelu_fast2(float2 x) {
    float2 er, e;
    erf_fast_parts2(x, er, e);
    const float2 h = __fmul2_rn(x, make_float2(0.5f, 0.5f));
    return __ffma2_rn(h, er, h);
}
// value and derivative from one erf evaluation
__device__ __forceinline__ void gelu_both_fast2(float2 x, float2& g, float2& gp) {
    float2 er, e;
    erf_fast_parts2(x, er, e);
    const float2 h = __fmul2_rn(x, make_float2(0.5f, 0.5f));
    g = __ffma2_rn(h, er, h);
    const float2 cdf = __ffma2_rn(er, make_float2(0.5f, 0.5f), make_float2(0.5f, 0.5f));
    gp = __ffma2_rn(__fmul2_rn(x, make_float2(0.39894228040143267794f, 0.39894228040143267794f)), e, cdf);
}
__device__ __forceinline__ float2 gelu_grad_fast2(float2 x) {
    float2 er, e;
    erf_fast_parts2(x, er, e);
    const float2 cdf = __ffma2_rn(er, make_float2(0.5f, 0.5f), make_float2(0.5f, 0.5f));
    return __ffma2_rn(__fmul2_rn(x, make_float2(0.39894228040143267794f, 0.39894228040143267794f)), e, cdf);
}

// storage-type dispatch: bf16 outputs take the MUFU-based approximation (error far below bf16 rounding), fp32
// (parity mode) keeps the libm erff / expf forms
template <typename T> __device__ __forceinline__ float gelu_for(float x) {
    if constexpr (sizeof(T) == 2) return gelu_fast(x);
    else return gelu_erf(x);
}
template <typename T> __device__ __forceinline__ float gelu_grad_for(float x) {
    if constexpr (sizeof(T) == 2) return gelu_grad_fast(x);
    else return gelu_erf_grad(x);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// block-wide sum for blockDim.x <= 1024; `red` is >= 32 floats of shared memory
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[wid] = v;
    __syncthreads();
    float r = (lane < nw) ? red[lane] : 0.f;
    r = warp_sum(r);
    return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[wid] = v;
    __syncthreads();
    float r = (lane < nw) ? red[lane] : -INFINITY;
    r = warp_max(r);
    return r;
}

// ---- programmatic dependent launch ------------------------------------------------------------------------------
// A kernel launched through launch_pdl may be scheduled while its predecessor on the stream is still draining (its
// launch latency and prologue overlap the predecessor's tail).  Such a kernel MUST call pdl_wait() before its first
// access to global memory written by earlier work; pdl_trigger() lets ITS dependents be scheduled early in turn.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() { pdl_trigger(); pdl_wait(); }

bool pdl_enabled();  // B200ASR_PDL=0 disables (A/B measurements)

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---- counter-based dropout masks (dropout.cu, LayerNorm forward with a fused dropout + residual add) ----------------
// Philox4x32-10 keyed by (seed, site) over the counter (element index / 8, step): one call = the 16-bit uniforms of 8
// consecutive elements, so an element's keep decision does not depend on the dtype's vector width.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}
struct DropKeys {
    uint2 key;
    uint32_t thresh16;  // keep iff u16 >= thresh16
    float scale;        // 1 / (1 - p)
    unsigned long long offset;
    const unsigned long long* dev_offset;  // optional device-side step counter added to offset
};
static inline DropKeys make_drop_keys(float p, unsigned long long seed, unsigned long long site, unsigned long long offset,
                                      const unsigned long long* dev_offset) {
    DropKeys k;
    const unsigned long long kk = seed ^ (site * 0x9E3779B97F4A7C15ull);
    k.key = make_uint2((uint32_t)kk, (uint32_t)(kk >> 32));
    k.thresh16 = (uint32_t)lrintf(p * 65536.f);
    k.scale = 1.0f / (1.0f - p);
    k.offset = offset;
    k.dev_offset = dev_offset;
    return k;
}
// 16-bit uniforms of the N (4 or 8) consecutive elements starting at element e0 (e0 % N == 0)
template <int N>
__device__ __forceinline__ void dropout_uniforms(unsigned long long e0, uint2 key, unsigned long long offset, uint32_t* u) {
    const unsigned long long blk = e0 >> 3;
    const uint4 r = philox4x32_10(make_uint4((uint32_t)blk, (uint32_t)(blk >> 32), (uint32_t)offset, (uint32_t)(offset >> 32)), key);
    if constexpr (N == 8) {
        u[0] = r.x & 0xffffu; u[1] = r.x >> 16; u[2] = r.y & 0xffffu; u[3] = r.y >> 16;
        u[4] = r.z & 0xffffu; u[5] = r.z >> 16; u[6] = r.w & 0xffffu; u[7] = r.w >> 16;
    } else {  // fp32: 4 elements per vector = the low or the high half of the 8-element block
        const bool hi = (e0 & 4) != 0;
        const uint32_t a = hi ? r.z : r.x, b = hi ? r.w : r.y;
        u[0] = a & 0xffffu; u[1] = a >> 16; u[2] = b & 0xffffu; u[3] = b >> 16;
    }
}

#define B200_DISPATCH_DTYPE(dtype, ...)                                        \
    do {                                                                       \
        if ((dtype) == B200ASR_F32) { using T = float; __VA_ARGS__ }           \
        else if ((dtype) == B200ASR_BF16) { using T = __nv_bfloat16; __VA_ARGS__ } \
        else { ::b200asr::set_error("unsupported dtype %d", (int)(dtype)); return B200ASR_ERR_BAD_ARG; } \
    } while (0)

}  // namespace b200asr
