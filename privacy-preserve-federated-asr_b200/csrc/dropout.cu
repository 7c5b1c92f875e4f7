// Element dropout (torch.nn.functional.dropout semantics: keep with probability 1-p, scale kept values by 1/(1-p)),
// used at the encoder's hidden_dropout / activation_dropout / feat_proj_dropout / final_dropout sites
// (HF modeling_data2vec_audio.py:160-172, 300-346, 359-386; federated/src/models.py:403).
//
// No mask is stored: forward and backward regenerate it from a counter-based generator (Philox4x32-10) keyed by
// (seed, site) with the counter (element index / 8, step).  One Philox call yields 8 x 16-bit uniforms = the keep
// decisions of 8 consecutive elements, so the mask of element e does not depend on the dtype's vector width.
// HBM-bound: read x [+ res], write y -> 2-3 x sizeof(T) bytes per element.
#include "common.cuh"

namespace b200asr {

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

template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(const T* x, const T* __restrict__ res, T* y /* may alias x */,
                                                      long long nvec, uint32_t thresh16, float scale,
                                                      uint2 key, unsigned long long offset,
                                                      const unsigned long long* __restrict__ dev_offset) {
    constexpr int N = Vec16<T>::N;
    if (dev_offset) offset += *dev_offset;  // per-step counter kept on the device: a captured launch draws a fresh mask
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        const unsigned long long e0 = (unsigned long long)i * N;
        const unsigned long long blk = e0 >> 3;
        const uint4 r = philox4x32_10(make_uint4((uint32_t)blk, (uint32_t)(blk >> 32), (uint32_t)offset,
                                                 (uint32_t)(offset >> 32)), key);
        uint32_t u[N];
        if constexpr (N == 8) {
            u[0] = r.x & 0xffffu; u[1] = r.x >> 16; u[2] = r.y & 0xffffu; u[3] = r.y >> 16;
            u[4] = r.z & 0xffffu; u[5] = r.z >> 16; u[6] = r.w & 0xffffu; u[7] = r.w >> 16;
        } else {  // fp32: 4 elements per vector = the low or the high half of the 8-element block
            const bool hi = (e0 & 4) != 0;
            const uint32_t a = hi ? r.z : r.x, b = hi ? r.w : r.y;
            u[0] = a & 0xffffu; u[1] = a >> 16; u[2] = b & 0xffffu; u[3] = b >> 16;
        }
        Vec16<T> v;
        float f[N], o[N];
        v.load(x + e0);
        v.unpack(f);
        if (res) {
            Vec16<T> rv;
            rv.load(res + e0);
            rv.unpack(o);
        } else {
#pragma unroll
            for (int j = 0; j < N; ++j) o[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < N; ++j) o[j] += (u[j] >= thresh16) ? f[j] * scale : 0.f;
        v.pack(o);
        v.store(y + e0);
    }
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

int b200asr_dropout(const void* x, const void* res, void* y, long long n, float p, unsigned long long seed,
                    unsigned long long site, unsigned long long offset, const unsigned long long* dev_offset, int dtype,
                    void* stream) {
    if (n == 0) return B200ASR_OK;
    B200_REQUIRE(x && y && n > 0, "dropout: bad argument");
    B200_REQUIRE(p >= 0.f && p < 1.f, "dropout: p=%f must be in [0, 1)", (double)p);
    B200_REQUIRE((((uintptr_t)x | (uintptr_t)y | (uintptr_t)res) & 15) == 0, "dropout: buffers must be 16-byte aligned");
    cudaStream_t st = as_stream(stream);
    const uint32_t thresh16 = (uint32_t)lrintf(p * 65536.f);  // keep iff u16 >= thresh16: P(drop) = thresh16 / 65536
    const float scale = 1.0f / (1.0f - p);
    const unsigned long long k = seed ^ (site * 0x9E3779B97F4A7C15ull);
    const uint2 key = make_uint2((uint32_t)k, (uint32_t)(k >> 32));
    B200_DISPATCH_DTYPE(dtype, {
        constexpr int N = Vec16<T>::N;
        B200_REQUIRE(n % N == 0, "dropout: n=%lld must be a multiple of %d", n, N);
        const long long nvec = n / N;
        long long grid = (nvec + 255) / 256;
        const long long cap = (long long)num_sms() * 16;
        if (grid > cap) grid = cap;
        dropout_kernel<T><<<(int)grid, 256, 0, st>>>((const T*)x, (const T*)res, (T*)y, nvec, thresh16, scale, key, offset,
                                                     dev_offset);
    });
    return check_launch("dropout");
}

}  // extern "C"
