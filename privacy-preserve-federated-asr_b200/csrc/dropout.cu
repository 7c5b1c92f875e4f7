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

template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(const T* x, const T* __restrict__ res, T* y /* may alias x */,
                                                      long long nvec, uint32_t thresh16, float scale,
                                                      uint2 key, unsigned long long offset,
                                                      const unsigned long long* __restrict__ dev_offset) {
    constexpr int N = Vec16<T>::N;
    if (dev_offset) offset += *dev_offset;  // per-step counter kept on the device: a captured launch draws a fresh mask
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
        const unsigned long long e0 = (unsigned long long)i * N;
        uint32_t u[N];
        dropout_uniforms<N>(e0, key, offset, u);
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
    const DropKeys dk = make_drop_keys(p, seed, site, offset, dev_offset);  // P(drop) = p rounded to 1/65536
    const uint32_t thresh16 = dk.thresh16;
    const float scale = dk.scale;
    const uint2 key = dk.key;
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
