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

// Keep BITS of `nsites` dropout sites (site0, site0 + stride, ...) over a row-major [rows, cols] tensor each, drawn ahead of
// their consumers (the decisions are data independent): bit i of word (m, cg) = "element (m, 32 cg + i) is kept", the same
// decision b200asr_dropout takes for the same (p, seed, site, offset) -- four Philox calls per word.  Words are stored
// column-group major, [site][cols / 32][mpad] (mpad >= rows): the consumer is the epilogue of the tcgen05 GEMM, where a
// lane owns one row of the tile and the 32 lanes of a warp read 32 consecutive words.
__global__ void __launch_bounds__(256) dropout_keep_bits_kernel(uint32_t* __restrict__ out, int nsites, int rows, int cols,
                                                                int mpad, uint32_t thresh16, unsigned long long seed,
                                                                unsigned long long site0, unsigned long long site_stride,
                                                                unsigned long long offset,
                                                                const unsigned long long* __restrict__ dev_offset) {
    if (dev_offset) offset += *dev_offset;
    const int cgs = cols >> 5;
    const long long wps = (long long)cgs * mpad;
    const long long total = wps * nsites;
    for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (long long)gridDim.x * blockDim.x) {
        const int s = (int)(w / wps);
        const long long r = w - (long long)s * wps;
        const int cg = (int)(r / mpad), m = (int)(r - (long long)cg * mpad);
        uint32_t bits = 0;
        if (m < rows) {
            const unsigned long long kk = seed ^ ((site0 + (unsigned long long)s * site_stride) * 0x9E3779B97F4A7C15ull);
            const uint2 key = make_uint2((uint32_t)kk, (uint32_t)(kk >> 32));
            const unsigned long long e0 = (unsigned long long)m * (unsigned long long)cols + 32ull * cg;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint32_t u[8];
                dropout_uniforms<8>(e0 + 8 * q, key, offset, u);
#pragma unroll
                for (int e = 0; e < 8; ++e) bits |= (u[e] >= thresh16 ? 1u : 0u) << (8 * q + e);
            }
        }
        out[w] = bits;
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

size_t b200asr_dropout_keep_words(int cols, int mpad) { return (size_t)(cols / 32) * (size_t)mpad; }

int b200asr_dropout_keep_bits(unsigned* out, int nsites, unsigned long long site0, unsigned long long site_stride, int rows,
                              int cols, int mpad, float p, unsigned long long seed, unsigned long long offset,
                              const unsigned long long* dev_offset, void* stream) {
    B200_REQUIRE(out && nsites >= 1 && rows >= 1, "dropout_keep_bits: bad argument");
    B200_REQUIRE(cols >= 32 && cols % 32 == 0 && mpad >= rows, "dropout_keep_bits: cols=%d must be a multiple of 32, mpad=%d >= rows=%d",
                 cols, mpad, rows);
    B200_REQUIRE(p >= 0.f && p < 1.f, "dropout_keep_bits: p=%f must be in [0, 1)", (double)p);
    const DropKeys dk = make_drop_keys(p, seed, site0, offset, dev_offset);
    const long long total = (long long)b200asr_dropout_keep_words(cols, mpad) * nsites;
    long long grid = (total + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    if (grid > cap) grid = cap;
    dropout_keep_bits_kernel<<<(int)grid, 256, 0, as_stream(stream)>>>(out, nsites, rows, cols, mpad, dk.thresh16, seed, site0,
                                                                      site_stride, offset, dev_offset);
    return check_launch("dropout_keep_bits");
}

}  // extern "C"
