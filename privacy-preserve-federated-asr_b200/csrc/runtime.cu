// Error plumbing + small elementwise helpers of the C ABI.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>

#include "common.cuh"

namespace b200asr {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
        return B200ASR_ERR_CUDA;
    }
    return B200ASR_OK;
}

bool pdl_enabled() {
    static const bool on = !(getenv("B200ASR_PDL") && atoi(getenv("B200ASR_PDL")) == 0);
    return on;
}

int num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
    }
    return n;
}

// ---------------------------------------------------------------------------
template <typename T>
__global__ void cast_from_f32_kernel(const float* __restrict__ src, T* __restrict__ dst, long long n) {
    long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const long long stride = (long long)gridDim.x * blockDim.x * 4;
    for (; i + 3 < n; i += stride) {
        float4 v = *reinterpret_cast<const float4*>(src + i);
        if constexpr (sizeof(T) == 2) {
            uint2 o;
            o.x = bf2_pack(v.x, v.y);
            o.y = bf2_pack(v.z, v.w);
            *reinterpret_cast<uint2*>(dst + i) = o;
        } else {
            *reinterpret_cast<float4*>(dst + i) = v;
        }
    }
    if (i < n && i + 3 >= n)
        for (long long j = i; j < n; ++j) dst[j] = from_f<T>(src[j]);
}

template <typename T>
__global__ void cast_to_f32_kernel(const T* __restrict__ src, float* __restrict__ dst, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) dst[i] = to_f<T>(src[i]);
}

template <typename T>
__global__ void scale_kernel(T* __restrict__ x, long long n, float hs, const float* __restrict__ ds) {
    const float s = hs * (ds ? *ds : 1.0f);
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) x[i] = from_f<T>(to_f<T>(x[i]) * s);
}

template <typename T>
__global__ void axpy_kernel(T* __restrict__ y, const T* __restrict__ x, float alpha, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n; i += stride) y[i] = from_f<T>(to_f<T>(y[i]) + alpha * to_f<T>(x[i]));
}

template <typename T>
__global__ void mask_frames_kernel(T* __restrict__ x, const int* __restrict__ lengths, int T_, int cols) {
    // grid: (frames, B); zero rows t >= lengths[b]
    const int b = blockIdx.y;
    const int len = lengths[b];
    for (int t = len + blockIdx.x; t < T_; t += gridDim.x) {
        T* row = x + ((long long)b * T_ + t) * cols;
        for (int c = threadIdx.x; c < cols; c += blockDim.x) row[c] = from_f<T>(0.f);
    }
}

static int grid_for(long long n, int per_block) {
    long long g = (n + per_block - 1) / per_block;
    long long cap = (long long)num_sms() * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

const char* b200asr_last_error(void) { return g_err; }
int b200asr_abi_version(void) { return B200ASR_ABI_VERSION; }
long long b200asr_launch_count(void) { return g_launches.load(); }

int b200asr_cast_f32(const float* src, void* dst, long long n, int dst_dtype, void* stream) {
    if (n <= 0) return B200ASR_OK;
    B200_REQUIRE(src && dst, "cast_f32: null pointer");
    B200_REQUIRE(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 7) == 0, "cast_f32: unaligned pointer");
    B200_DISPATCH_DTYPE(dst_dtype, {
        cast_from_f32_kernel<T><<<grid_for(n, 1024), 256, 0, as_stream(stream)>>>(src, (T*)dst, n);
    });
    return check_launch("cast_f32");
}

int b200asr_cast_to_f32(const void* src, float* dst, long long n, int src_dtype, void* stream) {
    if (n <= 0) return B200ASR_OK;
    B200_REQUIRE(src && dst, "cast_to_f32: null pointer");
    B200_DISPATCH_DTYPE(src_dtype, {
        cast_to_f32_kernel<T><<<grid_for(n, 256), 256, 0, as_stream(stream)>>>((const T*)src, dst, n);
    });
    return check_launch("cast_to_f32");
}

int b200asr_scale(void* x, long long n, float host_scale, const float* dev_scale, int dtype, void* stream) {
    if (n <= 0) return B200ASR_OK;
    B200_REQUIRE(x, "scale: null pointer");
    B200_DISPATCH_DTYPE(dtype, {
        scale_kernel<T><<<grid_for(n, 256), 256, 0, as_stream(stream)>>>((T*)x, n, host_scale, dev_scale);
    });
    return check_launch("scale");
}

int b200asr_axpy(void* y, const void* x, float alpha, long long n, int dtype, void* stream) {
    if (n <= 0) return B200ASR_OK;
    B200_REQUIRE(x && y, "axpy: null pointer");
    B200_DISPATCH_DTYPE(dtype, {
        axpy_kernel<T><<<grid_for(n, 256), 256, 0, as_stream(stream)>>>((T*)y, (const T*)x, alpha, n);
    });
    return check_launch("axpy");
}

int b200asr_mask_frames(void* x, const int* lengths, int B, int T_, int cols, int dtype, void* stream) {
    if (B <= 0 || T_ <= 0 || cols <= 0) return B200ASR_OK;
    B200_REQUIRE(x && lengths, "mask_frames: null pointer");
    dim3 grid(T_ < 64 ? T_ : 64, B);
    B200_DISPATCH_DTYPE(dtype, {
        mask_frames_kernel<T><<<grid, 128, 0, as_stream(stream)>>>((T*)x, lengths, T_, cols);
    });
    return check_launch("mask_frames");
}

}  // extern "C"
