// b200asr_gemm: argument validation + back-end selection.
//   bf16 operands, aligned            -> tcgen05 tensor cores (gemm_tcgen05.cu)
//   fp32 operands (parity mode)       -> fp32 FMA pipes      (gemm_simt.cu)
//   bf16 but TMA-ineligible (row pitch not a multiple of 16 B, N < 8: the 2-wide dementia head) -> fp32 FMA
//   pipes as well; those products are HBM-bound re-reads of the activations, not tensor-core work.
#include "common.cuh"
#include "gemm.h"

namespace b200asr {

int gemm_dispatch(const b200asr_gemm_args& a, cudaStream_t st) {
    if (a.dtype == B200ASR_BF16 && gemm_tc_eligible(a, nullptr)) return gemm_tc_launch(a, st);
    return gemm_simt_launch(a, st);
}

}  // namespace b200asr

using namespace b200asr;

extern "C" int b200asr_gemm(const b200asr_gemm_args* args, void* stream) {
    B200_REQUIRE(args, "gemm: null args");
    const b200asr_gemm_args& a = *args;
    B200_REQUIRE(a.A && a.B && a.C, "gemm: null operand");
    B200_REQUIRE(a.M >= 0 && a.N >= 0 && a.K >= 1, "gemm: bad shape M=%d N=%d K=%d", a.M, a.N, a.K);
    B200_REQUIRE(a.batch0 >= 1 && a.batch1 >= 1, "gemm: batch counts must be >= 1");
    B200_REQUIRE(a.dtype == B200ASR_F32 || a.dtype == B200ASR_BF16, "gemm: bad dtype %d", a.dtype);
    B200_REQUIRE(a.c_dtype == B200ASR_F32 || a.c_dtype == B200ASR_BF16, "gemm: bad c_dtype %d", a.c_dtype);
    B200_REQUIRE(!a.mul_gelu_grad || a.aux_in, "gemm: mul_gelu_grad needs aux_in");
    B200_REQUIRE(a.lda >= (a.a_kmajor ? 1 : a.M) && a.ldb >= (a.b_kmajor ? 1 : a.N) && a.ldc >= a.N,
                 "gemm: leading dimension too small (lda=%d ldb=%d ldc=%d)", a.lda, a.ldb, a.ldc);
    if (a.M == 0 || a.N == 0) return B200ASR_OK;
    return gemm_dispatch(a, as_stream(stream));
}

// test hook: force a back end (0 = simt, 1 = tcgen05) regardless of the dispatch rule
extern "C" int b200asr_gemm_backend(const b200asr_gemm_args* args, int backend, void* stream) {
    B200_REQUIRE(args, "gemm: null args");
    if (backend == 0) return gemm_simt_launch(*args, as_stream(stream));
    const char* why = nullptr;
    if (!gemm_tc_eligible(*args, &why)) {
        set_error("gemm: not eligible for tcgen05: %s", why ? why : "?");
        return B200ASR_ERR_UNSUPPORTED;
    }
    return gemm_tc_launch(*args, as_stream(stream));
}

// test / tuning hook: force the tcgen05 tile configuration (0 = cost model; else ncta * 1000 + bn, e.g. 2256 = CTA
// pair with 256 x 256 tiles, 1128 = single CTA 128 x 128).  Returns the previous setting.
extern "C" int b200asr_gemm_force_config(int cfg) { return gemm_tc_force_config(cfg); }
