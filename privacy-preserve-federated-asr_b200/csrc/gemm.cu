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

// y = drop(gelu(A B^T + bias)), aux_out = drop(gelu'(A B^T + bias)) with ONE mask: the FFN's activation dropout (HF:335-339)
// inside the W1 GEMM epilogue.  The backward GEMM multiplies by aux_out as it is (mul_gelu_grad = 2), so neither direction
// needs a dropout pass over the [M, 4 d] activations.  Mask = b200asr_dropout's for the same (p, seed, site, offset).
extern "C" int b200asr_gemm_dropout(const b200asr_gemm_args* args, float p, unsigned long long seed, unsigned long long site,
                                    unsigned long long offset, const unsigned long long* dev_offset, void* stream) {
    B200_REQUIRE(args, "gemm_dropout: null args");
    const b200asr_gemm_args& a = *args;
    B200_REQUIRE(a.A && a.B && a.C && a.aux_out, "gemm_dropout: null operand (aux_out is required)");
    B200_REQUIRE(a.act == B200ASR_ACT_GELU_SAVE_GRAD && !a.mul_gelu_grad && !a.residual,
                 "gemm_dropout: only the GELU_SAVE_GRAD epilogue carries the fused dropout");
    B200_REQUIRE(p >= 0.f && p < 1.f, "gemm_dropout: p=%f must be in [0, 1)", (double)p);
    B200_REQUIRE(a.batch0 == 1 && a.batch1 == 1 && a.ldc == a.N && a.N % 8 == 0,
                 "gemm_dropout: C must be one contiguous [M, N] matrix with N %% 8 == 0 (mask index = m * N + n)");
    B200_REQUIRE(a.M >= 0 && a.N >= 0 && a.K >= 1, "gemm_dropout: bad shape M=%d N=%d K=%d", a.M, a.N, a.K);
    const char* why = nullptr;
    if (a.c_dtype != B200ASR_BF16 || !gemm_tc_eligible(a, &why)) {
        set_error("gemm_dropout: needs the tcgen05 back end with bf16 output (%s); run b200asr_gemm + b200asr_dropout instead",
                  why ? why : "c_dtype is not bf16");
        return B200ASR_ERR_UNSUPPORTED;
    }
    if (a.M == 0 || a.N == 0) return B200ASR_OK;
    const DropKeys k = make_drop_keys(p, seed, site, offset, dev_offset);
    return gemm_tc_launch_dropout(a, k, nullptr, 0, as_stream(stream));
}

// The same with the keep decisions read from bits drawn ahead of time by b200asr_dropout_keep_bits (one launch per step for
// every block, off the critical path) instead of four Philox calls per 32 columns inside the epilogue.
extern "C" int b200asr_gemm_dropout_bits(const b200asr_gemm_args* args, const unsigned* keep_bits, int mpad, float p,
                                         void* stream) {
    B200_REQUIRE(args && keep_bits, "gemm_dropout_bits: null pointer");
    const b200asr_gemm_args& a = *args;
    B200_REQUIRE(a.A && a.B && a.C && a.aux_out, "gemm_dropout_bits: null operand (aux_out is required)");
    B200_REQUIRE(a.act == B200ASR_ACT_GELU_SAVE_GRAD && !a.mul_gelu_grad && !a.residual,
                 "gemm_dropout_bits: only the GELU_SAVE_GRAD epilogue carries the fused dropout");
    B200_REQUIRE(p >= 0.f && p < 1.f, "gemm_dropout_bits: p=%f must be in [0, 1)", (double)p);
    B200_REQUIRE(a.batch0 == 1 && a.batch1 == 1 && a.ldc == a.N && a.N % 64 == 0,
                 "gemm_dropout_bits: C must be one contiguous [M, N] matrix with N %% 64 == 0");
    B200_REQUIRE(a.M >= 0 && a.N >= 0 && a.K >= 1, "gemm_dropout_bits: bad shape M=%d N=%d K=%d", a.M, a.N, a.K);
    B200_REQUIRE(mpad >= (a.M + 511) / 512 * 512, "gemm_dropout_bits: mpad=%d must cover M=%d rounded up to 512 (the largest tile)",
                 mpad, a.M);
    const char* why = nullptr;
    if (a.c_dtype != B200ASR_BF16 || !gemm_tc_eligible(a, &why)) {
        set_error("gemm_dropout_bits: needs the tcgen05 back end with bf16 output (%s)", why ? why : "c_dtype is not bf16");
        return B200ASR_ERR_UNSUPPORTED;
    }
    if (a.M == 0 || a.N == 0) return B200ASR_OK;
    const DropKeys k = make_drop_keys(p, 0, 0, 0, nullptr);
    return gemm_tc_launch_dropout(a, k, keep_bits, mpad, as_stream(stream));
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
