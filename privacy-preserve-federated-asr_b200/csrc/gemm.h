// Internal: the two GEMM back ends behind b200asr_gemm.
#pragma once
#include <cuda_runtime.h>
#include "../../include/b200asr.h"

namespace b200asr {

// fp32 FMA pipes (parity mode; HBM-bound skinny / unaligned products)
int gemm_simt_launch(const b200asr_gemm_args& a, cudaStream_t st);
// tcgen05 tensor cores (bf16 operands, fp32 accumulate in TMEM)
bool gemm_tc_eligible(const b200asr_gemm_args& a, const char** why);
int gemm_tc_launch(const b200asr_gemm_args& a, cudaStream_t st);
int gemm_tc_force_config(int cfg);  // test hook: 0 = cost model, else ncta * 1000 + bn
// act = GELU_SAVE_GRAD with element dropout on both outputs (C and aux_out) in the epilogue
struct DropKeys;
// (keep_bits != null: the decisions are read from bits drawn ahead, [(n / 32) * mpad + m]; k then only carries 1 / (1 - p))
int gemm_tc_launch_dropout(const b200asr_gemm_args& a, const DropKeys& k, const unsigned* keep_bits, int mpad, cudaStream_t st);

// grouped positional conv (64 channels per group, bf16) on the tcgen05 GEMM kernel's conv addressing modes
int posconv_tc_taps(const void* x, const void* w_pack, const float* bias, void* y, int B, int T, int C, int groups, int k,
                    int tap_pad, cudaStream_t st);
int posconv_tc_wgrad(const void* x, const void* dy, float* dwt, int B, int T, int C, int groups, int k, int pad,
                     cudaStream_t st);

// internal convenience used by the composed attention path
int gemm_dispatch(const b200asr_gemm_args& a, cudaStream_t st);

}  // namespace b200asr
