// Internal: fused tcgen05 attention back end behind b200asr_mha_fwd / b200asr_mha_bwd (attn_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include "../../include/b200asr.h"

namespace b200asr {

// bf16, head_dim 64, 16-byte aligned packed projections
bool mha_tc_eligible(int dtype, int dh, const void* qkv);
int mha_fwd_tc_launch(const void* qkv, void* out, float* lse, const int* klen, int B, int T, int H, float scale,
                      cudaStream_t st);
size_t mha_bwd_tc_workspace_bytes(int B, int T, int H);
int mha_bwd_tc_launch(const void* qkv, const void* out, const void* dout, const float* lse, const int* klen, void* dqkv,
                      void* workspace, int B, int T, int H, float scale, cudaStream_t st);

}  // namespace b200asr
