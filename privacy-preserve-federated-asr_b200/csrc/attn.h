// Internal: fused tcgen05 attention back end behind b200asr_mha_fwd / b200asr_mha_bwd (attn_tc.cu).
#pragma once
#include <cuda_runtime.h>
#include "../../include/b200asr.h"

namespace b200asr {

// bf16, head_dim 64, 16-byte aligned packed projections
bool mha_tc_eligible(int dtype, int dh, const void* qkv);
// keep / p_drop: attention-probability dropout keep bits (attn_keep_mask_launch layout) or nullptr
int mha_fwd_tc_launch(const void* qkv, void* out, float* lse, const int* klen, const unsigned* keep, float p_drop, int B,
                      int T, int H, float scale, cudaStream_t st);
size_t mha_bwd_tc_workspace_bytes(int B, int T, int H);
int mha_bwd_tc_launch(const void* qkv, const void* out, const void* dout, const float* lse, const int* klen,
                      const unsigned* keep, float p_drop, void* dqkv, void* workspace, int B, int T, int H, float scale,
                      cudaStream_t st);

// keep-bit layout: word [((b H + h) KW + key / 32) TQ + query], bit key & 31; KW = ceil(T / 32), TQ = T rounded up to 128
int attn_keep_kw(int T);
int attn_keep_tq(int T);
int attn_keep_mask_launch(unsigned* keep, int n_sites, int B, int T, int H, float p, unsigned long long seed,
                          unsigned long long site0, unsigned long long site_stride, unsigned long long offset,
                          const unsigned long long* dev_offset, cudaStream_t st);

}  // namespace b200asr
