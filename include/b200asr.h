/* b200asr -- C ABI of the B200-native client-local ASR update + FedAvg hot path.
 *
 * Drop-in boundary.  The reference (JackingChen/Privacy-preserve-Federated-ASR)
 * is pure Python over torch/ATen + HF transformers; the interface each entry point replaces is the ATen /
 * transformers call the reference makes on the hot path, cited as <file>:<line> into /root/reference or
 * `HF:` = transformers/models/data2vec/modeling_data2vec_audio.py.
 *
 * Conventions
 *   - every function returns 0 (B200ASR_OK) or a negative B200ASR_ERR_*; b200asr_last_error() gives the text.
 *   - plain device pointers + sizes only; the library never allocates or frees user-visible memory.
 *     Scratch is a caller-provided `workspace` (+ a *_workspace_bytes query where its size is not obvious).
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*); no host synchronisation.
 *   - `dtype` selects the storage type of activations/weights: B200ASR_F32 or B200ASR_BF16.  Statistics,
 *     accumulators, losses, CTC state and FedAvg are always fp32.
 *   - matrices are row-major; `ld*` are element strides.
 */
#ifndef B200ASR_H_
#define B200ASR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ASR_OK 0
#define B200ASR_ERR_BAD_ARG (-1)
#define B200ASR_ERR_UNSUPPORTED (-2)
#define B200ASR_ERR_CUDA (-3)

#define B200ASR_F32 0
#define B200ASR_BF16 1

#define B200ASR_ABI_VERSION 2

/* activation codes */
#define B200ASR_ACT_NONE 0
#define B200ASR_ACT_GELU 1
#define B200ASR_ACT_GELU_SAVE_GRAD 2 /* GELU, and aux_out receives gelu'(pre-activation) instead of the pre-activation */

const char* b200asr_last_error(void);
int b200asr_abi_version(void);
/* number of kernel launches issued through this library since load (bench.py's gpu_launches) */
long long b200asr_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * LayerNorm over the last dim (+ optional residual add before, + optional GELU after).
 * Replaces torch.nn.functional.layer_norm / nn.LayerNorm at HF:70-74 (conv layers: LN(512)+GELU),
 * HF:160-172 (feature projection), HF:119-126 (pos-conv LN without affine + GELU), HF:381-383,
 * HF:313-346 (post-LN encoder block).
 *   s = x (+ res);  y = act(LN(s) * w + b)
 *   w, b   : fp32 [cols] or NULL (no affine)
 *   sum_out: NULL or [rows, cols] (dtype) receives s (saved for backward when res != NULL)
 *   mean, rstd: NULL or fp32 [rows]
 */
int b200asr_layernorm_fwd(const void* x, const void* res, const float* w, const float* b, void* y, void* sum_out,
                          float* mean, float* rstd, int rows, int cols, float eps, int act, int dtype, void* stream);
/* Backward of the above w.r.t. s (dx) and w, b.  `s` is the pre-norm input.  dw/db: fp32 [cols] or NULL.
 * dx_colsum: NULL or fp32 [cols] = column sums of dx (the bias gradient of the nn.Linear whose output + residual
 * fed this LayerNorm: HF:330-343 -- saves a separate reduction pass over dx).
 * workspace: fp32, >= b200asr_layernorm_bwd_workspace_bytes(cols) bytes (only used when dw/db/dx_colsum != NULL).
 * dres_add: NULL or [rows, cols] (dtype) added to dx (fuses the residual-branch gradient sum). */
size_t b200asr_layernorm_bwd_workspace_bytes(int cols);
int b200asr_layernorm_bwd(const void* dy, const void* s, const float* w, const float* b, const float* mean,
                          const float* rstd, const void* dres_add, void* dx, float* dw, float* db, float* dx_colsum,
                          void* workspace, size_t workspace_bytes, int rows, int cols, int act, int dtype, void* stream);
/* The same with the hidden dropout that sits in front of the residual sum (HF:310-312, 344-345: s = res + drop(o)):
 * dx continues along the residual path, dx_drop = dropout(dx) -- bit-identical to b200asr_dropout(dx) with the same
 * (p, seed, site, offset, dev_offset) -- is the gradient of the dropped nn.Linear output o, and dx_drop_colsum its bias
 * gradient.  Removes one pass over [rows, cols] and one column-sum launch per dropout site of the backward pass. */
int b200asr_layernorm_bwd_dropout(const void* dy, const void* s, const float* w, const float* b, const float* mean,
                                  const float* rstd, const void* dres_add, void* dx, void* dx_drop, float* dw, float* db,
                                  float* dx_drop_colsum, void* workspace, size_t workspace_bytes, int rows, int cols, int act,
                                  float p, unsigned long long seed, unsigned long long site, unsigned long long offset,
                                  const unsigned long long* dev_offset, int dtype, void* stream);

/* Column sums of a [rows, cols] matrix (bias gradients: what autograd does for nn.Linear.bias).
 * out fp32 [cols]; workspace >= b200asr_colsum_workspace_bytes(cols). */
size_t b200asr_colsum_workspace_bytes(int cols);
int b200asr_colsum(const void* x, float* out, void* workspace, size_t workspace_bytes, int rows, int cols, int ld,
                   int dtype, void* stream);
/* b200asr_colsum for callers that cannot provide a workspace and must not allocate (CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The partial sums go to a device array built into the library (2 MiB per GPU that loads it, held until exit), so the
 * call allocates nothing -- no cudaMalloc / cudaFree / allocating library routine.  Limit: cols <=
 * B200ASR_COLSUM_STATIC_MAX_COLS (also returned by b200asr_colsum_static_max_cols()); larger -> B200ASR_ERR_BAD_ARG with
 * a message.  Same result, bit for bit, as b200asr_colsum.  All calls on one GPU share the array (it keeps the last
 * call's partial sums; nothing in it has to be cleared, every call overwrites what it reads): calls that may overlap on
 * the GPU -- different streams without an ordering between them -- race; order them or use b200asr_colsum. */
#define B200ASR_COLSUM_STATIC_MAX_COLS 8192
int b200asr_colsum_static_max_cols(void);
int b200asr_colsum_static(const void* x, float* out, int rows, int cols, int ld, int dtype, void* stream);

/* Zero the frames t >= lengths[b] of x [B, T, cols] in place (HF:369-371 `hidden_states[~mask] = 0`). */
int b200asr_mask_frames(void* x, const int* lengths, int B, int T, int cols, int dtype, void* stream);

/* dtype conversion / copy helpers used by the host shim for weight shadows: dst[i] = (dtype) src_f32[i]. */
int b200asr_cast_f32(const float* src, void* dst, long long n, int dst_dtype, void* stream);
/* dst_f32[i] = (float) src[i] */
int b200asr_cast_to_f32(const void* src, float* dst, long long n, int src_dtype, void* stream);
/* x[i] *= host_scale * (*dev_scale)   (dev_scale may be NULL)  -- upstream-gradient scaling, GRL -lambda. */
int b200asr_scale(void* x, long long n, float host_scale, const float* dev_scale, int dtype, void* stream);

/* y[i] += alpha * x[i]  (gradient accumulation where two branches meet: h + pos_conv(h), HF:381). */
int b200asr_axpy(void* y, const void* x, float alpha, long long n, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Waveform feature extractor layer 0: Conv1d(1 -> C, k, stride) + LayerNorm(C, affine) + GELU, fused.
 * Replaces HF:51-75 for layer_id 0 (reached from federated/src/models.py:395, finetune_ASRs.py:231).
 *   x [B, L] fp32 waveform; w fp32 [C, k]; bias fp32 [C] or NULL; ln_w, ln_b fp32 [C]
 *   y [B, Tout, C] (dtype), channels-last, Tout = (L - k) / stride + 1
 */
int b200asr_conv0_ln_gelu_fwd(const float* x, const float* w, const float* bias, const float* ln_w,
                              const float* ln_b, void* y, int B, int L, int C, int k, int stride, float eps,
                              int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * General (batched) matrix product with fused epilogue -- every dense contraction of the path:
 * nn.Linear forward/backward (HF:160-172, 203-346; federated/src/models.py:292-296), conv layers 1-6 as
 * implicit GEMM over overlapping channels-last windows (HF:57-63), attention products in fp32 mode.
 *
 *   C[m, n] = epi( sum_k A(m, k) * B(n, k) )           m < M, n < N, k < K
 *   A(m, k) = a_kmajor ? A[m*lda + k] : A[k*lda + m]
 *   B(n, k) = b_kmajor ? B[n*ldb + k] : B[k*ldb + n]
 *   epi(v)  = v*alpha (+ bias[n]) -> (aux_out = that, if given) -> (GELU if act) (* gelu'(aux_in[m,n]) if
 *             mul_gelu_grad) (+ residual[m, n])
 *   act = B200ASR_ACT_GELU_SAVE_GRAD stores gelu'(v) in aux_out (both come out of one erf evaluation) and
 *   mul_gelu_grad = 2 multiplies by aux_in as it is: the backward GEMM epilogue then carries no transcendental math.
 * Two batch dims (batch0 fastest) with element strides per operand; batch counts of 1 for a plain GEMM.
 * dtype = storage of A and B (both the same).  BF16 runs on tcgen05 tensor cores (TMA -> smem -> UMMA ->
 * TMEM), F32 runs on FP32 FMA pipes (parity mode; no TF32).  c_dtype = storage of C, residual, aux.
 * bias is fp32 [N].
 */
typedef struct b200asr_gemm_args {
    const void* A;
    const void* B;
    void* C;
    const float* bias;     /* NULL or [N] */
    const void* residual;  /* NULL or [M, N] with ldc / batch strides of C */
    void* aux_out;         /* NULL or [M, N]: pre-activation copy (ldc, C strides) */
    const void* aux_in;    /* NULL or [M, N]: pre-activation for mul_gelu_grad (ldc, C strides) */
    int M, N, K;
    int lda, ldb, ldc;
    int a_kmajor, b_kmajor;
    int act;            /* B200ASR_ACT_* */
    int mul_gelu_grad;  /* 0; 1: * gelu'(aux_in); 2: * aux_in (already a derivative) */
    float alpha;
    int batch0, batch1;
    long long sa0, sa1, sb0, sb1, sc0, sc1; /* element strides of the batch dims */
    int dtype, c_dtype;
} b200asr_gemm_args;
int b200asr_gemm(const b200asr_gemm_args* args, void* stream);

/* b200asr_gemm with act = B200ASR_ACT_GELU_SAVE_GRAD and the FFN's activation dropout (HF:335-339,
 * `nn.Dropout(config.activation_dropout)` after the GELU) applied in the epilogue: C = drop(gelu(v)), aux_out =
 * drop(gelu'(v)) with the SAME keep decisions, which are those of b200asr_dropout for (p, seed, site, offset, dev_offset)
 * over the row-major [M, N] result.  The backward GEMM (mul_gelu_grad = 2 on aux_out) then needs no dropout pass either.
 * bf16 operands and output on the tcgen05 back end only (B200ASR_ERR_UNSUPPORTED otherwise: use b200asr_gemm followed
 * by b200asr_dropout); one batch, ldc = N, N % 8 == 0. */
int b200asr_gemm_dropout(const b200asr_gemm_args* args, float p, unsigned long long seed, unsigned long long site,
                         unsigned long long offset, const unsigned long long* dev_offset, void* stream);
/* Keep bits of `nsites` dropout sites (site0, site0 + site_stride, ...) over a row-major [rows, cols] tensor each, drawn
 * ahead of their consumer: bit (n & 31) of word out[(site * (cols / 32) + n / 32) * mpad + m] = "element (m, n) is kept" --
 * the decision b200asr_dropout takes for the same (p, seed, site, offset, dev_offset).  cols % 32 == 0, mpad >= rows;
 * b200asr_dropout_keep_words(cols, mpad) words per site.  One launch per training step serves the activation dropout
 * (HF:335-339) of every encoder block; b200asr_gemm_dropout_bits is b200asr_gemm_dropout reading these bits (mpad >= M
 * rounded up to 512, N % 64 == 0). */
size_t b200asr_dropout_keep_words(int cols, int mpad);
int b200asr_dropout_keep_bits(unsigned* out, int nsites, unsigned long long site0, unsigned long long site_stride, int rows,
                              int cols, int mpad, float p, unsigned long long seed, unsigned long long offset,
                              const unsigned long long* dev_offset, void* stream);
int b200asr_gemm_dropout_bits(const b200asr_gemm_args* args, const unsigned* keep_bits, int mpad, float p, void* stream);

/* test / tuning hook: force the tcgen05 tile configuration: 0 = built-in cost model, else ncta * 1000 + bn
 * (1064, 1128, 1256 single CTA; 2128, 2256 CTA pair = cluster of 2 with cta_group::2 MMA).  Returns the old value. */
int b200asr_gemm_force_config(int cfg);
/* diagnostics: while `buf` (device memory, >= 148 * 20 * 8 bytes) is set, every tensor-core GEMM launch records 20 u64
 * time stamps per CTA (entry, set-up done, dependency wait done, first operands landed, last MMA issued, last
 * accumulator complete, epilogue done, exit; SM clock cycles, slots 0 / 9 = %globaltimer ns); NULL stops. */
int b200asr_gemm_set_trace(void* buf);
/* diagnostics, same idea for the attention backward kernel: 24 u64 SM-clock stamps per CTA of its grid
 * (ceil(T/128) x H x B CTAs), slot layout next to the function in csrc/attn_tc.cu; NULL stops. */
int b200asr_attn_set_trace(void* buf);
/* diagnostics: 8 u64 SM-clock stamps of thread 0 per block of the LayerNorm backward grid (148 blocks); NULL stops. */
int b200asr_ln_set_trace(void* buf);
/* test hook: run b200asr_gemm on a forced back end (0 = fp32 FMA pipes, 1 = tcgen05 tensor cores). */
int b200asr_gemm_backend(const b200asr_gemm_args* args, int backend, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Row softmax with key-length masking, in place on scores [B, H, Tq, ldk] (HF:189-197 eager attention).
 * Keys k >= klen[b] get probability 0 (klen NULL = all valid).  Columns in [Tk, ldk) are zero-filled.
 * lse: NULL or fp32 [B, H, Tq] = log-sum-exp of each row. */
int b200asr_softmax_rows(void* s, const int* klen, float* lse, int B, int H, int Tq, int Tk, int ldk, int dtype,
                         void* stream);
/* Softmax backward in place: dp <- P * (dP - rowsum(dP * P)).  If lse != NULL, `p` holds the scaled logits S on
 * entry and is first overwritten with P = exp(S - lse) (keys >= klen[b] -> 0). */
int b200asr_softmax_rows_bwd(void* p, void* dp, const float* lse, const int* klen, int B, int H, int Tq, int Tk,
                             int ldk, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Multi-head self-attention, replaces Data2VecAudioAttention.forward's bmm -> softmax -> bmm
 * (HF:189-197 eager path == transformers 4.17).
 *   qkv [B, T, 3, H, dh] packed projections (q NOT pre-scaled; `scale` is applied to the logits)
 *   out [B, T, H*dh]; lse fp32 [B, H, T] (log-sum-exp of the scaled logits, saved for backward)
 *   klen int32 [B] or NULL: number of valid key frames per utterance.
 *   workspace >= b200asr_mha_fwd_workspace_bytes(...)
 */
size_t b200asr_mha_fwd_workspace_bytes(int B, int T, int H, int dh, int dtype);
int b200asr_mha_fwd(const void* qkv, void* out, float* lse, const int* klen, void* workspace, size_t workspace_bytes,
                    int B, int T, int H, int dh, float scale, int dtype, void* stream);
/* test hook: bf16 + head_dim 64 runs the fused tcgen05 flash-attention kernels; 0 forces the composed
 * (batched GEMM + row softmax) path that fp32 parity mode and other head sizes use.  Returns the old setting. */
int b200asr_mha_set_fused(int on);
/* dqkv [B, T, 3, H, dh] (fully written).  workspace >= b200asr_mha_bwd_workspace_bytes(...). */
size_t b200asr_mha_bwd_workspace_bytes(int B, int T, int H, int dh, int dtype);
int b200asr_mha_bwd(const void* qkv, const void* out, const void* dout, const float* lse, const int* klen,
                    void* dqkv, void* workspace, size_t workspace_bytes, int B, int T, int H, int dh, float scale,
                    int dtype, void* stream);
/* Where dQ of the fused backward is formed.  0 (default): the kernel stores dS^T (one TMA store per query block; the tile
 * stays L2-resident) and dQ = dS K is a batched GEMM -- fixed summation order, bit-reproducible, and the faster of the two
 * on B200 (72 vs 85 us per layer at B=8 x 10 s: the fp32 reduce-adds of the alternative are bound by the L2 atomic rate).
 * 1: inside the backward kernel -- dQ_i += dS_i K_j is a fifth MMA of the query-block loop, the partial tiles are added
 * into an fp32 accumulator with bulk reduce-add (cp.reduce.async.bulk .add.f32) and converted once; no dS^T round trip
 * through HBM; the key blocks of a head add in arrival order: reproducible to fp32 rounding, not bit for bit.
 * 2: measurement hook (as 1 without the reduce-adds: dQ is wrong).  Returns the old setting. */
int b200asr_mha_set_dq_fused(int on);
/* tuning hook: share of the softmax exponentials the fused attention kernels compute with a degree-3 polynomial on the FMA
 * pipe (relative error 1e-4, far below the bf16 rounding of the probabilities) instead of the MUFU: fwd / bwd in
 * {0 = none, 4 = 25 %, 2 = 50 %}, negative = leave unchanged.  Returns the previous setting (fwd * 16 + bwd). */
int b200asr_mha_set_exp2_poly(int fwd, int bwd);

/* Attention-probability dropout: `nn.functional.dropout(attn_weights, p=config.attention_dropout)` between the softmax and
 * the P V product (HF:189-197; Data2VecAudioConfig.attention_dropout = 0.1 in every reference training run, also inside the
 * frozen encoder of DACS stages 1-2, federated/src/update.py:433,650).
 * The keep decisions are data independent, so they are drawn OFF the attention kernels' critical path, for all blocks of a
 * step in one launch: keep[site][((b H + h) KW + key / 32) TQ + query] bit (key & 31), 1 = keep; KW = ceil(T / 32),
 * TQ = T rounded up to 128; b200asr_attn_keep_words(B, T, H) uint32 words per site.  Philox4x32-10 keyed by
 * (seed, site0 + site_stride * s) over (word, step = offset + *dev_offset); P(drop) = p rounded to 1 / 65536.
 * The fused kernels read one word per 32 probabilities; the forward keeps the softmax normalisation and the saved
 * log-sum-exp of the UNDROPPED probabilities, drops P before P V and scales the output by 1 / (1 - p); the backward
 * applies the same bits to P (for dV) and to dP (for dS).  The composed path applies them to the materialised planes. */
size_t b200asr_attn_keep_words(int B, int T, int H);
int b200asr_attn_dropout_mask(unsigned* keep, int n_sites, int B, int T, int H, float p, unsigned long long seed,
                              unsigned long long site0, unsigned long long site_stride, unsigned long long offset,
                              const unsigned long long* dev_offset, void* stream);
/* b200asr_mha_fwd / b200asr_mha_bwd with the keep bits of one site (keep == NULL: no dropout). */
int b200asr_mha_fwd_dropout(const void* qkv, void* out, float* lse, const int* klen, const unsigned* keep, float p_drop,
                            void* workspace, size_t workspace_bytes, int B, int T, int H, int dh, float scale, int dtype,
                            void* stream);
int b200asr_mha_bwd_dropout(const void* qkv, const void* out, const void* dout, const float* lse, const int* klen,
                            const unsigned* keep, float p_drop, void* dqkv, void* workspace, size_t workspace_bytes, int B,
                            int T, int H, int dh, float scale, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Grouped positional convolution (HF:89-128): y[b,t,g*Cg+o] = bias + sum_{j,i} W[g*Cg+o, i, j] *
 * x[b, t+j-k/2, g*Cg+i]  (zero padding; for even k the reference drops the last frame, which equals
 * computing only t < T).  Channels-last x, y [B, T, C] (dtype).
 * Weights are consumed in a packed layout produced once per weight update by b200asr_posconv_pack_weight
 * from the HF layout w [C, Cg, k] (dtype) into two buffers of C*Cg*k elements each (an internal layout: tap-major
 * K-major rows for the tcgen05 implicit GEMM when Cg == 64 and dtype is bf16, [groups][k][Cg][Cg] otherwise).
 */
int b200asr_posconv_pack_weight(const void* w, void* w_fwd, void* w_bwd, int C, int groups, int k, int dtype,
                                void* stream);
int b200asr_posconv_fwd(const void* x, const void* w_fwd, const float* bias, void* y, int B, int T, int C, int groups,
                        int k, int dtype, void* stream);
/* dx [B,T,C] (dtype) and dw fp32 [C, Cg, k] (HF layout; overwritten); dbias via b200asr_colsum.  Either output
 * may be NULL.  workspace >= b200asr_posconv_bwd_workspace_bytes(...) (tap-major staging of the tensor-core weight
 * gradient; only needed when dw != NULL). */
size_t b200asr_posconv_bwd_workspace_bytes(int C, int groups, int k);
int b200asr_posconv_bwd(const void* x, const void* w_bwd, const void* dy, void* dx, float* dw, void* workspace,
                        size_t workspace_bytes, int B, int T, int C, int groups, int k, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Wav2Vec2 / Hubert members of the family (centralized/finetune_ASRs.py:154-172, BASELINE configs[4]).
 *
 * conv layer 0 with GroupNorm(num_groups = C, affine) + GELU: replaces nn.Conv1d + nn.GroupNorm + gelu of
 * transformers wav2vec2/modeling_wav2vec2.py Wav2Vec2GroupNormConvLayer (feat_extract_norm = "group").  The
 * statistics are per (utterance, channel) over ALL Tout frames (padded frames included, as torch does).
 *   x [B, L] fp32; w fp32 [C, k]; bias fp32 [C] or NULL; gn_w, gn_b fp32 [C]; y [B, Tout, C] (dtype) channels-last
 *   workspace >= b200asr_conv0_gn_gelu_workspace_bytes(...)
 */
size_t b200asr_conv0_gn_gelu_workspace_bytes(int B, int L, int C, int k, int stride);
int b200asr_conv0_gn_gelu_fwd(const float* x, const float* w, const float* bias, const float* gn_w, const float* gn_b,
                              void* y, void* workspace, size_t workspace_bytes, int B, int L, int C, int k, int stride,
                              float eps, int dtype, void* stream);
/* torch.nn.utils.weight_norm(conv, name="weight", dim=2) of the positional conv filter
 * (wav2vec2/modeling_wav2vec2.py Wav2Vec2PositionalConvEmbedding.__init__): v fp32 [rows, k] (rows = C * C/groups,
 * tap fastest), g fp32 [k];  w[r][j] = g[j] * v[r][j] / norm[j],  norm[j] = ||v[:, j]||_2  (written to `norm`, fp32 [k]).
 * Backward: dv fp32 [rows, k], dg fp32 [k] from dw fp32 [rows, k].  workspace >= b200asr_weight_norm_workspace_bytes(k). */
size_t b200asr_weight_norm_workspace_bytes(int k);
int b200asr_weight_norm_fwd(const float* v, const float* g, float* w, float* norm, void* workspace, size_t workspace_bytes,
                            int rows, int k, void* stream);
int b200asr_weight_norm_bwd(const float* dw, const float* v, const float* g, const float* norm, float* dv, float* dg,
                            void* workspace, size_t workspace_bytes, int rows, int k, void* stream);
/* y = res + gelu(c) (res may be NULL) -- the activation after the weight-normed positional conv fused with
 * `hidden_states + position_embeddings` (Wav2Vec2Encoder.forward); backward dc = dy * gelu'(c).  n % (16/sizeof) == 0. */
int b200asr_gelu_add_fwd(const void* c, const void* res, void* y, long long n, int dtype, void* stream);
int b200asr_gelu_bwd(const void* dy, const void* c, void* dc, long long n, int dtype, void* stream);

/* b200asr_layernorm_fwd with an inverted dropout applied to x before the residual add:
 *   sum_out = res + dropout(x),  y = act(LN(sum_out) * w + b)
 * = `hidden_states = attn_residual + self.dropout(hidden_states); hidden_states = self.layer_norm(hidden_states)`
 * (HF modeling_data2vec_audio.py:321-324, 327-328) in one pass.  Mask keys as b200asr_dropout (below): the same
 * (seed, site, offset) give the same mask there, which is how the backward pass re-draws it.  cols % 8 == 0. */
int b200asr_layernorm_dropout_fwd(const void* x, const void* res, const float* w, const float* b, void* y, void* sum_out,
                                  float* mean, float* rstd, int rows, int cols, float eps, int act, float p,
                                  unsigned long long seed, unsigned long long site, unsigned long long offset,
                                  const unsigned long long* dev_offset, int dtype, void* stream);

/* y = dropout(x) and out[cols] = column sums of y in one pass over x [rows, cols] (contiguous): the backward of a
 * Linear whose output went through a dropout -- `d_out = dropout(ds)` feeds the data / weight gradients, its column sums
 * are the bias gradient.  Keys as b200asr_dropout; workspace as b200asr_colsum.  cols % 8 == 0. */
int b200asr_dropout_colsum(const void* x, void* y, float* out, void* workspace, size_t workspace_bytes, int rows, int cols,
                           float p, unsigned long long seed, unsigned long long site, unsigned long long offset,
                           const unsigned long long* dev_offset, int dtype, void* stream);

/* Element dropout, torch.nn.functional.dropout(x, p, training=True) semantics (HF modeling_data2vec_audio.py
 * hidden_dropout / activation_dropout / feat_proj_dropout :160-172, 300-346, 359-386; final_dropout
 * federated/src/models.py:403):  y = res + keep(x) / (1 - p)   (res may be NULL; y may alias x).
 * The mask is not stored: it is a pure function of (seed, site, offset + *dev_offset, element index), so the backward
 * pass calls the same entry point on dy with the same keys.  `dev_offset` (device u64, may be NULL) lets a launch
 * captured in a CUDA graph draw a fresh mask per replay.  n % (16/sizeof) == 0; P(drop) is p rounded to 1/65536. */
int b200asr_dropout(const void* x, const void* res, void* y, long long n, float p, unsigned long long seed,
                    unsigned long long site, unsigned long long offset, const unsigned long long* dev_offset, int dtype,
                    void* stream);

/* ------------------------------------------------------------------------------------------------
 * CTC: log_softmax(fp32) + alpha/beta recursions + gradient w.r.t. the logits, one launch for `n_heads`
 * logit tensors sharing targets.  Replaces nn.functional.log_softmax + nn.functional.ctc_loss at
 * federated/src/models.py:498-533, centralized/trainer_data2vec_2st.py:384-408, HF:897-908.
 *   logits [n_heads, B, T, V] (dtype); targets int32 flat (concatenated, as `labels.masked_select`);
 *   tgt_off/tgt_len/in_len int32 [B]; loss fp32 [n_heads, B] = -log p(target) per utterance (inf -> 0
 *   when zero_infinity); grad (dtype) [n_heads, B, T, V] = grad_scale[h, b] * d loss[h, b] / d logits, or NULL
 *   (forward only); grad_scale fp32 [n_heads, B] or NULL (= 1: reduction "sum").
 *   workspace >= b200asr_ctc_workspace_bytes(...).
 */
size_t b200asr_ctc_workspace_bytes(int n_heads, int B, int T, int max_tgt_len);
int b200asr_ctc_fwd_bwd(const void* logits, const int* targets, const int* tgt_off, const int* tgt_len,
                        const int* in_len, const float* grad_scale, float* loss, void* grad, void* workspace,
                        size_t workspace_bytes,
                        int n_heads, int B, int T, int V, int max_tgt_len, int blank, int zero_infinity, int dtype,
                        void* stream);

/* Greedy decode (finetune_ASRs.py:113-122 `map_to_result`, update.py:162-178): ids = argmax(logits, -1);
 * out_ids int32 [B, T] raw argmax; collapsed int32 [B, T] (repeat-collapsed, blank dropped, padded with -1);
 * collapsed_len int32 [B].  in_len NULL = all T frames. */
int b200asr_greedy_decode(const void* logits, const int* in_len, int* out_ids, int* collapsed, int* collapsed_len,
                          int B, int T, int V, int blank, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * DACS toggling network, fused: score split + TOGGLE_RATIO lerp + Gumbel noise + softmax(2)/tau + hard
 * straight-through mask + mask*hidden for both branches.  Replaces federated/src/models.py:428-455 +
 * gumbel_softmax :75-129 (centralized/trainer_data2vec_2st.py:314-341, Models.py:76-130).
 *   scores [M, 4d] = arbitrator(h); h [M, d]
 *   noise_lm / noise_ad: fp32 [M, d, 2] Gumbel samples g = -log(Exp(1)) or NULL -> drawn in-kernel (Philox,
 *   seed/offset).  Outputs (dtype): lm_masked, ad_masked [M, d]; soft_lm, soft_ad [M, d] = y_soft[...,0]
 *   (saved for backward); lm_mask, ad_mask [M, d] in {0,1} (optional, may be NULL).
 */
int b200asr_dacs_mask_fwd(const void* scores, const void* h, const float* noise_lm, const float* noise_ad,
                          void* lm_masked, void* ad_masked, void* soft_lm, void* soft_ad, void* lm_mask,
                          void* ad_mask, long long M, int d, float tau, float toggle_ratio,
                          unsigned long long seed, unsigned long long offset, int dtype, void* stream);
/* Same, with the Philox offset advanced by *dev_offset (device u64, may be NULL): the launch can be captured in a CUDA
 * graph and still draw fresh Gumbel noise on every replay (the caller bumps the counter with a device-side add). */
int b200asr_dacs_mask_fwd_ctr(const void* scores, const void* h, const float* noise_lm, const float* noise_ad,
                              void* lm_masked, void* ad_masked, void* soft_lm, void* soft_ad, void* lm_mask, void* ad_mask,
                              long long M, int d, float tau, float toggle_ratio, unsigned long long seed,
                              unsigned long long offset, const unsigned long long* dev_offset, int dtype, void* stream);
/* Straight-through backward: d_scores [M, 4d] from d_lm_masked, d_ad_masked; dh [M, d] or NULL. */
int b200asr_dacs_mask_bwd(const void* d_lm_masked, const void* d_ad_masked, const void* h, const void* soft_lm,
                          const void* soft_ad, const void* lm_mask, const void* ad_mask, void* d_scores, void* dh,
                          long long M, int d, float tau, float toggle_ratio, int dtype, void* stream);

/* Mean over the time axis of per-frame AD logits (models.py:468-476): x [B, T, C] -> out fp32 [B, C]
 * (mean over all T frames, as the reference does -- padded frames included). */
int b200asr_mean_time_fwd(const void* x, float* out, int B, int T, int C, int dtype, void* stream);
/* dx[b,t,c] = dout[b,c] * scale / T */
int b200asr_mean_time_bwd(const float* dout, void* dx, int B, int T, int C, float scale, int dtype, void* stream);

/* RecallLoss / CrossEntropy on [B, 2] mean AD logits (models.py:187-260, :535-541).
 * kind: 0 recall, 1 prec, 2 f1, 3 recall_ori, 4 prec_ori, 5 cel.  w fp32 host [C] (normalised inside).
 * loss fp32 [1]; dlogits fp32 [B, C] = d loss / d logits (NULL = forward only). */
int b200asr_ad_loss(const float* logits, const long long* target, float* loss, float* dlogits, int B, int C,
                    int kind, const float* w_host, void* stream);

/* CosFace / AM-softmax diversity term without the N x N gather (models.py:131-185, call :592-607):
 * rows of x1 [n, d] carry label 0, rows of x2 [n, d] label 1, fc_w fp32 [2, d] (not normalised, as the
 * reference's no-op loop leaves it).  loss fp32 [1] (mean over 2n rows).  dx1/dx2 (dtype) = d loss / dx * gscale
 * or NULL.  workspace: fp32 >= 2n floats. */
int b200asr_cosface_fwd_bwd(const void* x1, const void* x2, const float* fc_w, float* loss, void* dx1, void* dx2,
                            void* workspace, size_t workspace_bytes, long long n, int d, float s, float m,
                            int dtype, void* stream);

/* ------------------------------------------------------------------------------------------------
 * FedAvg (federated/src/utils.py:240-249 `average_weights`), flat fp32 parameter buffers.
 *   out[i] = (((in0[i] + in1[i]) + in2[i]) + ...) / K      when weights == NULL  (bit-exact with the
 *            reference's left-to-right sum then torch.div)
 *   out[i] = sum_j weights[j] * in_j[i]                     otherwise
 * inputs: HOST array of K device pointers (K <= 16); weights: HOST array or NULL.  out may alias inputs[0].
 */
int b200asr_fedavg_scale_accum(const float* const* inputs, const float* weights, float* out, int K, long long n,
                               void* stream);
/* One rank's share of an N-rank FedAvg before the sum all-reduce: buf[i] *= w  (in place). */
int b200asr_fedavg_prescale(float* buf, float w, long long n, void* stream);
/* Fused multi-GPU FedAvg over NVLink peer memory: every rank passes the HOST array of the `world` ranks'
 * device pointers to the same-sized flat buffer (peer-mapped, e.g. CUDA IPC), its own rank, and per-rank
 * weights (NULL = uniform 1/world).  Rank r reduces slice r of every peer (scale-accumulate straight from
 * peer memory) and writes the result back into every peer's slice.  Caller brackets with cross-rank barriers. */
int b200asr_fedavg_peer_reduce(float* const* peer_bufs, const float* weights, int world, int rank, long long n,
                               void* stream);
/* Same, but the kernel returns without touching any buffer when *abort_flag (device int, the barrier's timeout flag)
 * is non-zero: a straggler that missed the barrier must not receive -- or contribute to -- a torn average. */
int b200asr_fedavg_peer_reduce_guarded(float* const* peer_bufs, const float* weights, int world, int rank, long long n,
                                       const int* abort_flag, void* stream);

/* NVLS form of the same exchange (uniform weights): `multicast_buf` is the multicast address of the ranks' flat buffers
 * (one NVSwitch multicast object bound to every rank's allocation, e.g. torch symmetric memory's multicast_ptr).  Rank r
 * reduces slice r IN THE SWITCH (multimem.ld_reduce.add.v4.f32), divides by `world` and multicasts the result into every
 * rank's buffer (multimem.st): 4P/N bytes in + 4P/N bytes out per GPU at any N.  Same bracketing barriers as above;
 * equal to the sequential reference sum to fp32 rounding (the switch fixes the order), not bit for bit.
 * max_blocks > 0 caps the grid (leaves SMs free for work overlapped with the exchange). */
int b200asr_fedavg_nvls_reduce(float* multicast_buf, int world, int rank, long long n, const int* abort_flag, int max_blocks,
                               void* stream);

/* Grid cap (thread blocks) of the two exchange kernels above; 0 = full grid.  Returns the previous cap.  Used when the
 * exchange is overlapped with compute (layer-wise pipelined FedAvg): the links saturate with a fraction of the SMs. */
int b200asr_fedavg_set_max_blocks(int max_blocks);

/* Cross-rank barrier over peer memory for the fused kernel above: every rank passes the HOST array of the `world`
 * ranks' flag buffers (>= world unsigned each, zero-initialised, peer-mapped) and a barrier number `epoch` that grows
 * by one per call on every rank.  Stream-ordered; bounded spin (sets *timeout_flag, a device int, instead of hanging). */
int b200asr_fedavg_peer_barrier(unsigned* const* peer_flags, int world, int rank, unsigned epoch, int* timeout_flag,
                                void* stream);
/* Spin limit of the barrier in seconds of SM clock (default ~2 s).  A rank whose flag is already set publishes a poison
 * value at its next barrier, which sets the flag on every peer that meets it (the failure is visible on all ranks). */
int b200asr_fedavg_peer_barrier_set_timeout(double seconds);
/* Peer-visible device allocations (zero-filled cudaMalloc, exportable through CUDA IPC) and the IPC plumbing that
 * replaces the reference's pickle / deepcopy hops of the weights between client processes and the server
 * (federated/src/federated_main.py:124-141, federated/src/update.py:246-283).  Explicit create / destroy pairs. */
int b200asr_peer_alloc(void** ptr, size_t bytes);
int b200asr_peer_free(void* ptr);
int b200asr_ipc_export(const void* ptr, unsigned char handle[64]);
int b200asr_ipc_open(const unsigned char handle[64], void** ptr);
int b200asr_ipc_close(void* ptr);

/* ------------------------------------------------------------------------------------------------
 * Input pipeline on the device.  Replaces Wav2Vec2Processor's zero-mean / unit-variance normalisation
 * (centralized/finetune_ASRs.py:88-96: (x - mean) / sqrt(var + 1e-7) per utterance) and DataCollatorCTCWithPadding's
 * padding + attention mask (finetune_ASRs.py:59-86) on ONE packed buffer of raw utterances.
 *   packed fp32: the B utterances back to back; offsets int64 [B+1] (device) their boundaries
 *   input_values fp32 [B, Lmax] (zero padded), attention_mask int32 [B, Lmax] or NULL,
 *   frame_lengths int32 [B] or NULL = _get_feat_extract_output_lengths (conv_kernel / conv_stride: DEVICE int32 [n_conv])
 */
size_t b200asr_normalize_pad_workspace_bytes(int B, long long max_len);
int b200asr_normalize_pad(const float* packed, const long long* offsets, float* input_values, int* attention_mask,
                          int* frame_lengths, void* workspace, size_t workspace_bytes, int B, int Lmax,
                          const int* conv_kernel, const int* conv_stride, int n_conv, void* stream);
/* Word error rate on the device (federated/src/update.py:36-50 compute_metrics -> jiwer.wer on the decoded
 * strings).  words_from_ids: ids int32 [B, ld] (negative = padding), len [B] or NULL; a word is a maximal run of
 * ids != delim and is represented by a 64-bit hash (delim < 0: every id is a token = character error rate);
 * words u64 [B, ldw], nwords [B].  edit_distance: Levenshtein distance per pair of token rows -> dist int32 [B];
 * WER = sum(dist) / sum(ref word counts). */
int b200asr_words_from_ids(const int* ids, const int* len, int B, int ld, int delim, unsigned long long* words, int* nwords,
                           int ldw, void* stream);
int b200asr_edit_distance(const unsigned long long* hyp, const int* hyp_len, const unsigned long long* ref, const int* ref_len,
                          int* dist, int B, int ld_hyp, int ld_ref, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused AdamW on flat fp32 buffers (HF Trainer "adamw_hf" semantics = transformers.AdamW with
 * correct_bias=True, configured at centralized/finetune_ASRs.py:202-220, federated/src/update.py:657-678):
 *   g = grad * clip_coef,  clip_coef = min(1, max_norm / (sqrt(*gnorm_sq) + 1e-6)) (gnorm_sq NULL -> 1)
 *   m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; p -= lr * sqrt(1-b2^t)/(1-b1^t) * m / (sqrt(v) + eps)
 *   p -= lr * wd * p       (decoupled, applied after the Adam update as transformers.AdamW does)
 * shadow: NULL or bf16 [n] refreshed with the new p (weight shadow for the tensor-core GEMMs).
 */
int b200asr_adamw_step(float* p, const float* grad, float* m, float* v, void* shadow_bf16, long long n, float lr,
                       float beta1, float beta2, float eps, float weight_decay, int step, const float* gnorm_sq,
                       float max_norm, void* stream);
/* Same update with the per-step scalars read from device memory: hyper fp32 [2] = {lr, lr * sqrt(1-b2^t) / (1-b1^t)}.
 * No host value changes between steps, so the launch can be captured in a CUDA graph and replayed. */
int b200asr_adamw_step_dev(float* p, const float* grad, float* m, float* v, void* shadow_bf16, long long n,
                           const float* hyper, float beta1, float beta2, float eps, float weight_decay,
                           const float* gnorm_sq, float max_norm, void* stream);
/* Same update with PER-PARAMETER scalars (transformers.AdamW keeps state["step"] per parameter and skips parameters
 * whose grad is None -- a block dropped by LayerDrop (HF:392-404), an unused head): the flat buffer is described by
 * `chunks` (device int32 [nchunks, 4] = {first float4 index, float4 count, parameter index, 0}; a chunk never straddles
 * a parameter) and `hyper` (device fp32 [n_params, 2] = {lr, lr * sqrt(1-b2^t_p) / (1-b1^t_p)}); a NEGATIVE second entry
 * marks the parameter idle in this step: p / m / v / shadow are left untouched.  Graph-capturable. */
int b200asr_adamw_step_chunks(float* p, const float* grad, float* m, float* v, void* shadow_bf16, const int* chunks,
                              int nchunks, const float* hyper, float beta1, float beta2, float eps, float weight_decay,
                              const float* gnorm_sq, float max_norm, void* stream);
/* out[0] += sum(x^2)  (out must be zeroed by the caller for a fresh norm).  Deterministic (fixed-order combine).
 * b200asr_sumsq parks its per-block partials in module-global device memory: one launch in flight per device.
 * b200asr_sumsq_ws takes a caller-owned workspace (zero-initialised once) and may run on any number of streams. */
int b200asr_sumsq(const float* x, long long n, float* out, void* stream);
size_t b200asr_sumsq_workspace_bytes(void);
int b200asr_sumsq_ws(const float* x, long long n, float* out, void* workspace, size_t workspace_bytes, void* stream);

/* Frame lengths after the conv feature extractor, `_get_feat_extract_output_lengths` (HF:507-526, centralized/Models.py:249-271):
 * out[i] = the chain floor((len - k_j) / s_j) + 1 over the n_conv (<= 16) layers applied to lengths[i] (int64, e.g.
 * attention_mask.sum(-1)); floor division as torch.div(rounding_mode="floor").  conv_kernel / conv_stride are HOST arrays.
 * One launch instead of three element-wise kernels per layer on the critical path of every training step. */
int b200asr_conv_out_lengths(const long long* lengths, int* out, int n, const int* conv_kernel, const int* conv_stride,
                             int n_conv, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ASR_H_ */
