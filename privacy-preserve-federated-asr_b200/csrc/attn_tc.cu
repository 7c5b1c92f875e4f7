// Fused multi-head self-attention on the 5th-generation tensor cores (bf16, head_dim 64), replacing the
// bmm -> softmax -> bmm of Data2VecAudioAttention (HF:189-197) without ever materialising the [B,H,T,T]
// probabilities in HBM on the forward pass.
//
// Forward  (one CTA = 128 queries of one (utterance, head); keys stream through in blocks of 64):
//   warp 0      TMA producer: Q tile once, then a ring of {K_j, V_j} tiles (128B-swizzled boxes of the packed
//               [B,T,3,H,dh] projection buffer, out-of-range frames zero-filled by TMA)
//   warp 1      TMEM owner + single-thread tcgen05.mma issuer:  S_j = Q K_j^T  (double buffered in TMEM),
//               O += P_j V_j
//   warps 2..5  online softmax, one query row per thread (TMEM lane): tcgen05.ld S_j, running max / sum in the
//               log2 domain, P_j -> bf16 -> swizzled shared memory (the A operand of the P V product),
//               O rescale in TMEM only when a row maximum moved, final O / l -> bf16 -> HBM, log-sum-exp saved.
// Backward (one CTA = 128 keys of one (utterance, head); queries stream through in blocks of 64):
//   S^T = K_j Q_i^T and dP^T = V_j dO_i^T land in TMEM (three accumulator pairs: two blocks ahead of the element-wise warps); 8 element-wise warps rebuild
//   P^T = exp2(S^T c - lse) and dS^T = P^T (dP^T - D) scale, one key row per thread (no cross-thread reduction
//   is needed in this orientation), write both as bf16 A operands into swizzled shared memory and dS^T to HBM;
//   dV += P^T dO_i and dK += dS^T Q_i accumulate in TMEM over the whole query loop.  dQ = dS K is one batched
//   tcgen05 GEMM over the stored dS^T (it stays L2-resident between the two kernels at the sizes of this path).
#include <cuda.h>

#include "attn.h"
#include "common.cuh"
#include "gemm.h"
#include "tc_ptx.cuh"

namespace b200asr {
namespace attn {

using namespace tc;

constexpr int DH = 64;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

// ------------------------------------------------------------------------------------------ forward
constexpr int FQ = 128;       // queries per CTA (= TMEM lanes)
constexpr int FK = 64;        // keys per block
constexpr int F_STAGES = 3;
constexpr int F_SW = 8;         // softmax warps: two per TMEM lane quarter, each thread owns HALF a query row (32 of the
                                // block's 64 keys): the per-block chain (max, exp2, sum, pack) of a thread is latency bound
constexpr int F_THREADS = 64 + 32 * F_SW;
constexpr int F_Q_BYTES = FQ * DH * 2;        // 16 KB
constexpr int F_KV_BYTES = 2 * FK * DH * 2;   // K + V: 16 KB
constexpr int F_P_BYTES = FQ * FK * 2;        // 16 KB
constexpr int F_XCH_BYTES = 2 * FQ * 2 * 4;   // row maxima of the two half-row threads, double buffered by block parity
constexpr int F_SMEM = F_Q_BYTES + F_STAGES * F_KV_BYTES + F_P_BYTES + 1024 + 256 + F_XCH_BYTES;
constexpr int F_TMEM_COLS = 256;              // S[2] (2 x 64) + O (64) -> next power of two

struct FwdParams {
    __nv_bfloat16* out;
    float* lse;
    const int* klen;
    int B, T, H;
    float scale_log2;  // softmax scale * log2(e)
    // attention-probability dropout (HF:189-197 `nn.functional.dropout(attn_weights, p=self.dropout)`): keep bits drawn by
    // attn_keep_mask_kernel, bit (key & 31) of word [((b H + h) KW + key / 32) TQ + query]; dropped probabilities leave
    // the P V product, the row sum (softmax normalisation) and the saved log-sum-exp are those of the undropped softmax
    const uint32_t* keep;
    int KW, TQ;
    float inv_keep;    // 1 / (1 - p)
    unsigned long long* trace;   // diagnostics (b200asr_attn_set_trace): 24 time stamps per CTA, null in normal use
};
static unsigned long long* g_attn_trace = nullptr;
// share of the softmax exponentials computed by the FMA-pipe polynomial (exp2_poly2): 0 = none, 4 = every 4th pair (25 %),
// 2 = every 2nd (50 %); forward / backward kernel (b200asr_mha_set_exp2_poly)
static int g_exp2_poly_fwd = 0, g_exp2_poly_bwd = 0;
#define ATRACE(slot) do { if (p.trace) p.trace[(long long)((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 24 + (slot)] = (unsigned long long)clock64(); } while (0)

// PM: every PM-th pair of probabilities is computed by exp2_poly2 on the FMA pipe instead of two MUFU ex2 (0 = none): the
// exp2 / sum / pack phase of a key block was bound by the MUFU's 16 results per clock shared by the 16 softmax warps of an SM
template <bool DROP, int PM>
__global__ void __launch_bounds__(F_THREADS, 2)
mha_fwd_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                  const __grid_constant__ CUtensorMap map_v, const FwdParams p) {
    if (threadIdx.x == 0) ATRACE(0);
    pdl_enter();
    if (threadIdx.x == 0) ATRACE(1);  // launched with programmatic stream serialization: wait for the producers before touching global memory
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t sQ = base;
    const uint32_t sKV = sQ + F_Q_BYTES;
    const uint32_t sP = sKV + F_STAGES * F_KV_BYTES;
    const uint32_t bars = sP + F_P_BYTES;
    uint8_t* genP = gen + F_Q_BYTES + F_STAGES * F_KV_BYTES;
    const uint32_t q_full = bars;
    auto kv_full = [&](int s) { return bars + 8u * (1 + s); };
    auto kv_empty = [&](int s) { return bars + 8u * (1 + F_STAGES + s); };
    auto s_full = [&](int s) { return bars + 8u * (1 + 2 * F_STAGES + s); };
    auto s_free = [&](int s) { return bars + 8u * (3 + 2 * F_STAGES + s); };
    const uint32_t p_ready = bars + 8u * (5 + 2 * F_STAGES);
    const uint32_t pv_done = bars + 8u * (6 + 2 * F_STAGES);
    const uint32_t tmem_slot = bars + 8u * (7 + 2 * F_STAGES);
    volatile uint32_t* tmem_slot_gen =
        reinterpret_cast<volatile uint32_t*>(genP + F_P_BYTES + 8 * (7 + 2 * F_STAGES));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * FQ, h = blockIdx.y, b = blockIdx.z;
    const int T = p.T, H = p.H;
    const int kl = p.klen ? min(max(p.klen[b], 0), T) : T;
    const int nkb = (kl + FK - 1) / FK;
    const int D = H * DH;

    if (nkb == 0) {  // no valid key: probabilities are all zero (softmax_rows_kernel convention), lse = -inf
        if (warp >= 2 && warp < 6) {
            const int row = (warp & 3) * 32 + lane, t = q0 + row;
            if (t < T) {
                uint4* o = reinterpret_cast<uint4*>(p.out + ((long long)b * T + t) * D + h * DH);
#pragma unroll
                for (int c = 0; c < 8; ++c) o[c] = make_uint4(0, 0, 0, 0);
                p.lse[((long long)b * H + h) * T + t] = -INFINITY;
            }
        }
        return;
    }

    if (warp == 0 && elect_one_sync()) {
        tma_prefetch_desc(&map_q);
        tma_prefetch_desc(&map_k);
        tma_prefetch_desc(&map_v);
        mbar_init(q_full, 1);
        for (int s = 0; s < F_STAGES; ++s) {
            mbar_init(kv_full(s), 1);
            mbar_init(kv_empty(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(s_full(s), 1);
            mbar_init(s_free(s), F_SW);
        }
        mbar_init(p_ready, F_SW);
        mbar_init(pv_done, 1);
        fence_barrier_init();
        // the operand loads start now and overlap the TMEM allocation
        mbar_expect_tx(q_full, F_Q_BYTES);
        tma_load_4d(sQ, &map_q, q_full, 0, q0, h, b);
        for (int j = 0; j < nkb && j < F_STAGES; ++j) {
            mbar_expect_tx(kv_full(j), F_KV_BYTES);
            tma_load_4d(sKV + j * F_KV_BYTES, &map_k, kv_full(j), 0, j * FK, h, b);
            tma_load_4d(sKV + j * F_KV_BYTES + FK * DH * 2, &map_v, kv_full(j), 0, j * FK, h, b);
        }
    }
    if (warp == 1) tmem_alloc(tmem_slot, F_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_gen;
    const uint32_t tmem_o = tmem_base + 128;
    if (threadIdx.x == 0) ATRACE(2);

    if (warp == 0) {
        if (elect_one_sync()) {
            for (int j = F_STAGES; j < nkb; ++j) {   // (Q and the first F_STAGES key blocks were requested during set-up)
                const int s = j % F_STAGES;
                mbar_wait(kv_empty(s), ((j / F_STAGES) & 1) ^ 1u);
                mbar_expect_tx(kv_full(s), F_KV_BYTES);
                tma_load_4d(sKV + s * F_KV_BYTES, &map_k, kv_full(s), 0, j * FK, h, b);
                tma_load_4d(sKV + s * F_KV_BYTES + FK * DH * 2, &map_v, kv_full(s), 0, j * FK, h, b);
            }
        }
    } else if (warp == 1) {
        if (elect_one_sync()) {
            const uint32_t idesc_s = make_idesc_bf16(FQ, FK, 0, 0);
            const uint32_t idesc_pv = make_idesc_bf16(FQ, DH, 0, 1);
            auto issue_s = [&](int j) {
                const int s = j % F_STAGES;
                mbar_wait(kv_full(s), (j / F_STAGES) & 1);
                if (j >= 2) mbar_wait(s_free(j & 1), ((j >> 1) - 1) & 1);
                tc_fence_after();
                const uint32_t sk = sKV + s * F_KV_BYTES;
#pragma unroll
                for (int k = 0; k < DH / 16; ++k)
                    umma_bf16(tmem_base + (uint32_t)((j & 1) * FK), make_smem_desc(sQ + k * 32, 16, 1024),
                              make_smem_desc(sk + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
                umma_commit(s_full(j & 1));
            };
            mbar_wait(q_full, 0);
            ATRACE(3);
            issue_s(0);
            for (int j = 0; j < nkb; ++j) {
                if (j + 1 < nkb) issue_s(j + 1);
                mbar_wait(p_ready, j & 1);
                tc_fence_after();
                const int s = j % F_STAGES;
                const uint32_t sv = sKV + s * F_KV_BYTES + FK * DH * 2;
#pragma unroll
                for (int k = 0; k < FK / 16; ++k)
                    umma_bf16(tmem_o, make_smem_desc(sP + k * 32, 16, 1024), make_smem_desc(sv + k * 2048, 8192, 1024),
                              idesc_pv, (j > 0 || k > 0) ? 1u : 0u);
                umma_commit(kv_empty(s));
                umma_commit(pv_done);
            }
            ATRACE(4);
        }
    } else {
        constexpr int HC = FK / 2;   // key columns per thread and block (32)
        const int quarter = warp & 3, half = (warp - 2) >> 2;
        const int row = quarter * 32 + lane;
        const uint32_t tlane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const uint32_t tlane_o = tlane + 128 + (uint32_t)(half * 32);   // this thread's 32 of the 64 output columns
        float* xch = reinterpret_cast<float*>(genP + F_P_BYTES + 256);  // [parity][row][half]
        float m = -INFINITY, l = 0.f;   // running row maximum (common to both halves), this half's running sum
        const float c2 = p.scale_log2;
        // keep bits of this thread's query row: word kw = 2 j + half covers its 32 keys of block j (coalesced across the warp)
        const uint32_t* keep_row = nullptr;
        if constexpr (DROP) keep_row = p.keep + ((long long)b * H + h) * p.KW * (long long)p.TQ + (q0 + row);
        for (int j = 0; j < nkb; ++j) {
            const bool trc = threadIdx.x == 64 && j == 2;
            if (trc) ATRACE(8);
            uint32_t kbits = 0xffffffffu;
            if constexpr (DROP) {
                const int kw = 2 * j + half;
                if (kw < p.KW) kbits = __ldg(keep_row + (long long)kw * p.TQ);   // (requested before the wait on S_j)
            }
            mbar_wait(s_full(j & 1), (j >> 1) & 1);
            tc_fence_after();
            if (threadIdx.x == 64 && j == 3) ATRACE(15);
            if (trc) ATRACE(9);
            uint32_t r[HC];
            tmem_ld_32x32(tlane + (uint32_t)((j & 1) * FK + half * HC), r);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_free(j & 1));
            if (trc) ATRACE(10);
            const int nvalid = kl - j * FK - half * HC;  // valid keys among this thread's columns (may be <= 0)
            float mx0 = -INFINITY, mx1 = -INFINITY;
            if (nvalid >= HC) {
#pragma unroll
                for (int c = 0; c < HC; c += 2) {
                    mx0 = fmaxf(mx0, __uint_as_float(r[c]));
                    mx1 = fmaxf(mx1, __uint_as_float(r[c + 1]));
                }
            } else {
#pragma unroll
                for (int c = 0; c < HC; ++c) {
                    if (c >= nvalid) r[c] = __float_as_uint(-INFINITY);
                    mx0 = fmaxf(mx0, __uint_as_float(r[c]));
                }
            }
            // the other half of the row lives in the warp four further on (same lane): exchange the block maxima
            float* xr = xch + ((j & 1) * FQ + row) * 2;
            xr[half] = fmaxf(mx0, mx1);
            named_bar_sync(1 + quarter, 64);
            if (trc) ATRACE(11);
            const float mx = fmaxf(xr[0], xr[1]);   // >= one valid key in the block: finite
            // lazy rescaling: the reference maximum only moves when the block maximum exceeds it by more than 2^8 (the
            // probabilities then stay <= 256, exact enough in fp32 sums and bf16 operands); O in TMEM is rescaled only then
            const float mb = mx * c2;
            const float m_new = (mb > m + 8.0f) ? mb : m;
            const float alpha = fast_exp2(m - m_new);
            float sum0 = 0.f, sum1 = 0.f;
            uint32_t pk[HC / 2];
#pragma unroll
            for (int c = 0; c < HC; c += 2) {
                float p0, p1;
                if (PM > 0 && ((c >> 1) % (PM > 0 ? PM : 1)) == (PM > 0 ? PM : 1) - 1) {
                    const float2 pq = exp2_poly2(make_float2(fmaf(__uint_as_float(r[c]), c2, -m_new),
                                                             fmaf(__uint_as_float(r[c + 1]), c2, -m_new)));
                    p0 = pq.x; p1 = pq.y;
                } else {
                    p0 = fast_exp2(fmaf(__uint_as_float(r[c]), c2, -m_new));
                    p1 = fast_exp2(fmaf(__uint_as_float(r[c + 1]), c2, -m_new));
                }
                sum0 += p0;
                sum1 += p1;
                if constexpr (DROP) pk[c >> 1] = bf2_pack((kbits >> c) & 1u ? p0 : 0.f, (kbits >> (c + 1)) & 1u ? p1 : 0.f);
                else pk[c >> 1] = bf2_pack(p0, p1);
            }
            l = fmaf(l, alpha, sum0 + sum1);
            m = m_new;
            if (trc) ATRACE(12);
            if (j > 0) {
                mbar_wait(pv_done, (j - 1) & 1);
                tc_fence_after();
            }
            if (trc) ATRACE(13);
#pragma unroll
            for (int c = 0; c < HC / 8; ++c)
                *reinterpret_cast<uint4*>(genP + sw128_offset(row, half * (HC / 8) + c)) =
                    make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
            if (j > 0 && __any_sync(0xffffffffu, alpha != 1.0f)) {   // (both halves see the same alpha)
                uint32_t o[32];
                tmem_ld_32x32(tlane_o, o);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
                tmem_st_32x32(tlane_o, o);
                tmem_st_wait();
            }
            fence_proxy_async();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_ready);
            if (trc) ATRACE(14);
        }
        if (threadIdx.x == 64) ATRACE(5);
        // row sum = the two halves' sums (same running maximum)
        float* xl = xch + ((nkb & 1) * FQ + row) * 2;
        xl[half] = l;
        named_bar_sync(1 + quarter, 64);
        const float lt = xl[0] + xl[1];
        mbar_wait(pv_done, (nkb - 1) & 1);
        tc_fence_after();
        const int t = q0 + row;
        const float inv = DROP ? p.inv_keep / lt : 1.0f / lt;
        __nv_bfloat16* orow = p.out + ((long long)b * T + t) * D + h * DH + half * 32;
        {
            uint32_t o[32];
            tmem_ld_32x32(tlane_o, o);
            tmem_ld_wait();
            if (t < T) {
#pragma unroll
                for (int c = 0; c < 32; c += 8) {
                    uint4 v;
                    v.x = bf2_pack(__uint_as_float(o[c]) * inv, __uint_as_float(o[c + 1]) * inv);
                    v.y = bf2_pack(__uint_as_float(o[c + 2]) * inv, __uint_as_float(o[c + 3]) * inv);
                    v.z = bf2_pack(__uint_as_float(o[c + 4]) * inv, __uint_as_float(o[c + 5]) * inv);
                    v.w = bf2_pack(__uint_as_float(o[c + 6]) * inv, __uint_as_float(o[c + 7]) * inv);
                    *reinterpret_cast<uint4*>(orow + c) = v;
                }
            }
        }
        if (half == 0 && t < T) p.lse[((long long)b * H + h) * T + t] = (m + log2f(lt)) * LN2;
        if (threadIdx.x == 64) ATRACE(6);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, F_TMEM_COLS);
    }
    if (threadIdx.x == 0) ATRACE(7);
}

// ------------------------------------------------------------------------------------------ backward
constexpr int BK = 128;       // keys per CTA (= TMEM lanes)
constexpr int BQ = 64;        // queries per block
constexpr int B_STAGES = 5;    // Q_i + dO_i ring: S^T / dP^T run ahead of dV / dK, which still reads block i
constexpr int B_EW = 16;        // element-wise warps: four per TMEM lane quarter, 16 query columns each (the per-block chain
                                // of one thread is latency bound; twice the warps of the first version hide twice the latency)
constexpr int B_KV_BYTES = 2 * BK * DH * 2;   // K_j + V_j: 32 KB
constexpr int B_QD_BYTES = 2 * BQ * DH * 2;   // Q_i + dO_i: 16 KB
constexpr int B_PT_BYTES = BK * BQ * 2;       // one P^T (or dS^T) operand tile: 16 KB
constexpr int B_DQS_BYTES = BQ * DH * 4;      // fp32 staging tile of one block's dQ contribution: 16 KB
constexpr int B_TMEM_COLS = 512;

// Two builds of the kernel:
//   FDQ = false  dS^T goes to HBM (one TMA store per query block) and dQ = dS K is a batched GEMM afterwards: every dQ
//                element is one fixed-order sum -> bit-reproducible; three S^T / dP^T accumulator pairs in TMEM.
//   FDQ = true   dQ_i += dS_i K_j is a fifth MMA of the query-block loop (A = the dS^T tile read MN-major, B = K_j);
//                two drain warps move the 64 x 64 fp32 partial from TMEM through shared memory into the fp32 accumulator
//                [B,H,T,64] with ONE bulk reduce-add (cp.reduce.async.bulk .add.f32, 16 KB) per block.  No dS^T round trip
//                through HBM (64 MB per layer at B=8 x 10 s) and no extra GEMM; the T / 128 key CTAs of a head add in
//                arrival order, so dQ is reproducible only to fp32 rounding.  Two S^T / dP^T pairs (dQ takes 128 columns).
template <bool FDQ> struct BwdCfg {
    static constexpr int SDP = FDQ ? 2 : 3;
    static constexpr int THREADS = FDQ ? 32 * 22 : 96 + 32 * B_EW;   // FDQ: warps 20 / 21 drain dQ (TMEM lane quarters 0 / 1)
    static constexpr int COL_DP = 64 * SDP;        // S^T[s] at 64 s, dP^T[s] at COL_DP + 64 s
    static constexpr int COL_DV = 128 * SDP;
    static constexpr int COL_DK = COL_DV + 64;
    static constexpr int COL_DQ = COL_DK + 64;     // FDQ: dQ[2] (M = 128 accumulators, rows 0..63 used)
    static constexpr int NBAR = 10 + 2 * SDP + 2 * B_STAGES;
    // K,V | P^T[2], dS^T[2] | Q/dO ring | dQ staging | barriers | lse, D
    static constexpr int SMEM_FIXED = B_KV_BYTES + 4 * B_PT_BYTES + B_STAGES * B_QD_BYTES + (FDQ ? B_DQS_BYTES : 0) + 1024 + 256;
};
static_assert(BwdCfg<true>::COL_DQ + 128 <= B_TMEM_COLS && BwdCfg<false>::COL_DK + 64 <= B_TMEM_COLS, "TMEM columns");
static_assert(BwdCfg<false>::NBAR < 32 && BwdCfg<true>::NBAR < 32, "one barrier per lane, 256-byte barrier area");

struct BwdParams {
    const float* lse;   // [B,H,T] natural-log sum-exp of the scaled logits
    const float* dsum;  // [B,H,T] D = rowsum(dO * O)
    const int* klen;
    __nv_bfloat16* dqkv;  // [B,T,3,H,dh]: this kernel writes the K and V thirds
    __nv_bfloat16* dst;   // !FDQ: dS^T [B,H,T(key),ldq] (scaled by `scale`)
    float* dq_acc;        // FDQ: fp32 [B,H,T,64], zeroed by mha_bwd_dsum_kernel, converted by mha_bwd_dq_store_kernel
    int B, T, H, ldq;
    float scale, scale_log2;
    const uint32_t* keep;   // attention-probability dropout keep bits (see FwdParams) or unused
    int KW, TQ;
    float inv_keep;
    int skip_reduce;        // measurement hook (b200asr_mha_set_dq_fused(2)): run the fused-dQ kernel without its reduce-adds
    unsigned long long* trace;   // diagnostics (b200asr_attn_set_trace): 24 time stamps per CTA, null in normal use
};

__device__ __forceinline__ void bulk_reduce_add_f32(float* gdst, uint32_t ssrc, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                 ::"l"(gdst), "r"(ssrc), "r"(bytes) : "memory");
}

template <bool DROP, bool FDQ, int PM = 0>   // PM: share of the exponentials on the FMA pipe, as in the forward kernel
__global__ void __launch_bounds__(BwdCfg<FDQ>::THREADS, 1)
mha_bwd_tc_kernel(const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                  const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_do,
                  const __grid_constant__ CUtensorMap map_dst, const BwdParams p) {
    using Cfg = BwdCfg<FDQ>;
    constexpr int SDP = Cfg::SDP;
    if (threadIdx.x == 0) ATRACE(0);
    pdl_enter();  // launched with programmatic stream serialization: wait for the producers before touching global memory
    if (threadIdx.x == 0) ATRACE(1);
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t sK = base, sV = base + BK * DH * 2;
    const uint32_t sPT = base + B_KV_BYTES;             // PT[0], PT[1], DST[0], DST[1]
    uint8_t* genPT = gen + B_KV_BYTES;
    const uint32_t sQD = sPT + 4 * B_PT_BYTES;          // (the 16 KB after DST[1]: second, unused M block of the dQ MMA's A operand)
    const uint32_t sDQS = sQD + B_STAGES * B_QD_BYTES;
    constexpr int OFF_BARS = B_KV_BYTES + 4 * B_PT_BYTES + B_STAGES * B_QD_BYTES + (FDQ ? B_DQS_BYTES : 0);
    const uint32_t bars = base + OFF_BARS;
    const uint32_t kv_full = bars;
    auto qd_full = [&](int s) { return bars + 8u * (1 + s); };
    auto qd_empty = [&](int s) { return bars + 8u * (1 + B_STAGES + s); };
    auto sdp_full = [&](int s) { return bars + 8u * (1 + 2 * B_STAGES + s); };
    auto sdp_free = [&](int s) { return bars + 8u * (1 + SDP + 2 * B_STAGES + s); };
    auto pds_ready = [&](int s) { return bars + 8u * (1 + 2 * SDP + 2 * B_STAGES + s); };
    auto pds_free = [&](int s) { return bars + 8u * (3 + 2 * SDP + 2 * B_STAGES + s); };
    const uint32_t dkv_done = bars + 8u * (5 + 2 * SDP + 2 * B_STAGES);
    auto dq_full = [&](int s) { return bars + 8u * (6 + 2 * SDP + 2 * B_STAGES + s); };
    auto dq_free = [&](int s) { return bars + 8u * (8 + 2 * SDP + 2 * B_STAGES + s); };
    const uint32_t tmem_slot = bars + 8u * Cfg::NBAR;
    static_assert(8 * (Cfg::NBAR + 1) <= 256, "barrier area");
    volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen + OFF_BARS + 8 * Cfg::NBAR);
    // whole-row softmax statistics of this (utterance, head), staged once: lse * log2(e) then D, Tpad floats each
    // (Tpad = T rounded up to the 64-query block; the padding holds +inf / 0 so that out-of-range queries give P = 0)
    const int Tpad = ((p.T + BQ - 1) / BQ) * BQ;
    float* lse2_s = reinterpret_cast<float*>(gen + OFF_BARS + 256);
    float* dsum_s = lse2_s + Tpad;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k0 = blockIdx.x * BK, h = blockIdx.y, b = blockIdx.z;
    const int T = p.T, H = p.H, ldq = p.ldq;
    const int kl = p.klen ? min(max(p.klen[b], 0), T) : T;
    const int nqb = (T + BQ - 1) / BQ;
    const int D = H * DH;
    const long long bh = (long long)b * H + h;

    // this thread's share of the softmax statistics: the loads fly while the barriers / TMEM are set up
    float pre_lse = INFINITY, pre_d = 0.f;
    {
        const int q = threadIdx.x - 64;
        if (q >= 0 && q < T && k0 < kl) {
            pre_lse = __ldg(p.lse + bh * T + q) * LOG2E;
            pre_d = __ldg(p.dsum + bh * T + q);
        }
    }
    if (k0 >= kl) {  // fully masked key block: dK = dV = 0, no contribution to dQ (dS^T rows = 0)
        if (warp >= 2) {
            const int et = threadIdx.x - 64;  // first 256: two threads per key row
            const int row = et >> 1, half = et & 1, key = k0 + row;
            if (et < 2 * BK && key < T) {
                __nv_bfloat16* dk = p.dqkv + (((long long)b * T + key) * 3 + 1) * D + h * DH + half * 32;
                __nv_bfloat16* dv = dk + D;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    reinterpret_cast<uint4*>(dk)[c] = make_uint4(0, 0, 0, 0);
                    reinterpret_cast<uint4*>(dv)[c] = make_uint4(0, 0, 0, 0);
                }
                if constexpr (!FDQ) {
                    uint4* drow = reinterpret_cast<uint4*>(p.dst + (bh * T + key) * ldq);
                    for (int c = half; c < ldq / 8; c += 2) drow[c] = make_uint4(0, 0, 0, 0);
                }
            }
        }
        return;
    }

    if (warp == 0) {   // one barrier per lane (they are laid out back to back), then one lane starts the operand loads
        if (lane < Cfg::NBAR) {
            const int first_ew = 1 + SDP + 2 * B_STAGES, first_pf = 3 + 2 * SDP + 2 * B_STAGES;   // sdp_free, pds_free
            const int first_dqf = 8 + 2 * SDP + 2 * B_STAGES;                                     // dq_free
            // sdp_free / pds_ready: one arrival per element-wise warp; pds_free: the MMAs that read P^T / dS^T have
            // completed (+ the dS^T tile has been stored, !FDQ); dq_free: the two drain warps; everything else: one arrival
            uint32_t cnt = 1;
            if (lane >= first_ew && lane < first_pf) cnt = B_EW;
            else if (lane >= first_pf && lane < first_pf + 2) cnt = FDQ ? 1 : 2;
            else if (lane >= first_dqf && lane < first_dqf + 2) cnt = 2;
            mbar_init(bars + 8u * lane, cnt);
        }
        fence_barrier_init();
        __syncwarp();
    }
    if (warp == 0 && elect_one_sync()) {
        tma_prefetch_desc(&map_k);
        tma_prefetch_desc(&map_v);
        tma_prefetch_desc(&map_q);
        tma_prefetch_desc(&map_do);
        if constexpr (!FDQ) tma_prefetch_desc(&map_dst);
        // the operand loads start now: they overlap the TMEM allocation and the staging of the softmax statistics below
        mbar_expect_tx(kv_full, B_KV_BYTES);
        tma_load_4d(sK, &map_k, kv_full, 0, k0, h, b);
        tma_load_4d(sV, &map_v, kv_full, 0, k0, h, b);
        for (int i = 0; i < nqb && i < B_STAGES; ++i) {
            mbar_expect_tx(qd_full(i), B_QD_BYTES);
            tma_load_4d(sQD + i * B_QD_BYTES, &map_q, qd_full(i), 0, i * BQ, h, b);
            tma_load_4d(sQD + i * B_QD_BYTES + BQ * DH * 2, &map_do, qd_full(i), 0, i * BQ, h, b);
        }
    }
    if (warp == 1) tmem_alloc(tmem_slot, B_TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_gen;
    // the softmax statistics are only read by the element-wise warps: they stage them among themselves (named barrier)
    // while the MMA issuer already queues the first S^T / dP^T products
    if (warp >= 2 && warp < 2 + B_EW) {
        const int q0s = threadIdx.x - 64;
        if (q0s < Tpad) { lse2_s[q0s] = pre_lse; dsum_s[q0s] = pre_d; }   // (requested before the set-up barrier)
        for (int q = q0s + 32 * B_EW; q < Tpad; q += 32 * B_EW) {
            lse2_s[q] = (q < T) ? p.lse[bh * T + q] * LOG2E : INFINITY;
            dsum_s[q] = (q < T) ? p.dsum[bh * T + q] : 0.f;
        }
        named_bar_sync(1, 32 * B_EW);
    }
    if (threadIdx.x == 64) ATRACE(2);
    const uint32_t tm_dv = tmem_base + Cfg::COL_DV, tm_dk = tmem_base + Cfg::COL_DK;

    if (warp == 0) {
        if (elect_one_sync()) {
            for (int i = B_STAGES; i < nqb; ++i) {   // (K, V and the first B_STAGES blocks were requested during set-up)
                const int s = i % B_STAGES;
                mbar_wait(qd_empty(s), ((i / B_STAGES) & 1) ^ 1u);
                mbar_expect_tx(qd_full(s), B_QD_BYTES);
                tma_load_4d(sQD + s * B_QD_BYTES, &map_q, qd_full(s), 0, i * BQ, h, b);
                tma_load_4d(sQD + s * B_QD_BYTES + BQ * DH * 2, &map_do, qd_full(s), 0, i * BQ, h, b);
            }
        }
    } else if (warp == 1) {
        if (elect_one_sync()) {
            const uint32_t idesc_kk = make_idesc_bf16(BK, BQ, 0, 0);  // S^T, dP^T: both operands K-major
            const uint32_t idesc_km = make_idesc_bf16(BK, DH, 0, 1);  // dV, dK: B operand MN-major
            const uint32_t idesc_mm = make_idesc_bf16(128, DH, 1, 1); // dQ: A = dS^T tile read MN-major, B = K_j MN-major
            const uint64_t desc_k = make_smem_desc(sK, 16, 1024), desc_v = make_smem_desc(sV, 16, 1024);
            const uint64_t desc_k_mn = make_smem_desc(sK, 8192, 1024);   // K_j as the [dh x keys] B operand of dQ
            auto issue_sdp = [&](int i) {
                const int s = i % B_STAGES;
                const int a3 = i % SDP;
                mbar_wait(qd_full(s), (i / B_STAGES) & 1);
                if (i >= SDP) mbar_wait(sdp_free(a3), ((i / SDP) - 1) & 1);
                tc_fence_after();
                const uint32_t sq = sQD + s * B_QD_BYTES, sdo = sq + BQ * DH * 2;
                const uint32_t t_st = tmem_base + (uint32_t)(a3 * BQ), t_dp = tmem_base + Cfg::COL_DP + (uint32_t)(a3 * BQ);
                const uint64_t dq = make_smem_desc(sq, 16, 1024), ddo = make_smem_desc(sdo, 16, 1024);
#pragma unroll
                for (int k = 0; k < DH / 16; ++k) {
                    umma_bf16(t_st, smem_desc_advance(desc_k, 2 * k), smem_desc_advance(dq, 2 * k), idesc_kk, k > 0 ? 1u : 0u);
                    umma_bf16(t_dp, smem_desc_advance(desc_v, 2 * k), smem_desc_advance(ddo, 2 * k), idesc_kk, k > 0 ? 1u : 0u);
                }
                umma_commit(sdp_full(a3));
            };
            mbar_wait(kv_full, 0);
            ATRACE(3);
            // S^T / dP^T of block i + SDP - 1 are queued before this thread waits for the element-wise warps of block i:
            // the tensor pipe always has the next block's logits ready when they come back
            constexpr int LA = SDP - 1;
            for (int i = 0; i < LA && i < nqb; ++i) issue_sdp(i);
            for (int i = 0; i < nqb; ++i) {
                if (i + LA < nqb) issue_sdp(i + LA);
                if (i == 2) ATRACE(16);
                mbar_wait(pds_ready(i & 1), (i >> 1) & 1);
                tc_fence_after();
                if (i == 2) ATRACE(17);
                const int s = i % B_STAGES;
                const uint32_t sq = sQD + s * B_QD_BYTES, sdo = sq + BQ * DH * 2;
                const uint32_t spt = sPT + (i & 1) * B_PT_BYTES, sdst = sPT + (2 + (i & 1)) * B_PT_BYTES;
                const uint64_t dpt = make_smem_desc(spt, 16, 1024), ddst = make_smem_desc(sdst, 16, 1024);
                const uint64_t ddo = make_smem_desc(sdo, 8192, 1024), dq = make_smem_desc(sq, 8192, 1024);   // MN-major B
#pragma unroll
                for (int k = 0; k < BQ / 16; ++k) {
                    umma_bf16(tm_dv, smem_desc_advance(dpt, 2 * k), smem_desc_advance(ddo, 128 * k), idesc_km, (i > 0 || k > 0) ? 1u : 0u);
                    umma_bf16(tm_dk, smem_desc_advance(ddst, 2 * k), smem_desc_advance(dq, 128 * k), idesc_km, (i > 0 || k > 0) ? 1u : 0u);
                }
                if constexpr (FDQ) {
                    // dQ_i[q, n] = sum_key dS^T[key, q] K[key, n]: the dS^T tile [128 keys][64 queries] is the MN-major form of
                    // A = dS_i (M = queries, one 64-wide swizzle atom per M block; the second M block, LBO = 16 KB further,
                    // is whatever follows in shared memory: its 64 result rows land in TMEM lanes 64..127 and are never read)
                    const int slot = i & 1;
                    if (i >= 2) {
                        mbar_wait(dq_free(slot), ((i >> 1) - 1) & 1);
                        tc_fence_after();
                    }
                    const uint32_t t_dq = tmem_base + Cfg::COL_DQ + (uint32_t)(slot * DH);
                    const uint64_t da = make_smem_desc(sdst, B_PT_BYTES, 1024);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                        umma_bf16(t_dq, smem_desc_advance(da, 128 * k), smem_desc_advance(desc_k_mn, 128 * k), idesc_mm, k > 0 ? 1u : 0u);
                }
                umma_commit(qd_empty(s));
                umma_commit(pds_free(i & 1));
                if constexpr (FDQ) umma_commit(dq_full(i & 1));
                if (i == 2) ATRACE(18);
            }
            umma_commit(dkv_done);
            ATRACE(4);
        }
    } else if (!FDQ && warp == 2 + B_EW) {
        // dS^T (the A operand tile the element-wise warps just built, 128 keys x 64 queries, swizzle-128B) -> HBM as ONE TMA
        // store per query block.  Row-per-thread global stores from the element-wise warps touched 32 lines per instruction
        // and cost ~800 clk of every ~2500 clk block (tools/attn_trace.py).
        if (elect_one_sync()) {
            for (int i = 0; i < nqb; ++i) {
                const int slot = i & 1;
                mbar_wait(pds_ready(slot), (i >> 1) & 1);
                tma_store_4d(&map_dst, sPT + (2 + slot) * B_PT_BYTES, i * BQ, k0, (int)bh, 0);
                bulk_commit_group();
                bulk_wait_group_read<0>();
                mbar_arrive(pds_free(slot));
            }
        }
    } else if (FDQ && warp >= 20) {
        // dQ drain: warp 20 owns TMEM lanes 0..31 (queries 0..31 of the block), warp 21 lanes 32..63
        const int dw = warp - 20;
        const int row = dw * 32 + lane;
        float* stage = reinterpret_cast<float*>(gen + (sDQS - base));
        const uint32_t tl = tmem_base + ((uint32_t)(dw * 32) << 16) + Cfg::COL_DQ;
        const bool issuer = (dw == 0 && lane == 0);
        for (int i = 0; i < nqb; ++i) {
            const int slot = i & 1;
            mbar_wait(dq_full(slot), (i >> 1) & 1);
            tc_fence_after();
            uint32_t r0[32], r1[32];
            tmem_ld_32x32(tl + (uint32_t)(slot * DH), r0);
            tmem_ld_32x32(tl + (uint32_t)(slot * DH + 32), r1);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dq_free(slot));
            if (i > 0) {   // the previous block's reduce has finished reading the staging tile
                if (issuer) bulk_wait_group_read<0>();
                named_bar_sync(2, 64);
            }
            float4* dstrow = reinterpret_cast<float4*>(stage + row * DH);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                dstrow[c] = make_float4(__uint_as_float(r0[4 * c]), __uint_as_float(r0[4 * c + 1]), __uint_as_float(r0[4 * c + 2]),
                                        __uint_as_float(r0[4 * c + 3]));
                dstrow[8 + c] = make_float4(__uint_as_float(r1[4 * c]), __uint_as_float(r1[4 * c + 1]),
                                            __uint_as_float(r1[4 * c + 2]), __uint_as_float(r1[4 * c + 3]));
            }
            fence_proxy_async();
            named_bar_sync(2, 64);
            if (issuer && !p.skip_reduce) {
                const int rows = min(BQ, T - i * BQ);
                bulk_reduce_add_f32(p.dq_acc + (bh * T + (long long)i * BQ) * DH, sDQS, (uint32_t)rows * DH * 4u);
                bulk_commit_group();
            }
        }
        if (issuer) bulk_wait_group<0>();
    } else if (warp >= 2 && warp < 2 + B_EW) {
        constexpr int CG = 4 * BQ / B_EW;   // query columns per thread and block (16)
        static_assert(CG == 16 && DH == 64, "element-wise layout");
        const int quarter = warp & 3, cg = (warp - 2) >> 2;   // TMEM lane quarter, column group
        const int row = quarter * 32 + lane;
        const int key = k0 + row;
        const bool key_valid = key < kl;
        const uint32_t tlane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const float c2 = p.scale_log2, sc = p.scale;
        // keep bits: word row (k0 / 32 + quarter) holds this warp's 32 keys (bit = lane), 16 consecutive query words per block
        const uint32_t* keep_base = nullptr;
        bool keep_ok = false;
        float ik = 1.f;
        if constexpr (DROP) {
            const int kw = (k0 >> 5) + quarter;
            keep_ok = kw < p.KW;
            keep_base = p.keep + (bh * p.KW + (keep_ok ? kw : 0)) * (long long)p.TQ + cg * CG;
            ik = p.inv_keep;
        }
        for (int i = 0; i < nqb; ++i) {
            const int slot = i & 1, a3 = i % SDP;
            const bool trc = threadIdx.x == 64 && (i == 2 || i == 3);
            if (trc && i == 2) ATRACE(8);
            uint4 kb[CG / 4];
            if constexpr (DROP) {   // (requested before the wait on the accumulators; the same 64 bytes for the whole warp)
                const uint4* kp = reinterpret_cast<const uint4*>(keep_base + i * BQ);
#pragma unroll
                for (int c = 0; c < CG / 4; ++c) kb[c] = keep_ok ? __ldg(kp + c) : make_uint4(0, 0, 0, 0);
            }
            mbar_wait(sdp_full(a3), (i / SDP) & 1);
            tc_fence_after();
            if (trc) ATRACE(i == 2 ? 9 : 15);
            uint32_t s[CG], dp[CG];
            tmem_ld_32x16(tlane + (uint32_t)(a3 * BQ + cg * CG), s);
            tmem_ld_32x16(tlane + Cfg::COL_DP + (uint32_t)(a3 * BQ + cg * CG), dp);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sdp_free(a3));
            if (trc && i == 2) ATRACE(10);
            uint32_t pp[CG / 2], dd[CG / 2];
            const float4* l2v = reinterpret_cast<const float4*>(lse2_s + i * BQ + cg * CG);  // broadcast 16-byte reads
            const float4* dqv = reinterpret_cast<const float4*>(dsum_s + i * BQ + cg * CG);
#pragma unroll
            for (int c = 0; c < CG; c += 4) {
                const float4 l4 = l2v[c >> 2], d4 = dqv[c >> 2];
                float p0 = fast_exp2(fmaf(__uint_as_float(s[c]), c2, -l4.x));
                float p1 = fast_exp2(fmaf(__uint_as_float(s[c + 1]), c2, -l4.y));
                float p2, p3;
                if (PM == 2 || (PM == 4 && ((c >> 2) & 1))) {
                    const float2 pq = exp2_poly2(make_float2(fmaf(__uint_as_float(s[c + 2]), c2, -l4.z),
                                                             fmaf(__uint_as_float(s[c + 3]), c2, -l4.w)));
                    p2 = pq.x; p3 = pq.y;
                } else {
                    p2 = fast_exp2(fmaf(__uint_as_float(s[c + 2]), c2, -l4.z));
                    p3 = fast_exp2(fmaf(__uint_as_float(s[c + 3]), c2, -l4.w));
                }
                if (!key_valid) p0 = p1 = p2 = p3 = 0.f;
                if constexpr (DROP) {
                    // P_drop = P o M / (1 - p) feeds dV (the 1 / (1 - p) is applied once, when dV leaves TMEM);
                    // dS = P o (dP_drop o M / (1 - p) - D) with D = rowsum(dO o O) unchanged
                    const uint4 w4 = kb[c >> 2];
                    const bool k0b = (w4.x >> lane) & 1u, k1b = (w4.y >> lane) & 1u, k2b = (w4.z >> lane) & 1u, k3b = (w4.w >> lane) & 1u;
                    const float t0 = k0b ? fmaf(__uint_as_float(dp[c]), ik, -d4.x) : -d4.x;
                    const float t1 = k1b ? fmaf(__uint_as_float(dp[c + 1]), ik, -d4.y) : -d4.y;
                    const float t2 = k2b ? fmaf(__uint_as_float(dp[c + 2]), ik, -d4.z) : -d4.z;
                    const float t3 = k3b ? fmaf(__uint_as_float(dp[c + 3]), ik, -d4.w) : -d4.w;
                    dd[c >> 1] = bf2_pack(p0 * t0 * sc, p1 * t1 * sc);
                    dd[(c >> 1) + 1] = bf2_pack(p2 * t2 * sc, p3 * t3 * sc);
                    pp[c >> 1] = bf2_pack(k0b ? p0 : 0.f, k1b ? p1 : 0.f);
                    pp[(c >> 1) + 1] = bf2_pack(k2b ? p2 : 0.f, k3b ? p3 : 0.f);
                } else {
                    const float d0 = p0 * (__uint_as_float(dp[c]) - d4.x) * sc;
                    const float d1 = p1 * (__uint_as_float(dp[c + 1]) - d4.y) * sc;
                    const float d2 = p2 * (__uint_as_float(dp[c + 2]) - d4.z) * sc;
                    const float d3 = p3 * (__uint_as_float(dp[c + 3]) - d4.w) * sc;
                    pp[c >> 1] = bf2_pack(p0, p1);
                    pp[(c >> 1) + 1] = bf2_pack(p2, p3);
                    dd[c >> 1] = bf2_pack(d0, d1);
                    dd[(c >> 1) + 1] = bf2_pack(d2, d3);
                }
            }
            if (trc && i == 2) ATRACE(11);
            if (i >= 2) mbar_wait(pds_free(slot), ((i >> 1) - 1) & 1);
            if (trc && i == 2) ATRACE(12);
            uint8_t* gpt = genPT + slot * B_PT_BYTES;
            uint8_t* gds = genPT + (2 + slot) * B_PT_BYTES;
#pragma unroll
            for (int c = 0; c < CG / 8; ++c) {
                const uint32_t off = sw128_offset(row, cg * (CG / 8) + c);
                *reinterpret_cast<uint4*>(gpt + off) = make_uint4(pp[4 * c], pp[4 * c + 1], pp[4 * c + 2], pp[4 * c + 3]);
                *reinterpret_cast<uint4*>(gds + off) = make_uint4(dd[4 * c], dd[4 * c + 1], dd[4 * c + 2], dd[4 * c + 3]);
            }
            if (trc && i == 2) ATRACE(13);
            fence_proxy_async();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(pds_ready(slot));
            if (trc && i == 2) ATRACE(14);
        }
        if (threadIdx.x == 64) ATRACE(5);
        mbar_wait(dkv_done, 0);
        tc_fence_after();
        if (threadIdx.x == 64) ATRACE(6);
        uint32_t a[CG], c_[CG];   // this thread's 16 of the 64 head-dim columns of dV / dK
        tmem_ld_32x16(tlane + Cfg::COL_DV + (uint32_t)(cg * CG), a);   // dV
        tmem_ld_32x16(tlane + Cfg::COL_DK + (uint32_t)(cg * CG), c_);  // dK
        tmem_ld_wait();
        if (key < T) {
            __nv_bfloat16* dk = p.dqkv + (((long long)b * T + key) * 3 + 1) * D + h * DH + cg * CG;
            __nv_bfloat16* dv = dk + D;
#pragma unroll
            for (int c = 0; c < CG; c += 8) {
                uint4 v;
                v.x = bf2_pack(__uint_as_float(a[c]) * ik, __uint_as_float(a[c + 1]) * ik);
                v.y = bf2_pack(__uint_as_float(a[c + 2]) * ik, __uint_as_float(a[c + 3]) * ik);
                v.z = bf2_pack(__uint_as_float(a[c + 4]) * ik, __uint_as_float(a[c + 5]) * ik);
                v.w = bf2_pack(__uint_as_float(a[c + 6]) * ik, __uint_as_float(a[c + 7]) * ik);
                *reinterpret_cast<uint4*>(dv + c) = v;
                v.x = bf2_pack(__uint_as_float(c_[c]), __uint_as_float(c_[c + 1]));
                v.y = bf2_pack(__uint_as_float(c_[c + 2]), __uint_as_float(c_[c + 3]));
                v.z = bf2_pack(__uint_as_float(c_[c + 4]), __uint_as_float(c_[c + 5]));
                v.w = bf2_pack(__uint_as_float(c_[c + 6]), __uint_as_float(c_[c + 7]));
                *reinterpret_cast<uint4*>(dk + c) = v;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, B_TMEM_COLS);
    }
    if (threadIdx.x == 0) ATRACE(7);
}

// D[b,h,t] = sum_c dO[b,t,h,c] * O[b,t,h,c]   (one warp per frame, 8 lanes per head per 256-column sweep)
// zero_n4 > 0: also clears the fp32 dQ accumulator of the fused backward (zero_n4 float4), so that no memset launch is needed
__global__ void __launch_bounds__(256) mha_bwd_dsum_kernel(const __nv_bfloat16* __restrict__ o,
                                                           const __nv_bfloat16* __restrict__ dout,
                                                           float* __restrict__ dsum, int B, int T, int H,
                                                           float4* __restrict__ zero, long long zero_n4) {
    pdl_enter();  // launched with programmatic stream serialization: wait for the producers before touching global memory
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < zero_n4; i += (long long)gridDim.x * blockDim.x)
        zero[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int D = H * DH;
    for (long long f = warp; f < (long long)B * T; f += nwarps) {
        const int b = (int)(f / T), t = (int)(f % T);
        const __nv_bfloat16* op = o + f * D;
        const __nv_bfloat16* dp = dout + f * D;
        for (int c0 = 0; c0 < D; c0 += 256) {
            const int c = c0 + lane * 8;
            float acc = 0.f;
            if (c < D) {
                Vec16<__nv_bfloat16> a, g;
                float af[8], gf[8];
                a.load(op + c);
                g.load(dp + c);
                a.unpack(af);
                g.unpack(gf);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc = fmaf(af[j], gf[j], acc);
            }
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            acc += __shfl_xor_sync(0xffffffffu, acc, 4);
            if ((lane & 7) == 0 && c < D) dsum[((long long)b * H + c / DH) * T + t] = acc;
        }
    }
}

// dQ accumulator fp32 [B,H,T,64] -> bf16 Q third of dqkv [B,T,3,H,64]  (8 elements per thread)
__global__ void __launch_bounds__(256) mha_bwd_dq_store_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dqkv,
                                                               int B, int T, int H) {
    pdl_enter();
    const long long n = (long long)B * H * T * (DH / 8);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % (DH / 8));
        const long long r = i / (DH / 8);
        const int t = (int)(r % T);
        const long long bh = r / T;
        const int h = (int)(bh % H), b = (int)(bh / H);
        const float4 v0 = __ldcs(reinterpret_cast<const float4*>(acc + r * DH + c8 * 8));
        const float4 v1 = __ldcs(reinterpret_cast<const float4*>(acc + r * DH + c8 * 8 + 4));
        uint4 o;
        o.x = bf2_pack(v0.x, v0.y); o.y = bf2_pack(v0.z, v0.w); o.z = bf2_pack(v1.x, v1.y); o.w = bf2_pack(v1.z, v1.w);
        *reinterpret_cast<uint4*>(dqkv + (((long long)b * T + t) * 3) * (H * DH) + h * DH + c8 * 8) = o;
    }
}

// Keep bits of the attention-probability dropout for `n_sites` consecutive sites (encoder blocks) in one launch: one
// thread = one 32-bit word = the 32 keys [32 kw, 32 kw + 32) of one (site, utterance, head, query).  Philox4x32-10 keyed
// by (seed, site) over the counter (word index, sub-block 0..3, step): 4 calls = 32 16-bit uniforms; keep iff u >= p * 2^16.
__global__ void __launch_bounds__(256) attn_keep_mask_kernel(uint32_t* __restrict__ keep, long long words_per_site, int n_sites,
                                                             unsigned long long seed, unsigned long long site0,
                                                             unsigned long long site_stride, uint32_t thresh16,
                                                             unsigned long long offset,
                                                             const unsigned long long* __restrict__ dev_offset) {
    if (dev_offset) offset += *dev_offset;
    const long long total = words_per_site * n_sites;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long site_i = i / words_per_site, w = i - site_i * words_per_site;
        const unsigned long long kk = seed ^ ((site0 + site_i * site_stride) * 0x9E3779B97F4A7C15ull);
        const uint2 key = make_uint2((uint32_t)kk, (uint32_t)(kk >> 32));
        uint32_t bits = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const unsigned long long ctr = ((unsigned long long)w << 2) | (unsigned long long)q;
            const uint4 r = philox4x32_10(make_uint4((uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)offset, (uint32_t)(offset >> 32)), key);
            const uint32_t u[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                bits |= ((u[j] & 0xffffu) >= thresh16 ? 1u : 0u) << (8 * q + 2 * j);
                bits |= ((u[j] >> 16) >= thresh16 ? 1u : 0u) << (8 * q + 2 * j + 1);
            }
        }
        keep[i] = bits;
    }
}

static int make_head_map(CUtensorMap* map, const void* base, int T, int H, int B, long long row_stride_elems,
                         long long batch_stride_elems, int box_rows, const char* what) {
    const cuuint64_t dims[4] = {(cuuint64_t)DH, (cuuint64_t)T, (cuuint64_t)H, (cuuint64_t)B};
    const cuuint64_t strides[3] = {(cuuint64_t)row_stride_elems * 2, (cuuint64_t)DH * 2, (cuuint64_t)batch_stride_elems * 2};
    const cuuint32_t box[4] = {(cuuint32_t)DH, (cuuint32_t)box_rows, 1, 1};
    return encode_map_bf16_4d(map, base, dims, strides, box, what);
}

// b200asr_mha_set_dq_fused: 0 (default) = dS^T + batched GEMM, 1 = dQ inside the backward kernel (fp32 bulk reduce-add),
// 2 = as 1 with the reduce-adds skipped (wrong dQ; isolates the cost of the reduction in measurements)
static int g_dq_fused = 0;

template <typename K>
static int set_max_smem(K kernel, int bytes, const char* what) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) {
        set_error("%s: cudaFuncSetAttribute failed: %s", what, cudaGetErrorString(e));
        return B200ASR_ERR_CUDA;
    }
    return 0;
}

}  // namespace attn

bool mha_tc_eligible(int dtype, int dh, const void* qkv) {
    return dtype == B200ASR_BF16 && dh == attn::DH && (reinterpret_cast<uintptr_t>(qkv) & 15) == 0;
}

int attn_keep_kw(int T) { return (T + 31) / 32; }
int attn_keep_tq(int T) { return (T + 127) / 128 * 128; }

int attn_keep_mask_launch(unsigned* keep, int n_sites, int B, int T, int H, float p, unsigned long long seed,
                          unsigned long long site0, unsigned long long site_stride, unsigned long long offset,
                          const unsigned long long* dev_offset, cudaStream_t st) {
    const long long words = (long long)B * H * attn_keep_kw(T) * attn_keep_tq(T);
    const long long total = words * n_sites;
    long long grid = (total + 255) / 256;
    const long long cap = (long long)num_sms() * 16;
    if (grid > cap) grid = cap;
    attn::attn_keep_mask_kernel<<<(int)grid, 256, 0, st>>>(keep, words, n_sites, seed, site0, site_stride,
                                                           (uint32_t)lrintf(p * 65536.f), offset, dev_offset);
    return check_launch("attn_dropout_mask");
}

int mha_fwd_tc_launch(const void* qkv, void* out, float* lse, const int* klen, const unsigned* keep, float p_drop, int B,
                      int T, int H, float scale, cudaStream_t st) {
    using namespace attn;
    const int D = H * DH;
    const __nv_bfloat16* q = (const __nv_bfloat16*)qkv;
    CUtensorMap mq, mk, mv;
    int rc = make_head_map(&mq, q, T, H, B, 3LL * D, (long long)T * 3 * D, FQ, "mha_fwd(tcgen05)");
    if (rc) return rc;
    rc = make_head_map(&mk, q + D, T, H, B, 3LL * D, (long long)T * 3 * D, FK, "mha_fwd(tcgen05)");
    if (rc) return rc;
    rc = make_head_map(&mv, q + 2 * D, T, H, B, 3LL * D, (long long)T * 3 * D, FK, "mha_fwd(tcgen05)");
    if (rc) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        if ((rc = set_max_smem(mha_fwd_tc_kernel<false, 0>, F_SMEM, "mha_fwd(tcgen05)"))) return rc;
        if ((rc = set_max_smem(mha_fwd_tc_kernel<true, 0>, F_SMEM, "mha_fwd(tcgen05)"))) return rc;
        if ((rc = set_max_smem(mha_fwd_tc_kernel<false, 2>, F_SMEM, "mha_fwd(tcgen05)"))) return rc;
        if ((rc = set_max_smem(mha_fwd_tc_kernel<true, 2>, F_SMEM, "mha_fwd(tcgen05)"))) return rc;
        if ((rc = set_max_smem(mha_fwd_tc_kernel<false, 4>, F_SMEM, "mha_fwd(tcgen05)"))) return rc;
        if ((rc = set_max_smem(mha_fwd_tc_kernel<true, 4>, F_SMEM, "mha_fwd(tcgen05)"))) return rc;
        attr_set = true;
    }
    FwdParams p;
    p.out = (__nv_bfloat16*)out; p.lse = lse; p.klen = klen; p.B = B; p.T = T; p.H = H;
    p.scale_log2 = scale * LOG2E;
    p.keep = keep; p.KW = attn_keep_kw(T); p.TQ = attn_keep_tq(T); p.inv_keep = 1.0f / (1.0f - p_drop);
    p.trace = g_attn_trace;
    dim3 grid(ceil_div(T, FQ), H, B);
#define FWD_LAUNCH(DR, PMV) launch_pdl(mha_fwd_tc_kernel<DR, PMV>, grid, dim3(F_THREADS), (size_t)F_SMEM, st, mq, mk, mv, p)
    const int pm = g_exp2_poly_fwd;
    if (keep) { if (pm == 2) FWD_LAUNCH(true, 2); else if (pm == 4) FWD_LAUNCH(true, 4); else FWD_LAUNCH(true, 0); }
    else { if (pm == 2) FWD_LAUNCH(false, 2); else if (pm == 4) FWD_LAUNCH(false, 4); else FWD_LAUNCH(false, 0); }
#undef FWD_LAUNCH
    return check_launch("mha_fwd_tcgen05");
}

size_t mha_bwd_tc_workspace_bytes(int B, int T, int H) {
    const size_t ldq = (size_t)((T + 7) & ~7);
    const size_t dst = (size_t)B * H * T * ldq * 2;                 // dS^T (deterministic path)
    const size_t acc = (size_t)B * H * T * attn::DH * sizeof(float);  // fp32 dQ accumulator (fused path)
    return (dst > acc ? dst : acc) + (((size_t)B * H * T * 4 + 255) & ~(size_t)255);
}

template <bool DROP, bool FDQ, int PM = 0>
static int launch_bwd_kernel(dim3 grid, size_t smem, cudaStream_t st, const CUtensorMap& mk, const CUtensorMap& mv,
                             const CUtensorMap& mq, const CUtensorMap& mdo, const CUtensorMap& mdst, const attn::BwdParams& p) {
    static bool attr_set = false;
    if (!attr_set) {
        int rc = attn::set_max_smem(attn::mha_bwd_tc_kernel<DROP, FDQ, PM>, 227 * 1024, "mha_bwd(tcgen05)");
        if (rc) return rc;
        attr_set = true;
    }
    launch_pdl(attn::mha_bwd_tc_kernel<DROP, FDQ, PM>, grid, dim3(attn::BwdCfg<FDQ>::THREADS), smem, st, mk, mv, mq, mdo, mdst, p);
    return check_launch("mha_bwd_tcgen05");
}

int mha_bwd_tc_launch(const void* qkv, const void* out, const void* dout, const float* lse, const int* klen,
                      const unsigned* keep, float p_drop, void* dqkv, void* workspace, int B, int T, int H, float scale,
                      cudaStream_t st) {
    using namespace attn;
    const int D = H * DH;
    const int ldq = (T + 7) & ~7;
    const bool fdq = g_dq_fused != 0;
    const __nv_bfloat16* q = (const __nv_bfloat16*)qkv;
    float* dsum = (float*)workspace;
    void* big = (char*)workspace + (((size_t)B * H * T * 4 + 255) & ~(size_t)255);
    __nv_bfloat16* dst = (__nv_bfloat16*)big;
    float* dq_acc = (float*)big;
    {
        long long frames = (long long)B * T;
        long long g = (frames + 7) / 8;
        if (g > (long long)num_sms() * 8) g = (long long)num_sms() * 8;
        launch_pdl(mha_bwd_dsum_kernel, dim3((unsigned)g), dim3(256), 0, st, (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout,
                   dsum, B, T, H, (float4*)dq_acc, fdq ? (long long)B * H * T * DH / 4 : 0LL);
        int rc = check_launch("mha_bwd_dsum");
        if (rc) return rc;
    }
    CUtensorMap mk, mv, mq, mdo;
    int rc = make_head_map(&mk, q + D, T, H, B, 3LL * D, (long long)T * 3 * D, BK, "mha_bwd(tcgen05)");
    if (rc) return rc;
    rc = make_head_map(&mv, q + 2 * D, T, H, B, 3LL * D, (long long)T * 3 * D, BK, "mha_bwd(tcgen05)");
    if (rc) return rc;
    rc = make_head_map(&mq, q, T, H, B, 3LL * D, (long long)T * 3 * D, BQ, "mha_bwd(tcgen05)");
    if (rc) return rc;
    rc = make_head_map(&mdo, dout, T, H, B, (long long)D, (long long)T * D, BQ, "mha_bwd(tcgen05)");
    if (rc) return rc;
    BwdParams p;
    p.lse = lse; p.dsum = dsum; p.klen = klen; p.dqkv = (__nv_bfloat16*)dqkv; p.dst = dst; p.dq_acc = dq_acc;
    p.B = B; p.T = T; p.H = H; p.ldq = ldq; p.scale = scale; p.scale_log2 = scale * LOG2E;
    p.keep = keep; p.KW = attn_keep_kw(T); p.TQ = attn_keep_tq(T); p.inv_keep = keep ? 1.0f / (1.0f - p_drop) : 1.0f;
    p.skip_reduce = g_dq_fused == 2;
    p.trace = g_attn_trace;
    dim3 grid(ceil_div(T, BK), H, B);
    const int Tpad = ceil_div(T, BQ) * BQ;
    const size_t smem = (size_t)(fdq ? BwdCfg<true>::SMEM_FIXED : BwdCfg<false>::SMEM_FIXED) + 2 * (size_t)Tpad * sizeof(float);
    if (smem > 227 * 1024) {
        set_error("mha_bwd(tcgen05): T=%d too long for the staged softmax statistics", T);
        return B200ASR_ERR_UNSUPPORTED;
    }
    CUtensorMap mdst;
    {   // (only the deterministic path stores dS^T; the map is encoded either way so that both builds share a signature)
        const cuuint64_t dims[4] = {(cuuint64_t)ldq, (cuuint64_t)T, (cuuint64_t)B * H, 1};
        const cuuint64_t strides[3] = {(cuuint64_t)ldq * 2, (cuuint64_t)T * ldq * 2, (cuuint64_t)B * H * T * ldq * 2};
        const cuuint32_t box[4] = {BQ, BK, 1, 1};
        rc = encode_map_bf16_4d(&mdst, dst, dims, strides, box, "mha_bwd(tcgen05) dS^T");
        if (rc) return rc;
    }
    if (fdq) {
        rc = keep ? launch_bwd_kernel<true, true>(grid, smem, st, mk, mv, mq, mdo, mdst, p)
                  : launch_bwd_kernel<false, true>(grid, smem, st, mk, mv, mq, mdo, mdst, p);
        if (rc) return rc;
        long long n = (long long)B * H * T * (DH / 8);
        long long g = (n + 255) / 256;
        if (g > (long long)num_sms() * 16) g = (long long)num_sms() * 16;
        launch_pdl(mha_bwd_dq_store_kernel, dim3((unsigned)g), dim3(256), 0, st, (const float*)dq_acc, (__nv_bfloat16*)dqkv, B, T, H);
        return check_launch("mha_bwd_dq_store");
    }
    const int pm = g_exp2_poly_bwd;
    if (pm == 2)
        rc = keep ? launch_bwd_kernel<true, false, 2>(grid, smem, st, mk, mv, mq, mdo, mdst, p)
                  : launch_bwd_kernel<false, false, 2>(grid, smem, st, mk, mv, mq, mdo, mdst, p);
    else if (pm == 4)
        rc = keep ? launch_bwd_kernel<true, false, 4>(grid, smem, st, mk, mv, mq, mdo, mdst, p)
                  : launch_bwd_kernel<false, false, 4>(grid, smem, st, mk, mv, mq, mdo, mdst, p);
    else
        rc = keep ? launch_bwd_kernel<true, false>(grid, smem, st, mk, mv, mq, mdo, mdst, p)
                  : launch_bwd_kernel<false, false>(grid, smem, st, mk, mv, mq, mdo, mdst, p);
    if (rc) return rc;
    // dQ[q, n] = sum_key dS^T[key, q] K[key, n]   (dS already carries the softmax scale)
    b200asr_gemm_args g = {};
    g.dtype = B200ASR_BF16; g.c_dtype = B200ASR_BF16; g.batch0 = H; g.batch1 = B; g.alpha = 1.f;
    g.A = dst; g.B = q + D; g.C = dqkv;
    g.M = T; g.N = DH; g.K = T; g.lda = ldq; g.ldb = 3 * D; g.ldc = 3 * D; g.a_kmajor = 0; g.b_kmajor = 0;
    g.sa0 = (long long)T * ldq; g.sa1 = (long long)H * T * ldq; g.sb0 = DH; g.sb1 = (long long)T * 3 * D;
    g.sc0 = DH; g.sc1 = (long long)T * 3 * D;
    return gemm_dispatch(g, st);
}

}  // namespace b200asr

// diagnostics: while `buf` (device memory, 24 u64 per CTA of the backward grid) is set, mha_bwd records per-CTA SM-clock
// time stamps: 0 entry, 1 dependency wait done, 2 set-up done, 3 K/V landed, 4 last MMA issued, 5 element-wise loop done,
// 6 dK/dV complete, 7 exit; query block 2: 8..14 element-wise phases, 15 block 3 ready, 16..18 MMA issuer
extern "C" int b200asr_attn_set_trace(void* buf) {
    b200asr::attn::g_attn_trace = (unsigned long long*)buf;
    return 0;
}

// 0 (default): the backward stores dS^T and a batched GEMM produces dQ (bit-reproducible; measured faster);
// 1: dQ is accumulated inside the backward kernel (fp32 bulk reduce-add, reproducible to rounding); 2: measurement hook.
// Returns the previous setting.
// Share of the softmax exponentials of the fused attention kernels computed by a degree-3 polynomial on the FMA pipe
// instead of the MUFU: fwd / bwd in {0, 4 (every 4th pair: 25 %), 2 (50 %)}; a negative value keeps the current
// setting.  Returns the previous setting as fwd * 16 + bwd.
extern "C" int b200asr_mha_set_exp2_poly(int fwd, int bwd) {
    const int old = b200asr::attn::g_exp2_poly_fwd * 16 + b200asr::attn::g_exp2_poly_bwd;
    if (fwd >= 0) b200asr::attn::g_exp2_poly_fwd = (fwd == 2 || fwd == 4) ? fwd : 0;
    if (bwd >= 0) b200asr::attn::g_exp2_poly_bwd = (bwd == 2 || bwd == 4) ? bwd : 0;
    return old;
}

extern "C" int b200asr_mha_set_dq_fused(int on) {
    const int old = b200asr::attn::g_dq_fused;
    b200asr::attn::g_dq_fused = (on < 0 || on > 2) ? 0 : on;
    return old;
}
