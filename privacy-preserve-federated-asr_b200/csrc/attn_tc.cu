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
    unsigned long long* trace;   // diagnostics (b200asr_attn_set_trace): 24 time stamps per CTA, null in normal use
};
static unsigned long long* g_attn_trace = nullptr;
#define ATRACE(slot) do { if (p.trace) p.trace[(long long)((blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 24 + (slot)] = (unsigned long long)clock64(); } while (0)

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
        for (int j = 0; j < nkb; ++j) {
            const bool trc = threadIdx.x == 64 && j == 2;
            if (trc) ATRACE(8);
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
                const float p0 = fast_exp2(fmaf(__uint_as_float(r[c]), c2, -m_new));
                const float p1 = fast_exp2(fmaf(__uint_as_float(r[c + 1]), c2, -m_new));
                sum0 += p0;
                sum1 += p1;
                pk[c >> 1] = bf2_pack(p0, p1);
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
        const float inv = 1.0f / lt;
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
constexpr int B_STAGES = 5;    // Q_i + dO_i ring: S^T / dP^T run two blocks ahead of dV / dK, which still reads block i
constexpr int B_SDP = 3;       // S^T / dP^T accumulator pairs in TMEM
constexpr int B_EW = 16;        // element-wise warps: four per TMEM lane quarter, 16 query columns each (the per-block chain
                                // of one thread is latency bound; twice the warps of the first version hide twice the latency)
constexpr int B_THREADS = 96 + 32 * B_EW;  // TMA warp, MMA warp, element-wise warps, dS^T store warp
constexpr int B_KV_BYTES = 2 * BK * DH * 2;   // K_j + V_j: 32 KB
constexpr int B_QD_BYTES = 2 * BQ * DH * 2;   // Q_i + dO_i: 16 KB
constexpr int B_PT_BYTES = BK * BQ * 2;       // one P^T (or dS^T) operand tile: 16 KB
constexpr int B_SMEM_FIXED = B_KV_BYTES + B_STAGES * B_QD_BYTES + 4 * B_PT_BYTES + 1024 + 256;  // + 2 * Tpad floats (lse, D)
constexpr int B_TMEM_COLS = 512;  // S^T[3], dP^T[3] (6 x 64) + dV (64) + dK (64)

struct BwdParams {
    const float* lse;   // [B,H,T] natural-log sum-exp of the scaled logits
    const float* dsum;  // [B,H,T] D = rowsum(dO * O)
    const int* klen;
    __nv_bfloat16* dqkv;  // [B,T,3,H,dh]: this kernel writes the K and V thirds
    __nv_bfloat16* dst;   // dS^T [B,H,T(key),ldq] (scaled by `scale`)
    int B, T, H, ldq;
    float scale, scale_log2;
    unsigned long long* trace;   // diagnostics (b200asr_attn_set_trace): 24 time stamps per CTA, null in normal use
};

__global__ void __launch_bounds__(B_THREADS, 1)
mha_bwd_tc_kernel(const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v,
                  const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_do,
                  const __grid_constant__ CUtensorMap map_dst, const BwdParams p) {
    if (threadIdx.x == 0) ATRACE(0);
    pdl_enter();  // launched with programmatic stream serialization: wait for the producers before touching global memory
    if (threadIdx.x == 0) ATRACE(1);
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t sK = base, sV = base + BK * DH * 2;
    const uint32_t sQD = base + B_KV_BYTES;
    const uint32_t sPT = sQD + B_STAGES * B_QD_BYTES;   // PT[0], PT[1], DST[0], DST[1]
    uint8_t* genPT = gen + B_KV_BYTES + B_STAGES * B_QD_BYTES;
    const uint32_t bars = sPT + 4 * B_PT_BYTES;
    const uint32_t kv_full = bars;
    auto qd_full = [&](int s) { return bars + 8u * (1 + s); };
    auto qd_empty = [&](int s) { return bars + 8u * (1 + B_STAGES + s); };
    auto sdp_full = [&](int s) { return bars + 8u * (1 + 2 * B_STAGES + s); };
    auto sdp_free = [&](int s) { return bars + 8u * (1 + B_SDP + 2 * B_STAGES + s); };
    auto pds_ready = [&](int s) { return bars + 8u * (1 + 2 * B_SDP + 2 * B_STAGES + s); };
    auto pds_free = [&](int s) { return bars + 8u * (3 + 2 * B_SDP + 2 * B_STAGES + s); };
    const uint32_t dkv_done = bars + 8u * (5 + 2 * B_SDP + 2 * B_STAGES);
    const uint32_t tmem_slot = bars + 8u * (6 + 2 * B_SDP + 2 * B_STAGES);
    static_assert(8 * (7 + 2 * B_SDP + 2 * B_STAGES) <= 256, "barrier area");
    volatile uint32_t* tmem_slot_gen =
        reinterpret_cast<volatile uint32_t*>(genPT + 4 * B_PT_BYTES + 8 * (6 + 2 * B_SDP + 2 * B_STAGES));
    // whole-row softmax statistics of this (utterance, head), staged once: lse * log2(e) then D, Tpad floats each
    // (Tpad = T rounded up to the 64-query block; the padding holds +inf / 0 so that out-of-range queries give P = 0)
    const int Tpad = ((p.T + BQ - 1) / BQ) * BQ;
    float* lse2_s = reinterpret_cast<float*>(genPT + 4 * B_PT_BYTES + 256);
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
    if (k0 >= kl) {  // fully masked key block: dK = dV = 0, dS^T rows = 0
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
                uint4* drow = reinterpret_cast<uint4*>(p.dst + (bh * T + key) * ldq);
                for (int c = half; c < ldq / 8; c += 2) drow[c] = make_uint4(0, 0, 0, 0);
            }
        }
        return;
    }

    if (warp == 0) {   // one barrier per lane (they are laid out back to back), then one lane starts the operand loads
        constexpr int NBAR = 6 + 2 * B_SDP + 2 * B_STAGES;
        static_assert(NBAR <= 32, "one barrier per lane");
        if (lane < NBAR) {
            const int first_ew = 1 + B_SDP + 2 * B_STAGES, first_pf = 3 + 2 * B_SDP + 2 * B_STAGES;   // sdp_free, pds_free
            // sdp_free / pds_ready: one arrival per element-wise warp; pds_free: the MMAs that read P^T / dS^T have
            // completed + the dS^T tile has been stored; everything else: one arrival (+ transaction bytes)
            const uint32_t cnt = (lane >= first_ew && lane < first_pf) ? B_EW : (lane >= first_pf && lane < first_pf + 2) ? 2 : 1;
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
        tma_prefetch_desc(&map_dst);
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
    // TMEM columns: S^T[s] at 64 s, dP^T[s] at 192 + 64 s (s < 3), dV at 384, dK at 448
    const uint32_t tm_dv = tmem_base + 384, tm_dk = tmem_base + 448;

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
            const uint64_t desc_k = make_smem_desc(sK, 16, 1024), desc_v = make_smem_desc(sV, 16, 1024);
            auto issue_sdp = [&](int i) {
                const int s = i % B_STAGES;
                const int a3 = i % B_SDP;
                mbar_wait(qd_full(s), (i / B_STAGES) & 1);
                if (i >= B_SDP) mbar_wait(sdp_free(a3), ((i / B_SDP) - 1) & 1);
                tc_fence_after();
                const uint32_t sq = sQD + s * B_QD_BYTES, sdo = sq + BQ * DH * 2;
                const uint32_t t_st = tmem_base + (uint32_t)(a3 * BQ), t_dp = tmem_base + 192 + (uint32_t)(a3 * BQ);
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
            // S^T / dP^T of block i+2 are queued before this thread waits for the element-wise warps of block i: the
            // tensor pipe always has the next block's logits ready when they come back (three accumulator pairs)
            issue_sdp(0);
            if (nqb > 1) issue_sdp(1);
            for (int i = 0; i < nqb; ++i) {
                if (i + 2 < nqb) issue_sdp(i + 2);
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
                umma_commit(qd_empty(s));
                umma_commit(pds_free(i & 1));
                if (i == 2) ATRACE(18);
            }
            umma_commit(dkv_done);
            ATRACE(4);
        }
    } else if (warp == 2 + B_EW) {
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
    } else {
        constexpr int CG = 4 * BQ / B_EW;   // query columns per thread and block (16)
        static_assert(CG == 16 && DH == 64, "element-wise layout");
        const int quarter = warp & 3, cg = (warp - 2) >> 2;   // TMEM lane quarter, column group
        const int row = quarter * 32 + lane;
        const int key = k0 + row;
        const bool key_valid = key < kl;
        const uint32_t tlane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const float c2 = p.scale_log2, sc = p.scale;
        for (int i = 0; i < nqb; ++i) {
            const int slot = i & 1, a3 = i % B_SDP;
            const bool trc = threadIdx.x == 64 && (i == 2 || i == 3);
            if (trc && i == 2) ATRACE(8);
            mbar_wait(sdp_full(a3), (i / B_SDP) & 1);
            tc_fence_after();
            if (trc) ATRACE(i == 2 ? 9 : 15);
            uint32_t s[CG], dp[CG];
            tmem_ld_32x16(tlane + (uint32_t)(a3 * BQ + cg * CG), s);
            tmem_ld_32x16(tlane + 192 + (uint32_t)(a3 * BQ + cg * CG), dp);
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
                float p2 = fast_exp2(fmaf(__uint_as_float(s[c + 2]), c2, -l4.z));
                float p3 = fast_exp2(fmaf(__uint_as_float(s[c + 3]), c2, -l4.w));
                if (!key_valid) p0 = p1 = p2 = p3 = 0.f;
                const float d0 = p0 * (__uint_as_float(dp[c]) - d4.x) * sc;
                const float d1 = p1 * (__uint_as_float(dp[c + 1]) - d4.y) * sc;
                const float d2 = p2 * (__uint_as_float(dp[c + 2]) - d4.z) * sc;
                const float d3 = p3 * (__uint_as_float(dp[c + 3]) - d4.w) * sc;
                pp[c >> 1] = bf2_pack(p0, p1);
                pp[(c >> 1) + 1] = bf2_pack(p2, p3);
                dd[c >> 1] = bf2_pack(d0, d1);
                dd[(c >> 1) + 1] = bf2_pack(d2, d3);
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
        tmem_ld_32x16(tlane + 384 + (uint32_t)(cg * CG), a);   // dV
        tmem_ld_32x16(tlane + 448 + (uint32_t)(cg * CG), c_);  // dK
        tmem_ld_wait();
        if (key < T) {
            __nv_bfloat16* dk = p.dqkv + (((long long)b * T + key) * 3 + 1) * D + h * DH + cg * CG;
            __nv_bfloat16* dv = dk + D;
#pragma unroll
            for (int c = 0; c < CG; c += 8) {
                uint4 v;
                v.x = bf2_pack(__uint_as_float(a[c]), __uint_as_float(a[c + 1]));
                v.y = bf2_pack(__uint_as_float(a[c + 2]), __uint_as_float(a[c + 3]));
                v.z = bf2_pack(__uint_as_float(a[c + 4]), __uint_as_float(a[c + 5]));
                v.w = bf2_pack(__uint_as_float(a[c + 6]), __uint_as_float(a[c + 7]));
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
__global__ void __launch_bounds__(256) mha_bwd_dsum_kernel(const __nv_bfloat16* __restrict__ o,
                                                           const __nv_bfloat16* __restrict__ dout,
                                                           float* __restrict__ dsum, int B, int T, int H) {
    pdl_enter();  // launched with programmatic stream serialization: wait for the producers before touching global memory
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

static int make_head_map(CUtensorMap* map, const void* base, int T, int H, int B, long long row_stride_elems,
                         long long batch_stride_elems, int box_rows, const char* what) {
    const cuuint64_t dims[4] = {(cuuint64_t)DH, (cuuint64_t)T, (cuuint64_t)H, (cuuint64_t)B};
    const cuuint64_t strides[3] = {(cuuint64_t)row_stride_elems * 2, (cuuint64_t)DH * 2, (cuuint64_t)batch_stride_elems * 2};
    const cuuint32_t box[4] = {(cuuint32_t)DH, (cuuint32_t)box_rows, 1, 1};
    return encode_map_bf16_4d(map, base, dims, strides, box, what);
}

}  // namespace attn

bool mha_tc_eligible(int dtype, int dh, const void* qkv) {
    return dtype == B200ASR_BF16 && dh == attn::DH && (reinterpret_cast<uintptr_t>(qkv) & 15) == 0;
}

int mha_fwd_tc_launch(const void* qkv, void* out, float* lse, const int* klen, int B, int T, int H, float scale,
                      cudaStream_t st) {
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
        cudaError_t e = cudaFuncSetAttribute(mha_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM);
        if (e != cudaSuccess) {
            set_error("mha_fwd(tcgen05): cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
            return B200ASR_ERR_CUDA;
        }
        attr_set = true;
    }
    FwdParams p;
    p.out = (__nv_bfloat16*)out; p.lse = lse; p.klen = klen; p.B = B; p.T = T; p.H = H;
    p.scale_log2 = scale * LOG2E;
    p.trace = g_attn_trace;
    dim3 grid(ceil_div(T, FQ), H, B);
    launch_pdl(mha_fwd_tc_kernel, grid, dim3(F_THREADS), (size_t)F_SMEM, st, mq, mk, mv, p);
    return check_launch("mha_fwd_tcgen05");
}

size_t mha_bwd_tc_workspace_bytes(int B, int T, int H) {
    const size_t ldq = (size_t)((T + 7) & ~7);
    return (size_t)B * H * T * ldq * 2 + (((size_t)B * H * T * 4 + 255) & ~(size_t)255);
}

int mha_bwd_tc_launch(const void* qkv, const void* out, const void* dout, const float* lse, const int* klen, void* dqkv,
                      void* workspace, int B, int T, int H, float scale, cudaStream_t st) {
    using namespace attn;
    const int D = H * DH;
    const int ldq = (T + 7) & ~7;
    const __nv_bfloat16* q = (const __nv_bfloat16*)qkv;
    float* dsum = (float*)workspace;
    __nv_bfloat16* dst = (__nv_bfloat16*)((char*)workspace + (((size_t)B * H * T * 4 + 255) & ~(size_t)255));
    {
        long long frames = (long long)B * T;
        long long g = (frames + 7) / 8;
        if (g > (long long)num_sms() * 8) g = (long long)num_sms() * 8;
        launch_pdl(mha_bwd_dsum_kernel, dim3((unsigned)g), dim3(256), 0, st, (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, dsum, B, T, H);
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
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(mha_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) {
            set_error("mha_bwd(tcgen05): cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
            return B200ASR_ERR_CUDA;
        }
        attr_set = true;
    }
    BwdParams p;
    p.lse = lse; p.dsum = dsum; p.klen = klen; p.dqkv = (__nv_bfloat16*)dqkv; p.dst = dst;
    p.B = B; p.T = T; p.H = H; p.ldq = ldq; p.scale = scale; p.scale_log2 = scale * LOG2E;
    p.trace = g_attn_trace;
    dim3 grid(ceil_div(T, BK), H, B);
    const int Tpad = ceil_div(T, BQ) * BQ;
    const size_t smem = (size_t)B_SMEM_FIXED + 2 * (size_t)Tpad * sizeof(float);
    if (smem > 227 * 1024) {
        set_error("mha_bwd(tcgen05): T=%d too long for the staged softmax statistics", T);
        return B200ASR_ERR_UNSUPPORTED;
    }
    CUtensorMap mdst;
    {
        const cuuint64_t dims[4] = {(cuuint64_t)ldq, (cuuint64_t)T, (cuuint64_t)B * H, 1};
        const cuuint64_t strides[3] = {(cuuint64_t)ldq * 2, (cuuint64_t)T * ldq * 2, (cuuint64_t)B * H * T * ldq * 2};
        const cuuint32_t box[4] = {BQ, BK, 1, 1};
        rc = encode_map_bf16_4d(&mdst, dst, dims, strides, box, "mha_bwd(tcgen05) dS^T");
        if (rc) return rc;
    }
    launch_pdl(mha_bwd_tc_kernel, grid, dim3(B_THREADS), smem, st, mk, mv, mq, mdo, mdst, p);
    rc = check_launch("mha_bwd_tcgen05");
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
