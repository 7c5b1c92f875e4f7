// bf16 GEMM on the 5th-generation tensor cores: TMA (128B-swizzled tiles) -> shared memory ring ->
// tcgen05.mma (single issuing thread, fp32 accumulators in TMEM, double buffered) -> tcgen05.ld epilogue.
// Persistent, warp specialised: warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..5 =
// epilogue (one TMEM lane quarter each).  Either operand may be K-major or MN-major in memory, so that
// nn.Linear forward (NT), dgrad (NN) and wgrad (TN) all run without materialised transposes.
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <type_traits>

#include "common.cuh"
#include "gemm.h"
#include "tc_ptx.cuh"

namespace b200asr {

namespace tc {

static unsigned long long* g_gemm_trace = nullptr;  // b200asr_gemm_set_trace

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;   // 64 bf16 = 128 bytes = one swizzle-128B row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 320;  // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter: column halves)
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB

// NCTA = 1: one CTA per 128 x BN tile.  NCTA = 2: a CTA pair (cluster of 2, tcgen05 cta_group::2) per 256 x BN tile; each
// CTA stages its own 128 rows of A and HALF of the B tile, so the L2 -> shared-memory traffic per FLOP drops by 1.5x
// (BN = 256) against the single-CTA 128 x 256 tile -- the measured limiter of this kernel (profiles/: ~6.3 KB/clk chip
// wide L2 read cap reached at ~950 TFLOP/s).
// NCTA = 4: a cluster of TWO pairs on vertically adjacent 256 x BN tiles (a 512 x BN super tile).  Both pairs need the same
// B tile: every CTA loads one quarter of it and TMA-multicasts it to the CTA of the other pair that holds the same B half,
// so the L2 -> shared-memory traffic per FLOP drops by another 25% (A 16 KB + B 8 KB per CTA and 64-deep K block).
template <int BN, int NCTA = 1> struct Cfg {
    static constexpr int NP = NCTA >= 2 ? 2 : 1;  // CTAs per MMA (cta_group)
    static constexpr int B_STAGE_BYTES = (BN / NP) * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    // Staged epilogue: every epilogue warp owns 8 KB of shared memory -- two chunks of 32 rows x 128 bytes (64 bf16 / 32 fp32
    // columns, swizzle-128B).  Results are TMA-stored from there (full coalesced rows, asynchronous); a residual / gelu'
    // operand is TMA-prefetched into the same chunk a whole tile ahead and overwritten in place by the result; the
    // forward FFN GEMM uses the pair as (result, gelu').  Measured before the staging (profiles/
    // r01_gemm_epilogue_ablation.txt): the row-per-thread global accesses cost 9 us of a 34 us QKV GEMM and 19 us of a
    // 52 us FFN data-gradient GEMM; per-CTA time stamps (tools/gemm_trace.py, profiles/r01_gemm_trace.txt) then showed the
    // accumulator drain itself at 25% of a single-tile GEMM, which is what the 64-column chunks, the compile-time
    // epilogue modes and the register-held bias slice are for.
    static constexpr bool EPI_TMA = !(NCTA == 1 && BN == 256);  // 48 KB stages leave no room there
    static constexpr int EPI_WARP_BYTES = 8192;
    static constexpr int EPI_BYTES = EPI_TMA ? 8 * EPI_WARP_BYTES : 0;
    static constexpr int STAGES = (NCTA >= 2) ? (BN == 256 ? 5 : 6) : ((BN == 256) ? 4 : (BN == 128 ? 5 : 6));
    static constexpr int TMEM_COLS = (NCTA >= 2) ? 512 : ((2 * BN < 32) ? 32 : 2 * BN);  // two accumulator buffers
    static constexpr int BAR_BYTES = 512;  // full/empty[STAGES], tmem_full/empty[2], tmem slot, epilogue in-barriers[32]
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 /*align slack*/ + BAR_BYTES;
    static_assert(SMEM_BYTES <= 232448, "shared memory budget");
    static_assert(8 * (2 * STAGES + 4) + 8 + 32 * 8 <= BAR_BYTES, "barrier area");
};

struct Params {
    void* C;
    const float* bias;
    const void* residual;
    void* aux_out;
    const void* aux_in;
    int M, N, K;
    int ldc;
    int a_mn, b_mn;  // 1 = MN-major operand
    int act, mul_gelu_grad;
    float alpha;
    int batch0, batch1;
    long long sc0, sc1;
    int tiles_m, tiles_n;
    long long num_tiles;
    // operand addressing mode of the TMA producer
    //   0  plain GEMM
    //   1  convolution taps (grouped positional conv forward / data gradient): K-block kb is filter tap kb; the A rows
    //      of that block are the activation frames m0 + kb - tap_pad (out-of-range frames zero-filled by TMA)
    //   2  convolution weight gradient: both operands MN-major, K runs over (utterance, 64-frame block); N-block nb is
    //      filter tap nb, whose B rows are the activation frames t0 + nb - tap_pad
    int mode, tap_pad, tblocks, num_kb;
    long long bias_sb0;  // bias element stride per batch0 index (grouped conv: channels per group)
    int epi_tma;         // 1: staged epilogue (TMA store of C [+ second tensor]); 0: direct row-per-thread accesses
    int x_kind;          // second epilogue tensor of the staged path: 0 none, 1 aux_out (store), 2 residual, 3 aux_in (gelu')
    int epi_mode;        // staged path: 8 * ACTM + XK (see the epilogue)
    int dbg;             // experiment knobs (B200ASR_GEMM_DBG): 1 skip GELU math, 2 skip aux/residual traffic, 4 skip the C store
    // element dropout fused into the GELU_SAVE_GRAD epilogue (b200asr_gemm_dropout): both outputs are multiplied by
    // keep / (1 - p) with the keep decisions of dropout.cu (Philox keyed by (seed, site) over (element index / 8, step));
    // drop_on = 0 otherwise
    int drop_on;
    uint32_t drop_thresh16;
    float drop_scale;
    uint2 drop_key;
    unsigned long long drop_offset;
    const unsigned long long* drop_dev_offset;
    // ... or read from keep bits drawn ahead of the GEMM (b200asr_dropout_keep_bits: word [(n / 32) * drop_mpad + m], bit n & 31):
    // the four Philox calls per 32 columns are 45 % on top of the epilogue's own work, and the epilogue bounds the K = 1024 GEMMs
    const uint32_t* drop_bits;
    int drop_mpad;
    unsigned long long* trace;  // diagnostics (b200asr_gemm_set_trace): 12 time stamps per CTA, null in normal use
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// slot layout: 0 globaltimer at entry, 1 clock at entry, 2 after set-up, 3 after griddepcontrol.wait, 4 first operand stage
// landed (MMA issuer), 5 last MMA issued, 6 last accumulator complete (epilogue warp 2), 7 epilogue warp 2 done, 8 clock at
// exit, 9 globaltimer at exit
#define TRACE(slot) do { if (p.trace) p.trace[(long long)blockIdx.x * 20 + (slot)] = (unsigned long long)clock64(); } while (0)

template <typename TC> __device__ __forceinline__ void store_row_chunk(TC* dst, const float* v, int ncols_valid, bool vec_ok);
template <> __device__ __forceinline__ void store_row_chunk<float>(float* dst, const float* v, int n, bool vec_ok) {
    if (n >= 32 && vec_ok) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < n) dst[j] = v[j];
    }
}
template <> __device__ __forceinline__ void store_row_chunk<__nv_bfloat16>(__nv_bfloat16* dst, const float* v, int n, bool vec_ok) {
    if (n >= 32 && vec_ok) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
            uint4 o;
            o.x = bf2_pack(v[j], v[j + 1]);
            o.y = bf2_pack(v[j + 2], v[j + 3]);
            o.z = bf2_pack(v[j + 4], v[j + 5]);
            o.w = bf2_pack(v[j + 6], v[j + 7]);
            *reinterpret_cast<uint4*>(dst + j) = o;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < n) dst[j] = __float2bfloat16_rn(v[j]);
    }
}
template <typename TC> __device__ __forceinline__ void load_row_chunk(const TC* src, float* v, int n, bool vec_ok);
template <> __device__ __forceinline__ void load_row_chunk<float>(const float* src, float* v, int n, bool vec_ok) {
    if (n >= 32 && vec_ok) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const float4 t = *reinterpret_cast<const float4*>(src + j);
            v[j] = t.x; v[j + 1] = t.y; v[j + 2] = t.z; v[j + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = (j < n) ? src[j] : 0.f;
    }
}
template <> __device__ __forceinline__ void load_row_chunk<__nv_bfloat16>(const __nv_bfloat16* src, float* v, int n, bool vec_ok) {
    if (n >= 32 && vec_ok) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
            const uint4 t = *reinterpret_cast<const uint4*>(src + j);
            bf2_unpack(t.x, v[j], v[j + 1]);
            bf2_unpack(t.y, v[j + 2], v[j + 3]);
            bf2_unpack(t.z, v[j + 4], v[j + 5]);
            bf2_unpack(t.w, v[j + 6], v[j + 7]);
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = (j < n) ? __bfloat162float(src[j]) : 0.f;
    }
}

template <int BN, typename TC, int NCTA>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
               const __grid_constant__ CUtensorMap tma_c, const __grid_constant__ CUtensorMap tma_x, const Params p) {
    using C_ = Cfg<BN, NCTA>;
    constexpr int NP = C_::NP;
    constexpr bool PAIR = NCTA >= 2;        // tcgen05 cta_group::2
    constexpr bool MC = NCTA == 4;          // two pairs per cluster sharing the B tile by TMA multicast
    constexpr int TILE_M = BLOCK_M * NP;    // rows of one (pair) tile
    constexpr int BNH = BN / NP;            // B rows held by this CTA
    const uint32_t crank = PAIR ? cluster_ctarank() : 0u;  // rank in the cluster
    const uint32_t rank = crank & 1u;       // rank in the pair: 0 = pair leader (issues the MMAs)
    const uint32_t pair = crank >> 1;       // which pair of the cluster (NCTA = 4)
    const uint32_t leader = crank & ~1u;    // cluster rank of this pair's leader
    // worker = the unit that walks the tile list: a CTA, a pair, or (NCTA = 4) a cluster of two pairs whose "tile" is the
    // vertical pair of 256-row tiles (pair p takes the p-th); tiles_m then counts super rows
    const int worker = (NCTA == 4) ? (int)(blockIdx.x >> 2) : PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int nworkers = (NCTA == 4) ? (int)(gridDim.x >> 2) : PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment: the 128B swizzle pattern is a function of the absolute shared address
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t epi_smem = smem_base + C_::STAGES * C_::STAGE_BYTES;
    const uint32_t bars = epi_smem + C_::EPI_BYTES;
    // barrier layout (8 bytes each): full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], then tmem ptr
    auto full_bar = [&](int s) { return bars + 8u * s; };
    auto empty_bar = [&](int s) { return bars + 8u * (C_::STAGES + s); };
    auto tfull_bar = [&](int s) { return bars + 8u * (2 * C_::STAGES + s); };
    auto tempty_bar = [&](int s) { return bars + 8u * (2 * C_::STAGES + 2 + s); };
    const uint32_t tmem_slot = bars + 8u * (2 * C_::STAGES + 4);
    auto in_bar = [&](int ew, int i) { return bars + 8u * (2 * C_::STAGES + 5) + 8u * (4 * ew + i); };
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + C_::STAGES * C_::STAGE_BYTES + C_::EPI_BYTES + 8 * (2 * C_::STAGES + 4));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (p.trace && threadIdx.x == 0) { p.trace[(long long)blockIdx.x * 20] = globaltimer_ns(); TRACE(1); }

    if (warp == 0) {   // one pipeline barrier per lane: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2] lie back to back
        static_assert(2 * C_::STAGES + 4 <= 32, "one barrier per lane");
        if (lane < 2 * C_::STAGES + 4) {
            // empty (NCTA = 4): both pairs' MMAs must have drained the slot (multicast target);
            // tmem_empty: one arrival per epilogue warp (of both CTAs of a pair)
            const uint32_t cnt = (lane >= C_::STAGES && lane < 2 * C_::STAGES) ? (MC ? 2u : 1u)
                                 : (lane >= 2 * C_::STAGES + 2) ? (uint32_t)(8 * NP) : 1u;
            mbar_init(bars + 8u * lane, cnt);
        }
        fence_barrier_init();
        if (lane == 0) {
            tma_prefetch_desc(&tma_a);
            tma_prefetch_desc(&tma_b);
            if (p.epi_tma) {
                tma_prefetch_desc(&tma_c);
                if (p.x_kind) tma_prefetch_desc(&tma_x);
            }
        }
    }
    if (warp == 2) {   // the epilogue warps' operand-load barriers, one per lane
        mbar_init(in_bar(lane >> 2, lane & 3), 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        if constexpr (PAIR) tmem_alloc_2sm(tmem_slot, C_::TMEM_COLS);
        else tmem_alloc(tmem_slot, C_::TMEM_COLS);
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // the peers' barriers are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_gen;
    // Programmatic dependent launch: everything above (barrier init, TMEM allocation, descriptor prefetch) may overlap the
    // tail of the kernel this one depends on; nothing below may touch global memory before that kernel has completed.
    if (threadIdx.x == 0) TRACE(2);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0) TRACE(3);

    const int num_kb = p.num_kb;
    const int tiles_per_batch = p.tiles_m * p.tiles_n;   // (32-bit tile arithmetic: the host rejects tile counts beyond 2^31)
    const int num_tiles_i = (int)p.num_tiles;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one_sync()) {   // (not `lane == 0`: elect.sync keeps the region warp-uniform for the compiler, so the
            // TMA / MMA instructions take their uniform-register operands without a per-instruction election loop)
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = worker; tile < num_tiles_i; tile += nworkers) {
                const int z = tile / tiles_per_batch;
                const int rem = tile - z * tiles_per_batch;
                const int mb = rem / p.tiles_n, nb = rem % p.tiles_n;  // n fastest: neighbours share the A tile in L2
                const int b0 = z % p.batch0, b1 = z / p.batch0;
                const int m0 = (MC ? (2 * mb + (int)pair) : mb) * TILE_M + (int)rank * BLOCK_M, n0 = nb * BN + (int)rank * BNH;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    const uint32_t sa = smem_base + stage * C_::STAGE_BYTES;
                    const uint32_t sb = sa + A_STAGE_BYTES;
                    const int k0 = kb * BLOCK_K;
                    if constexpr (PAIR) {
                        // the loads of both CTAs complete on the pair LEADER's full barrier, which expects the pair's bytes
                        const uint32_t fb = mapa_shared(full_bar(stage), leader);
                        if (rank == 0) mbar_expect_tx(full_bar(stage), 2 * C_::STAGE_BYTES);
                        if (!p.a_mn) {
                            tma_load_4d_2sm(sa, &tma_a, fb, k0, m0, b0, b1);
                        } else {
#pragma unroll
                            for (int c = 0; c < BLOCK_M / 64; ++c)
                                tma_load_4d_2sm(sa + c * (64 * BLOCK_K * 2), &tma_a, fb, m0 + 64 * c, k0, b0, b1);
                        }
                        if constexpr (MC) {
                            // this CTA fetches quarter `pair` (64 of the BNH = 128 rows) of its B half and multicasts it to
                            // the CTA holding the same half in the other pair (cluster ranks rank and rank + 2); the bytes
                            // are credited to the leader of each destination's pair (peer bit of the barrier address cleared)
                            static_assert(!MC || BNH == 128, "multicast variant: 256-wide tiles");
                            const uint16_t mask = (uint16_t)((1u << rank) | (1u << (rank + 2)));
                            const uint32_t fbm = full_bar(stage) & 0xFEFFFFFFu;
                            const uint32_t dst = sb + pair * (64 * BLOCK_K * 2);
                            if (!p.b_mn) tma_load_4d_2sm_mc(dst, &tma_b, fbm, mask, k0, n0 + 64 * (int)pair, b0, b1);
                            else tma_load_4d_2sm_mc(dst, &tma_b, fbm, mask, n0 + 64 * (int)pair, k0, b0, b1);
                        } else if (!p.b_mn) {
                            tma_load_4d_2sm(sb, &tma_b, fb, k0, n0, b0, b1);
                        } else {
#pragma unroll
                            for (int c = 0; c < BNH / 64; ++c)
                                tma_load_4d_2sm(sb + c * (64 * BLOCK_K * 2), &tma_b, fb, n0 + 64 * c, k0, b0, b1);
                        }
                        if (++stage == C_::STAGES) { stage = 0; phase ^= 1u; }
                        continue;
                    }
                    mbar_expect_tx(full_bar(stage), C_::STAGE_BYTES);
                    if (p.mode != 0) {
                        if (p.mode == 1) {
                            tma_load_4d(sa, &tma_a, full_bar(stage), 0, m0 + kb - p.tap_pad, b0, b1);
                            tma_load_4d(sb, &tma_b, full_bar(stage), k0, n0, b0, 0);
                        } else {
                            const int bb = kb / p.tblocks, t0 = (kb % p.tblocks) * BLOCK_K;
#pragma unroll
                            for (int c = 0; c < BLOCK_M / 64; ++c)
                                tma_load_4d(sa + c * (64 * BLOCK_K * 2), &tma_a, full_bar(stage), m0 + 64 * c, t0, b0, bb);
                            // N-block nb holds the BN / 64 filter taps nb * (BN / 64) + c: each is the activation window
                            // shifted by its tap, one 64-channel box per tap (taps beyond the filter: columns >= N, never stored)
#pragma unroll
                            for (int c = 0; c < BN / 64; ++c)
                                tma_load_4d(sb + c * (64 * BLOCK_K * 2), &tma_b, full_bar(stage), 0,
                                            t0 + nb * (BN / 64) + c - p.tap_pad, b0, bb);
                        }
                        if (++stage == C_::STAGES) { stage = 0; phase ^= 1u; }
                        continue;
                    }
                    if (!p.a_mn) {
                        tma_load_4d(sa, &tma_a, full_bar(stage), k0, m0, b0, b1);
                    } else {
#pragma unroll
                        for (int c = 0; c < BLOCK_M / 64; ++c)
                            tma_load_4d(sa + c * (64 * BLOCK_K * 2), &tma_a, full_bar(stage), m0 + 64 * c, k0, b0, b1);
                    }
                    if (!p.b_mn) {
                        tma_load_4d(sb, &tma_b, full_bar(stage), k0, n0, b0, b1);
                    } else {
#pragma unroll
                        for (int c = 0; c < BN / 64; ++c)
                            tma_load_4d(sb + c * (64 * BLOCK_K * 2), &tma_b, full_bar(stage), n0 + 64 * c, k0, b0, b1);
                    }
                    if (++stage == C_::STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (rank == 0 && elect_one_sync()) {
            // instruction descriptor: D fp32, A/B bf16, majors, N>>3, M>>4 (cute::UMMA::InstrDescriptor)
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                                   ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
            // K-major SW128: 8-row groups 1024 B apart (SBO); LBO unused (1).  +32 B per UMMA_K inside the atom.
            // MN-major SW128: 64-element chunks along M/N are one TMA box (64 k-rows x 128 B = 8 KB) apart (LBO),
            //                 8-k-row groups 1024 B apart (SBO).  +2048 B per UMMA_K.
            const uint32_t a_lbo = p.a_mn ? 64 * BLOCK_K * 2 : 16, a_sbo = 1024;
            const uint32_t b_lbo = p.b_mn ? 64 * BLOCK_K * 2 : 16, b_sbo = 1024;
            const uint32_t a_kstep = p.a_mn ? (UMMA_K * 128) : (UMMA_K * 2);
            const uint32_t b_kstep = p.b_mn ? (UMMA_K * 128) : (UMMA_K * 2);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = worker; tile < num_tiles_i; tile += nworkers) {
                mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    if (kb == 0 && tile == worker) TRACE(4);
                    const uint32_t sa = smem_base + stage * C_::STAGE_BYTES;
                    const uint32_t sb = sa + A_STAGE_BYTES;
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        const uint64_t adesc = make_smem_desc(sa + k * a_kstep, a_lbo, a_sbo);
                        const uint64_t bdesc = make_smem_desc(sb + k * b_kstep, b_lbo, b_sbo);
                        if constexpr (PAIR) umma_bf16_2sm(tmem_d, adesc, bdesc, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                        else umma_bf16(tmem_d, adesc, bdesc, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    }
                    // frees the smem slot (in both CTAs of a pair) once these MMAs have read it
                    if constexpr (MC) umma_commit_2sm(empty_bar(stage), 0xF);   // all four CTAs: the slot is a multicast target
                    else if constexpr (PAIR) umma_commit_2sm(empty_bar(stage), 3);
                    else umma_commit(empty_bar(stage));
                    if (++stage == C_::STAGES) { stage = 0; phase ^= 1u; }
                }
                // accumulator complete -> epilogue warps (of both CTAs)
                if constexpr (PAIR) umma_commit_2sm(tfull_bar(acc), (uint16_t)(3u << leader));
                else umma_commit(tfull_bar(acc));
                TRACE(5);
                if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int quarter = warp & 3;  // TMEM lanes [32*quarter, 32*quarter+32) are accessible to this warp
        const int chalf = (warp - 2) >> 2;  // which half of the tile's columns this warp drains
        constexpr int CHUNKS = BN / 64;     // 32-column chunks per warp
        int acc = 0;
        uint32_t acc_phase = 0;
        TC* Cb = (TC*)p.C;
        const TC* Rb = (const TC*)p.residual;
        TC* AOb = (TC*)p.aux_out;
        const TC* AIb = (const TC*)p.aux_in;
        const bool vec_ok = (p.ldc % (16 / (int)sizeof(TC)) == 0) && ((p.sc0 | p.sc1) % (16 / (int)sizeof(TC)) == 0) &&
                            ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0) &&
                            (!p.residual || (reinterpret_cast<uintptr_t>(p.residual) & 15) == 0) &&
                            (!p.aux_out || (reinterpret_cast<uintptr_t>(p.aux_out) & 15) == 0) &&
                            (!p.aux_in || (reinterpret_cast<uintptr_t>(p.aux_in) & 15) == 0);
        bool staged = false;
        if constexpr (C_::EPI_TMA) staged = p.epi_tma != 0;
        if (staged) {
            if constexpr (C_::EPI_TMA) {
                // ---------- staged epilogue: TMEM -> registers -> swizzled shared chunk -> TMA store ----------
                // A chunk is 32 rows x 128 bytes (64 bf16 / 32 fp32 columns, swizzle-128B): two (bf16) or one (fp32)
                // 32-column TMEM reads feed one fence + one TMA store.  The math is specialised at compile time per
                // epilogue mode (ACTM, XK) so that every chunk is straight-line code:
                //   ACTM 1: GELU on the result.   XK 0: no second tensor, 1: second output = the pre-activation,
                //   2: result GELU, second output gelu' (forward of the FFN), 3: + residual, 4: * stored gelu',
                //   5: * gelu'(stored pre-activation).
                constexpr int ES = (int)sizeof(TC);
                constexpr int GC = (ES == 2) ? 2 : 1;              // 32-column accumulator groups per chunk
                constexpr int CW = 32 * GC;                         // chunk width in columns
                constexpr int CHUNK_BYTES = 4096;
                constexpr bool SOLO = (BN / 2) < CW;                // BN = 64, bf16: the tile is one chunk wide
                constexpr int CPT = SOLO ? 1 : (BN / 2) / CW;       // chunks per tile drained by this warp
                constexpr int NSLOT = C_::EPI_WARP_BYTES / CHUNK_BYTES;
                static_assert(NSLOT == 2, "two staging chunks per epilogue warp");
                const int ew = warp - 2;
                const uint32_t my = epi_smem + (uint32_t)ew * C_::EPI_WARP_BYTES;
                uint8_t* my_gen = smem_gen + C_::STAGES * C_::STAGE_BYTES + ew * C_::EPI_WARP_BYTES;
                const int sw = lane & 7;                            // swizzle-128B: 16-byte unit u of row r sits at u ^ (r & 7)
                const int ntiles = num_tiles_i, tpb = tiles_per_batch;
                const int col0 = SOLO ? 0 : chalf * CPT * CW;       // first tile column of this warp's part
                const bool works = !SOLO || chalf == 0;

                // first row / first column of this warp's part of a tile (32-bit arithmetic: once per tile)
                auto tile_coords = [&](int tile, int& m0w, int& nbase, int& b0, int& b1) {
                    const int z = tile / tpb, rem = tile - z * tpb;
                    const int mb = rem / p.tiles_n, nb = rem - mb * p.tiles_n;
                    b1 = z / p.batch0; b0 = z - b1 * p.batch0;
                    m0w = (MC ? (2 * mb + (int)pair) : mb) * TILE_M + (int)rank * BLOCK_M + quarter * 32;
                    nbase = nb * BN + col0;
                };
                // lane 0: fetch the second-tensor chunk of work item `it` (= CPT * tile ordinal + chunk) into slot it & 1
                auto issue_in = [&](int it) {
                    const int tile = worker + (it / CPT) * nworkers;
                    if (tile >= ntiles) return;
                    int m0w, nbase, b0, b1;
                    tile_coords(tile, m0w, nbase, b0, b1);
                    const int n = nbase + (it % CPT) * CW;
                    if (n >= p.N || m0w >= p.M) return;
                    const int sl = it & 1;
                    mbar_expect_tx(in_bar(ew, sl), CHUNK_BYTES);
                    tma_load_4d(my + sl * CHUNK_BYTES, &tma_x, in_bar(ew, sl), n, m0w, b0, b1);
                };

                auto run = [&](auto actm_c, auto xk_c) {
                    constexpr int ACTM = decltype(actm_c)::value;
                    constexpr int XK = decltype(xk_c)::value;
                    constexpr bool HAS_IN = XK >= 3 && XK <= 5;   // operand chunk TMA-loaded into the slot, result written in place
                    constexpr bool TWO_OUT = XK == 1 || XK == 2 || XK == 6;   // result in slot 0, second output in slot 1
                    [[maybe_unused]] unsigned long long drop_off = 0;
                    if constexpr (XK == 6) drop_off = p.drop_offset + (p.drop_dev_offset ? *p.drop_dev_offset : 0ull);
                    int acc = 0;
                    uint32_t acc_phase = 0, in_phase = 0;
                    if (HAS_IN && works && elect_one_sync()) { issue_in(0); issue_in(1); }
                    int item = 0;
                    for (int tile = worker; tile < ntiles; tile += nworkers) {
                        int m0w, nbase, b0, b1;
                        tile_coords(tile, m0w, nbase, b0, b1);
                        // this warp's bias slice: four columns per lane, broadcast by shuffles chunk by chunk (a global
                        // load per chunk costs an L2 round trip on the critical path of the accumulator drain)
                        const float* bpt = p.bias ? p.bias + (long long)b0 * p.bias_sb0 + nbase : nullptr;
                        const bool bias_fast = bpt && works && (p.N - nbase >= CPT * CW) && (reinterpret_cast<uintptr_t>(bpt) & 15) == 0;
                        float4 bq = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (bias_fast && lane * 4 < CPT * CW) bq = __ldg(reinterpret_cast<const float4*>(bpt) + lane);
                        mbar_wait(tfull_bar(acc), acc_phase);
                        tc_fence_after();
                        if (threadIdx.x == 64) TRACE(6);
                        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN + col0);
                        auto release_acc = [&]() {   // this warp has read its part of the accumulator: the MMA issuer may reuse it
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) {
                                if constexpr (PAIR) mbar_arrive_cluster(mapa_shared(tempty_bar(acc), leader));
                                else mbar_arrive(tempty_bar(acc));
                            }
                        };
                        if (!works) {
                            release_acc();
                            if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
                            continue;
                        }
#pragma unroll 1
                        for (int cc = 0; cc < CPT; ++cc, ++item) {
                            const int n = nbase + cc * CW;
                            if (n >= p.N || m0w >= p.M) {  // chunk entirely outside C: nothing staged, nothing waited for
                                if (cc == CPT - 1) release_acc();
                                if (HAS_IN && item >= 1 && elect_one_sync()) {
                                    bulk_wait_group_read<0>();
                                    issue_in(item + 1);
                                }
                                continue;
                            }
                            const int sl = TWO_OUT ? 0 : (item & 1);
                            uint8_t* orow = my_gen + sl * CHUNK_BYTES + lane * 128;      // result row (= operand row when HAS_IN)
                            [[maybe_unused]] uint8_t* xrow = my_gen + (TWO_OUT ? 1 : sl) * CHUNK_BYTES + lane * 128;
                            if constexpr (HAS_IN) {
                                mbar_wait(in_bar(ew, sl), (in_phase >> sl) & 1u);   // operand landed (issued after the slot's last store)
                                in_phase ^= 1u << sl;
                            } else {
                                if (lane == 0) {   // the stores that last read this chunk's slot(s) are done with them
                                    if (TWO_OUT) bulk_wait_group_read<0>();
                                    else bulk_wait_group_read<1>();
                                }
                                __syncwarp();
                            }
                            uint32_t rr[GC][32];   // the chunk's accumulator columns: all TMEM reads in flight, one wait
#pragma unroll
                            for (int g = 0; g < GC; ++g) tmem_ld_32x32(taddr + cc * CW + g * 32, rr[g]);
                            tmem_ld_wait();
                            if (cc == CPT - 1) release_acc();
#pragma unroll
                            for (int g = 0; g < GC; ++g) {
                                const uint32_t* r = rr[g];
                                float2 v2[16];   // the 32 accumulator columns of this lane's row, as pairs for the packed fp32 pipe
                                float* v = reinterpret_cast<float*>(v2);
                                const float2 al = make_float2(p.alpha, p.alpha);
                                if (bias_fast) {
                                    const int src0 = (cc * CW + g * 32) >> 2;
#pragma unroll
                                    for (int j = 0; j < 32; j += 4) {
                                        const int src = src0 + (j >> 2);
                                        const float2 b01 = make_float2(__shfl_sync(0xffffffffu, bq.x, src), __shfl_sync(0xffffffffu, bq.y, src));
                                        const float2 b23 = make_float2(__shfl_sync(0xffffffffu, bq.z, src), __shfl_sync(0xffffffffu, bq.w, src));
                                        v2[j / 2] = __ffma2_rn(make_float2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])), al, b01);
                                        v2[j / 2 + 1] = __ffma2_rn(make_float2(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])), al, b23);
                                    }
                                } else {
#pragma unroll
                                    for (int j = 0; j < 32; j += 2)
                                        v2[j / 2] = __fmul2_rn(make_float2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])), al);
                                    if (bpt) {
                                        const int nvalid = p.N - (n + g * 32);
                                        const float* bp = bpt + cc * CW + g * 32;
#pragma unroll
                                        for (int j = 0; j < 32; ++j)
                                            if (j < nvalid) v[j] += __ldg(bp + j);
                                    }
                                }
                                if constexpr (ES == 2) {
                                    auto pack8 = [](const float* f) {
                                        uint4 o;
                                        o.x = bf2_pack(f[0], f[1]); o.y = bf2_pack(f[2], f[3]);
                                        o.z = bf2_pack(f[4], f[5]); o.w = bf2_pack(f[6], f[7]);
                                        return o;
                                    };
                                    if constexpr (XK == 1) {          // second output: the pre-activation
#pragma unroll
                                        for (int q = 0; q < 4; ++q)
                                            *reinterpret_cast<uint4*>(xrow + (((g * 4 + q) ^ sw) << 4)) = pack8(v + 8 * q);
                                    }
                                    if constexpr (XK == 2 || XK == 6) {   // result gelu, second output gelu' (6: both * keep / (1 - p))
                                        float2 g2[16];
#pragma unroll
                                        for (int j = 0; j < 16; ++j) {
                                            float2 gv;
                                            gelu_both_fast2(v2[j], gv, g2[j]);
                                            v2[j] = gv;
                                        }
                                        if constexpr (XK == 6) {
                                            // this lane's 32 columns of row m0w + lane = 4 Philox blocks of 8 elements of the
                                            // row-major [M, N] result (ldc = N, N % 8 == 0: checked by the caller)
                                            if (p.drop_bits) {
                                                const uint32_t kb = __ldg(p.drop_bits + (long long)((n + g * 32) >> 5) * p.drop_mpad + (m0w + lane));
#pragma unroll
                                                for (int j = 0; j < 16; ++j) {
                                                    const float2 mk = make_float2((kb >> (2 * j)) & 1u ? p.drop_scale : 0.f,
                                                                                  (kb >> (2 * j + 1)) & 1u ? p.drop_scale : 0.f);
                                                    v2[j] = __fmul2_rn(v2[j], mk);
                                                    g2[j] = __fmul2_rn(g2[j], mk);
                                                }
                                            } else {
                                                const unsigned long long e0 =
                                                    (unsigned long long)(m0w + lane) * (unsigned long long)p.ldc + (unsigned long long)(n + g * 32);
#pragma unroll
                                                for (int q = 0; q < 4; ++q) {
                                                    uint32_t u[8];
                                                    dropout_uniforms<8>(e0 + 8 * q, p.drop_key, drop_off, u);
#pragma unroll
                                                    for (int e = 0; e < 8; e += 2) {
                                                        const float2 mk = make_float2(u[e] >= p.drop_thresh16 ? p.drop_scale : 0.f,
                                                                                      u[e + 1] >= p.drop_thresh16 ? p.drop_scale : 0.f);
                                                        v2[4 * q + e / 2] = __fmul2_rn(v2[4 * q + e / 2], mk);
                                                        g2[4 * q + e / 2] = __fmul2_rn(g2[4 * q + e / 2], mk);
                                                    }
                                                }
                                            }
                                        }
                                        const float* gf = reinterpret_cast<const float*>(g2);
#pragma unroll
                                        for (int q = 0; q < 4; ++q)
                                            *reinterpret_cast<uint4*>(xrow + (((g * 4 + q) ^ sw) << 4)) = pack8(gf + 8 * q);
                                    } else if constexpr (ACTM == 1) {
#pragma unroll
                                        for (int j = 0; j < 16; ++j) v2[j] = gelu_fast2(v2[j]);
                                    }
                                    if constexpr (HAS_IN) {
                                        uint4 t[4];
#pragma unroll
                                        for (int q = 0; q < 4; ++q) t[q] = *reinterpret_cast<const uint4*>(xrow + (((g * 4 + q) ^ sw) << 4));
#pragma unroll
                                        for (int q = 0; q < 4; ++q) {
                                            float u[8];
                                            bf2_unpack(t[q].x, u[0], u[1]); bf2_unpack(t[q].y, u[2], u[3]);
                                            bf2_unpack(t[q].z, u[4], u[5]); bf2_unpack(t[q].w, u[6], u[7]);
#pragma unroll
                                            for (int e = 0; e < 8; e += 2) {
                                                const float2 uu = make_float2(u[e], u[e + 1]);
                                                float2& d = v2[4 * q + e / 2];
                                                if constexpr (XK == 3) d = __fadd2_rn(d, uu);
                                                else if constexpr (XK == 4) d = __fmul2_rn(d, uu);
                                                else d = __fmul2_rn(d, gelu_grad_fast2(uu));
                                            }
                                        }
                                    }
#pragma unroll
                                    for (int q = 0; q < 4; ++q)
                                        *reinterpret_cast<uint4*>(orow + (((g * 4 + q) ^ sw) << 4)) = pack8(v + 8 * q);
                                } else {
                                    if constexpr (ACTM == 1) {
#pragma unroll
                                        for (int j = 0; j < 16; ++j) v2[j] = gelu_fast2(v2[j]);
                                    }
#pragma unroll
                                    for (int q = 0; q < 8; ++q)
                                        *reinterpret_cast<float4*>(orow + ((q ^ sw) << 4)) =
                                            make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                                }
                            }
                            fence_proxy_async();
                            __syncwarp();
                            if (elect_one_sync()) {
                                tma_store_4d(&tma_c, my + sl * CHUNK_BYTES, n, m0w, b0, b1);
                                if constexpr (TWO_OUT) tma_store_4d(&tma_x, my + CHUNK_BYTES, n, m0w, b0, b1);
                                bulk_commit_group();
                                if (HAS_IN && item >= 1) {
                                    // the store of chunk item-1 has read its slot: refill it with the operand of chunk item+1
                                    bulk_wait_group_read<1>();
                                    issue_in(item + 1);
                                }
                            }
                            if (cc == 0 && threadIdx.x == 64) TRACE(10);
                        }
                        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
                    }
                };
                using I0 = std::integral_constant<int, 0>;
                using I1 = std::integral_constant<int, 1>;
                if constexpr (ES == 4) {
                    if (p.epi_mode & 8) run(I1{}, I0{});
                    else run(I0{}, I0{});
                } else {
                    switch (p.epi_mode) {
                        case 0: run(I0{}, I0{}); break;
                        case 8: run(I1{}, I0{}); break;
                        case 1: run(I0{}, I1{}); break;
                        case 9: run(I1{}, I1{}); break;
                        case 2: run(I0{}, std::integral_constant<int, 2>{}); break;
                        case 6: run(I0{}, std::integral_constant<int, 6>{}); break;
                        case 3: run(I0{}, std::integral_constant<int, 3>{}); break;
                        case 11: run(I1{}, std::integral_constant<int, 3>{}); break;
                        case 4: run(I0{}, std::integral_constant<int, 4>{}); break;
                        default: run(I0{}, std::integral_constant<int, 5>{}); break;
                    }
                }
                if (threadIdx.x == 64) TRACE(11);
                if (lane == 0) bulk_wait_group_read<0>();   // shared memory may be released; the writes complete with the grid
                if (threadIdx.x == 64) TRACE(7);
            }
        } else {
        for (int tile = worker; tile < num_tiles_i; tile += nworkers) {
                const int z = tile / tiles_per_batch;
                const int rem = tile - z * tiles_per_batch;
                const int mb = rem / p.tiles_n, nb = rem % p.tiles_n;
                const int b0 = z % p.batch0, b1 = z / p.batch0;
                const int m = (MC ? (2 * mb + (int)pair) : mb) * TILE_M + (int)rank * BLOCK_M + quarter * 32 + lane;
                const int n0 = nb * BN;
                const long long row_off = (long long)b0 * p.sc0 + (long long)b1 * p.sc1 + (long long)m * p.ldc;
                mbar_wait(tfull_bar(acc), acc_phase);
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
    #pragma unroll 1
                for (int c = chalf * CHUNKS; c < (chalf + 1) * CHUNKS; ++c) {
                    uint32_t r[32];
                    tmem_ld_32x32(taddr + c * 32, r);
                    tmem_ld_wait();
                    const int n = n0 + c * 32;
                    const int nvalid = p.N - n;  // may be <= 0 or > 32
                    if (m < p.M && nvalid > 0) {
                        float v[32];
    #pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
                        if (p.bias) {
                            const float* bp = p.bias + (long long)b0 * p.bias_sb0 + n;
                            if (nvalid >= 32 && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
    #pragma unroll
                                for (int j = 0; j < 32; j += 4) {
                                    const float4 bv = __ldg(reinterpret_cast<const float4*>(bp + j));
                                    v[j] += bv.x; v[j + 1] += bv.y; v[j + 2] += bv.z; v[j + 3] += bv.w;
                                }
                            } else {
    #pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (j < nvalid) v[j] += __ldg(bp + j);
                            }
                        }
                        float mk[32];   // fused element dropout (drop_on): keep / (1 - p) per column of this row chunk
                        if (p.drop_on && p.drop_bits) {
                            const uint32_t kb = __ldg(p.drop_bits + (long long)(n >> 5) * p.drop_mpad + m);
    #pragma unroll
                            for (int j = 0; j < 32; ++j) mk[j] = (kb >> j) & 1u ? p.drop_scale : 0.f;
                        } else if (p.drop_on) {
                            const unsigned long long doff = p.drop_offset + (p.drop_dev_offset ? *p.drop_dev_offset : 0ull);
                            const unsigned long long e0 = (unsigned long long)m * (unsigned long long)p.ldc + (unsigned long long)n;
    #pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint32_t u[8];
                                dropout_uniforms<8>(e0 + 8 * q, p.drop_key, doff, u);
    #pragma unroll
                                for (int e = 0; e < 8; ++e) mk[8 * q + e] = u[e] >= p.drop_thresh16 ? p.drop_scale : 0.f;
                            }
                        }
                        if (AOb && !(p.dbg & 2)) {
                            if (p.act == B200ASR_ACT_GELU_SAVE_GRAD) {
                                float gp[32];
    #pragma unroll
                                for (int j = 0; j < 32; ++j) gp[j] = gelu_grad_fast(v[j]);
                                if (p.drop_on) {
    #pragma unroll
                                    for (int j = 0; j < 32; ++j) gp[j] *= mk[j];
                                }
                                store_row_chunk<TC>(AOb + row_off + n, gp, nvalid, vec_ok);
                            } else {
                                store_row_chunk<TC>(AOb + row_off + n, v, nvalid, vec_ok);
                            }
                        }
                        if (p.act != B200ASR_ACT_NONE && !(p.dbg & 1)) {
    #pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = gelu_fast(v[j]);
                        }
                        if (p.drop_on) {
    #pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] *= mk[j];
                        }
                        if (p.mul_gelu_grad && !(p.dbg & 2)) {
                            float u[32];
                            load_row_chunk<TC>(AIb + row_off + n, u, nvalid, vec_ok);
                            if (p.mul_gelu_grad == 2) {
    #pragma unroll
                                for (int j = 0; j < 32; ++j) v[j] *= u[j];
                            } else if (!(p.dbg & 1)) {
    #pragma unroll
                                for (int j = 0; j < 32; ++j) v[j] *= gelu_grad_fast(u[j]);
                            } else {
    #pragma unroll
                                for (int j = 0; j < 32; ++j) v[j] += u[j];
                            }
                        }
                        if (Rb && !(p.dbg & 2)) {
                            float u[32];
                            load_row_chunk<TC>(Rb + row_off + n, u, nvalid, vec_ok);
    #pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] += u[j];
                        }
                        if (!(p.dbg & 4) || v[0] == 123.456f) store_row_chunk<TC>(Cb + row_off + n, v, nvalid, vec_ok);
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (PAIR) mbar_arrive_cluster(mapa_shared(tempty_bar(acc), leader));
                    else mbar_arrive(tempty_bar(acc));
                }
                if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // no CTA of the cluster exits while a peer may still signal / read it
    if (warp == 1) {
        tc_fence_after();
        if constexpr (PAIR) tmem_dealloc_2sm(tmem_base, C_::TMEM_COLS);
        else tmem_dealloc(tmem_base, C_::TMEM_COLS);
    }
    if (p.trace && threadIdx.x == 0) { TRACE(8); p.trace[(long long)blockIdx.x * 20 + 9] = globaltimer_ns(); }
}

// ---------------------------------------------------------------- host side
EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(f);
    });
    return fn;
}

int encode_map_bf16_4d(CUtensorMap* map, const void* base, const cuuint64_t dims[4], const cuuint64_t strides_bytes[3],
                       const cuuint32_t box[4], const char* what) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("%s: cuTensorMapEncodeTiled entry point unavailable", what);
        return B200ASR_ERR_CUDA;
    }
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides_bytes, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("%s: cuTensorMapEncodeTiled failed (%d) dims=(%llu,%llu,%llu,%llu) strides=(%llu,%llu,%llu)", what, (int)r,
                  (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2],
                  (unsigned long long)dims[3], (unsigned long long)strides_bytes[0], (unsigned long long)strides_bytes[1],
                  (unsigned long long)strides_bytes[2]);
        return B200ASR_ERR_CUDA;
    }
    return B200ASR_OK;
}

// operand matrix X(r, k): kmajor -> X[r*ld + k], else X[k*ld + r];  r < R, k < K; batched with strides s0, s1.
static int make_operand_map(CUtensorMap* map, const void* base, int R, int K, int ld, int kmajor, int box_r, int batch0,
                            int batch1, long long s0, long long s1) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("gemm(tcgen05): cuTensorMapEncodeTiled entry point unavailable");
        return B200ASR_ERR_CUDA;
    }
    cuuint64_t dims[4];
    cuuint64_t strides[3];
    cuuint32_t box[4];
    cuuint32_t estr[4] = {1, 1, 1, 1};
    const long long row_bytes = (long long)ld * 2;
    if (kmajor) {
        dims[0] = (cuuint64_t)K; dims[1] = (cuuint64_t)R;
        box[0] = BLOCK_K; box[1] = (cuuint32_t)box_r;
    } else {
        dims[0] = (cuuint64_t)R; dims[1] = (cuuint64_t)K;
        box[0] = 64; box[1] = BLOCK_K;
    }
    dims[2] = (cuuint64_t)batch0; dims[3] = (cuuint64_t)batch1;
    box[2] = 1; box[3] = 1;
    strides[0] = (cuuint64_t)row_bytes;
    // unit batch dims still need a legal (multiple-of-16) stride
    strides[1] = (cuuint64_t)((batch0 > 1 ? s0 : 8) * 2);
    strides[2] = (cuuint64_t)((batch1 > 1 ? s1 : 8) * 2);
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("gemm(tcgen05): cuTensorMapEncodeTiled failed (%d) R=%d K=%d ld=%d kmajor=%d b=(%d,%d) s=(%lld,%lld)", (int)r,
                  R, K, ld, kmajor, batch0, batch1, s0, s1);
        return B200ASR_ERR_CUDA;
    }
    return B200ASR_OK;
}

template <int BN, typename TC, int NCTA = 1>
static int launch_kernel(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& mx,
                         const Params& p, cudaStream_t st) {
    using C_ = Cfg<BN, NCTA>;
    if (p.num_tiles <= 0 || p.num_tiles > 0x7fffffffLL / 8) {   // the kernel walks the tile list with 32-bit indices
        set_error("gemm(tcgen05): %lld tiles are outside the supported range", p.num_tiles);
        return B200ASR_ERR_UNSUPPORTED;
    }
    static bool attr_set = false;  // per instantiation
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN, TC, NCTA>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             C_::SMEM_BYTES);
        if (e != cudaSuccess) {
            set_error("gemm(tcgen05): cudaFuncSetAttribute failed: %s", cudaGetErrorString(e));
            return B200ASR_ERR_CUDA;
        }
        attr_set = true;
    }
    long long workers = num_sms() / NCTA;
    if (NCTA == 4) {  // clusters of 4 must fit inside a GPC: ask how many can be co-resident (persistent kernel, one wave)
        static int max_clusters = -1;
        if (max_clusters < 0) {
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(NCTA * (num_sms() / NCTA));
            q.blockDim = dim3(NUM_THREADS);
            q.dynamicSmemBytes = C_::SMEM_BYTES;
            cudaLaunchAttribute qa[1];
            qa[0].id = cudaLaunchAttributeClusterDimension;
            qa[0].val.clusterDim.x = NCTA; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
            q.attrs = qa; q.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, gemm_tc_kernel<BN, TC, NCTA>, &q) != cudaSuccess || n <= 0) {
                (void)cudaGetLastError();
                n = num_sms() / NCTA;
            }
            max_clusters = n;
        }
        if (workers > max_clusters) workers = max_clusters;
    }
    const long long grid = (p.num_tiles < workers ? p.num_tiles : workers) * NCTA;
    static const int use_pdl = getenv("B200ASR_PDL") ? atoi(getenv("B200ASR_PDL")) : 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = C_::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (NCTA > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = NCTA;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    if (use_pdl) {  // the kernel's prologue may start while its predecessor on the stream drains (griddepcontrol.wait inside)
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_tc_kernel<BN, TC, NCTA>, ma, mb, mc, mx, p);
    if (e != cudaSuccess) {
        set_error("gemm(tcgen05): launch failed: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return B200ASR_ERR_CUDA;
    }
    return check_launch("gemm_tcgen05");
}

// 32 x 32 chunk map over an epilogue tensor X[b1][b0][m][n] (strides sc1, sc0, ld; same geometry as C).  Returns false
// when TMA cannot address it (alignment), in which case the launch falls back to the direct epilogue.
template <typename TC>
static bool make_epilogue_map(CUtensorMap* map, const void* base, int M, int N, int ld, int batch0, int batch1, long long s0,
                              long long s1) {
    EncodeTiledFn enc = get_encode_fn();
    constexpr int ES = (int)sizeof(TC);
    if (!enc || !base || ((uintptr_t)base & 15) || ((long long)ld * ES) % 16) return false;
    if ((batch0 > 1 && (s0 * ES) % 16) || (batch1 > 1 && (s1 * ES) % 16)) return false;
    const cuuint64_t dims[4] = {(cuuint64_t)N, (cuuint64_t)M, (cuuint64_t)batch0, (cuuint64_t)batch1};
    // unit batch dims still need a legal (multiple-of-16-byte) stride
    const cuuint64_t strides[3] = {(cuuint64_t)ld * ES, (cuuint64_t)(batch0 > 1 ? s0 * ES : 16),
                                   (cuuint64_t)(batch1 > 1 ? s1 * ES : 16)};
    const cuuint32_t box[4] = {(cuuint32_t)(128 / ES), 32, 1, 1};   // one staged chunk: 32 rows of 128 bytes
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, ES == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
                     const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

// keys of the element dropout fused into the next launch's GELU_SAVE_GRAD epilogue (set by gemm_tc_launch_dropout around
// its call; thread-local: the argument struct of the C ABI stays as it is)
struct EpiDrop {
    DropKeys k;
    const uint32_t* bits;   // keep bits drawn ahead (b200asr_dropout_keep_bits) or null: Philox in the epilogue
    int mpad;
};
static thread_local const EpiDrop* t_epi_drop = nullptr;

template <int BN, typename TC, int NCTA = 1>
static int launch(const b200asr_gemm_args& a, cudaStream_t st) {
    CUtensorMap ma, mb, mc, mx;
    memset(&mc, 0, sizeof(mc));
    memset(&mx, 0, sizeof(mx));
    int rc = make_operand_map(&ma, a.A, a.M, a.K, a.lda, a.a_kmajor, BLOCK_M, a.batch0, a.batch1, a.sa0, a.sa1);
    if (rc) return rc;
    // B box rows: the whole tile (single CTA), this CTA's half (pair), or one multicast quarter (two pairs)
    rc = make_operand_map(&mb, a.B, a.N, a.K, a.ldb, a.b_kmajor, NCTA == 4 ? 64 : BN / NCTA, a.batch0, a.batch1, a.sb0, a.sb1);
    if (rc) return rc;
    Params p;
    p.C = a.C; p.bias = a.bias; p.residual = a.residual; p.aux_out = a.aux_out; p.aux_in = a.aux_in;
    p.M = a.M; p.N = a.N; p.K = a.K; p.ldc = a.ldc;
    p.a_mn = a.a_kmajor ? 0 : 1; p.b_mn = a.b_kmajor ? 0 : 1;
    p.act = a.act; p.mul_gelu_grad = a.mul_gelu_grad; p.alpha = a.alpha;
    p.batch0 = a.batch0; p.batch1 = a.batch1; p.sc0 = a.sc0; p.sc1 = a.sc1;
    p.tiles_m = ceil_div(a.M, BLOCK_M * NCTA);  // NCTA = 4: super rows of 512 (two vertically adjacent pair tiles)
    p.tiles_n = ceil_div(a.N, BN);
    p.num_tiles = (long long)p.tiles_m * p.tiles_n * a.batch0 * a.batch1;
    p.mode = 0; p.tap_pad = 0; p.tblocks = 1; p.num_kb = ceil_div(a.K, BLOCK_K); p.bias_sb0 = 0;
    static const int dbg = getenv("B200ASR_GEMM_DBG") ? atoi(getenv("B200ASR_GEMM_DBG")) : 0;
    p.dbg = dbg;
    p.trace = g_gemm_trace;
    // staged epilogue when the kernel variant has the shared memory for it and at most ONE second tensor is involved
    p.epi_tma = 0; p.x_kind = 0; p.epi_mode = 0;
    p.drop_on = 0; p.drop_thresh16 = 0; p.drop_scale = 1.f; p.drop_key = make_uint2(0, 0); p.drop_offset = 0;
    p.drop_dev_offset = nullptr; p.drop_bits = nullptr; p.drop_mpad = 0;
    if (t_epi_drop) {
        const DropKeys& k = t_epi_drop->k;
        p.drop_on = 1; p.drop_thresh16 = k.thresh16; p.drop_scale = k.scale; p.drop_key = k.key;
        p.drop_offset = k.offset; p.drop_dev_offset = k.dev_offset;
        p.drop_bits = t_epi_drop->bits; p.drop_mpad = t_epi_drop->mpad;
    }
    if (Cfg<BN, NCTA>::EPI_TMA && !(dbg & 8)) {
        const void* xs[3] = {a.aux_out, a.residual, a.mul_gelu_grad ? a.aux_in : nullptr};
        int nx = 0, kind = 0;
        for (int i = 0; i < 3; ++i)
            if (xs[i]) { ++nx; kind = i + 1; }
        const bool f32c = sizeof(TC) == 4;
        // epilogue mode 8 * ACTM + XK; combinations outside the specialised list take the direct path
        const bool gelu = a.act == B200ASR_ACT_GELU || a.act == B200ASR_ACT_GELU_SAVE_GRAD;
        int mode = -1;
        if (nx == 0) mode = gelu ? 8 : 0;
        else if (nx == 1 && kind == 1) mode = a.act == B200ASR_ACT_GELU_SAVE_GRAD ? (p.drop_on ? 6 : 2) : (a.act == B200ASR_ACT_GELU ? 9 : 1);
        else if (nx == 1 && kind == 2) mode = gelu ? 11 : 3;
        else if (nx == 1 && kind == 3 && !gelu) mode = a.mul_gelu_grad == 2 ? 4 : 5;
        p.epi_mode = mode < 0 ? 0 : mode;
        if (mode >= 0 && nx <= 1 && !(f32c && nx) && make_epilogue_map<TC>(&mc, a.C, a.M, a.N, a.ldc, a.batch0, a.batch1, a.sc0, a.sc1) &&
            (nx == 0 || make_epilogue_map<TC>(&mx, xs[kind - 1], a.M, a.N, a.ldc, a.batch0, a.batch1, a.sc0, a.sc1))) {
            p.epi_tma = 1;
            p.x_kind = kind;
        }
    }
    return launch_kernel<BN, TC, NCTA>(ma, mb, mc, mx, p, st);
}

// ---------------------------------------------------------------- grouped positional conv on the same kernel
// x, y channels-last [B, T, C] bf16, 64 channels per group.  w_pack [groups][64 n][k][64 c] (K-major rows of k*64).
//   forward:       n = out channel o, c = in channel i,  tap_pad = k/2,         w_pack[g][o][j][i] = W[g*64+o][i][j]
//   data gradient: n = in channel i,  c = out channel o, tap_pad = k-1-k/2,     w_pack[g][i][j][o] = W[g*64+o][i][k-1-j]
static int posconv_taps_launch(const void* x, const void* w_pack, const float* bias, void* y, int B, int T, int C,
                               int groups, int k, int tap_pad, cudaStream_t st) {
    CUtensorMap ma, mb;
    {
        const cuuint64_t dims[4] = {64, (cuuint64_t)T, (cuuint64_t)groups, (cuuint64_t)B};
        const cuuint64_t strides[3] = {(cuuint64_t)C * 2, 128, (cuuint64_t)T * C * 2};
        const cuuint32_t box[4] = {64, BLOCK_M, 1, 1};
        int rc = encode_map_bf16_4d(&ma, x, dims, strides, box, "posconv(tcgen05) A");
        if (rc) return rc;
    }
    {
        const cuuint64_t dims[4] = {(cuuint64_t)k * 64, 64, (cuuint64_t)groups, 1};
        const cuuint64_t strides[3] = {(cuuint64_t)k * 64 * 2, (cuuint64_t)64 * k * 64 * 2, 16};
        const cuuint32_t box[4] = {64, 64, 1, 1};
        int rc = encode_map_bf16_4d(&mb, w_pack, dims, strides, box, "posconv(tcgen05) B");
        if (rc) return rc;
    }
    Params p = {};
    p.C = y; p.bias = bias; p.M = T; p.N = 64; p.K = k * 64; p.ldc = C; p.alpha = 1.f;
    p.batch0 = groups; p.batch1 = B; p.sc0 = 64; p.sc1 = (long long)T * C;
    p.tiles_m = ceil_div(T, BLOCK_M); p.tiles_n = 1;
    p.num_tiles = (long long)p.tiles_m * groups * B;
    p.mode = 1; p.tap_pad = tap_pad; p.tblocks = 1; p.num_kb = k; p.bias_sb0 = 64;
    return launch_kernel<64, __nv_bfloat16>(ma, mb, ma, ma, p, st);  // epi_tma = 0: the epilogue maps are unused
}

// dwt[g*64+o][j][i] = sum_{b,t} dy[b,t,g*64+o] * x[b, t+j-pad, g*64+i]   (fp32, tap-major; permuted by the caller)
static int posconv_wgrad_launch(const void* x, const void* dy, float* dwt, int B, int T, int C, int groups, int k, int pad,
                                cudaStream_t st) {
    CUtensorMap ma, mb;
    const cuuint64_t dims[4] = {64, (cuuint64_t)T, (cuuint64_t)groups, (cuuint64_t)B};
    const cuuint64_t strides[3] = {(cuuint64_t)C * 2, 128, (cuuint64_t)T * C * 2};
    const cuuint32_t box[4] = {64, BLOCK_K, 1, 1};
    int rc = encode_map_bf16_4d(&ma, dy, dims, strides, box, "posconv_wgrad(tcgen05) A");
    if (rc) return rc;
    rc = encode_map_bf16_4d(&mb, x, dims, strides, box, "posconv_wgrad(tcgen05) B");
    if (rc) return rc;
    Params p = {};
    p.C = dwt; p.M = 64; p.N = k * 64; p.K = B * T; p.ldc = k * 64; p.alpha = 1.f;
    p.a_mn = 1; p.b_mn = 1;
    p.batch0 = groups; p.batch1 = 1; p.sc0 = (long long)64 * k * 64; p.sc1 = 0;
    // four taps per tile (BN = 256): the dy tile is fetched once per four taps instead of once per tap -- the one-tap tiles
    // were bound by the L2 -> shared-memory fill (69 us per layer at 10 % of the tensor peak)
    p.tiles_m = 1; p.tiles_n = ceil_div(k, 4);
    p.num_tiles = (long long)p.tiles_n * groups;
    p.mode = 2; p.tap_pad = pad; p.tblocks = ceil_div(T, BLOCK_K); p.num_kb = B * p.tblocks; p.bias_sb0 = 0;
    return launch_kernel<256, float>(ma, mb, ma, ma, p, st);  // epi_tma = 0
}

}  // namespace tc

bool gemm_tc_eligible(const b200asr_gemm_args& a, const char** why) {
    auto fail = [&](const char* w) { if (why) *why = w; return false; };
    if (a.dtype != B200ASR_BF16) return fail("operands are not bf16");
    if (a.M < 1 || a.N < 8 || a.K < 8) return fail("N or K below 8");
    if (((uintptr_t)a.A & 15) || ((uintptr_t)a.B & 15)) return fail("operand base not 16-byte aligned");
    if (a.lda % 8 || a.ldb % 8) return fail("leading dimension not a multiple of 8 elements");
    if ((a.batch0 > 1 && (a.sa0 % 8 || a.sb0 % 8)) || (a.batch1 > 1 && (a.sa1 % 8 || a.sb1 % 8)))
        return fail("batch stride not a multiple of 8 elements");
    return true;
}

static int g_force_cfg = 0;  // test hook: 0 = cost model, else (ncta * 1000 + bn)

int gemm_tc_force_config(int cfg) {
    const int old = g_force_cfg;
    g_force_cfg = cfg;
    return old;
}

// Tile shape / CTA-pair selection by a two-term cost model (cycles): tensor-pipe time of the busiest worker
// (waves x 2*BN cycles per 64-deep K block) against the chip-wide L2 -> shared-memory fill time (bytes of every
// staged operand tile / ~6.3 KB per clock, measured: profiles/r01_gemm_ncu_*.txt), plus a per-tile epilogue term.
static void choose_config(const b200asr_gemm_args& a, int* ncta_out, int* bn_out) {
    const long long batches = (long long)a.batch0 * a.batch1;
    const int num_kb = ceil_div(a.K, tc::BLOCK_K);
    const int sms = num_sms();
    double best = 1e30;
    int best_ncta = 1, best_bn = 128;
    // The multicast cluster variant (NCTA = 4) is correct on every operand layout (tests force it) but measured no faster
    // than the pair tile on the model's shapes and 0.8% slower in the full step (profiles/r01_gemm_multicast_ab.txt): the
    // pair kernel is not bound by L2 read traffic after all.  Opt-in via B200ASR_GEMM_MC=1.
    static const int allow4 = getenv("B200ASR_GEMM_MC") ? atoi(getenv("B200ASR_GEMM_MC")) : 0;
    for (int ncta = 1; ncta <= 4; ncta *= 2) {
        if (ncta >= 2 && a.M <= 128) continue;
        if (ncta == 4 && (!allow4 || a.M <= 256)) continue;
        for (int bn = 64; bn <= 256; bn *= 2) {
            if (ncta >= 2 && bn == 64) continue;
            if (ncta == 4 && bn != 256) continue;
            if (bn > 64 && a.N <= bn / 2) continue;  // more than half of the tile would be padding
            const long long tiles = (long long)ceil_div(a.M, 128 * ncta) * ceil_div(a.N, bn) * batches;
            const long long workers = ncta == 4 ? 32 : sms / ncta;  // clusters of 4: GPC granularity leaves ~32 co-resident
            const long long waves = (tiles + workers - 1) / workers;
            const double mma = (double)waves * 2.0 * bn * num_kb;
            // bytes fetched from L2 per CTA and K block: its 128 A rows + its share of the B tile (half of it in a pair,
            // a quarter in the multicast cluster of two pairs)
            const double bytes = (double)tiles * ncta * (128.0 + (double)bn / ncta) * 128.0 * num_kb;
            const double l2 = bytes / 6300.0;
            const double epi = (double)waves * (600.0 + 8.0 * bn);
            const double cost = (mma > l2 ? mma : l2) + epi;
            if (cost < best) { best = cost; best_ncta = ncta; best_bn = bn; }
        }
    }
    *ncta_out = best_ncta;
    *bn_out = best_bn;
}

int gemm_tc_launch(const b200asr_gemm_args& a, cudaStream_t st) {
    int ncta, bn;
    choose_config(a, &ncta, &bn);
    if (g_force_cfg) { ncta = g_force_cfg / 1000; bn = g_force_cfg % 1000; }
    const bool f32c = a.c_dtype == B200ASR_F32;
    if (ncta == 4) return f32c ? tc::launch<256, float, 4>(a, st) : tc::launch<256, __nv_bfloat16, 4>(a, st);
    if (ncta == 2) {
        if (bn == 256) return f32c ? tc::launch<256, float, 2>(a, st) : tc::launch<256, __nv_bfloat16, 2>(a, st);
        return f32c ? tc::launch<128, float, 2>(a, st) : tc::launch<128, __nv_bfloat16, 2>(a, st);
    }
    switch (bn) {
        case 64: return f32c ? tc::launch<64, float>(a, st) : tc::launch<64, __nv_bfloat16>(a, st);
        case 256: return f32c ? tc::launch<256, float>(a, st) : tc::launch<256, __nv_bfloat16>(a, st);
        default: return f32c ? tc::launch<128, float>(a, st) : tc::launch<128, __nv_bfloat16>(a, st);
    }
}

int gemm_tc_launch_dropout(const b200asr_gemm_args& a, const DropKeys& k, const unsigned* keep_bits, int mpad, cudaStream_t st) {
    const tc::EpiDrop ed{k, keep_bits, mpad};
    tc::t_epi_drop = &ed;
    const int rc = gemm_tc_launch(a, st);
    tc::t_epi_drop = nullptr;
    return rc;
}

}  // namespace b200asr

namespace b200asr {
int posconv_tc_taps(const void* x, const void* w_pack, const float* bias, void* y, int B, int T, int C, int groups, int k,
                    int tap_pad, cudaStream_t st) {
    return tc::posconv_taps_launch(x, w_pack, bias, y, B, T, C, groups, k, tap_pad, st);
}
int posconv_tc_wgrad(const void* x, const void* dy, float* dwt, int B, int T, int C, int groups, int k, int pad,
                     cudaStream_t st) {
    return tc::posconv_wgrad_launch(x, dy, dwt, B, T, C, groups, k, pad, st);
}
}  // namespace b200asr

// diagnostics: while `buf` (device memory, 20 u64 per CTA of the largest grid = 148 * 20 * 8 bytes) is set, every tcgen05 GEMM
// launch records per-CTA time stamps (slot layout next to TRACE above); pass null to stop
extern "C" int b200asr_gemm_set_trace(void* buf) {
    b200asr::tc::g_gemm_trace = (unsigned long long*)buf;
    return 0;
}
