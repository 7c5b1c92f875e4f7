#!/usr/bin/env python
"""Back-to-back timing (CUDA events, rotating buffers > L2) of the non-GEMM kernels at the configs[1] sizes
(B=8 x 10 s: 3992 tokens x 1024): LayerNorm fwd/bwd, colsum, attention fwd/bwd, CTC, AdamW, conv0, pos-conv.
Prints us per launch, achieved GB/s or TFLOP/s and the fraction of the measured peak."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch  # noqa: E402

from b200asr import ops  # noqa: E402
from b200asr._lib import ACT_GELU, ACT_NONE  # noqa: E402

dev = "cuda"
try:
    PK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    PK = {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0}
HBM, TF = PK["hbm_gbs"], PK["bf16_tflops_sustained"]
M, D, F, B, T, H = 3992, 1024, 4096, 8, 499, 16
NROT = 6
bf = torch.bfloat16


def rnd(*s, dtype=bf, scale=1.0):
    return (torch.randn(*s, device=dev) * scale).to(dtype)


def bench(name, mk, nbytes=None, flops=None, iters=30):
    sets = [mk() for _ in range(NROT)]
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for s in sets:
            s()
    torch.cuda.synchronize()
    # the launches are replayed from a CUDA graph so that the host-side cost of a call (allocations, ctypes) is not
    # what gets timed
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for i in range(iters):
            sets[i % NROT]()
    graph.replay()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    graph.replay()
    t1.record()
    torch.cuda.synchronize()
    us = t0.elapsed_time(t1) * 1e3 / iters
    extra = ""
    if nbytes:
        gbs = nbytes / us / 1e3
        extra = f"{gbs:8.0f} GB/s ({gbs / HBM:5.1%} of HBM peak)"
    if flops:
        tf = flops / us / 1e6
        extra = f"{tf:8.1f} TFLOP/s ({tf / TF:5.1%} of bf16 peak)"
    print(f"{name:52s} {us:8.1f} us  {extra}", flush=True)
    return us


def ln_fwd(cols, rows, res, act, affine=True):
    def mk():
        x = rnd(rows, cols); r = rnd(rows, cols) if res else None
        w = torch.ones(cols, device=dev) if affine else None
        b = torch.zeros(cols, device=dev) if affine else None
        return lambda: ops.layernorm_fwd(x, w, b, 1e-5, res=r, act=act, want_sum=res)
    return mk


def ln_bwd(cols, rows, affine, cs, act=ACT_NONE):
    def mk():
        x = rnd(rows, cols); dy = rnd(rows, cols)
        w = torch.ones(cols, device=dev) if affine else None
        b = torch.zeros(cols, device=dev) if affine else None
        mean = torch.zeros(rows, device=dev); rstd = torch.ones(rows, device=dev)
        dw = torch.empty(cols, device=dev); db = torch.empty(cols, device=dev)
        dxs = torch.empty(cols, device=dev) if cs else None
        return lambda: ops.layernorm_bwd(dy, x, w, b, mean, rstd, act=act, dw_out=dw if affine else None,
                                         db_out=db if affine else None, dxsum_out=dxs)
    return mk


def ln_bwd_drop(cols, rows):
    """LayerNorm backward + the hidden dropout's backward + the dropped Linear's bias gradient in one pass."""
    def mk():
        x = rnd(rows, cols); dy = rnd(rows, cols)
        w = torch.ones(cols, device=dev); b = torch.zeros(cols, device=dev)
        mean = torch.zeros(rows, device=dev); rstd = torch.ones(rows, device=dev)
        dw = torch.empty(cols, device=dev); db = torch.empty(cols, device=dev); dxs = torch.empty(cols, device=dev)
        return lambda: ops.layernorm_bwd(dy, x, w, b, mean, rstd, dw_out=dw, db_out=db, dxsum_out=dxs, drop=(0.1, 1234, 16, 7))
    return mk


def w1_gemm(drop):
    """The FFN's first GEMM [M, d] x [4d, d]^T with GELU + saved gelu' epilogue, without / with the activation dropout
    (drop = "philox": decisions computed in the epilogue, "bits": read from bits drawn ahead)."""
    def mk():
        from b200asr._lib import ACT_GELU_SAVE_GRAD
        x = rnd(M, D); w = rnd(F, D); b = torch.zeros(F, device=dev)
        d = None
        if drop == "bits":
            d = (0.1, ops.dropout_keep_bits(1, 17, 4, M, F, 0.1, 1234, 7, device=dev)[0], ops.dropout_keep_mpad(M))
        elif drop:
            d = (0.1, 1234, 17, 7)
        return lambda: ops.linear_fwd(x, w, b, act=ACT_GELU_SAVE_GRAD, drop=d)
    return mk


def keep_bits_all_blocks():
    def mk():
        out = ops.dropout_keep_bits(24, 17, 4, M, F, 0.1, 1234, 7, device=dev)
        return lambda: ops.dropout_keep_bits(24, 17, 4, M, F, 0.1, 1234, 7, out=out)
    return mk


def colsum(cols):
    def mk():
        x = rnd(M, cols); out = torch.empty(cols, device=dev)
        return lambda: ops.colsum(x, out=out)
    return mk


def dropout(cols, res):
    def mk():
        x = rnd(M, cols); r = rnd(M, cols) if res else None
        y = torch.empty_like(x)
        return lambda: ops.dropout(x, 0.1, 1234, 16, 7, residual=r, out=y)
    return mk


def mha_f():
    def mk():
        qkv = rnd(M, 3 * D)
        return lambda: ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)
    return mk


def mha_b():
    def mk():
        qkv = rnd(M, 3 * D); do = rnd(M, D, scale=0.1)
        out, lse = ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)
        return lambda: ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
    return mk


def mha_drop(kind, dq_fused=1):
    """attention with attention_dropout 0.1 keep bits; kind = "fwd" | "bwd" | "mask" (the keep-bit draw for 24 blocks)."""
    from b200asr._lib import lib

    def mk():
        if kind == "mask":
            keep24 = torch.empty(24, ops.attn_keep_words(B, T, H), dtype=torch.int32, device=dev)
            return lambda: ops.attn_dropout_mask(24, B, T, H, 0.1, 1234, 19, 4, 5, out=keep24)
        qkv = rnd(M, 3 * D); do = rnd(M, D, scale=0.1)
        keep = ops.attn_dropout_mask(1, B, T, H, 0.1, 1234, 19, 4, 5)[0]
        if kind == "fwd":
            return lambda: ops.mha_fwd(qkv, None, B, T, H, 64, 0.125, keep=keep, p_drop=0.1)
        out, lse = ops.mha_fwd(qkv, None, B, T, H, 64, 0.125, keep=keep, p_drop=0.1)

        def run():
            old = lib.b200asr_mha_set_dq_fused(dq_fused)
            ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125, keep=keep, p_drop=0.1)
            lib.b200asr_mha_set_dq_fused(old)
        return run
    return mk


def mha_poly(kind, share):
    """fused attention with `share` (4 = 25 %, 2 = 50 %) of the exponentials on the FMA pipe (b200asr_mha_set_exp2_poly)."""
    from b200asr._lib import lib

    def mk():
        qkv = rnd(M, 3 * D); do = rnd(M, D, scale=0.1)
        out, lse = ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)

        def run():
            old = lib.b200asr_mha_set_exp2_poly(share, share)
            if kind == "fwd":
                ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)
            else:
                ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
            lib.b200asr_mha_set_exp2_poly(old >> 4, old & 15)
        return run
    return mk


def mha_b_mode(dq_fused):
    from b200asr._lib import lib

    def mk():
        qkv = rnd(M, 3 * D); do = rnd(M, D, scale=0.1)
        out, lse = ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)

        def run():
            old = lib.b200asr_mha_set_dq_fused(dq_fused)
            ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
            lib.b200asr_mha_set_dq_fused(old)
        return run
    return mk


def dacs(kind):
    def mk():
        scores = rnd(M, 4 * D); h = rnd(M, D)
        if kind == "fwd":
            return lambda: ops.dacs_mask_fwd(scores, h, 1.0, 0.0, seed=1, offset=3)
        lm, ad, sl, sa, ml, ma = ops.dacs_mask_fwd(scores, h, 1.0, 0.0, seed=1, offset=3)
        dl, da = rnd(M, D), rnd(M, D)
        return lambda: ops.dacs_mask_bwd(dl, da, h, sl, sa, ml, ma, 1.0, 0.0, want_dh=False)
    return mk


def ctc():
    def mk():
        logits = rnd(1, B, T, 32)
        tl = torch.randint(20, 121, (B,), device=dev, dtype=torch.int32)
        tg = torch.randint(1, 32, (B * 120,), device=dev, dtype=torch.int32)
        off = torch.arange(B, device=dev, dtype=torch.int32) * 120
        il = torch.full((B,), T, device=dev, dtype=torch.int32)
        return lambda: ops.ctc_fwd_bwd(logits, tg, off, tl, il, 120, 0, True)
    return mk


def adamw(n):
    def mk():
        p, g, m, v = (torch.randn(n, device=dev) * 0.01 for _ in range(4))
        v.abs_()
        sh = torch.empty(n, device=dev, dtype=bf)
        gn = torch.ones(1, device=dev)
        return lambda: ops.adamw_step_(p, g, m, v, sh, 1e-5, 0.9, 0.999, 1e-8, 0.005, 3, gn, 1.0)
    return mk


def conv0():
    def mk():
        x = torch.randn(B, 160000, device=dev); w = torch.randn(512, 1, 10, device=dev)
        lw = torch.ones(512, device=dev); lb = torch.zeros(512, device=dev)
        return lambda: ops.conv0_ln_gelu(x, w, None, lw, lb, 5, 1e-5, bf)
    return mk


def posconv(kind):
    def mk():
        x = rnd(B, T, D); dy = rnd(B, T, D); w = rnd(D, 64, 19, scale=0.03)
        wf, wb = ops.posconv_pack(w, 16)
        bias = torch.zeros(D, device=dev)
        if kind == "fwd":
            return lambda: ops.posconv_fwd(x, wf, bias, 16, 19)
        if kind == "dgrad":
            return lambda: ops.posconv_bwd(x, wb, dy, 16, 19, want_dw=False)
        return lambda: ops.posconv_bwd(x, wb, dy, 16, 19, want_dx=False)
    return mk


e = 2  # bytes per bf16
tot = 0.0
tot += 48 * bench("ln_fwd   [3992,1024] affine          (x48)", ln_fwd(D, M, False, ACT_NONE), nbytes=2 * M * D * e)
tot += 7 * bench("ln_fwd   [8*T_l,512] affine+gelu conv (x7, avg rows)", ln_fwd(512, 8 * 9070, False, ACT_GELU), nbytes=2 * 8 * 9070 * 512 * e)
tot += 5 * bench("ln_fwd   [3992,1024] no affine + gelu (x5)", ln_fwd(D, M, False, ACT_GELU, affine=False), nbytes=2 * M * D * e)
tot += 48 * bench("ln_bwd   [3992,1024] affine + colsum  (x48)", ln_bwd(D, M, True, True), nbytes=3 * M * D * e)
tot += 5 * bench("ln_bwd   [3992,1024] no affine gelu cs (x5)", ln_bwd(D, M, False, True, ACT_GELU), nbytes=3 * M * D * e)
tot += 24 * bench("colsum   [3992,4096]                 (x24)", colsum(F), nbytes=M * F * e)
tot += 24 * bench("colsum   [3992,3072]                 (x24)", colsum(3 * D), nbytes=M * 3 * D * e)
bench("dropout [3992,1024] + residual (x96 when p > 0)", dropout(D, True), nbytes=3 * M * D * e)
bench("dropout [3992,1024]            (x52 when p > 0)", dropout(D, False), nbytes=2 * M * D * e)
bench("dropout [3992,4096] activation (fp32 mode only: fused below)", dropout(F, False), nbytes=2 * M * F * e)
bench("ln_bwd   [3992,1024] + hidden dropout bwd + colsum (x48 when p > 0)", ln_bwd_drop(D, M), nbytes=4 * M * D * e)
bench("gemm W1  [3992,1024]x[4096,1024] gelu + gelu'        (x24)", w1_gemm(False), flops=2.0 * M * D * F)
bench("gemm W1  ... + activation dropout, Philox in the epilogue", w1_gemm("philox"), flops=2.0 * M * D * F)
bench("gemm W1  ... + activation dropout from keep bits     (x24 when p > 0)", w1_gemm("bits"), flops=2.0 * M * D * F)
bench("activation keep bits, 24 blocks in one launch (x1, side stream)", keep_bits_all_blocks(), nbytes=24 * M * F // 8)
tot += 24 * bench("mha_fwd  B8 T499 H16 dh64            (x24)", mha_f(), flops=4.0 * B * H * T * T * 64)
tot += 24 * bench("mha_bwd  B8 T499 H16 dh64 (+dsum,dQ) (x24)", mha_b(), flops=10.0 * B * H * T * T * 64)
for share, pct in ((0, 0), (4, 25), (2, 50)):
    bench(f"mha_fwd  ... {pct:2d} % of the exp2 as FMA-pipe polynomial", mha_poly("fwd", share), flops=4.0 * B * H * T * T * 64)
    bench(f"mha_bwd  ... {pct:2d} % of the exp2 as FMA-pipe polynomial", mha_poly("bwd", share), flops=10.0 * B * H * T * T * 64)
bench("mha_bwd  ... dQ in-kernel, fp32 bulk reduce-add", mha_b_mode(1), flops=10.0 * B * H * T * T * 64)
bench("mha_bwd  ... dQ in-kernel, reduce-adds skipped (hook)", mha_b_mode(2), flops=10.0 * B * H * T * T * 64)
bench("mha_fwd  ... + attention dropout 0.1 keep bits", mha_drop("fwd"), flops=4.0 * B * H * T * T * 64)
bench("mha_bwd  ... + attention dropout 0.1 keep bits", mha_drop("bwd"), flops=10.0 * B * H * T * T * 64)
bench("attn keep bits, 24 blocks in one launch (x1)", mha_drop("mask"), nbytes=24 * 4.0 * B * H * 16 * 512, iters=6)
# DACS toggling network (configs[2]): read scores [M,4d] + h, write 2 masked states + 2 soft + 2 masks = 18 KB / token fwd
bench("dacs_mask_fwd [3992,1024] in-kernel gumbel (x1)", dacs("fwd"), nbytes=M * D * e * (4 + 1 + 6), iters=10)
bench("dacs_mask_bwd [3992,1024]                  (x1)", dacs("bwd"), nbytes=M * D * e * (2 + 1 + 4 + 4), iters=10)
tot += bench("ctc      B8 T499 V32 L<=120 fwd+bwd   (x1)", ctc(), iters=10)
tot += bench("adamw    309M params (+clip,+shadow)  (x1)", adamw(309_101_600 // 4 * 4), nbytes=309_101_600 * 30.0, iters=6)
tot += bench("conv0    B8 x 160000 -> [8,31999,512] (x1)", conv0(), nbytes=8 * 160000 * 4 + 8 * 31999 * 512 * e, iters=10)
tot += 5 * bench("posconv fwd   k19 g16                (x5)", posconv("fwd"), flops=2.0 * M * D * 64 * 19)
tot += 5 * bench("posconv dgrad k19 g16                (x5)", posconv("dgrad"), flops=2.0 * M * D * 64 * 19)
tot += 5 * bench("posconv wgrad k19 g16 (+permute)     (x5)", posconv("wgrad"), flops=2.0 * M * D * 64 * 19)
print(f"sum over one training step (multiplicities above): {tot / 1e3:.2f} ms")
