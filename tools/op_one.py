#!/usr/bin/env python
"""A few launches of the non-GEMM kernels at configs[1] sizes, for `ncu --set full -k regex:...` captures."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch
from b200asr import ops
M, D, B, T, H = 3992, 1024, 8, 499, 16
bf = torch.bfloat16
r = lambda *s, sc=1.0: (torch.randn(*s, device="cuda") * sc).to(bf)
qkv, do = r(M, 3 * D), r(M, D, sc=0.1)
x, dy = r(M, D), r(M, D)
w, b = torch.ones(D, device="cuda"), torch.zeros(D, device="cuda")
dw, db, dxs = (torch.empty(D, device="cuda") for _ in range(3))
for _ in range(4):
    out, lse = ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)
    ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
    y, _, mean, rstd = ops.layernorm_fwd(x, w, b, 1e-5)
    ops.layernorm_bwd(dy, x, w, b, mean, rstd, dw_out=dw, db_out=db, dxsum_out=dxs)
torch.cuda.synchronize()
print("ok")
wav = torch.randn(8, 160000, device="cuda")
cw, lw, lb = torch.randn(512, 1, 10, device="cuda") * 0.4, torch.ones(512, device="cuda"), torch.zeros(512, device="cuda")
for _ in range(3):
    ops.conv0_ln_gelu(wav, cw, None, lw, lb, 5, 1e-5, bf)
torch.cuda.synchronize()
print("ok2")

# CTC (B=8, T=499, V=32, <= 120 labels), DACS toggling masks (in-kernel Gumbel noise), dropout, attention with keep bits
logits = r(1, B, T, 32)
tl = torch.randint(20, 121, (B,), device="cuda", dtype=torch.int32)
tg = torch.randint(1, 32, (B * 120,), device="cuda", dtype=torch.int32)
off = torch.arange(B, device="cuda", dtype=torch.int32) * 120
il = torch.full((B,), T, device="cuda", dtype=torch.int32)
scores, hid = r(M, 4 * D), r(M, D)
keep24 = ops.attn_dropout_mask(24, B, T, H, 0.1, 1234, 19, 4, 5)
for _ in range(3):
    ops.ctc_fwd_bwd(logits, tg, off, tl, il, 120, 0, True)
    lm, ad, sl, sa, ml, ma = ops.dacs_mask_fwd(scores, hid, 1.0, 0.0, seed=1, offset=3)
    ops.dacs_mask_bwd(x, dy, hid, sl, sa, ml, ma, 1.0, 0.0, want_dh=False)
    ops.dropout(x, 0.1, 1234, 16, 7, residual=dy)
    ops.attn_dropout_mask(24, B, T, H, 0.1, 1234, 19, 4, 5, out=keep24)
    ops.dropout_keep_bits(24, 17, 4, M, 4 * D, 0.1, 1234, 5, device="cuda")
    out, lse = ops.mha_fwd(qkv, None, B, T, H, 64, 0.125, keep=keep24[0], p_drop=0.1)
    ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125, keep=keep24[0], p_drop=0.1)
torch.cuda.synchronize()
print("ok3")
