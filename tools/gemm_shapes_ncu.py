#!/usr/bin/env python
"""The encoder's GEMM shapes, one launch each after a warm-up launch (for `ncu --set full -k regex:gemm_tc_kernel`): forward
QKV / O / FFN1 (gelu + gelu', without and with the fused activation dropout) / FFN2, the FFN2 data gradient (* gelu') and the
O / W1 weight gradients at M = 3992 tokens.  Prints the launch order so the ncu rows can be named."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch
from b200asr import ops
from b200asr._lib import ACT_GELU_SAVE_GRAD
M, D, F = 3992, 1024, 4096
r = lambda *s: (torch.randn(*s, device="cuda") * 0.1).bfloat16()
x, xf = r(M, D), r(M, F)
wqkv, wo, w1, w2 = r(3 * D, D), r(D, D), r(F, D), r(D, F)
bq, bo, b1 = torch.zeros(3 * D, device="cuda"), torch.zeros(D, device="cuda"), torch.zeros(F, device="cuda")
cases = [
    ("fwd QKV   [M,3072] K=1024 +bias", lambda: ops.linear_fwd(x, wqkv, bq)),
    ("fwd O     [M,1024] K=1024 +bias", lambda: ops.linear_fwd(x, wo, bo)),
    ("fwd FFN1  [M,4096] K=1024 +bias, gelu + gelu'", lambda: ops.linear_fwd(x, w1, b1, act=ACT_GELU_SAVE_GRAD)),
    ("fwd FFN1  ... + activation dropout (Philox in the epilogue)", lambda: ops.linear_fwd(x, w1, b1, act=ACT_GELU_SAVE_GRAD, drop=(0.1, 1, 17, 3))),
    ("fwd FFN2  [M,1024] K=4096 +bias", lambda: ops.linear_fwd(xf, w2, bo)),
    ("dgrad FFN2 [M,4096] K=1024 * gelu'", lambda: ops.linear_dgrad(x, w2, gelu_grad=xf)),
    ("dgrad FFN1 [M,1024] K=4096 + residual", lambda: ops.linear_dgrad(xf, w1, residual=x)),
    ("wgrad O   [1024,1024] K=3992 fp32", lambda: ops.linear_wgrad(x, x)),
    ("wgrad W1  [4096,1024] K=3992 fp32", lambda: ops.linear_wgrad(xf, x)),
]
for name, fn in cases:
    fn()
torch.cuda.synchronize()
torch.cuda.profiler.start()
for i, (name, fn) in enumerate(cases):
    fn()
    print(i, name)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
