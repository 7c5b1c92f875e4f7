#!/usr/bin/env python
"""Per-shape timing of the tcgen05 GEMM on the shapes the data2vec-large update issues (B=8 x 10 s: M = 3992
tokens).  CUDA events over `iters` back-to-back launches on rotating buffers (> L2), TFLOP/s vs MEASURED_PEAKS."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch  # noqa: E402

from b200asr import ops  # noqa: E402
from b200asr._lib import ACT_GELU, ACT_GELU_SAVE_GRAD, ACT_NONE  # noqa: E402

dev = "cuda"
M = 3992
NROT = 6


def bench(name, mk, flops, iters=20):
    sets = [mk() for _ in range(NROT)]
    for s in sets[:2]:
        s()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(iters):
        sets[i % NROT]()
    t1.record()
    torch.cuda.synchronize()
    us = t0.elapsed_time(t1) * 1e3 / iters
    print(f"{name:44s} {us:8.1f} us  {flops / us / 1e6:7.1f} TFLOP/s", flush=True)
    return us


def rnd(*shape, dtype=torch.bfloat16):
    return (torch.randn(*shape, device=dev) * 0.1).to(dtype)


def fwd(N, K, act=ACT_NONE, residual=False, preact=False):
    def mk():
        x, w, b = rnd(M, K), rnd(N, K), rnd(N, dtype=torch.float32)
        y = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        r = rnd(M, N) if residual else None
        pre = torch.empty_like(y) if preact else None
        return lambda: ops.gemm_raw(x, w, y, M, N, K, K, K, N, True, True, bias=b, residual=r, aux_out=pre, act=act)
    return mk


def dgrad(Nout, Kin, gelu=False, residual=False):
    def mk():
        dy, w = rnd(M, Nout), rnd(Nout, Kin)
        dx = torch.empty(M, Kin, dtype=torch.bfloat16, device=dev)
        r = rnd(M, Kin) if residual else None
        u = rnd(M, Kin) if gelu else None
        return lambda: ops.gemm_raw(dy, w, dx, M, Kin, Nout, Nout, Kin, Kin, True, False, residual=r, aux_in=u, mul_gelu_grad=2 if gelu else 0)
    return mk


def wgrad(Nout, Kin):
    def mk():
        dy, x = rnd(M, Nout), rnd(M, Kin)
        dw = torch.empty(Nout, Kin, dtype=torch.float32, device=dev)
        return lambda: ops.gemm_raw(dy, x, dw, Nout, Kin, M, Nout, Kin, Kin, False, False)
    return mk


def conv(Tin, k, B=8, C=512):
    Tout = (Tin - k) // 2 + 1
    def mk():
        h, w = rnd(B, Tin, C), rnd(C, k * C)
        y = torch.empty(B, Tout, C, dtype=torch.bfloat16, device=dev)
        return lambda: ops.gemm_raw(h, w, y, Tout, C, k * C, 2 * C, k * C, C, True, True, batch0=1, batch1=B,
                                    sa=(0, Tin * C), sb=(0, 0), sc=(0, Tout * C))
    return mk, 2.0 * B * Tout * C * k * C


SWEEP = "--sweep" in sys.argv
from b200asr._lib import lib  # noqa: E402

tot = 0.0
for name, mk, fl, mult in [
    ("fwd  QKV   N=3072 K=1024 +bias", fwd(3072, 1024), 2.0 * M * 3072 * 1024, 24),
    ("fwd  O     N=1024 K=1024 +bias+res", fwd(1024, 1024, residual=True), 2.0 * M * 1024 * 1024, 24),
    ("fwd  FFN1  N=4096 K=1024 +bias+gelu+gelu'", fwd(4096, 1024, act=ACT_GELU_SAVE_GRAD, preact=True), 2.0 * M * 4096 * 1024, 24),
    ("fwd  FFN2  N=1024 K=4096 +bias+res", fwd(1024, 4096, residual=True), 2.0 * M * 4096 * 1024, 24),
    ("dgrad FFN2 -> [M,4096] K=1024 *gelu'", dgrad(1024, 4096, gelu=True), 2.0 * M * 4096 * 1024, 24),
    ("dgrad FFN1 -> [M,1024] K=4096 +res", dgrad(4096, 1024, residual=True), 2.0 * M * 4096 * 1024, 24),
    ("dgrad O    -> [M,1024] K=1024", dgrad(1024, 1024), 2.0 * M * 1024 * 1024, 24),
    ("dgrad QKV  -> [M,1024] K=3072 +res", dgrad(3072, 1024, residual=True), 2.0 * M * 3072 * 1024, 24),
    ("wgrad QKV  [3072,1024] K=3992 f32", wgrad(3072, 1024), 2.0 * M * 3072 * 1024, 24),
    ("wgrad O    [1024,1024] K=3992 f32", wgrad(1024, 1024), 2.0 * M * 1024 * 1024, 24),
    ("wgrad FFN1 [4096,1024] K=3992 f32", wgrad(4096, 1024), 2.0 * M * 4096 * 1024, 24),
    ("wgrad FFN2 [1024,4096] K=3992 f32", wgrad(1024, 4096), 2.0 * M * 4096 * 1024, 24),
]:
    us = bench(name, mk, fl)
    tot += us * mult
    if SWEEP:
        for cfg in (2256, 4256):
            lib.b200asr_gemm_force_config(cfg)
            bench(f"    forced {cfg}", mk, fl)
        lib.b200asr_gemm_force_config(0)
print(f"encoder linears per step (24 layers): {tot / 1e3:.2f} ms; at measured sustained peak: "
      f"{24 * 12 * 2.0 * M * 1024 * 1024 * 1.0 / 1e12:.2f}.. see DESIGN")
ctot = 0.0
for Tin, k in [(31999, 3), (15999, 3), (7999, 3), (3999, 3), (1999, 2), (999, 2)]:
    mk, fl = conv(Tin, k)
    ctot += bench(f"conv  Tin={Tin} k={k} (B=8, implicit GEMM)", mk, fl, iters=10)
print(f"conv layers 1-6 per step: {ctot / 1e3:.2f} ms")
