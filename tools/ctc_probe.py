#!/usr/bin/env python
"""CTC kernel phase probe: fwd+bwd vs fwd only, long vs short targets.  python tools/ctc_probe.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch
from b200asr import ops
dev = "cuda"
B, T = 8, 499
def bench(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(iters): fn()
    t1.record(); torch.cuda.synchronize()
    return t0.elapsed_time(t1) / iters * 1e3
logits = (torch.randn(1, B, T, 32, device=dev)).bfloat16()
for lo, hi, mt in ((20, 121, 120), (1, 2, 120), (1, 2, 1)):
    tl = torch.randint(lo, hi, (B,), device=dev, dtype=torch.int32)
    tg = torch.randint(1, 32, (B * 120,), device=dev, dtype=torch.int32)
    off = torch.arange(B, device=dev, dtype=torch.int32) * 120
    il = torch.full((B,), T, device=dev, dtype=torch.int32)
    for wg in (True, False):
        us = bench(lambda: ops.ctc_fwd_bwd(logits, tg, off, tl, il, mt, 0, True, want_grad=wg))
        print(f"targets {lo}..{hi - 1} (max_tgt {mt}) want_grad={wg}: {us:.1f} us")
