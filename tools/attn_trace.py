#!/usr/bin/env python
"""Where the cycles of the attention backward kernel go: per-CTA SM-clock stamps (b200asr_attn_set_trace), medians over
the CTAs of one launch.  python tools/attn_trace.py [B T H]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch
from b200asr import ops
from b200asr._lib import lib, check

FWD = "--fwd" in sys.argv   # trace the forward kernel instead
sys.argv = [a for a in sys.argv if a != "--fwd"]
B, T, H = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (8, 499, 16)
D = H * 64
M = B * T
dev = "cuda"
qkv = (torch.randn(M, 3 * D, device=dev) * 0.5).bfloat16(); do = (torch.randn(M, D, device=dev) * 0.1).bfloat16()
out, lse = ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)
nc = ((T + 127) // 128) * H * B
if FWD:
    tr = torch.zeros(nc, 24, dtype=torch.int64, device=dev)
    run = lambda: ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)
    for _ in range(3): run()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(10): run()
    t1.record(); torch.cuda.synchronize()
    print(f"B={B} T={T} H={H}: mha_fwd {t0.elapsed_time(t1) * 100:.1f} us per call, {nc} CTAs (2 per SM)")
    check(lib.b200asr_attn_set_trace(tr.data_ptr()), "set_trace")
    run()
    check(lib.b200asr_attn_set_trace(None), "set_trace")
    torch.cuda.synchronize()
    a = tr.cpu().double()
    a = a[a[:, 2] != 0]
    def med(x): return f"median {x.median():7.0f}  min {x.min():7.0f}  max {x.max():7.0f} clk"
    print("CTA lifetime                 ", med(a[:, 7] - a[:, 0]))
    print("dependency wait              ", med(a[:, 1] - a[:, 0]))
    print("set-up                       ", med(a[:, 2] - a[:, 1]))
    print("Q landed after set-up        ", med(a[:, 3] - a[:, 2]))
    print("MMA issue loop               ", med(a[:, 4] - a[:, 3]))
    print("softmax loop end after set-up", med(a[:, 5] - a[:, 2]))
    print("output drain                 ", med(a[:, 6] - a[:, 5]))
    print("teardown                     ", med(a[:, 7] - a[:, 6]))
    print("key block 2, softmax warp 2:")
    ph = ["wait S", "TMEM read + release", "max + exchange barrier", "exp2 / sum / pack", "wait PV(j-1) done", "P store (+O rescale) + arrive"]
    for k, n in enumerate(ph):
        print(f"   {n:30s}", med(a[:, 9 + k] - a[:, 8 + k]))
    print("   block period (S ready 2 -> 3) ", med(a[:, 15] - a[:, 9]))
    sys.exit(0)
tr = torch.zeros(nc, 24, dtype=torch.int64, device=dev)
for _ in range(3):
    ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for _ in range(10):
    ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
t1.record()
torch.cuda.synchronize()
print(f"B={B} T={T} H={H}: mha_bwd (dsum + main + dQ GEMM) {t0.elapsed_time(t1) * 100:.1f} us per call, {nc} CTAs")
check(lib.b200asr_attn_set_trace(tr.data_ptr()), "set_trace")
ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
check(lib.b200asr_attn_set_trace(None), "set_trace")
torch.cuda.synchronize()
a = tr.cpu().double()
a = a[a[:, 2] != 0]
def med(x): return f"median {x.median():7.0f}  min {x.min():7.0f}  max {x.max():7.0f} clk"
print("CTA lifetime                      ", med(a[:, 7] - a[:, 0]))
print("dependency wait                   ", med(a[:, 1] - a[:, 0]))
print("set-up (barriers, TMEM, lse/D)    ", med(a[:, 2] - a[:, 1]))
print("K/V landed after set-up           ", med(a[:, 3] - a[:, 2]))
print("MMA issue loop (all query blocks) ", med(a[:, 4] - a[:, 3]))
print("element-wise loop end after set-up", med(a[:, 5] - a[:, 2]))
print("wait for dK/dV after the loop     ", med(a[:, 6] - a[:, 5]))
print("dK/dV drain + teardown            ", med(a[:, 7] - a[:, 6]))
print("query block 2, element-wise warp 2:")
ph = ["wait S/dP", "TMEM read + release", "exp2 / dS math", "wait P/dS slot free", "smem + global writes", "fence + arrive",
      "-> block 3 S/dP ready"]
for k, n in enumerate(ph):
    print(f"   {n:24s}", med(a[:, 9 + k] - a[:, 8 + k]))
print("   block period (S/dP ready 2 -> 3)", med(a[:, 15] - a[:, 9]))
print("query block 2, MMA issuer: wait P/dS ready", med(a[:, 17] - a[:, 16]), "| issue dV/dK MMAs", med(a[:, 18] - a[:, 17]))
