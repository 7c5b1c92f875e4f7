#!/usr/bin/env python
"""Where the cycles of one tcgen05 GEMM launch go: per-CTA time stamps (b200asr_gemm_set_trace) of a GEMM launched back to
back with itself (programmatic dependent launch, steady state), medians over CTAs of the last launch.
  python tools/gemm_trace.py N K [M] [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch
from b200asr import ops
from b200asr._lib import lib, check

WGRAD = "--wgrad" in sys.argv   # fp32 weight-gradient GEMM dW[N, K] = dy[M, N]^T x[M, K] instead of the bf16 linear
argv = [a for a in sys.argv if a != "--wgrad"]
N, K = int(argv[1]), int(argv[2])
M = int(argv[3]) if len(argv) > 3 else 3992
reps = int(argv[4]) if len(argv) > 4 else 20
x = (torch.randn(M, K, device="cuda") * 0.1).bfloat16(); w = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
b = torch.zeros(N, device="cuda"); y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
res = torch.randn(M, N, device="cuda").bfloat16()
tr = torch.zeros(reps, 148 * 20, dtype=torch.int64, device="cuda")
if WGRAD:
    dy = (torch.randn(M, N, device="cuda") * 0.1).bfloat16(); dw = torch.empty(N, K, device="cuda")
    run = lambda: ops.linear_wgrad(dy, x, out=dw)
else:
    run = lambda: ops.gemm_raw(x, w, y, M, N, K, K, K, N, True, True, bias=b, residual=res)
for _ in range(5):
    run()
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for i in range(reps):
    check(lib.b200asr_gemm_set_trace(tr[i].data_ptr()), "set_trace")
    run()
t1.record()
check(lib.b200asr_gemm_set_trace(None), "set_trace")
torch.cuda.synchronize()
print(f"{'wgrad fp32 ' if WGRAD else ''}N={N} K={K} M={M}: {t0.elapsed_time(t1) / reps * 1e3:.1f} us per launch (back to back)")
t = tr.cpu().view(reps, 148, 20)
names = ["setup", "dep wait", "first stage landed", "mainloop issue", "acc complete after last issue", "epilogue", "teardown"]
for r in (reps // 2, reps - 1):
    a = t[r]
    used = a[:, 1] != 0
    a = a[used].double()
    seg = torch.stack([a[:, 2] - a[:, 1], a[:, 3] - a[:, 2], a[:, 4] - a[:, 3], a[:, 5] - a[:, 4], a[:, 6] - a[:, 5],
                       a[:, 7] - a[:, 6], a[:, 8] - a[:, 7]], 1)
    lead = a[:, 4] != 0
    print(f"launch {r}: {int(used.sum())} CTAs; wall (globaltimer) first entry -> last exit {(a[:, 9].max() - a[:, 0].min()) / 1e3:.2f} us; "
          f"entry spread {(a[:, 0].max() - a[:, 0].min()) / 1e3:.2f} us; exit spread {(a[:, 9].max() - a[:, 9].min()) / 1e3:.2f} us")
    if r > 0:
        prev = t[r - 1][t[r - 1][:, 1] != 0].double()
        print(f"   previous launch last exit -> this first entry {(a[:, 0].min() - prev[:, 9].max()) / 1e3:.2f} us, "
              f"-> median entry {(a[:, 0].median() - prev[:, 9].max()) / 1e3:.2f} us")
    print(f"   CTA lifetime median {(a[:, 8] - a[:, 1]).median():.0f} clk")
    print(f"   epilogue split: first chunk {(a[:, 10] - a[:, 6]).median():.0f}, chunks 2.. {(a[:, 11] - a[:, 10]).median():.0f}, "
          f"final store wait {(a[:, 7] - a[:, 11]).median():.0f} clk (medians)")
    for j, n in enumerate(names):
        col = seg[:, j]
        if j in (2, 3, 4):
            col = col[lead]   # stamps 4/5 exist in pair leaders only
            if j == 4: col = (a[:, 6] - a[:, 5])[lead]
        print(f"   {n:32s} median {col.median():8.0f}  min {col.min():8.0f}  max {col.max():8.0f} clk")
