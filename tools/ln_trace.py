#!/usr/bin/env python
"""Where the cycles of the LayerNorm backward kernel go: per-block SM-clock stamps (b200asr_ln_set_trace).
  python tools/ln_trace.py [rows cols]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch
from b200asr import ops
from b200asr._lib import lib, check

rows, cols = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3992, 1024)
dev = "cuda"
x = (torch.randn(rows, cols, device=dev)).bfloat16(); dy = (torch.randn(rows, cols, device=dev) * 0.1).bfloat16()
w = torch.ones(cols, device=dev); b = torch.zeros(cols, device=dev)
mean = torch.zeros(rows, device=dev); rstd = torch.ones(rows, device=dev)
dw = torch.empty(cols, device=dev); db = torch.empty(cols, device=dev); dxs = torch.empty(cols, device=dev)
run = lambda: ops.layernorm_bwd(dy, x, w, b, mean, rstd, dw_out=dw, db_out=db, dxsum_out=dxs)
for _ in range(3): run()
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for _ in range(20): run()
t1.record(); torch.cuda.synchronize()
print(f"layernorm_bwd [{rows},{cols}] bf16 (+ affine grads + column sums, incl. reduce_partials): {t0.elapsed_time(t1) * 50:.1f} us per call")
tr = torch.zeros(148, 8, dtype=torch.int64, device=dev)
check(lib.b200asr_ln_set_trace(tr.data_ptr()), "set_trace")
run()
torch.cuda.synchronize()
check(lib.b200asr_ln_set_trace(None), "set_trace")
a = tr.cpu().double()
a = a[a[:, 0] != 0]
def med(x): return f"median {x.median():7.0f}  min {x.min():7.0f}  max {x.max():7.0f} clk"
print(f"{a.shape[0]} blocks")
print("block lifetime             ", med(a[:, 6] - a[:, 0]))
print("dependency wait            ", med(a[:, 1] - a[:, 0]))
print("params + first rows request", med(a[:, 2] - a[:, 1]))
print("first row landed + reduced ", med(a[:, 3] - a[:, 2]))
print("remaining rows             ", med(a[:, 4] - a[:, 3]))
print("block barrier              ", med(a[:, 5] - a[:, 4]))
print("group turns + partial write", med(a[:, 6] - a[:, 5]))
