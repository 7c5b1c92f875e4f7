#!/usr/bin/env python
"""One GEMM shape launched a few times (for `ncu --set full -k regex:gemm_tc`): python tools/gemm_one.py N K [M]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch
from b200asr import ops
N, K = int(sys.argv[1]), int(sys.argv[2])
M = int(sys.argv[3]) if len(sys.argv) > 3 else 3992
x = (torch.randn(M, K, device="cuda") * 0.1).bfloat16(); w = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
b = torch.zeros(N, device="cuda"); y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(8):
    ops.gemm_raw(x, w, y, M, N, K, K, K, N, True, True, bias=b)
torch.cuda.synchronize()
print("ok")
