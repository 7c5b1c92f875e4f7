#!/usr/bin/env python
"""Probe the tcgen05 GEMM for one operand-major combination (kk, km, mk, mm: A then B; k = K-major, m = MN-major)
over several shapes; prints the error of each instead of asserting (debug aid for GPU bring-up)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"))
import torch  # noqa: E402

from b200asr import ops  # noqa: E402

combo = sys.argv[1] if len(sys.argv) > 1 else "kk"
a_k, b_k = combo[0] == "k", combo[1] == "k"
shapes = [(128, 128, 64), (128, 64, 64), (128, 256, 64), (128, 128, 256), (256, 256, 512), (200, 72, 136),
          (499, 1024, 1024), (3992, 1024, 1024), (3992, 4096, 1024), (3992, 1024, 4096)]
if len(sys.argv) > 2:
    shapes = shapes[: int(sys.argv[2])]
for (M, N, K) in shapes:
    g = torch.Generator().manual_seed(M + N + K)
    A = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16)
    B = (torch.randn(N, K, generator=g) * 0.5).to(torch.bfloat16)
    ref = A.float() @ B.float().t()
    Mp, Np = (M + 7) // 8 * 8, (N + 7) // 8 * 8
    if a_k:
        Ad, lda = A.cuda(), K
    else:
        t = torch.zeros(K, Mp, dtype=torch.bfloat16); t[:, :M] = A.t(); Ad, lda = t.cuda(), Mp
    if b_k:
        Bd, ldb = B.cuda(), K
    else:
        t = torch.zeros(K, Np, dtype=torch.bfloat16); t[:, :N] = B.t(); Bd, ldb = t.cuda(), Np
    C = torch.zeros(M, N, dtype=torch.float32, device="cuda")
    try:
        ops.gemm_raw(Ad, Bd, C, M, N, K, lda, ldb, N, a_k, b_k, backend=1)
        torch.cuda.synchronize()
        err = float((C.cpu() - ref).abs().max() / ref.abs().max())
        # timing
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(10):
            ops.gemm_raw(Ad, Bd, C, M, N, K, lda, ldb, N, a_k, b_k, backend=1)
        t1.record(); torch.cuda.synchronize()
        ms = t0.elapsed_time(t1) / 10
        print(f"{combo} M={M} N={N} K={K}: rel-inf err {err:.3e}  {ms*1e3:.1f} us  {2*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)
        if err > 1e-2:
            d = (C.cpu() - ref).abs()
            bad = (d > 1e-2 * ref.abs().max()).nonzero()
            print("   first bad idx", bad[:5].tolist(), "n_bad", len(bad), "C", C.cpu()[bad[0][0], bad[0][1]].item(),
                  "ref", ref[bad[0][0], bad[0][1]].item(), flush=True)
    except Exception as e:  # noqa
        print(f"{combo} M={M} N={N} K={K}: EXCEPTION {type(e).__name__}: {e}", flush=True)
        break
