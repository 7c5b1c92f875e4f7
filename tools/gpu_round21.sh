#!/bin/bash
mkdir -p gpurun_out
for d in 0 16 32 48 64; do
  echo "=== B200ASR_GEMM_DBG=$d"
  B200ASR_GEMM_DBG=$d timeout 300 python tools/gemm_bench.py 2>&1 | grep -E "fwd  QKV|fwd  FFN1|dgrad FFN2|fwd  O|wgrad FFN1"
done > gpurun_out/gemm_dbg2.log 2>&1
cat gpurun_out/gemm_dbg2.log
