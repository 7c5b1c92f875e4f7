#!/bin/bash
# GEMM per-CTA time-stamp traces (tools/gemm_trace.py) of the four encoder shapes
mkdir -p gpurun_out
{
for s in "1024 4096" "4096 1024"; do
  timeout 120 python tools/gemm_trace.py $s
done
} > gpurun_out/gemm_trace33.txt 2>&1
tail -60 gpurun_out/gemm_trace33.txt
