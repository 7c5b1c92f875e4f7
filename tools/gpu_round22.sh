#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/gemm_bench.py > gpurun_out/gemm_bench22.log 2>&1; echo "bench rc=$?"; cat gpurun_out/gemm_bench22.log | tail -25
timeout 240 python -m pytest tests/test_gpu_kernels.py -q -x -k "gemm or linear" > gpurun_out/t_gemm.log 2>&1; echo "gemm tests rc=$?"; tail -n 15 gpurun_out/t_gemm.log | cut -c1-250
