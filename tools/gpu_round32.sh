#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/gemm_one.py 1024 4096 > gpurun_out/plain_g1.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 5 -c 1 -f -o gpurun_out/prof_gemm32_ffn2 python tools/gemm_one.py 1024 4096 > gpurun_out/ncu_g1.log 2>&1; echo "ncu1 rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 5 -c 1 -f -o gpurun_out/prof_gemm32_ffn1 python tools/gemm_one.py 4096 1024 > gpurun_out/ncu_g2.log 2>&1; echo "ncu2 rc=$?"
