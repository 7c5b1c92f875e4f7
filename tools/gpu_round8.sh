#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/gemm_one.py 1024 4096 > gpurun_out/plain_g1.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 4 -c 2 -f -o gpurun_out/prof_gemm_n1024_k4096 python tools/gemm_one.py 1024 4096 > gpurun_out/ncu_g1.log 2>&1; echo "ncu g1 rc=$?"
timeout 300 python tools/gemm_one.py 4096 1024 > gpurun_out/plain_g2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 4 -c 2 -f -o gpurun_out/prof_gemm_n4096_k1024 python tools/gemm_one.py 4096 1024 > gpurun_out/ncu_g2.log 2>&1; echo "ncu g2 rc=$?"
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches2.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"; tail -n 2 gpurun_out/ncu.log | cut -c1-300
