#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches3.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"; tail -n 2 gpurun_out/ncu.log | cut -c1-200
timeout 600 python tools/gemm_bench.py --sweep > gpurun_out/gemm_sweep.log 2>&1; echo "sweep rc=$?"; cat gpurun_out/gemm_sweep.log
