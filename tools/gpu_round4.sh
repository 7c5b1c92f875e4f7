#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops2.py -q -x -k "mha" > gpurun_out/t_mha.log 2>&1; echo "mha tests rc=$?"; tail -n 15 gpurun_out/t_mha.log
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -x > gpurun_out/t_k.log 2>&1; echo "kernel tests rc=$?"; tail -n 5 gpurun_out/t_k.log
timeout 600 python tools/gemm_bench.py > gpurun_out/gemm_bench.log 2>&1; echo "gemm_bench rc=$?"; cat gpurun_out/gemm_bench.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench4.log 2> gpurun_out/bench4.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench4.log; tail -n 5 gpurun_out/bench4.err
