#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops2.py -q -x -k "posconv" > gpurun_out/t_pc.log 2>&1; echo "posconv tests rc=$?"; tail -n 15 gpurun_out/t_pc.log
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -n 5 gpurun_out/t_all.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench5.log 2> gpurun_out/bench5.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench5.log; tail -n 5 gpurun_out/bench5.err
