#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -n 30 gpurun_out/t_all.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench6.log 2> gpurun_out/bench6.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench6.log; tail -n 5 gpurun_out/bench6.err
