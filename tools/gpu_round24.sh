#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_ops2.py -q -x > gpurun_out/t_model.log 2>&1; echo "model tests rc=$?"; tail -n 4 gpurun_out/t_model.log | cut -c1-250
timeout 900 python tools/config_bench.py --only c3 > gpurun_out/config_bench_c3.jsonl 2> gpurun_out/config_bench.err; echo "config bench rc=$?"; cat gpurun_out/config_bench_c3.jsonl | cut -c1-400; tail -n 5 gpurun_out/config_bench.err | cut -c1-300
