#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -x -k "forced_tile" > gpurun_out/t_pair.log 2>&1; echo "pair tests rc=$?"; tail -n 15 gpurun_out/t_pair.log
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -n 5 gpurun_out/t_all.log
timeout 600 python tools/gemm_bench.py > gpurun_out/gemm_bench10.log 2>&1; echo "gemm_bench rc=$?"; cat gpurun_out/gemm_bench10.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench10.log 2> gpurun_out/bench10.err; echo "bench rc=$?"; python -c "
import json;d=json.loads(open('gpurun_out/bench10.log').read().strip().splitlines()[-1]);print('ms/step',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'launches',d['gpu_launches_per_step'],'host ms',d.get('host_enqueue_ms_per_step'),'loss',d['e2e']['last_loss'])"; tail -n 3 gpurun_out/bench10.err
