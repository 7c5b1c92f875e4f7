#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -n 8 gpurun_out/t_all.log
timeout 600 python tools/gemm_bench.py > gpurun_out/gemm_bench7.log 2>&1; echo "gemm_bench rc=$?"; cat gpurun_out/gemm_bench7.log
B200ASR_SIDE_STREAM=0 timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench7_noside.log 2> gpurun_out/bench7.err; echo "bench(noside) rc=$?"; python -c "
import json;d=json.loads(open('gpurun_out/bench7_noside.log').read().strip().splitlines()[-1]);print('noside ms/step',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'launches',d['gpu_launches_per_step'],'gemm frac',d['roofline']['frac'])"
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench7.log 2> gpurun_out/bench7.err; echo "bench rc=$?"; python -c "
import json;d=json.loads(open('gpurun_out/bench7.log').read().strip().splitlines()[-1]);print('side ms/step',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'launches',d['gpu_launches_per_step'],'gemm frac',d['roofline']['frac'], d['e2e']['last_loss'])"; tail -n 3 gpurun_out/bench7.err
