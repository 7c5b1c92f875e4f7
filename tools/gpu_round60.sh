#!/bin/bash
# GEMM set-up: one barrier per lane, 32-bit tile walk
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --ignore=tests/test_gpu_multi.py > gpurun_out/t60.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/t60.log
timeout 300 python tools/op_bench.py > gpurun_out/op_bench60.txt 2>&1; grep -i "mha\|ln_\|colsum" gpurun_out/op_bench60.txt
timeout 120 python tools/gemm_trace.py 1024 4096 > gpurun_out/gemm_trace60.txt 2>&1; grep -A12 "^launch 19" gpurun_out/gemm_trace60.txt
timeout 300 python tools/gemm_bench.py > gpurun_out/gemm_bench60.txt 2>&1; grep "encoder linears" gpurun_out/gemm_bench60.txt
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench60.json 2> gpurun_out/bench60.err; python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench60.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'ms_per_step')}, d['e2e']['value'], d['roofline']['frac'], d.get('clocks'))
PY
