#!/bin/bash
# reduce_partials with 32 warps (one batch of loads)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --ignore=tests/test_gpu_multi.py > gpurun_out/t47.log 2>&1; echo "gpu tests rc=$?"; tail -3 gpurun_out/t47.log
timeout 300 python tools/op_bench.py > gpurun_out/op_bench47.txt 2>&1; grep -i "mha\|ln_\|colsum" gpurun_out/op_bench47.txt
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench47.json 2> gpurun_out/bench47.err; python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench47.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'ms_per_step')}, d['e2e']['value'], d['roofline']['frac'], d.get('clocks'))
PY
