#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/gemm_bench.py > gpurun_out/gemm_bench23.log 2>&1; echo "gemm bench rc=$?"; grep -E "fwd|dgrad|wgrad|encoder" gpurun_out/gemm_bench23.log
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -n 6 gpurun_out/t_all.log | cut -c1-250
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench23.json 2> gpurun_out/bench23.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench23.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench23.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['ms_per_step'], 'roofline', d['roofline']['achieved'], d['roofline']['frac'], d['roofline']['other'])
PY
