#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -n 3 gpurun_out/t_all.log | cut -c1-200
for pdl in 1; do
B200ASR_PDL=$pdl timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench28_$pdl.json 2> gpurun_out/bench28_$pdl.err; echo "bench pdl=$pdl rc=$?"; tail -n 3 gpurun_out/bench28_$pdl.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench28_$pdl.json').read().strip().splitlines()[-1])
print('pdl=$pdl', {k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['ms_per_step'], 'roofline', round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), 'loss', d['e2e']['last_loss'])
PY
done
