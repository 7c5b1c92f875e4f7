#!/bin/bash
mkdir -p gpurun_out
for mc in 0 1; do
B200ASR_GEMM_MC=$mc timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench30_$mc.json 2> gpurun_out/bench30_$mc.err; echo "bench mc=$mc rc=$?"; tail -n 3 gpurun_out/bench30_$mc.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench30_$mc.json').read().strip().splitlines()[-1])
print('mc=$mc', {k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e']['ms_per_step'], 'roofline', round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), 'loss', d['e2e']['last_loss'])
PY
done
