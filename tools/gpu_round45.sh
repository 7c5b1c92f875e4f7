#!/bin/bash
# A/B: number of side streams for the weight-gradient kernels (same box, back to back)
mkdir -p gpurun_out
for n in 1 3 2 1 3; do
  B200ASR_SIDE_STREAMS=$n timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench45_$n.json 2> gpurun_out/bench45_$n.err
  python - <<PY
import json
d = json.loads(open('gpurun_out/bench45_$n.json').read().strip().splitlines()[-1])
print("side streams $n:", d['ms_per_step'], d['value'], d['e2e']['value'])
PY
done
timeout 900 python -m pytest tests -m gpu -x -q -k "train or model or graph or step" --ignore=tests/test_gpu_multi.py 2>&1 | tail -2
