#!/bin/bash
# N=2 check of HEAD (dropout default in the bench, per-parameter AdamW): multi-GPU FedAvg tests + the scaling bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu > gpurun_out/t65_multi.log 2>&1; echo "multi tests rc=$?"; tail -n 2 gpurun_out/t65_multi.log | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench65_n2.json 2> gpurun_out/bench65_n2.err; echo "bench n2 rc=$?"; tail -n 1 gpurun_out/bench65_n2.json | cut -c1-250
python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench65_n2.json').read().strip().splitlines()[-1])
print({k: d.get(k) for k in ('value', 'ms_per_step', 'fedavg_agg_ms', 'fedavg_agg_alone')}, d['config'].get('fedavg_impl'), d['e2e']['value'], d['clocks'])
PY
