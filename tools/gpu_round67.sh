#!/bin/bash
# N=8 scaling line of HEAD (dropout 0.1 default, FedAvg auto)
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench67_n8.json 2> gpurun_out/bench67_n8.err; echo "bench n8 rc=$?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench67_n8.json').read().strip().splitlines()[-1])
print({k: d.get(k) for k in ('value', 'ms_per_step', 'fedavg_agg_ms', 'fedavg_agg_alone')}, d['config'].get('fedavg_impl'), d['e2e']['value'], d['clocks'])
PY
