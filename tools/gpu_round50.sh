#!/bin/bash
# v15 scaling on one 8-GPU box: FedAvg round bench at N = 8 and N = 4 (--fedavg auto)
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2981$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench50_n$n.json 2> gpurun_out/bench50_n$n.err; echo "bench n$n rc=$?"; tail -n 2 gpurun_out/bench50_n$n.err | cut -c1-300
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench50_n$n.json').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('n_gpus','value','ms_per_step')}, d['e2e']['ms_per_step'], {k: v for k, v in d['config'].items() if k.startswith('fedavg_') and not isinstance(v, str) or k == 'fedavg_impl'})
except Exception as e: print('no json', e)
PY
done
