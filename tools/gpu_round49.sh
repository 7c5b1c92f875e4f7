#!/bin/bash
# v15 multi-GPU: peer FedAvg parity tests + bench at N=2 (run with gpurun --gpus 2), --fedavg auto
mkdir -p gpurun_out
n=${1:-2}
nvidia-smi --query-gpu=name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_gpu_multi.py -q -x -m gpu > gpurun_out/t49_multi.log 2>&1; echo "multi tests rc=$?"; tail -n 3 gpurun_out/t49_multi.log | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2971$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench49_n$n.json 2> gpurun_out/bench49_n$n.err; echo "bench n$n rc=$?"; tail -n 3 gpurun_out/bench49_n$n.err | cut -c1-300
python - <<PY
import json
d = json.loads(open('gpurun_out/bench49_n$n.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'ms_per_step', 'n_gpus')}, d['e2e']['value'], {k: v for k, v in d['config'].items() if 'fedavg' in k})
PY
