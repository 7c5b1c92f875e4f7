#!/bin/bash
# 2-GPU round: fused peer-memory FedAvg parity + the FedAvg round bench (peer kernel vs NCCL)
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
nvidia-smi topo -m 2>/dev/null | head -6
timeout 600 python -m pytest tests/test_gpu_multi.py -q -x > gpurun_out/t_multi.log 2>&1; echo "multi rc=$?"; tail -n 15 gpurun_out/t_multi.log | cut -c1-250
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench20_n2.json 2> gpurun_out/bench20_n2.err; echo "bench n2 rc=$?"; tail -n 5 gpurun_out/bench20_n2.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/bench20_n2.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','fedavg_agg_ms','fedavg_agg_alone','n_gpus')}, d['e2e']['ms_per_step'])
except Exception as e: print('no json', e)
PY
