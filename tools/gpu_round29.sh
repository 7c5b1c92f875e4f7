#!/bin/bash
# 8-GPU round: fused peer FedAvg parity at world 4 + the FedAvg round bench at N = 8 (and N = 4)
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 300 python -m pytest tests/test_gpu_multi.py -q -x > gpurun_out/t_multi8.log 2>&1; echo "multi rc=$?"; tail -n 3 gpurun_out/t_multi8.log | cut -c1-200
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench29_n$n.json 2> gpurun_out/bench29_n$n.err; echo "bench n$n rc=$?"; tail -n 3 gpurun_out/bench29_n$n.err | cut -c1-300
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench29_n$n.json').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('n_gpus','value','ms_per_step','fedavg_agg_ms','fedavg_agg_alone','fedavg_nvlink_gbs_per_dir')}, d['e2e']['ms_per_step'])
except Exception as e: print('no json', e)
PY
done
