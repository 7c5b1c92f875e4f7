#!/bin/bash
# tools/scale_variants.sh TAG N variant...   -- bench.py at N GPUs with several FedAvg settings (variant = name:flags,
# flags separated by '+', e.g. nvls:--fedavg+nvls)
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
for v in "$@"; do
  name=${v%%:*}; flags=${v#*:}; [ "$flags" = "$v" ] && flags=""
  flags=${flags//+/ }
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 \
    bench.py --gpus $N --steps 20 --warmup 3 $flags > gpurun_out/${TAG}_n${N}_${name}.json 2> gpurun_out/${TAG}_n${N}_${name}.err
  echo "$name rc=$?"
done
