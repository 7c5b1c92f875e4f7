mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops2.py -q -k "ctc" > gpurun_out/r2h_ctc.log 2>&1; echo ctc rc=$?
timeout 300 python tools/op_bench.py 2>&1 | grep -E "ctc" > gpurun_out/r2h_op.txt
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --steps 20 --warmup 3 "$@" > gpurun_out/r2h_${tag}.json 2> gpurun_out/r2h_${tag}.err; echo "$tag rc=$?"; }
run b128g2
run b296g1 --fedavg-blocks 296 --fedavg-group 1
run b64g4 --fedavg-blocks 64 --fedavg-group 4
run b0g3 --fedavg-blocks 0 --fedavg-group 3
run b296g24 --fedavg-blocks 296 --fedavg-group 24
run nolw --no-layerwise
