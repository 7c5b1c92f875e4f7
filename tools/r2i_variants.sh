mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_graph_layerdrop.py -q > gpurun_out/r2i_graph.log 2>&1; echo graph rc=$?
run() { tag=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --steps 20 --warmup 3 "$@" > gpurun_out/r2i_${tag}.json 2> gpurun_out/r2i_${tag}.err; echo "$tag rc=$?"; }
run b128g2
run b296g1 --fedavg-blocks 296 --fedavg-group 1
run b128g4 --fedavg-blocks 128 --fedavg-group 4
run b128g2_nochunk --no-chunked-adamw
run b128g2_nvls --fedavg nvls
