#!/bin/bash
# One parameterised GPU-box script (replaces the per-round scratch scripts):  tools/gpu_run.sh TAG STAGE [STAGE ...]
# Stages write their logs to gpurun_out/<TAG>_<stage>.*; every stage runs under its own timeout so that a hung kernel
# cannot hold the box.
TAG=$1; shift
mkdir -p gpurun_out
for stage in "$@"; do
  case $stage in
    newtests)  timeout 900 python -m pytest tests/test_gpu_attn_dropout.py tests/test_gpu_graph_layerdrop.py -q -x -s > gpurun_out/${TAG}_newtests.log 2>&1; echo "newtests rc=$?";;
    newtests_all)  timeout 1200 python -m pytest tests/test_gpu_attn_dropout.py tests/test_gpu_graph_layerdrop.py -q -s > gpurun_out/${TAG}_newtests.log 2>&1; echo "newtests rc=$?";;
    large)     timeout 1500 python -m pytest tests/test_gpu_large.py -q -s > gpurun_out/${TAG}_large.log 2>&1; echo "large rc=$?";;
    alltests)  timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_alltests.log 2>&1; echo "alltests rc=$?";;
    alltests_noexit)  timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_alltests.log 2>&1; echo "alltests rc=$?";;
    smoke)     timeout 600 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?";;
    bench)     timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?";;
    bench_p0)  timeout 900 python bench.py --steps 20 --warmup 3 --dropout 0 --layerdrop 0 --no-cpu-baseline > gpurun_out/${TAG}_bench_p0.json 2> gpurun_out/${TAG}_bench_p0.err; echo "bench_p0 rc=$?";;
    bench_ref) timeout 1500 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "bench_ref rc=$?";;
    opbench)   timeout 900 python tools/op_bench.py > gpurun_out/${TAG}_op_bench.txt 2>&1; echo "opbench rc=$?";;
    launches)  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/${TAG}_ncu.log 2>&1; echo "launches rc=$?";;
    probe*)    N=${stage#probe}; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 tools/nvls_probe.py > gpurun_out/${TAG}_probe${N}.json 2> gpurun_out/${TAG}_probe${N}.err; echo "probe$N rc=$?";;
    multitest) timeout 900 python -m pytest tests/test_gpu_multi.py -q -s > gpurun_out/${TAG}_multitest.log 2>&1; echo "multitest rc=$?";;
    mbench_noov*) N=${stage#mbench_noov}; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 20 --warmup 3 --no-overlap > gpurun_out/${TAG}_bench_noov_n${N}.json 2> gpurun_out/${TAG}_bench_noov_n${N}.err; echo "mbench_noov$N rc=$?";;
    mbench*)   N=${stage#mbench}; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err; echo "mbench$N rc=$?";;
    # one ncu use per box call, each after its own command exited 0 without ncu
    ncu_launches)  # launch list of the bench's timed steps
               B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --profile-range"
               timeout 600 $B > gpurun_out/${TAG}_ncu_plain1.log 2>&1 &&
               timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/${TAG}_ncu1.log 2>&1
               echo "ncu launches rc=$?";;
    ncu_gemm)  # DRAM traffic and tensor-pipe share of every GEMM launch of one eager step
               timeout 600 python tools/step_one.py > gpurun_out/${TAG}_ncu_plain2.log 2>&1 &&
               timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active -k regex:gemm_tc_kernel --clock-control none --csv --log-file gpurun_out/${TAG}_gemm_traffic.csv python tools/step_one.py > gpurun_out/${TAG}_ncu2.log 2>&1
               echo "ncu gemm rc=$?";;
    ncu_ops)   # full sets of the memory-bound / latency-bound kernels and of the attention kernels
               timeout 600 python tools/op_one.py > gpurun_out/${TAG}_ncu_plain3.log 2>&1 &&
               timeout 1500 ncu --set full --clock-control none -k "regex:conv0_ln_gelu|ln_fwd|ln_bwd|ctc_kernel|ctc_grad|dacs_mask|mha_fwd_tc|mha_bwd_tc|attn_keep|dropout_kernel|dropout_keep|dsum" -c 60 -o /tmp/${TAG}_prof_ops python tools/op_one.py > gpurun_out/${TAG}_ncu3.log 2>&1
               echo "ncu ops rc=$?"
               # the report itself is too large to travel back (64 MiB limit): export the raw metric page as CSV on the box
               ncu -i /tmp/${TAG}_prof_ops.ncu-rep --page raw --csv > gpurun_out/${TAG}_prof_ops_raw.csv 2> gpurun_out/${TAG}_ncu3_export.log
               echo "ncu export rc=$?";;
    ncu_gemm_full)  # --set full of the encoder's GEMM shapes, one launch each
               timeout 300 python tools/gemm_shapes_ncu.py > gpurun_out/${TAG}_ncu_plain4.log 2>&1 &&
               timeout 900 ncu --profile-from-start off --set full --clock-control none -k regex:gemm_tc_kernel -o /tmp/${TAG}_prof_gemm python tools/gemm_shapes_ncu.py > gpurun_out/${TAG}_ncu4.log 2>&1
               echo "ncu gemm full rc=$?"
               ncu -i /tmp/${TAG}_prof_gemm.ncu-rep --page raw --csv > gpurun_out/${TAG}_prof_gemm_raw.csv 2> gpurun_out/${TAG}_ncu4_export.log
               echo "ncu export rc=$?";;
    *) echo "unknown stage $stage";;
  esac
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
