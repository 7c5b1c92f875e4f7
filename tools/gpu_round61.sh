#!/bin/bash
# v16 evidence round: GPU parity tests, smoke, both bench arms, ncu launch list + GEMM dram traffic + one --set full capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv,noheader
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/t61_all.log 2>&1; echo "all tests rc=$?"; tail -n 3 gpurun_out/t61_all.log | cut -c1-300
timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke61.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke61.log
timeout 400 python bench.py > gpurun_out/bench61.json 2> gpurun_out/bench61.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench61.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench61_ref.json 2> gpurun_out/bench61_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench61_ref.json
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches61.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu61.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:gemm_tc_kernel -s 900 -c 320 --csv --log-file gpurun_out/gemm_traffic61.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu61b.log 2>&1; echo "ncu traffic rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 1000 -c 4 -f -o gpurun_out/prof_gemm61 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu61c.log 2>&1; echo "ncu full rc=$?"
timeout 200 python tools/op_bench.py > gpurun_out/op_bench61.txt 2>&1; echo "op bench rc=$?"
ls -la gpurun_out | grep 61
