#!/bin/bash
# v15 evidence round: GPU parity tests, smoke, both bench arms, ncu launch list + GEMM dram traffic + one --set full capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv,noheader
timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/t48_all.log 2>&1; echo "all tests rc=$?"; tail -n 3 gpurun_out/t48_all.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke48.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke48.log
timeout 900 python bench.py > gpurun_out/bench48.json 2> gpurun_out/bench48.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench48.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench48_ref.json 2> gpurun_out/bench48_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench48_ref.json
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches48.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu48.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:gemm_tc_kernel -s 900 -c 320 --csv --log-file gpurun_out/gemm_traffic48.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu48b.log 2>&1; echo "ncu traffic rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 1000 -c 6 -f -o gpurun_out/prof_gemm48 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu48c.log 2>&1; echo "ncu full rc=$?"
timeout 300 python tools/config_bench.py > gpurun_out/config_bench48.jsonl 2> gpurun_out/config_bench48.err; echo "config bench rc=$?"; cut -c1-200 gpurun_out/config_bench48.jsonl
ls -la gpurun_out | grep 48
