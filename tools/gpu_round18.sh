#!/bin/bash
# full verification round: GPU parity tests, smoke, both bench arms, ncu launch list + GEMM dram traffic + one --set full capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv,noheader
nproc
timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"; tail -n 6 gpurun_out/t_all.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench18.json 2> gpurun_out/bench18.err; echo "bench rc=$?"; tail -n 5 gpurun_out/bench18.err; cut -c1-600 gpurun_out/bench18.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench18_ref.json 2> gpurun_out/bench18_ref.err; echo "ref rc=$?"; cat gpurun_out/bench18_ref.json | cut -c1-400
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/plain.log 2>&1 || exit 1
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches18.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:gemm_tc_kernel -s 900 -c 320 --csv --log-file gpurun_out/gemm_traffic18.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu2.log 2>&1; echo "ncu traffic rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 1000 -c 6 -f -o gpurun_out/prof_gemm18 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu3.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | head -30
