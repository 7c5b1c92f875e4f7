#!/bin/bash
# v17 evidence round (LayerDrop + element dropout): GPU parity tests, smoke, both bench arms, ncu launch list, op bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/t64_all.log 2>&1; echo "all tests rc=$?"; tail -n 3 gpurun_out/t64_all.log | cut -c1-300
timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke64.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke64.log
timeout 400 python bench.py > gpurun_out/bench64.json 2> gpurun_out/bench64.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench64.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench64_ref.json 2> gpurun_out/bench64_ref.err; echo "ref rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches64.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu64.log 2>&1; echo "ncu list rc=$?"
timeout 200 python tools/op_bench.py > gpurun_out/op_bench64.txt 2>&1; echo "op bench rc=$?"; grep dropout gpurun_out/op_bench64.txt
ls -la gpurun_out | grep 64
