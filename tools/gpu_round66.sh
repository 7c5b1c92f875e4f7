#!/bin/bash
# v17b final check: whole GPU suite, smoke, default bench (+ dropout 0 for the parity setting), launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/t66_all.log 2>&1; echo "all tests rc=$?"; tail -n 2 gpurun_out/t66_all.log | cut -c1-300
timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke66.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke66.log
timeout 400 python bench.py > gpurun_out/bench66_final.json 2> gpurun_out/bench66_final.err; echo "bench rc=$?"; cut -c1-260 gpurun_out/bench66_final.json
timeout 300 python bench.py --no-cpu-baseline --dropout 0 > gpurun_out/bench66_p0.json 2> gpurun_out/bench66_p0.err; echo "bench p0 rc=$?"; cut -c100-260 gpurun_out/bench66_p0.json
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches66.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu66.log 2>&1; echo "ncu list rc=$?"
