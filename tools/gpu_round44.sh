#!/bin/bash
# v14 evidence: full GPU tests, bench, ncu launch list of one eager step
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --ignore=tests/test_gpu_multi.py > gpurun_out/t44.log 2>&1; echo "gpu tests rc=$?"; tail -2 gpurun_out/t44.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench44.json 2> gpurun_out/bench44.err; tail -c 600 gpurun_out/bench44.json
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches44.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu44.log 2>&1; echo "ncu list rc=$?"
python tools/launch_summary.py gpurun_out/launches44.csv 4 > gpurun_out/launches44_summary.txt 2>&1; head -40 gpurun_out/launches44_summary.txt
