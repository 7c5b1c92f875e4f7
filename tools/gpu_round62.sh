#!/bin/bash
# HEAD check after the LayerDrop / per-parameter AdamW change: whole GPU suite, smoke, configuration bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/t62_all.log 2>&1; echo "all tests rc=$?"; tail -n 3 gpurun_out/t62_all.log | cut -c1-300
timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke62.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke62.log
timeout 300 python tools/config_bench.py > gpurun_out/config_bench62.jsonl 2> gpurun_out/config_bench62.err; echo "config bench rc=$?"; cut -c1-260 gpurun_out/config_bench62.jsonl
