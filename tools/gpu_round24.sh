#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/config_bench.py > gpurun_out/config_bench.jsonl 2> gpurun_out/config_bench.err; echo "config bench rc=$?"; cat gpurun_out/config_bench.jsonl | cut -c1-400; tail -n 5 gpurun_out/config_bench.err | cut -c1-300
