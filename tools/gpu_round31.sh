#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_pipeline.py -q -x > gpurun_out/t_pipe.log 2>&1; echo "pipeline tests rc=$?"; tail -n 25 gpurun_out/t_pipe.log | cut -c1-220
