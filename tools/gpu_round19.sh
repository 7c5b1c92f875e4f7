#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_variants.py -q -x > gpurun_out/t_var.log 2>&1; echo "variants rc=$?"; tail -n 40 gpurun_out/t_var.log | cut -c1-250
