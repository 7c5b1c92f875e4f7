#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_ops2.py tests/test_gpu_model.py -q -x > gpurun_out/t_k.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/t_k.log | cut -c1-250
timeout 300 python tools/op_bench.py > gpurun_out/op_bench26.log 2>&1; echo "op bench rc=$?"; grep -E "ln_|conv0|ctc|sum over" gpurun_out/op_bench26.log | cut -c1-160
