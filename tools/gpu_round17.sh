#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/op_one.py > gpurun_out/plain_op.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"mha_bwd_tc|mha_fwd_tc|ln_bwd_kernel|ln_fwd_kernel" -s 8 -c 4 -f -o gpurun_out/prof_ops python tools/op_one.py > gpurun_out/ncu_ops.log 2>&1; echo "ncu ops rc=$?"; tail -3 gpurun_out/ncu_ops.log
