#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/op_one.py > gpurun_out/plain_op.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"ln_bwd_kernel|conv0_ln_gelu|ln_fwd_kernel" -s 5 -c 3 -f -o gpurun_out/prof_ln27 python tools/op_one.py > gpurun_out/ncu_ln.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_ln.log
