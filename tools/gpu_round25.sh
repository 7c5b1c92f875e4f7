#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/op_bench.py > gpurun_out/op_bench25.log 2>&1; echo "op bench rc=$?"; cat gpurun_out/op_bench25.log | cut -c1-160
timeout 120 python tools/op_one.py > gpurun_out/plain_op.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mha_bwd_tc|mha_fwd_tc" -s 4 -c 2 -f -o gpurun_out/prof_attn25 python tools/op_one.py > gpurun_out/ncu_attn.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_attn.log
