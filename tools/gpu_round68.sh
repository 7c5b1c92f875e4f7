#!/bin/bash
# ncu --set full of the dropout kernels (stand-alone pass and the LayerNorm forward with fused dropout + residual)
mkdir -p gpurun_out
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"dropout_kernel|ln_fwd_kernel" -s 120 -c 8 -f -o gpurun_out/prof_dropout68 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --eager > gpurun_out/ncu68.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/prof_dropout68.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__grid_size,sm__warps_active.avg.pct_of_peak_sustained_active > gpurun_out/prof_dropout68.csv 2>/dev/null; echo "csv rc=$?"; wc -l gpurun_out/prof_dropout68.csv
