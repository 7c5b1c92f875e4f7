#!/bin/bash
# staged epilogue v3 (64-column chunks, compile-time epilogue modes, bias slice in registers): parity + trace + bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_ops2.py -m gpu -x -q > gpurun_out/t35_gemm.log 2>&1; echo "gemm tests rc=$?"; tail -3 gpurun_out/t35_gemm.log
{
for s in "1024 4096" "4096 1024" "1024 1024"; do
  timeout 120 python tools/gemm_trace.py $s
done
} > gpurun_out/gemm_trace35.txt 2>&1
grep -A3 "^launch 19\|^N=" gpurun_out/gemm_trace35.txt | grep -v "previous"
grep -A12 "^launch 19" gpurun_out/gemm_trace35.txt | head -13
timeout 300 python tools/gemm_bench.py > gpurun_out/gemm_bench35.txt 2>&1; tail -25 gpurun_out/gemm_bench35.txt
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench35.json 2> gpurun_out/bench35.err; python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench35.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'ms_per_step')}, d['e2e']['value'], d['roofline']['frac'], d.get('clocks'))
PY
