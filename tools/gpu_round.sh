#!/bin/bash
# One gpurun call: kernel parity tests, tcgen05 probes per operand-major combo (each in its own process so a
# fault in one does not hide the others), model parity.  Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.log 2>&1
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -k "not tcgen05" > gpurun_out/t1.log 2>&1; echo "t1 rc=$?"
for c in kk km mk mm; do timeout 180 python tools/gemm_probe.py $c > gpurun_out/probe_$c.log 2>&1; echo "probe $c rc=$?"; done
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -k "tcgen05" > gpurun_out/t1b.log 2>&1; echo "t1b rc=$?"
timeout 900 python -m pytest tests/test_gpu_ops2.py -q > gpurun_out/t2.log 2>&1; echo "t2 rc=$?"
timeout 900 python -m pytest tests/test_gpu_model.py -q > gpurun_out/t3.log 2>&1; echo "t3 rc=$?"
for f in t1 probe_kk probe_km probe_mk probe_mm t1b t2 t3; do echo "=== $f"; tail -n 12 gpurun_out/$f.log; done
