#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_model.py -q -x -k "graphed" > gpurun_out/t_graph.log 2>&1; echo "graph test rc=$?"; tail -n 25 gpurun_out/t_graph.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench11.log 2> gpurun_out/bench11.err; echo "bench rc=$?"; python -c "
import json;d=json.loads(open('gpurun_out/bench11.log').read().strip().splitlines()[-1]);print('ms/step',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'launches',d['gpu_launches_per_step'],'host ms',d.get('host_enqueue_ms_per_step'),d.get('host_enqueue_ms_per_step_eager'),'loss',d['e2e']['last_loss'], d['step_mode'])"; tail -n 12 gpurun_out/bench11.err
