#!/bin/bash
# element dropout: cost of the reference-default regularisation in the timed step; kernel bandwidth
mkdir -p gpurun_out
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench63_p0.json 2> gpurun_out/bench63_p0.err; echo "p0 rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --dropout 0.1 > gpurun_out/bench63_p01.json 2> gpurun_out/bench63_p01.err; echo "p0.1 rc=$?"; tail -3 gpurun_out/bench63_p01.err
python - <<'PY'
import json, torch, sys
sys.path.insert(0, 'privacy-preserve-federated-asr_b200')
for f in ('gpurun_out/bench63_p0.json', 'gpurun_out/bench63_p01.json'):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches_per_step'], d['roofline']['frac'], d['clocks'])
    except Exception as e:
        print(f, 'ERR', e)
from b200asr import ops
out = []
for shape, dt in (((3992, 1024), torch.bfloat16), ((3992, 4096), torch.bfloat16), ((3992, 1024), torch.float32)):
    xs = [torch.randn(shape, device='cuda').to(dt) for _ in range(40)]   # rotate > L2
    rs = [torch.randn(shape, device='cuda').to(dt) for _ in range(40)]
    for res in (False, True):
        for _ in range(2):
            for x, r in zip(xs, rs):
                ops.dropout(x, 0.1, 1, 2, 3, residual=r if res else None, out=x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for x, r in zip(xs, rs):
            ops.dropout(x, 0.1, 1, 2, 3, residual=r if res else None, out=x)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / len(xs)
        nbytes = xs[0].numel() * xs[0].element_size() * (3 if res else 2)
        line = f"dropout {tuple(shape)} {str(dt).split('.')[-1]} residual={res}: {us:.1f} us  {nbytes / us / 1e3:.0f} GB/s"
        print(line); out.append(line)
open('gpurun_out/dropout_bw63.txt', 'w').write("\n".join(out) + "\n")
PY
