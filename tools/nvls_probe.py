#!/usr/bin/env python
"""torchrun --nproc-per-node N tools/nvls_probe.py : does torch symmetric memory give a multicast address on this box, is
the NVLS FedAvg kernel correct, and how fast is it against the peer kernel and NCCL on the real 1.24 GB buffer?"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "privacy-preserve-federated-asr_b200")]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
out = {"world": world}
try:
    from b200asr.fedavg import FlatParams, SymmFedAvg, fedavg_allreduce
    n = int(os.environ.get("PROBE_N", 309_101_600))
    p = torch.nn.Parameter(torch.empty(n, device=dev))
    fp = FlatParams([("w", p)], peer="symm", with_shadow=True)
    agg = SymmFedAvg([fp])
    out["has_multicast"] = agg.has_multicast
    g = torch.Generator(device=dev).manual_seed(rank + 1)
    vals = torch.randn(n, device=dev, generator=g) * (1 + rank)

    def expected():
        e = vals.clone()
        dist.all_reduce(e)
        return e / world

    exp = expected()
    for impl in (["nvls"] if agg.has_multicast else []) + ["peer"]:
        fp.flat.copy_(vals)
        torch.cuda.synchronize(); dist.barrier()
        agg(impl=impl)
        torch.cuda.synchronize(); dist.barrier()
        agg.check()
        err = float((fp.flat[:n] - exp).abs().max() / exp.abs().max())
        out[impl + "_rel_err"] = err
        # every rank must hold the same bits
        h = fp.flat[:n].double().sum().reshape(1)
        hs = [torch.zeros_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
        out[impl + "_same_on_all_ranks"] = bool(all(float(x) == float(hs[0]) for x in hs))

    def timeit(fn, iters=8):
        fn(); torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(); dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    if agg.has_multicast:
        out["nvls_ms"] = timeit(lambda: agg(impl="nvls", mark=False))
        out["nvls_ms_64blocks"] = timeit(lambda: agg(impl="nvls", mark=False, max_blocks=64))
        out["nvls_ms_296blocks"] = timeit(lambda: agg(impl="nvls", mark=False, max_blocks=296))
    out["peer_ms"] = timeit(lambda: agg(impl="peer", mark=False))
    out["nccl_ms"] = timeit(lambda: fedavg_allreduce(fp.flat))
    out["shadow_refresh_ms"] = timeit(lambda: fp.refresh_shadow())
    out["bytes"] = 4 * n
except Exception as e:
    import traceback
    out["error"] = f"{type(e).__name__}: {e}"
    out["trace"] = traceback.format_exc()[-1500:]
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
