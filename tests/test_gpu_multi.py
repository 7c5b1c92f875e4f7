"""Multi-GPU FedAvg: one federated client per process / GPU.  The fused peer-memory kernel
(PeerFedAvg: scale + reduce-scatter + all-gather over NVLink, barriers over peer flags) must reproduce the reference's
`average_weights` (federated/src/utils.py:240-249) bit for bit in the uniform case.  Needs >= 2 GPUs:
  gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu
The process group is gloo on purpose: torch.distributed only carries the 64-byte IPC handles, never the weights."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
    pytest.skip("needs >= 2 CUDA devices", allow_module_level=True)

import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


def _client_params(rank, seed=0):
    g = torch.Generator().manual_seed(seed + 17 * rank)
    shapes = {"enc.w": (257, 129), "enc.b": (129,), "head.w": (31, 64), "tiny": (3,), "big.w": (1024, 513)}
    return {k: torch.randn(*s, generator=g) * (1 + rank) for k, s in shapes.items()}


def _worker(rank, world, port, weighted, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "privacy-preserve-federated-asr_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from b200asr.fedavg import FlatParams, PeerFedAvg
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        params = {k: torch.nn.Parameter(v.cuda()) for k, v in _client_params(rank).items()}
        fp = FlatParams(list(params.items()), peer=True)
        agg = PeerFedAvg([fp])
        w = [0.1 + 0.2 * r for r in range(world)] if weighted else None
        if w is not None:
            tot = sum(w)
            w = [x / tot for x in w]
        for rounds in range(3):  # repeated rounds exercise the barrier epochs; after round 1 everyone holds the average
            agg(w)
            if rounds == 0:
                first = {k: p.detach().cpu().clone() for k, p in params.items()}
        torch.cuda.synchronize()
        agg.check()
        torch.save({"first": first, "last": {k: p.detach().cpu() for k, p in params.items()}},
                   os.path.join(out_dir, f"rank{rank}.pt"))
        dist.barrier()
        agg.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("weighted", [False, True])
def test_peer_fedavg_matches_average_weights(tmp_path, weighted):
    world = 2 if torch.cuda.device_count() < 4 else 4
    port = 29500 + (os.getpid() % 500) + (7 if weighted else 0)
    mp.spawn(_worker, args=(world, port, weighted, str(tmp_path)), nprocs=world, join=True)
    clients = [_client_params(r) for r in range(world)]
    res = [torch.load(os.path.join(tmp_path, f"rank{r}.pt")) for r in range(world)]
    for k in clients[0]:
        if weighted:
            w = [0.1 + 0.2 * r for r in range(world)]
            tot = sum(w)
            ref = sum((x / tot) * c[k] for x, c in zip(w, clients))
        else:  # the reference: sequential += then torch.div(., N)
            ref = clients[0][k].clone()
            for c in clients[1:]:
                ref += c[k]
            ref = torch.div(ref, world)
        for r in range(world):
            got = res[r]["first"][k]
            if weighted:
                assert (got - ref).abs().max() <= 1e-6 * max(1.0, float(ref.abs().max())), (k, r)
            else:
                assert torch.equal(got, ref), (k, r)  # bit-exact (BASELINE asks for <= 1e-6)
            assert torch.equal(res[r]["first"][k], res[0]["first"][k])  # every client holds the same model
        if not weighted:  # averaging identical models is the identity
            assert torch.equal(res[0]["last"][k], res[0]["first"][k]), k


def _symm_worker(rank, world, port, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "privacy-preserve-federated-asr_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from b200asr.fedavg import FlatParams, SymmFedAvg
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        params = {k: torch.nn.Parameter(v.cuda()) for k, v in _client_params(rank).items()}
        fp = FlatParams(list(params.items()), peer="symm", with_shadow=True)
        agg = SymmFedAvg([fp])
        start = fp.flat.clone()
        out = {"has_multicast": agg.has_multicast}
        for impl in (["nvls"] if agg.has_multicast else []) + ["peer"]:
            fp.flat.copy_(start)
            torch.cuda.synchronize()
            dist.barrier()
            agg(impl=impl)
            torch.cuda.synchronize()
            agg.check()
            out[impl] = {k: p.detach().cpu().clone() for k, p in params.items()}
            out[impl + "_shadow_ok"] = bool(torch.equal(fp.shadow, fp.flat.to(torch.bfloat16)))   # refreshed by mark_updated
            dist.barrier()
            # the same exchange chunk by chunk on a comm stream (PipelinedFedAvg: capped grid, per-chunk barrier + shadow slice)
            from b200asr.fedavg import PipelinedFedAvg
            n = fp.numel
            cuts = [0, (n // 3) // 4 * 4, (2 * n // 3) // 4 * 4, n]
            pipe = PipelinedFedAvg([fp], [[(0, cuts[i], cuts[i + 1])] for i in range(3)], agg=agg, max_blocks=8)
            fp.flat.copy_(start)
            fp.shadow.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            comm = torch.cuda.Stream()
            evs = pipe.launch(comm, impl)
            torch.cuda.current_stream().wait_event(evs[-1])
            torch.cuda.synchronize()
            agg.check()
            out["pipe_" + impl] = {k: p.detach().cpu().clone() for k, p in params.items()}
            out["pipe_" + impl + "_shadow_ok"] = bool(torch.equal(fp.shadow, fp.flat.to(torch.bfloat16)))
            dist.barrier()
        torch.save(out, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_symm_fedavg_nvls_and_peer_match_average_weights(tmp_path):
    """FlatParams over torch symmetric memory: the own NVLS kernel (multimem.ld_reduce in the switch + multimem.st) equals
    the reference's sequential mean to <= 1e-6 (BASELINE.json) on every rank, the peer kernel on the same buffers is
    bit-exact; the bf16 shadow follows."""
    world = min(torch.cuda.device_count(), 8)
    port = 29100 + (os.getpid() % 500)
    mp.spawn(_symm_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    clients = [_client_params(r) for r in range(world)]
    res = [torch.load(os.path.join(tmp_path, f"rank{r}.pt")) for r in range(world)]
    impls = (["nvls"] if res[0]["has_multicast"] else []) + ["peer"]
    print(f"[symm fedavg] world {world}, multicast {res[0]['has_multicast']}, impls {impls}")
    for k in clients[0]:
        ref = clients[0][k].clone()
        for c in clients[1:]:
            ref += c[k]
        ref = torch.div(ref, world)
        for impl in impls:
            for r in range(world):
                got = res[r][impl][k]
                if impl == "peer":
                    assert torch.equal(got, ref), (k, r)
                else:
                    assert (got - ref).abs().max() <= 1e-6 * max(1.0, float(ref.abs().max())), (impl, k, r)
                assert torch.equal(got, res[0][impl][k])   # every client holds the same bits
                assert res[r][impl + "_shadow_ok"]
                assert torch.equal(res[r]["pipe_" + impl][k], got), (impl, k, r)   # chunked == whole-buffer exchange
                assert res[r]["pipe_" + impl + "_shadow_ok"]
