#!/usr/bin/env python
"""bench.py -- the headline measurement (BASELINE.json): training audio-seconds / second of the data2vec-audio-large
CTC client-local update on B200, and the FedAvg round at 1/2/4/8 GPUs (one federated client per GPU).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" = one client-local update on one batch of synthetic audio (configs[1]: batch 8 x 10 s at 16 kHz, bf16):
zero_grad -> forward (frozen conv stack, 24-layer encoder, lm_head, CTC) -> backward -> fused AdamW; with N > 1 every
step ends with the FedAvg of the trainable parameters across the N clients (NCCL all-reduce over NVLink).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for what every key means.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "privacy-preserve-federated-asr_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

METRIC = "train audio-sec/sec (data2vec-large CTC) + FedAvg round s at 1/2/4/8 B200"
UNIT = "audio-s/s"
BATCH, SECONDS, SR = 8, 10, 16000
WORKLOAD = "data2vec-audio-large CTC fine-tune step, batch 8 x 10 s synthetic 16 kHz audio, bf16 (BASELINE configs[1])"


def large_config(**kw):
    from types import SimpleNamespace
    c = dict(vocab_size=32, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, intermediate_size=4096,
             conv_dim=(512,) * 7, conv_kernel=(10, 3, 3, 3, 3, 2, 2), conv_stride=(5, 2, 2, 2, 2, 2, 2), conv_bias=False,
             num_conv_pos_embeddings=5, conv_pos_kernel_size=19, num_conv_pos_embedding_groups=16, layer_norm_eps=1e-5,
             pad_token_id=0, ctc_loss_reduction="sum", ctc_zero_infinity=True, mask_time_prob=0.0,
             feat_extract_norm="layer", do_stable_layer_norm=False)
    c.update(kw)
    return SimpleNamespace(**c)


def synth_batch(batch, samples, seed, frames, pin=True):
    """Synthetic normalised 16 kHz audio: full-length utterances, attention mask supplied (as the
    training collator does), 20..120 labels per utterance, ids 1..31."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch, samples, generator=g)
    x = (x - x.mean(-1, keepdim=True)) / torch.sqrt(x.var(-1, unbiased=False, keepdim=True) + 1e-7)
    mask = torch.ones(batch, samples, dtype=torch.int32)
    hi = max(2, min(120, frames // 2))
    lo = min(20, hi - 1)
    n = torch.randint(lo, hi + 1, (batch,), generator=g)
    labels = torch.full((batch, int(n.max())), -100, dtype=torch.long)
    for b in range(batch):
        labels[b, : int(n[b])] = torch.randint(1, 32, (int(n[b]),), generator=g)
    out = dict(input_values=x, attention_mask=mask, labels=labels)
    if pin and torch.cuda.is_available():
        out = {k: v.pin_memory() for k, v in out.items()}
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.f.name)
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2]))
                for name, v in zip(names, r[5:9]):
                    if v.strip().lower() == "active":
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


# ======================================================================================= CPU reference arm
def cpu_reference_step_time(steps, warmup, seconds=2.0, batch=1):
    """The reference's CPU path for this metric: the oracle port (oracle/data2vec_ref.py, a restatement of
    HF Data2VecAudioForCTC + F.ctc_loss as the reference runs them) doing forward + backward + AdamW in fp32 on all
    host cores, on a BOUNDED sample of the workload (full-size model, `batch` x `seconds` s of audio)."""
    from oracle import data2vec_ref as R
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = R.ArchConfig()
    sd = R.init_state_dict(cfg, seed=0, dacs=False)
    params = {k: v.requires_grad_(not k.startswith("data2vec_audio.feature_extractor")) for k, v in sd.items()}
    opt = torch.optim.AdamW([p for p in params.values() if p.requires_grad], lr=1e-5, weight_decay=0.005)
    samples = int(seconds * SR)
    b = R.synth_batch(cfg, batch, samples, seed=1234, ragged=False, max_label=120)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        loss, _ = R.ctc_forward(params, cfg, b["input_values"], b["attention_mask"], b["labels"])
        loss.backward()
        torch.nn.utils.clip_grad_norm_([p for p in params.values() if p.requires_grad], 1.0)
        opt.step()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    t = statistics.median(times)
    return dict(value=batch * seconds / t, unit=UNIT, cores=os.cpu_count() or 1, kind="port",
                sample=f"full-size model, batch {batch} x {seconds:g} s, fp32 fwd+bwd+AdamW, median of {steps} step(s), "
                       f"{t:.2f} s/step", ms_per_step=t * 1e3)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 3)), max(0, min(args.warmup, 1))
    cb = cpu_reference_step_time(steps, warmup)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": cb["sample"]},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ======================================================================================= B200 arm
def run_b200(args):
    import torch.distributed as dist

    from b200asr import _lib, modeling as M, ops
    from b200asr.fedavg import PeerFedAvg, fedavg_allreduce
    from b200asr.optim import FusedAdamW

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the b200 arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    pd = float(args.dropout)
    cfg = large_config(hidden_dropout=pd, activation_dropout=pd, final_dropout=pd, feat_proj_dropout=0.0,
                       attention_dropout=0.0, layerdrop=0.0)
    torch.manual_seed(0)  # identical initial weights on every client
    model = M.Data2VecAudioForCTC(cfg, compute_dtype=torch.bfloat16)
    model.freeze_feature_encoder()
    model.validate_labels = False  # the reference's labels.max() check costs a host sync per step; inputs are synthetic
    model.to(dev).train()
    use_peer = world > 1 and args.fedavg in ("peer", "auto")
    opt = FusedAdamW(model.named_parameters(), lr=1e-5, weight_decay=0.005, max_grad_norm=1.0, warmup_steps=1000,
                     peer=use_peer)
    peer_agg = PeerFedAvg([g["fp"] for g in opt.groups]) if use_peer else None
    n_train = sum(g["fp"].numel for g in opt.groups)
    samples = SECONDS * SR
    frames = int(M.conv_out_lengths(cfg, torch.tensor(samples)))
    # a few distinct host batches per client (seed offset by rank: every client has its own shard)
    host = [synth_batch(BATCH, samples, 1234 + 100 * rank + i, frames) for i in range(4)]
    resident = [{k: v.to(dev) for k, v in hb.items()} for hb in host]

    def eager_step(batch):
        opt.zero_grad()
        out = model(batch["input_values"], batch["attention_mask"], labels=batch["labels"])
        out.loss.backward()
        opt.step()
        return out.loss

    # the call a user makes: the whole update captured once as a CUDA graph, replayed per batch (b200asr/graph.py)
    graphed = None
    if not args.eager:
        from b200asr.graph import GraphedTrainStep
        graphed = GraphedTrainStep(model, opt, resident[0], max_label_len=120)

    def local_step(batch):
        return graphed(batch) if graphed is not None else eager_step(batch)

    def fedavg_nccl():
        for g in opt.groups:
            fedavg_allreduce(g["fp"].flat)
            g["fp"].mark_updated()

    chosen = {"impl": "peer" if use_peer else "nccl"}

    def fedavg():
        if world > 1:
            if chosen["impl"] == "peer":   # fused peer-memory kernel: scale + reduce-scatter + all-gather over NVLink
                peer_agg()
            else:                          # our scale kernel + NCCL all-reduce (NVLS in-switch reduction at 8 GPUs)
                fedavg_nccl()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    copy_stream = torch.cuda.Stream()

    def timed(nsteps, e2e):
        """CUDA-event time of nsteps steps (max over ranks), barrier + synchronize on both sides."""
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        agg = []
        t0.record()
        last = None
        main = torch.cuda.current_stream()

        def stage(i):
            # pinned host batch -> device on the copy stream (every step's copy is inside the timed region; from the second
            # step on it runs while the previous step computes, as a prefetching data loader would do it)
            with torch.cuda.stream(copy_stream):
                b = {k: v.to(dev, non_blocking=True) for k, v in host[i % len(host)].items()}
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            for v in b.values():
                v.record_stream(main)
            return b, ev

        nxt = stage(0) if e2e else None
        for i in range(nsteps):
            if e2e:  # the call a user makes: pinned host batch -> device, update, loss read back
                batch, ev = nxt
                main.wait_event(ev)
            else:
                batch = resident[i % len(resident)]
            loss = local_step(batch)
            if world > 1:
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record(); fedavg(); a1.record()
                agg.append((a0, a1))
            if e2e:
                if i + 1 < nsteps:
                    nxt = stage(i + 1)
                last = float(loss)  # device -> host read of the step's result
        t1.record()
        barrier()
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        agg_ms = torch.tensor([sum(a.elapsed_time(b) for a, b in agg) / max(1, len(agg))], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(agg_ms, op=dist.ReduceOp.MAX)
        return float(ms) / nsteps, float(agg_ms), last

    def time_alone(fn, iters=5):
        barrier()
        fn(); barrier()
        o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        o0.record()
        for _ in range(iters):
            fn()
        o1.record()
        barrier()
        t = torch.tensor([o0.elapsed_time(o1) / iters], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    for i in range(args.warmup):
        local_step(resident[i % len(resident)])
        fedavg()
    agg_alone = {}
    if world > 1:  # both aggregation paths timed alone on the real buffers (max over ranks: every rank sees the same numbers)
        agg_alone["nccl_allreduce_ms"] = time_alone(fedavg_nccl)
        if peer_agg is not None:
            agg_alone["peer_kernel_ms"] = time_alone(peer_agg)
            peer_agg.check()
            if args.fedavg == "auto":
                chosen["impl"] = "peer" if agg_alone["peer_kernel_ms"] <= agg_alone["nccl_allreduce_ms"] else "nccl"
    clocks = ClockSampler(local_rank) if rank == 0 else None
    n0 = _lib.launch_count()
    ms_step, agg_ms, _ = timed(args.steps, e2e=False)
    launches = (_lib.launch_count() - n0) // args.steps
    if graphed is not None:  # replays launch the captured kernels without passing through the host wrappers
        launches = graphed.launches_per_replay
    clk = clocks.stop() if clocks else None
    ms_e2e, _, last_loss = timed(args.steps, e2e=True)

    # ---- the other FedAvg implementation, timed alone on identical buffers (every rank holds the same average by now)
    # ---- host enqueue time of one step (no synchronisation inside: pure Python + launch cost; the GPU queue holds a step)
    torch.cuda.synchronize()
    h0 = time.perf_counter()
    local_step(resident[0])
    host_ms = (time.perf_counter() - h0) * 1e3
    torch.cuda.synchronize()
    h0 = time.perf_counter()
    eager_step(resident[0])
    host_eager_ms = (time.perf_counter() - h0) * 1e3
    torch.cuda.synchronize()

    # ---- roofline leg: one extra instrumented step, every GEMM / attention call bracketed by CUDA events
    eager_step(resident[0])  # untimed: the eager path's workspaces are allocated here, not inside an event bracket
    torch.cuda.synchronize()
    ops.PROFILE = []
    eager_step(resident[0])
    torch.cuda.synchronize()
    prof, ops.PROFILE = ops.PROFILE, None
    kinds = {}
    for kind, e0, e1, fl in prof:
        k = kinds.setdefault(kind, [0, 0.0, 0.0])
        k[0] += 1; k[1] += e0.elapsed_time(e1); k[2] += fl
    peaks, src = measured_peaks()
    gm = kinds.get("gemm", [0, 1e-9, 0.0])
    achieved = gm[2] / (gm[1] * 1e-3) / 1e12
    peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0)))
    traffic = None
    try:  # per-launch DRAM traffic of this kernel from the committed ncu capture of this same command (profiles/)
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_ncu_gemm_traffic_v16.json")))["dram_bytes_per_launch"]
    except Exception:
        pass
    roofline = {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05 bf16 GEMM: QKV/O/FFN linears fwd+dgrad+wgrad, conv layers 1-6)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": "profiles/r01_ncu_gemm_traffic_v16.json (ncu dram__bytes_read.sum + dram__bytes_write.sum, mean over one step's launches)",
                "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({src})", "launches": gm[0],
                "avg_launch_us": gm[1] * 1e3 / max(1, gm[0]), "share_of_step": gm[1] / ms_step,
                "other": {k: {"launches": v[0], "ms": v[1], "tflops": v[2] / (v[1] * 1e-3) / 1e12} for k, v in kinds.items()
                          if k != "gemm"}}

    if rank == 0:
        audio_s = BATCH * SECONDS * world
        cb = cpu_reference_step_time(steps=1, warmup=0) if world == 1 and not args.no_cpu_baseline else None
        algo_tflop = 8.376  # conv + 3 x rest, B=8 x 10 s, frozen conv stack, no recompute
        line = {
            "metric": METRIC, "value": audio_s / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "global_batch": BATCH * world, "audio_s_per_step": audio_s,
                       "clients": world, "trainable_params": n_train, "frames_per_utt": frames,
                       "regularisation": (f"hidden / activation / final dropout {pd:g} applied in the timed step (masks drawn "
                                          "in-kernel, fresh per CUDA-graph replay); " if pd > 0 else "element dropout 0; ")
                                         + "attention-probability dropout not applied; layerdrop 0 = all 24 blocks run "
                                           "every step (the reference default 0.1 would skip 10% of them)",
                       "l2_policy": "inputs+weights+saved activations per step (>3 GB) exceed the 126 MB L2; no flush",
                       "e2e_input_pipeline": "each step's pinned-host -> device copy runs on a copy stream, issued while the previous step computes; the loss is read back every step",
                       "fedavg_impl": chosen["impl"] if world > 1 else None,
                       "fedavg": (("uniform FedAvg of all trainable params after every local step: fused peer-memory kernel "
                                   "(scale + reduce-scatter + all-gather over NVLink, no NCCL on the data path)") if chosen["impl"] == "peer" else
                                  "uniform FedAvg of all trainable params after every local step (scale kernel + NCCL all-reduce)")
                                 if world > 1 else "n/a at 1 GPU"},
            "e2e": {"value": audio_s / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": sum(v.numel() * v.element_size() for v in host[0].values()),
                    "d2h_bytes_per_step": 4, "last_loss": last_loss},
            "gpu_launches": int(launches) * args.steps,
            "gpu_launches_per_step": int(launches),
            "host_enqueue_ms_per_step": host_ms, "host_enqueue_ms_per_step_eager": host_eager_ms,
            "step_mode": "cuda_graph" if graphed is not None else "eager",
            "step_tflops": algo_tflop * world / (ms_step * 1e-3),
            "step_frac_of_tensor_peak": algo_tflop / (ms_step * 1e-3) / peak,
            "fedavg_round_s": ms_step * 1e-3, "fedavg_agg_ms": agg_ms if world > 1 else None,
            "fedavg_agg_alone": agg_alone or None,  # each implementation timed back to back without the local step
            # per GPU and direction: its own remote reads + the peers' writes into it = 2 * 4P (N-1)/N bytes
            "fedavg_nvlink_gbs_per_dir": (2 * 4 * n_train * (world - 1) / world / (agg_alone["peer_kernel_ms"] * 1e-3) / 1e9)
                                         if agg_alone.get("peer_kernel_ms") else None,
            "fedavg_bytes_per_client": 4 * n_train,
            "clocks": clk, "roofline": roofline,
        }
        if cb is not None:
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fedavg", default="auto", choices=["auto", "peer", "nccl"],
                    help="N>1 aggregation: fused peer-memory kernel, scale kernel + NCCL all-reduce, or whichever measures faster in warm-up")
    ap.add_argument("--dropout", type=float, default=0.1,
                    help="hidden_dropout = activation_dropout = final_dropout = P in the timed training step "
                         "(0.1 = the Data2VecAudioConfig defaults the reference fine-tunes with; 0 = the parity setting)")
    ap.add_argument("--eager", action="store_true", help="issue every launch from Python instead of replaying the CUDA graph")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        if args.warmup < 3:
            args.warmup = 3
        run_b200(args)


if __name__ == "__main__":
    main()
