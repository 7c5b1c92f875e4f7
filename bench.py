#!/usr/bin/env python
"""bench.py -- the headline measurement (BASELINE.json): training audio-seconds / second of the data2vec-audio-large
CTC client-local update on B200, and the FedAvg round at 1/2/4/8 GPUs (one federated client per GPU).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" = one client-local update on one batch of synthetic audio (configs[1]: batch 8 x 10 s at 16 kHz, bf16) with the
reference's real training configuration (Data2VecAudioConfig defaults, which centralized/finetune_ASRs.py:160-164 keeps:
hidden / activation / attention / final dropout 0.1, layerdrop 0.1): zero_grad -> forward (frozen conv stack, 24-layer
encoder, lm_head, CTC) -> backward -> fused AdamW; with N > 1 every step ends with the FedAvg of the trainable parameters
across the N clients.  Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for what every key means.

Legs of the b200 line (all measured in this process, on this box):
  value / ms_per_step   inputs resident in HBM, CUDA events, max over ranks
  e2e                   pinned host batch -> device copy and loss read-back inside the timed region
  roofline              the dominant kernel (tcgen05 GEMM): algorithmic FLOPs / CUDA-event time per launch
  torch_gpu_baseline    the oracle port (the reference's model restated on stock PyTorch ops) on THIS GPU under bf16
                        autocast + torch.optim.AdamW(fused=True), same batch, same regularisation: "the real bar"
  cpu_baseline          the oracle port on the host cores, bounded sample
  other_configs         C1 greedy inference / C3 DACS stage 2 / C5 wav2vec2-large 16 x 30 s at full size
`--impl reference` times the reference's CPU path (oracle port, all host cores) at the metric's own configuration.
"""
from __future__ import annotations

import argparse
import hashlib
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
N_LAYERS = 24
COIN_SEED = 20260  # host LayerDrop coins (np.random.uniform, as the pinned transformers 4.17): same stream in every arm


def large_config(**kw):
    from types import SimpleNamespace
    c = dict(vocab_size=32, hidden_size=1024, num_hidden_layers=N_LAYERS, num_attention_heads=16, intermediate_size=4096,
             conv_dim=(512,) * 7, conv_kernel=(10, 3, 3, 3, 3, 2, 2), conv_stride=(5, 2, 2, 2, 2, 2, 2), conv_bias=False,
             num_conv_pos_embeddings=5, conv_pos_kernel_size=19, num_conv_pos_embedding_groups=16, layer_norm_eps=1e-5,
             pad_token_id=0, ctc_loss_reduction="sum", ctc_zero_infinity=True, mask_time_prob=0.0,
             feat_extract_norm="layer", do_stable_layer_norm=False)
    c.update(kw)
    return SimpleNamespace(**c)


def regularisation(pd: float, ld: float) -> dict:
    return dict(hidden_dropout=pd, activation_dropout=pd, attention_dropout=pd, final_dropout=pd, feat_proj_dropout=0.0,
                layerdrop=ld)


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


def coin_stream(steps, p, seed=COIN_SEED):
    """Per step, the blocks LayerDrop skips (HF:392-404: one uniform per block per step, skipped below p)."""
    import numpy as np
    rs = np.random.RandomState(seed)
    return [tuple(l for l in range(N_LAYERS) if rs.uniform(0, 1) < p) if p > 0 else () for _ in range(steps)]


# ======================================================================================= the oracle port as a baseline
def oracle_step_time(device, batch, seconds, steps, warmup, pd, ld, autocast):
    """forward + backward + clip + AdamW of the full-size model through the oracle port (oracle/data2vec_ref.py: the HF
    Data2VecAudioForCTC arithmetic on stock torch ops).  CPU: fp32 on all host cores.  cuda: bf16 autocast, fused AdamW,
    CUDA-event timing.  Same regularisation as the product arm: torch dropout at the HF sites, LayerDrop coins."""
    import torch.nn.functional as F
    from oracle import data2vec_ref as R
    cfg = R.ArchConfig()
    sd = R.init_state_dict(cfg, seed=0, dacs=False)
    params = {k: v.to(device).requires_grad_(not k.startswith("data2vec_audio.feature_extractor")) for k, v in sd.items()}
    train = [p for p in params.values() if p.requires_grad]
    opt = torch.optim.AdamW(train, lr=1e-5, weight_decay=0.005, fused=(device != "cpu"))
    samples = int(seconds * SR)
    frames = int(R.conv_out_lengths(cfg, torch.tensor([samples]))[0])
    hb = synth_batch(batch, samples, 1234, frames, pin=False)
    x, am, lab = hb["input_values"].to(device), hb["attention_mask"].long().to(device), hb["labels"].to(device)
    coins = coin_stream(warmup + steps, ld)
    drop = (lambda site, t: F.dropout(t, pd)) if pd > 0 else (lambda site, t: t)
    times = []
    for i in range(warmup + steps):
        if device == "cpu":
            t0 = time.perf_counter()
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            loss, _ = R.ctc_forward(params, cfg, x, am, lab, skip_layers=set(coins[i]), drop=drop)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(train, 1.0)
        opt.step()
        if device == "cpu":
            dt = time.perf_counter() - t0
        else:
            e1.record()
            torch.cuda.synchronize()
            dt = e0.elapsed_time(e1) * 1e-3
        if i >= warmup:
            times.append(dt)
    t = statistics.median(times)
    return dict(value=batch * seconds / t, ms_per_step=t * 1e3, last_loss=float(loss.detach()), steps=steps, warmup=warmup)


def cpu_reference(steps, warmup, batch, seconds, pd, ld):
    torch.set_num_threads(os.cpu_count() or 1)
    r = oracle_step_time("cpu", batch, seconds, steps, warmup, 0.0, ld, autocast=False)
    r.update(unit=UNIT, cores=os.cpu_count() or 1, kind="port",
             sample=f"full-size model, batch {batch} x {seconds:g} s, fp32 fwd+bwd+clip+AdamW, layerdrop {ld:g} (same coin stream), "
                    f"element dropout not applied (cost-neutral on the CPU), {warmup} warm-up + median of {steps} step(s), "
                    f"{r['ms_per_step'] / 1e3:.2f} s/step")
    return r


def run_reference(args):
    """The reference arm: the reference's CPU implementation of the path (oracle port -- the reference has no native code
    and is not a package) on all host cores AT THE METRIC'S CONFIGURATION (batch 8 x 10 s), 1 warm-up +
    up to 3 timed steps (~0.5 min each on 16 cores)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 3)), 1
    cb = cpu_reference(steps, warmup, args.ref_batch, args.ref_seconds, args.dropout, args.layerdrop)
    same = args.ref_batch == BATCH and args.ref_seconds == SECONDS
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if same else WORKLOAD + f" -- BOUNDED SAMPLE batch {args.ref_batch} x {args.ref_seconds:g} s",
                       "global_batch": args.ref_batch, "audio_s_per_step": args.ref_batch * args.ref_seconds,
                       "sample": cb["sample"]},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    args._emit(json.dumps(line))


# ======================================================================================= other configurations (N = 1)
def other_configs(dev, steps=4, warmup=2):
    """C1 / C3 / C5 of BASELINE.json at full size, measured by the same process (tools/config_bench.py holds the longer
    version).  Parity-test cases, not the headline: a few steps each."""
    from types import SimpleNamespace

    from b200asr import modeling as M
    from b200asr.graph import GraphedTrainStep
    from b200asr.optim import FusedAdamW
    out = {}

    def timed(fn, n, w):
        for _ in range(w):
            fn()
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(n):
            fn()
        t1.record()
        torch.cuda.synchronize()
        return t0.elapsed_time(t1) / n

    def batch_of(cfg, B, seconds):
        frames = int(M.conv_out_lengths(cfg, torch.tensor(seconds * SR)))
        hb = synth_batch(B, seconds * SR, 1234, frames, pin=False)
        hb["dementia_labels"] = (torch.rand(B, generator=torch.Generator().manual_seed(1237)) < 0.5).long()
        return {k: v.to(dev) for k, v in hb.items()}

    torch.manual_seed(0)
    cfg = large_config()
    model = M.Data2VecAudioForCTC(cfg, compute_dtype=torch.bfloat16).to(dev).eval()
    b = batch_of(cfg, 1, 5)

    from b200asr.graph import GraphedInference
    ginfer = GraphedInference(model, b["input_values"])
    ms = timed(lambda: ginfer(b["input_values"]), 4 * steps, warmup)

    @torch.no_grad()
    def infer():
        o = model(b["input_values"])
        return model.greedy_decode(o.logits)
    ms_eager = timed(infer, 4 * steps, warmup)
    out["C1"] = dict(workload="greedy CTC inference, batch 1 x 5 s (forward + argmax / collapse as one CUDA graph, id lists "
                              "read back to the host every call)",
                     ms_per_step=ms, audio_s_per_s=5.0 / (ms * 1e-3), ms_per_step_eager_launches=ms_eager)
    del ginfer
    del model

    cfg = large_config(**regularisation(0.1, 0.0))
    args = SimpleNamespace(LAMBDA=0.5, TOGGLE_RATIO=0, GS_TAU=1, AD_loss="recall", W_LOSS=None, STAGE=2)
    model = M.DacsData2VecAudioForCTC(cfg, args, compute_dtype=torch.bfloat16)
    model.validate_labels = False
    model.to(dev).train()
    b = batch_of(cfg, 8, 10)
    opt = FusedAdamW(model.named_parameters(), lr=1e-5, weight_decay=0.005, max_grad_norm=1.0)
    gstep = GraphedTrainStep(model, opt, b, max_label_len=120)
    ms = timed(lambda: gstep(b), 2 * steps, warmup)
    out["C3"] = dict(workload="DACS stage-2 step (gumbel_softmax GS_TAU=1, GRL AD head, 2 x CTC, CosFace), batch 8 x 10 s, frozen "
                              "encoder forward in training mode (dropouts 0.1 incl. attention) + arbitrator update, one CUDA graph",
                     ms_per_step=ms, audio_s_per_s=80.0 / (ms * 1e-3))
    del model, opt, gstep
    torch.cuda.empty_cache()

    cfg = large_config(feat_extract_norm="layer", conv_bias=True, do_stable_layer_norm=True, num_conv_pos_embeddings=128,
                       **regularisation(0.1, 0.0))
    model = M.Wav2Vec2ForCTC(cfg, compute_dtype=torch.bfloat16)
    model.freeze_feature_encoder()
    model.validate_labels = False
    model.to(dev).train()
    b = batch_of(cfg, 16, 30)
    b.pop("dementia_labels")
    opt = FusedAdamW(model.named_parameters(), lr=1e-5, weight_decay=0.005, max_grad_norm=1.0)

    def step():
        opt.zero_grad()
        model(**b).loss.backward()
        opt.step()
    ms = timed(step, steps, warmup)
    out["C5"] = dict(workload="wav2vec2-large-960h-lv60 architecture CTC step, batch 16 x 30 s (T = 1499), dropouts 0.1 incl. attention, "
                              "eager launches", ms_per_step=ms, audio_s_per_s=480.0 / (ms * 1e-3),
                     peak_mem_gb=torch.cuda.max_memory_allocated() / 2 ** 30)
    del model, opt
    torch.cuda.empty_cache()
    return out


# ======================================================================================= B200 arm
def run_b200(args):
    import numpy as np
    import torch.distributed as dist

    from b200asr import _lib, modeling as M, ops
    from b200asr.fedavg import PeerFedAvg, PipelinedFedAvg, SymmFedAvg, fedavg_allreduce
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

    pd, ld = float(args.dropout), float(args.layerdrop)
    cfg = large_config(**regularisation(pd, ld))
    torch.manual_seed(0)  # identical initial weights on every client
    model = M.Data2VecAudioForCTC(cfg, compute_dtype=torch.bfloat16)
    model.freeze_feature_encoder()
    model.validate_labels = False  # the reference's labels.max() check costs a host sync per step; inputs are synthetic
    model.to(dev).train()
    # N > 1: the flat parameter buffers live in memory every client can reach.  Preferred: torch symmetric memory (peer
    # pointers AND one NVSwitch multicast address: our NVLS kernel + our peer kernel on the same buffers); fallback: plain
    # cudaMalloc exported over CUDA IPC (peer kernel only); "nccl": ordinary allocations.
    use_peer = world > 1 and args.fedavg in ("peer", "auto", "nvls")
    opt, peer_agg, symm_note = None, None, None
    if use_peer and args.fedavg in ("auto", "nvls") and not args.no_symm:
        try:
            opt = FusedAdamW(model.named_parameters(), lr=1e-5, weight_decay=0.005, max_grad_norm=1.0, warmup_steps=1000,
                             peer="symm")
            peer_agg = SymmFedAvg([g["fp"] for g in opt.groups])
            symm_note = "symmetric memory, multicast " + ("available" if peer_agg.has_multicast else "NOT available")
        except Exception as e:   # no symmetric-memory support on this box: CUDA IPC buffers + the peer kernel
            symm_note = f"symmetric memory unavailable ({type(e).__name__}: {str(e)[:120]})"
            opt, peer_agg = None, None
    if opt is None:
        opt = FusedAdamW(model.named_parameters(), lr=1e-5, weight_decay=0.005, max_grad_norm=1.0, warmup_steps=1000,
                         peer=use_peer)
        peer_agg = PeerFedAvg([g["fp"] for g in opt.groups]) if use_peer else None
    is_symm = isinstance(peer_agg, SymmFedAvg)
    n_train = sum(g["fp"].numel for g in opt.groups)
    samples = SECONDS * SR
    frames = int(M.conv_out_lengths(cfg, torch.tensor(samples)))
    # a few distinct host batches per client (seed offset by rank: every client has its own shard)
    host = [synth_batch(BATCH, samples, 1234 + 100 * rank + i, frames) for i in range(4)]
    resident = [{k: v.to(dev) for k, v in hb.items()} for hb in host]
    np.random.seed(COIN_SEED + rank)   # LayerDrop coins of this client

    def eager_step(batch):
        opt.zero_grad()
        out = model(batch["input_values"], batch["attention_mask"], labels=batch["labels"])
        out.loss.backward()
        opt.step()
        return out.loss

    # the call a user makes: the whole update captured once as CUDA graphs, replayed per batch (b200asr/graph.py)
    graphed = None
    if not args.eager:
        from b200asr.graph import GraphedTrainStep
        # N > 1: the frozen conv stack is a graph of its own, so that the next round's begin() can run under the aggregation
        layerwise = world > 1 and not args.no_overlap and not args.no_layerwise
        graphed = GraphedTrainStep(model, opt, resident[0], max_label_len=120, split_conv=(world > 1 and not args.no_overlap),
                                   defer_adamw=(layerwise and not args.no_chunked_adamw),
                                   per_block_graphs=(layerwise or args.per_block_graphs))
    skipped_blocks = [0]
    # N > 1 with per-block graphs: layer-wise FedAvg on a comm stream, each block's graph of the NEXT round waits only for
    # its own parameters (fedavg.PipelinedFedAvg) -- the aggregation hides behind the forward pass
    pipe = None
    if world > 1 and graphed is not None and graphed.segmented and not args.no_overlap and not args.no_layerwise:
        ch = graphed.fedavg_chunks()   # [non-block parameters, block 0, block 1, ...]; --fedavg-group blocks per exchange
        G = max(1, args.fedavg_group)
        grouped = [ch[0]] + [sum(ch[1 + i:1 + i + G], []) for i in range(0, len(ch) - 1, G)]
        pipe = PipelinedFedAvg([g["fp"] for g in opt.groups], grouped, agg=peer_agg, max_blocks=args.fedavg_blocks or 128)
        pipe_event_of_block = [0] + [1 + i // G for i in range(len(ch) - 1)]   # chunk index per graph chunk
        cp = graphed.fedavg_chunk_params()
        grouped_cp = [cp[0]] + [sum(cp[1 + i:1 + i + G], []) for i in range(0, len(cp) - 1, G)]

        def update_chunk(c):   # the AdamW launch(es) of one exchange chunk, on the main stream
            for gi, lo, hi in grouped_cp[c]:
                opt.adamw_params(gi, lo, hi)

    def local_step(batch):
        if graphed is None:
            loss = eager_step(batch)
        else:
            loss = graphed(batch)
        skipped_blocks[0] += len(model.backbone.last_skipped_layers)
        return loss

    def fedavg_nccl():
        for g in opt.groups:
            fedavg_allreduce(g["fp"].flat)
            g["fp"].mark_updated()

    chosen = {"impl": ("nvls" if (is_symm and peer_agg.has_multicast) else "peer") if use_peer else "nccl"}
    if args.fedavg in ("peer", "nvls") and use_peer:
        chosen["impl"] = args.fedavg if (args.fedavg == "peer" or (is_symm and peer_agg.has_multicast)) else "peer"

    def fedavg_impl(impl):
        if impl == "nccl":                 # our scale kernel + NCCL all-reduce
            fedavg_nccl()
        elif is_symm:                      # "nvls": in-switch reduction + multicast; "peer": fused peer-memory kernel
            peer_agg(impl=impl, max_blocks=(args.nvls_blocks if impl == "nvls" else 0))
        else:                              # fused peer-memory kernel over CUDA IPC buffers
            peer_agg()

    def fedavg():
        if world > 1:
            fedavg_impl(chosen["impl"])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    copy_stream = torch.cuda.Stream()
    comm_stream = torch.cuda.Stream(priority=-1)   # high priority: the exchange kernels' blocks go first whenever an SM frees up

    def timed(nsteps, e2e):
        """CUDA-event time of nsteps steps (max over ranks), barrier + synchronize on both sides."""
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        agg = []
        launches = 0
        skipped_blocks[0] = 0
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

        # loss read-back (e2e): every step's loss goes device -> pinned host inside the timed region, asynchronously; the host
        # looks at it one step later, so that it never stalls the launch of the next step (what a logging training loop does)
        pinned = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)] if e2e else None
        pending = None
        nxt = stage(0) if e2e else None
        pipelined = world > 1 and graphed is not None and graphed.split_conv

        def get_batch(i):
            nonlocal nxt
            if e2e:  # the call a user makes: pinned host batch -> device, update, loss read back
                b, ev = nxt
                main.wait_event(ev)
                if i + 1 < nsteps:
                    nxt = stage(i + 1)
                return b
            return resident[i % len(resident)]

        if pipelined:
            graphed.begin(get_batch(0))
        evs = None
        for i in range(nsteps):
            if pipe is not None:
                # round i: every block's graph waits only for the average of ITS parameters from round i - 1; the
                # aggregation of round i then starts on the comm stream and runs under the forward pass of round i + 1
                loss = graphed.finish(wait_events=evs, adamw=not graphed.defer_adamw)
                skipped_blocks[0] += len(model.backbone.last_skipped_layers)
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record(main)
                ge = pipe.launch(comm_stream, chosen["impl"], update_chunk=update_chunk if graphed.defer_adamw else None)
                evs = [ge[j] for j in pipe_event_of_block]   # one event per graph chunk (blocks of a group share theirs)
                a1.record(comm_stream)
                agg.append((a0, a1))
                if i + 1 < nsteps:
                    graphed.begin(get_batch(i + 1))
            elif pipelined:
                # round i: the weight-dependent part of the update, then the aggregation on the comm stream while the main
                # stream already runs the frozen conv stack of round i + 1; the main stream rejoins before it reads weights
                loss = graphed.finish()
                skipped_blocks[0] += len(model.backbone.last_skipped_layers)
                comm_stream.wait_stream(main)
                with torch.cuda.stream(comm_stream):
                    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a0.record(comm_stream); fedavg(); a1.record(comm_stream)
                agg.append((a0, a1))
                if i + 1 < nsteps:
                    graphed.begin(get_batch(i + 1))
                main.wait_stream(comm_stream)
            else:
                loss = local_step(get_batch(i))
                if world > 1:
                    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a0.record(); fedavg(); a1.record()
                    agg.append((a0, a1))
            if graphed is not None:
                launches += graphed.launches_last_replay if (graphed.segmented or graphed.split_conv) else graphed.launches_per_replay
            if e2e:
                pinned[i & 1].copy_(loss.detach().reshape(1), non_blocking=True)   # device -> host read of the step's result
                ev = torch.cuda.Event()
                ev.record()
                if pending is not None:
                    pending[0].synchronize()
                    last = float(pending[1])
                pending = (ev, pinned[i & 1])
        if e2e and pending is not None:
            pending[0].synchronize()
            last = float(pending[1])
        main.wait_stream(comm_stream)   # the last round's aggregation is part of the timed region
        t1.record()
        barrier()
        if peer_agg is not None:
            peer_agg.check()   # a barrier that timed out inside the timed region voids the round: raise, do not report
        ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
        agg_ms = torch.tensor([sum(a.elapsed_time(b) for a, b in agg) / max(1, len(agg))], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(agg_ms, op=dist.ReduceOp.MAX)
        return float(ms) / nsteps, float(agg_ms), last, launches, skipped_blocks[0]

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
            agg_alone["peer_kernel_ms"] = time_alone(lambda: fedavg_impl("peer"))
            if is_symm and peer_agg.has_multicast:
                agg_alone["nvls_kernel_ms"] = time_alone(lambda: fedavg_impl("nvls"))
            peer_agg.check()
            if args.fedavg == "auto":
                cand = {"nccl": agg_alone["nccl_allreduce_ms"], "peer": agg_alone["peer_kernel_ms"]}
                if "nvls_kernel_ms" in agg_alone:
                    cand["nvls"] = agg_alone["nvls_kernel_ms"]
                chosen["impl"] = min(cand, key=cand.get)
    if pipe is not None:
        # grid cap of the overlapped exchange kernels, measured per implementation (N = 2 / 4 / 8 sweeps in profiles/
        # r02_scaling.txt): the in-switch reduction is link bound with 32 CTAs and every further CTA only takes SM time from
        # the forward pass; the peer kernel needs 128 CTAs of P2P loads in flight
        if args.fedavg_blocks <= 0:
            args.fedavg_blocks = 32 if chosen["impl"] == "nvls" else 128
        pipe.max_blocks = args.fedavg_blocks
    clocks = ClockSampler(local_rank) if rank == 0 else None
    n0 = _lib.launch_count()
    if args.profile_range:   # `ncu --profile-from-start off`: exactly the timed steps are captured
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    ms_step, agg_ms, _, launches, skipped = timed(args.steps, e2e=False)
    if args.profile_range:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    if graphed is None:
        launches = _lib.launch_count() - n0
    clk = clocks.stop() if clocks else None
    ms_e2e, _, last_loss, _, skipped_e2e = timed(args.steps, e2e=True)

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
    # the eager path is host bound (Python enqueues a step more slowly than the GPU executes it), which would put the host's
    # time between "record start event" and "launch" into every bracket: park the GPU on a 60 ms spin first, so that the
    # whole step is queued before its first kernel runs and the brackets hold device time only
    # ... and keep the weight-gradient GEMMs on the main stream for this one step: a bracket around a kernel that shares the
    # SMs with a kernel of the other stream would count the shared time twice
    side_was, ops.SIDE_STREAM_ENABLED = ops.SIDE_STREAM_ENABLED, False
    torch.cuda._sleep(int(0.060 * 1.9e9))
    eager_step(resident[0])
    torch.cuda.synchronize()
    ops.SIDE_STREAM_ENABLED = side_was
    prof, ops.PROFILE = ops.PROFILE, None
    kinds = {}
    for kind, e0, e1, fl in prof:
        k = kinds.setdefault(kind, [0, 0.0, 0.0])
        k[0] += 1; k[1] += e0.elapsed_time(e1); k[2] += fl
    peaks, src = measured_peaks()
    gm = kinds.get("gemm", [0, 1e-9, 0.0])
    achieved = gm[2] / (gm[1] * 1e-3) / 1e12
    peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0)))
    # per-launch DRAM traffic of this kernel: ncu capture of THIS build (the JSON records the sha of the GEMM source it was
    # taken from; a capture of another build is not reported)
    traffic, traffic_src = None, "no ncu capture of this build (profiles/r02_ncu_gemm_traffic.json absent or from another source sha)"
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_gemm_traffic.json")))
        sha = hashlib.sha256(open(os.path.join(PKG, "csrc", "gemm_tcgen05.cu"), "rb").read()).hexdigest()[:16]
        if tj.get("gemm_source_sha16") == sha:
            traffic = tj["dram_bytes_per_launch"]
            traffic_src = ("profiles/r02_ncu_gemm_traffic.json (ncu dram__bytes_read.sum + dram__bytes_write.sum, mean over one "
                           f"step's launches, GEMM source sha {sha})")
    except Exception:
        pass
    roofline = {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05 bf16 GEMM: QKV/O/FFN linears fwd+dgrad+wgrad, conv layers 1-6)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": traffic_src,
                "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({src})", "launches": gm[0],
                "avg_launch_us": gm[1] * 1e3 / max(1, gm[0]),
                "other": {k: {"launches": v[0], "ms": v[1], "tflops": v[2] / (v[1] * 1e-3) / 1e12} for k, v in kinds.items()
                          if k != "gemm"}}

    if rank == 0:
        audio_s = BATCH * SECONDS * world
        # algorithmic work: 8.376 TFLOP with all 24 blocks; a block = 3 x its forward linears + 3.5 x attention
        blk = (3 * 2 * frames * BATCH * (4 * 1024 ** 2 + 2 * 1024 * 4096) + 3.5 * 4 * BATCH * 16 * frames ** 2 * 64) * 1e-12
        frac_run = 1.0 - skipped / (args.steps * N_LAYERS)
        algo_tflop = 8.376 - N_LAYERS * blk * (1.0 - frac_run)
        line = {
            "metric": METRIC, "value": audio_s / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "global_batch": BATCH * world, "audio_s_per_step": audio_s,
                       "clients": world, "trainable_params": n_train, "frames_per_utt": frames,
                       "regularisation": (f"the reference's training configuration (Data2VecAudioConfig defaults kept by finetune_ASRs.py:160-164): "
                                          f"hidden / activation / attention / final dropout {pd:g} (masks drawn on the device, fresh per "
                                          f"replay), feat_proj_dropout 0, layerdrop {ld:g} (host coin per block per step: "
                                          f"{skipped} of {args.steps * N_LAYERS} blocks skipped in the timed steps)"),
                       "l2_policy": "inputs+weights+saved activations per step (>3 GB) exceed the 126 MB L2; no flush",
                       "e2e_input_pipeline": "each step's pinned-host -> device copy runs on a copy stream, issued while the previous step computes; every step's loss is copied to pinned host memory inside the timed region and read by the host one step later",
                       "fedavg_impl": chosen["impl"] if world > 1 else None,
                       "fedavg_buffers": symm_note if world > 1 else None,
                       "fedavg_overlap": (("layer-wise pipelined: chunk = one encoder block, exchange + barrier + bf16 shadow refresh per chunk "
                                           f"on a comm stream (exchange grid capped at {args.fedavg_blocks} blocks); every block's CUDA graph "
                                           "of the next round waits only for its own chunk, so the aggregation runs under the next forward "
                                           "pass" + ("; AdamW is launched chunk by chunk too and chunk c's exchange starts while chunk c + 1 "
                                                     "is still being updated" if (graphed is not None and graphed.defer_adamw) else "")
                                           + "; fedavg_agg_ms is the span of the aggregation, not its exposed part") if pipe is not None else
                                          ("the aggregation runs on a comm stream while the main stream replays the next round's frozen "
                                           "conv-stack graph" if (world > 1 and graphed is not None and graphed.split_conv) else None)),
                       "fedavg": ({"peer": "uniform FedAvg of all trainable params after every local step: fused peer-memory kernel "
                                           "(scale + reduce-scatter + all-gather over NVLink, no NCCL on the data path)",
                                   "nvls": "uniform FedAvg of all trainable params after every local step: own NVLS kernel (multimem.ld_reduce "
                                           "in the NVSwitch + multimem.st multicast of the average, 4P/N bytes each way per GPU)",
                                   "nccl": "uniform FedAvg of all trainable params after every local step (scale kernel + NCCL all-reduce)"}
                                  [chosen["impl"]]) if world > 1 else "n/a at 1 GPU"},
            "e2e": {"value": audio_s / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": sum(v.numel() * v.element_size() for v in host[0].values()),
                    "d2h_bytes_per_step": 4, "last_loss": last_loss, "blocks_skipped": skipped_e2e},
            "gpu_launches": int(launches),
            "gpu_launches_per_step": int(launches) / args.steps,
            "host_enqueue_ms_per_step": host_ms, "host_enqueue_ms_per_step_eager": host_eager_ms,
            "step_mode": ("cuda_graph" + (" (one graph per block and direction: LayerDrop)" if graphed.segmented else ""))
                         if graphed is not None else "eager",
            "step_tflops": algo_tflop * world / (ms_step * 1e-3),
            "step_frac_of_tensor_peak": algo_tflop / (ms_step * 1e-3) / peak,
            "algorithmic_tflop_per_step": algo_tflop,
            "fedavg_round_s": ms_step * 1e-3, "fedavg_agg_ms": agg_ms if world > 1 else None,
            "fedavg_agg_alone": agg_alone or None,  # each implementation timed back to back without the local step
            # per GPU and direction: its own remote reads + the peers' writes into it = 2 * 4P (N-1)/N bytes
            "fedavg_nvlink_gbs_per_dir": (2 * 4 * n_train * (world - 1) / world / (agg_alone["peer_kernel_ms"] * 1e-3) / 1e9)
                                         if agg_alone.get("peer_kernel_ms") else None,
            "fedavg_bytes_per_client": 4 * n_train,
            "clocks": clk, "roofline": roofline,
        }
        if world == 1 and not args.no_extras:
            del graphed
            model.to("cpu")
            torch.cuda.empty_cache()
            # "the real bar" (BASELINE.md 3): the same model through stock PyTorch kernels on this GPU
            try:
                tb = oracle_step_time("cuda", BATCH, SECONDS, steps=6, warmup=2, pd=pd, ld=ld, autocast=True)
                line["torch_gpu_baseline"] = dict(
                    value=tb["value"], unit=UNIT, ms_per_step=tb["ms_per_step"], ratio=tb["ms_per_step"] / ms_step,
                    what=("oracle port (HF Data2VecAudioForCTC arithmetic on stock torch ops: cuDNN conv, cuBLAS, eager softmax "
                          "attention, F.ctc_loss) under torch.autocast(bfloat16) + clip_grad_norm_ + torch.optim.AdamW(fused=True), same "
                          "batch 8 x 10 s, same dropouts and LayerDrop coin stream, CUDA events, 2 warm-up + median of 6 steps"),
                    last_loss=tb["last_loss"])
            except Exception as e:  # the baseline leg must never take the product's number down with it
                line["torch_gpu_baseline"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
            torch.cuda.empty_cache()
            try:
                line["other_configs"] = other_configs(dev)
            except Exception as e:
                line["other_configs"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_reference(steps=2, warmup=1, batch=2, seconds=SECONDS, pd=pd, ld=ld)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        args._emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the torch_gpu_baseline and other_configs legs (profiling runs)")
    ap.add_argument("--fedavg", default="auto", choices=["auto", "peer", "nccl", "nvls"],
                    help="N>1 aggregation: own NVLS kernel (in-switch reduction), fused peer-memory kernel, scale kernel + NCCL "
                         "all-reduce, or whichever measures fastest in warm-up")
    ap.add_argument("--no-symm", action="store_true", help="do not try torch symmetric memory for the flat buffers (CUDA IPC + peer kernel)")
    ap.add_argument("--no-overlap", action="store_true", help="N>1: do not overlap the aggregation with the next round's conv stack")
    ap.add_argument("--nvls-blocks", type=int, default=0, help="cap of the NVLS kernel's grid (0 = full grid)")
    ap.add_argument("--per-block-graphs", action="store_true",
                    help="cut the captured update at the block boundaries even with layerdrop 0 (measures what the cuts cost)")
    ap.add_argument("--profile-range", action="store_true", help="cudaProfilerStart/Stop around the timed steps (ncu --profile-from-start off)")
    ap.add_argument("--no-chunked-adamw", action="store_true", help="layer-wise FedAvg: keep AdamW inside the last graph instead of chunk by chunk")
    ap.add_argument("--no-layerwise", action="store_true", help="N>1: aggregate the whole buffer at once instead of block by block")
    ap.add_argument("--fedavg-blocks", type=int, default=0,
                    help="grid cap of the exchange kernels in the layer-wise pipelined FedAvg (0 = 32 for nvls, 128 for peer / nccl)")
    ap.add_argument("--fedavg-group", type=int, default=2, help="encoder blocks per exchange of the layer-wise pipelined FedAvg")
    ap.add_argument("--dropout", type=float, default=0.1,
                    help="hidden = activation = attention = final dropout P in the timed training step (0.1 = the "
                         "Data2VecAudioConfig defaults the reference fine-tunes with; 0 = the parity setting)")
    ap.add_argument("--layerdrop", type=float, default=0.1, help="LayerDrop probability (0.1 = the reference's default)")
    ap.add_argument("--ref-batch", type=int, default=BATCH, help="--impl reference: utterances per step (8 = the metric's config)")
    ap.add_argument("--ref-seconds", type=float, default=SECONDS)
    ap.add_argument("--eager", action="store_true", help="issue every launch from Python instead of replaying the CUDA graph")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON: anything a library prints on the way (NCCL's version banner, warnings) is
    # sent to stderr by pointing fd 1 there until the line is ready
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    args._emit = lambda text: (sys.stdout.flush(), os.dup2(real_stdout, 1), print(text, flush=True))
    if args.impl == "reference":
        run_reference(args)
    else:
        if args.warmup < 3:
            args.warmup = 3
        run_b200(args)


if __name__ == "__main__":
    main()
