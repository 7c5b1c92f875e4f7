#!/usr/bin/env python
"""Full-size timing of the BASELINE.json configurations that bench.py does not headline (they are parity-test cases
there): C1 greedy CTC inference, C3 DACS stage-2 step, C5 wav2vec2-large / hubert-large steps at batch 16 x 30 s.
Synthetic 16 kHz audio, random-init weights, bf16 compute, CUDA events after warm-up; one JSON line per configuration.

  python tools/config_bench.py [--only c1,c3,c5w,c5h] [--steps 5] [--warmup 3]
"""
import argparse
import json
import os
import sys
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "privacy-preserve-federated-asr_b200")]
import torch  # noqa: E402

from b200asr import modeling as M  # noqa: E402
from b200asr.optim import FusedAdamW  # noqa: E402
from bench import large_config, synth_batch  # noqa: E402

SR = 16000
dev = torch.device("cuda", 0)


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        fn()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / steps


def batch_of(cfg, B, seconds, seed=1234):
    frames = int(M.conv_out_lengths(cfg, torch.tensor(seconds * SR)))
    hb = synth_batch(B, seconds * SR, seed, frames)
    hb["dementia_labels"] = (torch.rand(B, generator=torch.Generator().manual_seed(seed + 3)) < 0.5).long()
    return {k: v.to(dev) for k, v in hb.items()}, frames


def train_step_fn(model, batch, extra=()):
    opt = FusedAdamW(model.named_parameters(), lr=1e-5, weight_decay=0.005, max_grad_norm=1.0)
    kw = {k: batch[k] for k in extra}

    def step():
        opt.zero_grad()
        out = model(batch["input_values"], batch["attention_mask"], labels=batch["labels"], **kw)
        out.loss.backward()
        opt.step()
    return step, sum(g["fp"].numel for g in opt.groups)


def emit(name, workload, ms, audio_s, **kw):
    print(json.dumps(dict(config=name, workload=workload, ms_per_step=ms, audio_s_per_s=audio_s / (ms * 1e-3), dtype="bf16",
                          data="synthetic", **kw)), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="c1,c3,c5w,c5h")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    which = set(a.only.split(","))
    torch.manual_seed(0)

    if "c1" in which:  # configs[0]: greedy CTC inference, batch 1 x 5 s
        cfg = large_config()
        model = M.Data2VecAudioForCTC(cfg, compute_dtype=torch.bfloat16).to(dev).eval()
        b, frames = batch_of(cfg, 1, 5)

        @torch.no_grad()
        def infer():
            out = model(b["input_values"])
            return model.greedy_decode(out.logits)
        ms = timed(infer, a.steps * 4, a.warmup)
        emit("C1", "data2vec-audio-large greedy CTC inference, batch 1 x 5 s (forward + argmax/collapse on device + id lists on host)",
             ms, 5.0, frames=frames)
        del model

    if "c3" in which:  # configs[2]: DACS stage 2, batch 8 x 10 s
        cfg = large_config()
        args = SimpleNamespace(LAMBDA=0.5, TOGGLE_RATIO=0, GS_TAU=1, AD_loss="recall", W_LOSS=None, STAGE=2)
        model = M.DacsData2VecAudioForCTC(cfg, args, compute_dtype=torch.bfloat16)
        model.validate_labels = False
        model.to(dev).train()
        b, frames = batch_of(cfg, 8, 10)
        step, ntrain = train_step_fn(model, b, extra=("dementia_labels",))
        ms = timed(step, a.steps, a.warmup)
        emit("C3", "DACS stage-2 toggling network (gumbel_softmax GS_TAU=1, GRL AD head, 2 x CTC, CosFace), batch 8 x 10 s, "
                   "frozen encoder forward + arbitrator update", ms, 80.0, trainable_params=ntrain, frames=frames, mode="eager")
        from b200asr.graph import GraphedTrainStep
        opt = FusedAdamW(model.named_parameters(), lr=1e-5, weight_decay=0.005, max_grad_norm=1.0)
        gb = {k: b[k] for k in ("input_values", "attention_mask", "labels", "dementia_labels")}
        gstep = GraphedTrainStep(model, opt, gb, max_label_len=120)
        c0 = int(model.gumbel_counter)
        ms = timed(lambda: gstep(gb), a.steps * 2, a.warmup)
        assert int(model.gumbel_counter) == c0 + a.steps * 2 + a.warmup  # every replay advanced the noise counter
        emit("C3", "same, the whole update replayed as one CUDA graph (Gumbel noise offset read from a device counter)",
             ms, 80.0, trainable_params=ntrain, frames=frames, mode="cuda_graph")
        del model, step, gstep

    for key, cls, name, kw in (
            ("c5w", M.Wav2Vec2ForCTC, "wav2vec2-large-960h-lv60 arch (layer-norm conv stack + bias, stable LayerNorm, weight-normed pos conv k=128)",
             dict(feat_extract_norm="layer", conv_bias=True, do_stable_layer_norm=True, num_conv_pos_embeddings=128)),
            ("c5h", M.HubertForCTC, "hubert-large-ls960-ft arch (layer-norm conv stack + bias, stable LayerNorm, weight-normed pos conv k=128)",
             dict(feat_extract_norm="layer", conv_bias=True, do_stable_layer_norm=True, num_conv_pos_embeddings=128,
                  feat_proj_layer_norm=True))):
        if key not in which:
            continue
        cfg = large_config(**kw)
        model = cls(cfg, compute_dtype=torch.bfloat16)
        model.freeze_feature_encoder()
        model.validate_labels = False
        model.to(dev).train()
        b, frames = batch_of(cfg, 16, 30)
        step, ntrain = train_step_fn(model, b)
        ms = timed(step, a.steps, a.warmup)
        emit("C5/" + key[2:], name + ", CTC fine-tune step, batch 16 x 30 s", ms, 480.0, trainable_params=ntrain, frames=frames,
             peak_mem_gb=torch.cuda.max_memory_allocated() / 2**30)
        del model, step
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
