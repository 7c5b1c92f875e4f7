#!/usr/bin/env python
"""One eager client-local update of the full-size model (BASELINE configs[1]: batch 8 x 10 s, bf16, dropouts 0.1 incl.
attention, layerdrop 0 so that every block runs) between cudaProfilerStart / Stop, after two warm-up updates -- the
command behind the kernel-filtered ncu captures (`ncu --profile-from-start off -k regex:...`)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "privacy-preserve-federated-asr_b200")]
import torch  # noqa: E402

from b200asr import modeling as M  # noqa: E402
from b200asr.optim import FusedAdamW  # noqa: E402
from bench import BATCH, SECONDS, SR, large_config, regularisation, synth_batch  # noqa: E402

dev = torch.device("cuda", 0)
cfg = large_config(**regularisation(0.1, 0.0))
torch.manual_seed(0)
model = M.Data2VecAudioForCTC(cfg, compute_dtype=torch.bfloat16)
model.freeze_feature_encoder()
model.validate_labels = False
model.to(dev).train()
opt = FusedAdamW(model.named_parameters(), lr=1e-5, weight_decay=0.005, max_grad_norm=1.0)
frames = int(M.conv_out_lengths(cfg, torch.tensor(SECONDS * SR)))
b = {k: v.to(dev) for k, v in synth_batch(BATCH, SECONDS * SR, 1234, frames, pin=False).items()}


def step():
    opt.zero_grad()
    out = model(b["input_values"], b["attention_mask"], labels=b["labels"])
    out.loss.backward()
    opt.step()
    return out.loss


for _ in range(2):
    step()
torch.cuda.synchronize()
torch.cuda.profiler.start()
loss = step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok", float(loss))
