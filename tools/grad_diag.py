#!/usr/bin/env python
"""Diagnostic: every gradient of the full-size ragged C2 batch, CUDA path (fp32 / bf16) vs the oracle port on the same GPU
(fp32, TF32 off) -- worst parameters first, plus d loss / d logits per utterance."""
import os
import sys
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "privacy-preserve-federated-asr_b200"), os.path.join(ROOT, "tests")]
import torch  # noqa: E402

from b200asr import modeling as M  # noqa: E402
from oracle import data2vec_ref as R  # noqa: E402
import large_cases as LC  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
DEV = "cuda"
name = sys.argv[1] if len(sys.argv) > 1 else "large_c2_train"
dtype = torch.bfloat16 if (len(sys.argv) > 2 and sys.argv[2] == "bf16") else torch.float32
g, cfg, sd, b = LC.setup(os.path.join(ROOT, "tests", "golden"), name)
x, am, lab = b["input_values"].to(DEV), b["attention_mask"].to(DEV), b["labels"].to(DEV)
print("lengths", b["lengths"].tolist(), "frames", R.conv_out_lengths(cfg, b["lengths"]).tolist(), "labels", (b["labels"] >= 0).sum(-1).tolist())
fe = cfg.prefix + "feature_extractor"
sdg = {k: v.to(DEV).requires_grad_(not k.startswith(fe)) for k, v in sd.items()}
col = []
loss_ref, logits_ref = R.ctc_forward(sdg, cfg, x, am.long(), lab, collect=col)
logits_ref.retain_grad()
for h in col:
    h.retain_grad()
loss_ref.backward()
ns = SimpleNamespace(**{k: (tuple(v) if isinstance(v, (list, tuple)) else v) for k, v in cfg.__dict__.items()}) \
    if cfg.model_type == "data2vec" else SimpleNamespace(**cfg.to_hf_kwargs())
cls = {"data2vec": M.Data2VecAudioForCTC, "wav2vec2": M.Wav2Vec2ForCTC, "hubert": M.HubertForCTC}[cfg.model_type]
model = cls(ns, compute_dtype=dtype)
model.load_state_dict(sd)
model.freeze_feature_encoder()
model.to(DEV).train()
out = model(input_values=x, attention_mask=am, labels=lab, output_hidden_states=(len(sys.argv) > 3))
out.logits.retain_grad()
out.loss.backward()
torch.cuda.synchronize()


def rel(a, r):
    return float((a.float() - r.float()).abs().max() / max(float(r.float().abs().max()), 1e-30))


print("loss", float(out.loss), float(loss_ref), "logits rel", rel(out.logits, logits_ref.detach()))
dl, dr = out.logits.grad.float(), logits_ref.grad
for i in range(dl.shape[0]):
    print(f"  dlogits utt {i}: rel {rel(dl[i], dr[i]):.3e}  max|ref| {float(dr[i].abs().max()):.3e}  colsum ours {dl[i].sum(0)[:4].tolist()} ref {dr[i].sum(0)[:4].tolist()}")
params = dict(model.named_parameters())
rows = []
for k, v in sdg.items():
    if v.grad is None or params[k].grad is None:
        continue
    rows.append((rel(params[k].grad, v.grad), abs(float(params[k].grad.float().norm()) - float(v.grad.norm())) / max(float(v.grad.norm()), 1e-30), k))
rows.sort(reverse=True)
for e, n, k in rows[:40]:
    print(f"  {e:.3e} (norm {n:.3e})  {k}")
print("  ... best:", rows[-1])
