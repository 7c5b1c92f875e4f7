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
loss_ref.backward(retain_graph=True)
ns = SimpleNamespace(**{k: (tuple(v) if isinstance(v, (list, tuple)) else v) for k, v in cfg.__dict__.items()}) \
    if cfg.model_type == "data2vec" else SimpleNamespace(**cfg.to_hf_kwargs())
cls = {"data2vec": M.Data2VecAudioForCTC, "wav2vec2": M.Wav2Vec2ForCTC, "hubert": M.HubertForCTC}[cfg.model_type]
model = cls(ns, compute_dtype=dtype)
model.load_state_dict(sd)
model.freeze_feature_encoder()
model.to(DEV).train()
out = model(input_values=x, attention_mask=am, labels=lab, output_hidden_states=True)
out.logits.retain_grad()
for h in out.hidden_states:
    h.retain_grad()
out.loss.backward()
torch.cuda.synchronize()


def l2(a, r):
    a, r = a.double(), r.double()
    return float((a - r).norm() / max(float(r.norm()), 1e-30))


# the oracle's CTC on OUR logits: separates the loss kernel (bf16 in / out) from the deviation of the forward pass
lg = out.logits.detach().float().requires_grad_(True)
lens = R.conv_out_lengths(cfg, b["lengths"]).to(DEV)
lp = torch.log_softmax(lg, -1).transpose(0, 1)
tl = (lab >= 0).sum(-1)
tg = torch.cat([lab[i, : int(tl[i])] for i in range(lab.shape[0])])
l_own = torch.nn.functional.ctc_loss(lp, tg, lens, tl, blank=cfg.pad_token_id, reduction="sum", zero_infinity=False)
l_own.backward()
print("CTC on our logits (fp32 torch): loss", float(l_own), "ours", float(out.loss), " dlogits l2 (ours vs torch on the same logits)",
      l2(out.logits.grad, lg.grad), " dlogits l2 (ours vs oracle end to end)", l2(out.logits.grad, logits_ref.grad))
# the oracle's fp32 backward pass driven by OUR d loss / d logits: separates the backward kernels from the effect of the
# forward deviation on the CTC posteriors
g_ref = {k: v.grad.clone() for k, v in sdg.items() if v.grad is not None}
for v in sdg.values():
    v.grad = None
logits_ref.backward(out.logits.grad.float())
import numpy as np
pm = dict(model.named_parameters())
rm, em = [], []
for k, v in sdg.items():
    if v.grad is None or pm[k].grad is None or k.endswith("k_proj.bias"):
        continue
    rm.append(float(v.grad.norm()) / float(g_ref[k].norm()))
    em.append(l2(pm[k].grad, v.grad))
qq = lambda a: [round(float(np.quantile(a, t)), 4) for t in (0.1, 0.5, 0.9)]
print("oracle backward of OUR dlogits: |g_mixed| / |g_oracle| (10 / 50 / 90 %)", qq(rm), "  l2 error of our g against g_mixed", qq(em))
for k, v in sdg.items():
    if k in g_ref:
        v.grad = g_ref[k]
if dtype == torch.bfloat16:   # the yardstick: the oracle port under stock bf16 autocast
    sdy = {k: v.detach().clone().requires_grad_(v.requires_grad) for k, v in sdg.items()}
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss_y, logits_y = R.ctc_forward(sdy, cfg, x, am.long(), lab)
    logits_y.retain_grad()
    loss_y.backward()
    print("stock bf16 autocast: loss", float(loss_y), " logits l2 vs oracle", l2(logits_y.detach(), logits_ref.detach()),
          " dlogits l2 vs oracle", l2(logits_y.grad, logits_ref.grad), " ours vs stock: logits", l2(out.logits.detach(), logits_y.detach()),
          "dlogits", l2(out.logits.grad, logits_y.grad))
    print("  |dlogits|: ours", float(out.logits.grad.float().norm()), "stock", float(logits_y.grad.float().norm()), "oracle", float(logits_ref.grad.norm()))
    import numpy as np
    pm = dict(model.named_parameters())
    ro, ry, eo, ey = [], [], [], []
    for k, v in sdg.items():
        if v.grad is None or pm[k].grad is None or k.endswith("k_proj.bias"):
            continue
        n = float(v.grad.norm())
        ro.append(float(pm[k].grad.float().norm()) / n); ry.append(float(sdy[k].grad.float().norm()) / n)
        eo.append(l2(pm[k].grad, v.grad)); ey.append(l2(sdy[k].grad, v.grad))
    q = lambda a: [round(float(np.quantile(a, t)), 4) for t in (0.1, 0.5, 0.9)]
    print("  per-parameter |g| / |g_oracle| (10 / 50 / 90 %): ours", q(ro), "stock", q(ry))
    print("  per-parameter l2 error of g (10 / 50 / 90 %): ours", q(eo), "stock", q(ey))
if len(col) == len(out.hidden_states) and all(h.grad is not None for h in out.hidden_states):
    for i in range(0, len(col), max(1, len(col) // 12)):
        print(f"  d loss / d hidden[{i}]: l2 {l2(out.hidden_states[i].grad, col[i].grad):.3e}   hidden l2 {l2(out.hidden_states[i].detach(), col[i].detach()):.3e}"
              f"   |h| rms {float(col[i].detach().pow(2).mean().sqrt()):.3e}  |dh| rms {float(col[i].grad.pow(2).mean().sqrt()):.3e}")
else:
    print("hidden state lists differ:", len(col), len(out.hidden_states))


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
