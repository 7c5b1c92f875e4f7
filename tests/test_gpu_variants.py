"""GPU parity of the Wav2Vec2 / Hubert members of the family (BASELINE configs[4]; centralized/finetune_ASRs.py:154-172)
against fixtures produced by HF Wav2Vec2ForCTC / HubertForCTC (tests/golden/make_golden.py::VARIANTS) and the oracle:
GroupNorm conv stack, weight-normed positional conv (even kernel), pre-LN blocks + final LayerNorm."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from b200asr import modeling as M  # noqa: E402
from b200asr import ops  # noqa: E402
from b200asr.optim import FusedAdamW  # noqa: E402
from oracle import data2vec_ref as R  # noqa: E402

from test_oracle_golden import VARIANTS, variant_cfg  # noqa: E402

DEV = "cuda"
TOL = {torch.float32: 1e-3, torch.bfloat16: 2e-2}  # BASELINE.json: loss / logits relative tolerance
MODEL = {"wav2vec2": M.Wav2Vec2ForCTC, "hubert": M.HubertForCTC}


def hf_ns(cfg):
    """The HF-style config object the reference would hand to the model (num_conv_pos_embeddings = kernel size)."""
    return SimpleNamespace(**cfg.to_hf_kwargs())


def relinf(a, b):
    a = torch.as_tensor(a).float().cpu()
    b = torch.as_tensor(b).float().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-6))


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("name", list(VARIANTS))
def test_variant_ctc_vs_hf_golden(golden_dir, name, dtype):
    g = load(golden_dir, name)
    cfg = variant_cfg(name)
    seed, batch, samples, use_mask = int(g["seed"]), int(g["batch"]), int(g["samples"]), bool(g["use_mask"])
    sd = R.init_state_dict(cfg, seed=seed, dacs=False)
    model = MODEL[cfg.model_type](hf_ns(cfg), compute_dtype=dtype)
    model.load_state_dict(sd, strict=True)
    model.freeze_feature_encoder()
    model.to(DEV).train()
    b = R.synth_batch(cfg, batch, samples, seed=1234 + seed, ragged=use_mask, max_label=12)
    out = model(input_values=b["input_values"].to(DEV), attention_mask=b["attention_mask"].to(DEV) if use_mask else None,
                labels=b["labels"].to(DEV), output_hidden_states=True)
    out.loss.backward()
    tol = TOL[dtype]
    assert len(out.hidden_states) == cfg.num_hidden_layers + 1
    assert relinf(out.hidden_states[0], g["hidden_first"]) < tol
    assert relinf(out.hidden_states[-1], g["hidden_last"]) < tol
    assert relinf(out.logits, g["logits"]) < tol
    assert abs(float(out.loss) - float(g["loss"])) / abs(float(g["loss"])) < tol
    if dtype == torch.float32:
        ids, seqs = model.greedy_decode(out.logits.detach())
        assert (ids.cpu().numpy() == g["argmax"]).all()
        _, ref_seqs = R.greedy_ids(torch.from_numpy(g["logits"]))
        assert seqs == ref_seqs
    params = dict(model.named_parameters())
    gt = 2e-3 if dtype == torch.float32 else 6e-2
    for n in g["grad_names"]:
        assert relinf(params[str(n)].grad, g["grad:" + str(n)]) < gt, n
    have = {n for n, p in params.items() if p.grad is not None}
    assert have == {str(n) for n in g["gnorm_names"]}
    for n, v in zip(g["gnorm_names"], g["gnorm_values"]):
        if str(n).endswith("k_proj.bias"):
            continue  # zero in exact arithmetic (softmax shift invariance)
        got = float(params[str(n)].grad.float().norm())
        assert abs(got - v) <= gt * max(v, 1e-3), (n, got, v)


@pytest.mark.parametrize("name", ["w2v2_group_postln", "hubert_stable"])
def test_variant_loads_new_weight_norm_keys_and_trains_with_flat_params(name):
    """state_dicts saved by newer transformers (parametrizations.weight.original0/1) load; gradients delivered straight
    into the FlatParams sinks equal the autograd-delivered ones."""
    cfg = variant_cfg(name)
    sd = R.init_state_dict(cfg, seed=3, dacs=False)
    new_sd = {k.replace("conv.weight_g", "conv.parametrizations.weight.original0")
               .replace("conv.weight_v", "conv.parametrizations.weight.original1"): v for k, v in sd.items()}
    b = R.synth_batch(cfg, 2, 6000, seed=99, ragged=True, max_label=12)
    grads = []
    for flat in (False, True):
        model = MODEL[cfg.model_type](hf_ns(cfg), compute_dtype=torch.float32)
        model.load_state_dict(new_sd, strict=True)
        model.freeze_feature_encoder()
        model.to(DEV).train()
        if flat:
            opt = FusedAdamW(model.named_parameters(), lr=0.0, weight_decay=0.0, max_grad_norm=None)
            opt.zero_grad()
        out = model(b["input_values"].to(DEV), b["attention_mask"].to(DEV), labels=b["labels"].to(DEV))
        out.loss.backward()
        torch.cuda.synchronize()
        grads.append({n: p.grad.detach().float().clone() for n, p in model.named_parameters() if p.grad is not None})
    assert grads[0].keys() == grads[1].keys()
    for n in grads[0]:
        assert relinf(grads[1][n], grads[0][n]) < 1e-5, n


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,L,C,bias", [(2, 4000, 64, False), (3, 16000, 512, True), (1, 405, 128, False)])
def test_conv0_groupnorm_gelu(B, L, C, bias, dtype):
    g = torch.Generator().manual_seed(L + C)
    x = torch.randn(B, L, generator=g)
    x[-1, L // 2:] = 0.0  # a padded tail takes part in the statistics, as in torch
    w = torch.randn(C, 1, 10, generator=g) * 0.45
    bs = torch.randn(C, generator=g) * 0.1 if bias else None
    gw, gb = 1 + 0.1 * torch.randn(C, generator=g), 0.1 * torch.randn(C, generator=g)
    ref = F.gelu(F.group_norm(F.conv1d(x[:, None], w, bs, stride=5), C, gw, gb, 1e-5)).transpose(1, 2)
    y = ops.conv0_gn_gelu(x.to(DEV), w.to(DEV), None if bs is None else bs.to(DEV), gw.to(DEV), gb.to(DEV), 5, 1e-5, dtype)
    assert y.shape == ref.shape
    assert relinf(y, ref) < (2e-5 if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("rows,k", [(128 * 64, 16), (1024 * 64, 128), (77, 5)])
def test_weight_norm_fwd_bwd(rows, k):
    g = torch.Generator().manual_seed(rows + k)
    v = (torch.randn(rows, 1, k, generator=g) * 0.02).requires_grad_()
    gg = (torch.rand(1, 1, k, generator=g) + 0.5).requires_grad_()
    w_ref = gg * v / v.norm(dim=(0, 1), keepdim=True)
    dw = torch.randn(rows, 1, k, generator=g)
    w_ref.backward(dw)
    w, norm = ops.weight_norm_fwd(v.detach().to(DEV), gg.detach().to(DEV))
    assert relinf(w, w_ref.detach()) < 1e-5
    assert relinf(norm, v.detach().norm(dim=(0, 1))) < 1e-5
    dv, dg = ops.weight_norm_bwd(dw.to(DEV), v.detach().to(DEV), gg.detach().to(DEV), norm)
    assert relinf(dv, v.grad) < 2e-5
    assert relinf(dg, gg.grad) < 2e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gelu_add_fwd_bwd(dtype):
    g = torch.Generator().manual_seed(11)
    c = (torch.randn(37, 256, generator=g) * 2).to(dtype)
    r = torch.randn(37, 256, generator=g).to(dtype)
    dy = torch.randn(37, 256, generator=g).to(dtype)
    cf = c.float().requires_grad_()
    ref = r.float() + F.gelu(cf)
    ref.backward(dy.float())
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert relinf(ops.gelu_add(c.to(DEV), r.to(DEV)), ref.detach()) < tol
    assert relinf(ops.gelu_add(c.to(DEV)), F.gelu(c.float())) < tol
    assert relinf(ops.gelu_bwd(dy.to(DEV), c.to(DEV)), cf.grad) < tol
