#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the REFERENCE's own code in the build container.

Needs /root/reference (read-only mount) and the image's transformers; it therefore runs
only here, never on the GPU box -- the .npz files it writes are committed.

  python tests/golden/make_golden.py

What is executed:
  * HF `Data2VecAudioForCTC` (what centralized/finetune_ASRs.py:164 instantiates), eager attention
  * /root/reference/federated/src/models.py::Data2VecAudioForCTC for STAGE 0/1/2 (imported as is);
    its Gumbel noise comes from `Tensor.exponential_`, which is patched to replay a seeded
    sample queue so that the oracle/kernels can be fed the very same noise
  * /root/reference/federated/src/utils.py:240-249 `average_weights`, exec'd from the sliced source
    (the module itself cannot be imported: librosa + env-dependent np.load at import)
Weights come from oracle.data2vec_ref.init_state_dict (seeded, own RNG stream) and are loaded
into the reference modules with load_state_dict(strict) -- so fixtures only hold outputs.
"""
import copy
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
REF = "/root/reference"
sys.path.insert(0, os.path.join(REF, "federated", "src"))

from oracle import data2vec_ref as R  # noqa: E402
from oracle import dacs_ref as D  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
torch.set_num_threads(8)


def hf_config(cfg):
    from transformers import Data2VecAudioConfig
    return Data2VecAudioConfig(**cfg.to_hf_kwargs())


def grads_of(model, names):
    return {n: p.grad.detach().clone() for n, p in model.named_parameters() if n in names and p.grad is not None}


def case_plain_ctc(tag, cfg, batch, samples, seed, use_mask):
    from transformers import Data2VecAudioForCTC
    sd = R.init_state_dict(cfg, seed=seed, dacs=False)
    m = Data2VecAudioForCTC(hf_config(cfg))
    m.load_state_dict(sd, strict=True)
    m.freeze_feature_encoder()
    m.train()
    b = R.synth_batch(cfg, batch, samples, seed=1234 + seed, ragged=use_mask, max_label=12)
    out = m(input_values=b["input_values"], attention_mask=b["attention_mask"] if use_mask else None,
            labels=b["labels"], output_hidden_states=True)
    out.loss.backward()
    keep = ["lm_head.weight", "lm_head.bias",
            "data2vec_audio.encoder.layers.0.attention.q_proj.weight",
            "data2vec_audio.encoder.layers.0.attention.v_proj.bias",
            "data2vec_audio.encoder.layers.1.feed_forward.intermediate_dense.weight",
            "data2vec_audio.encoder.layers.1.final_layer_norm.weight",
            "data2vec_audio.encoder.layer_norm.bias",
            "data2vec_audio.encoder.pos_conv_embed.layers.0.conv.weight",
            "data2vec_audio.encoder.pos_conv_embed.layers.1.conv.bias",
            "data2vec_audio.feature_projection.projection.weight",
            "data2vec_audio.feature_projection.layer_norm.weight"]
    g = grads_of(m, keep)
    gnorm = {n: float(p.grad.norm()) for n, p in m.named_parameters() if p.grad is not None}
    ids = out.logits.argmax(-1)
    np.savez_compressed(
        os.path.join(OUT, f"{tag}.npz"),
        seed=seed, batch=batch, samples=samples, use_mask=int(use_mask),
        loss=out.loss.detach().numpy(), logits=out.logits.detach().numpy(), argmax=ids.numpy(),
        hidden_last=out.hidden_states[-1].detach().numpy(), hidden_first=out.hidden_states[0].detach().numpy(),
        grad_names=np.array(list(g.keys())), **{"grad:" + k: v.numpy() for k, v in g.items()},
        gnorm_names=np.array(list(gnorm.keys())), gnorm_values=np.array(list(gnorm.values())),
    )
    print(tag, "loss", float(out.loss), "logits", tuple(out.logits.shape))


def case_layerdrop(tag, cfg, batch, samples, seed, layerdrop, coin_seed):
    """HF Data2VecAudioForCTC in training mode with LayerDrop (HF:392-404; every dropout 0).  The per-layer coins
    are host RNG draws (`torch.rand([])` in the image's transformers, `np.random.uniform` in the pinned 4.17); the
    fixture stores the coins and which blocks were skipped (read off the hidden states: a skipped block leaves its
    input unchanged) so the oracle and the CUDA path can be driven with the same outcome."""
    from transformers import Data2VecAudioForCTC
    sd = R.init_state_dict(cfg, seed=seed, dacs=False)
    hc = hf_config(cfg)
    hc.layerdrop = layerdrop
    m = Data2VecAudioForCTC(hc)
    m.load_state_dict(sd, strict=True)
    m.freeze_feature_encoder()
    m.train()
    b = R.synth_batch(cfg, batch, samples, seed=1234 + seed, ragged=True, max_label=12)
    torch.manual_seed(coin_seed)
    coins = np.array([float(torch.rand([])) for _ in range(cfg.num_hidden_layers)])
    torch.manual_seed(coin_seed)
    out = m(input_values=b["input_values"], attention_mask=b["attention_mask"], labels=b["labels"],
            output_hidden_states=True)
    out.loss.backward()
    hs = out.hidden_states
    skipped = np.array([l for l in range(cfg.num_hidden_layers) if torch.equal(hs[l], hs[l + 1])], dtype=np.int64)
    assert list(skipped) == [l for l in range(cfg.num_hidden_layers) if coins[l] < layerdrop], (skipped, coins)
    assert 0 < len(skipped) < cfg.num_hidden_layers
    g = {n: p.grad.detach().clone() for n, p in m.named_parameters() if p.grad is not None}
    none_grad = [n for n, p in m.named_parameters() if p.requires_grad and p.grad is None]
    keep = ["lm_head.weight", "data2vec_audio.feature_projection.projection.weight",
            "data2vec_audio.encoder.pos_conv_embed.layers.0.conv.weight"]
    kept_layers = [l for l in range(cfg.num_hidden_layers) if l not in set(skipped)]
    keep += [f"data2vec_audio.encoder.layers.{l}.attention.q_proj.weight" for l in kept_layers]
    keep += [f"data2vec_audio.encoder.layers.{l}.feed_forward.output_dense.bias" for l in kept_layers]
    np.savez_compressed(
        os.path.join(OUT, f"{tag}.npz"),
        seed=seed, batch=batch, samples=samples, layerdrop=layerdrop, coin_seed=coin_seed, coins=coins,
        skipped=skipped, num_hidden_layers=cfg.num_hidden_layers, none_grad_names=np.array(none_grad),
        loss=out.loss.detach().numpy(), logits=out.logits.detach().numpy(), argmax=out.logits.argmax(-1).numpy(),
        hidden_last=hs[-1].detach().numpy(),
        grad_names=np.array(keep), **{"grad:" + k: g[k].numpy() for k in keep},
        gnorm_names=np.array(list(g.keys())), gnorm_values=np.array([float(v.norm()) for v in g.values()]),
    )
    print(tag, "loss", float(out.loss), "skipped", skipped.tolist(), "coins", coins.round(3).tolist())


def case_dropout(tag, cfg, batch, samples, seed, ps, mask_seed):
    """HF Data2VecAudioForCTC in training mode with element dropout (attention_dropout 0).  torch.nn.Dropout.forward
    is replaced by oracle.data2vec_ref.call_order_mask (k-th call -> generator seeded mask_seed + k), so the fixture
    pins WHERE each dropout sits and with which probability; the oracle replays the same rule."""
    from transformers import Data2VecAudioForCTC
    sd = R.init_state_dict(cfg, seed=seed, dacs=False)
    hc = hf_config(cfg)
    for k, v in ps.items():
        setattr(hc, k, v)
    m = Data2VecAudioForCTC(hc)
    m.load_state_dict(sd, strict=True)
    m.freeze_feature_encoder()
    m.train()
    b = R.synth_batch(cfg, batch, samples, seed=1234 + seed, ragged=True, max_label=12)
    calls = {"k": 0, "ps": []}
    orig = torch.nn.Dropout.forward

    def patched(self, x):
        k = calls["k"]
        calls["k"] += 1
        calls["ps"].append(self.p)
        return R.call_order_mask(x, self.p, mask_seed, k)

    torch.nn.Dropout.forward = patched
    try:
        out = m(input_values=b["input_values"], attention_mask=b["attention_mask"], labels=b["labels"],
                output_hidden_states=True)
    finally:
        torch.nn.Dropout.forward = orig
    out.loss.backward()
    g = {n: p.grad.detach().clone() for n, p in m.named_parameters() if p.grad is not None}
    keep = ["lm_head.weight", "data2vec_audio.feature_projection.projection.weight",
            "data2vec_audio.encoder.layers.0.attention.out_proj.bias",
            "data2vec_audio.encoder.layers.0.feed_forward.intermediate_dense.weight",
            "data2vec_audio.encoder.layers.1.feed_forward.output_dense.bias",
            "data2vec_audio.encoder.layers.1.attention.q_proj.weight"]
    np.savez_compressed(
        os.path.join(OUT, f"{tag}.npz"),
        seed=seed, batch=batch, samples=samples, mask_seed=mask_seed, n_dropout_calls=calls["k"],
        dropout_call_ps=np.array(calls["ps"]), p_names=np.array(list(ps.keys())), p_values=np.array(list(ps.values())),
        loss=out.loss.detach().numpy(), logits=out.logits.detach().numpy(),
        hidden_last=out.hidden_states[-1].detach().numpy(),
        grad_names=np.array(keep), **{"grad:" + k: g[k].numpy() for k in keep},
        gnorm_names=np.array(list(g.keys())), gnorm_values=np.array([float(v.norm()) for v in g.values()]),
    )
    print(tag, "loss", float(out.loss.detach()), "dropout calls", calls["k"], calls["ps"])


DROPOUT_PS = dict(hidden_dropout=0.1, activation_dropout=0.15, feat_proj_dropout=0.05, final_dropout=0.2)
ATTN_DROPOUT_PS = dict(DROPOUT_PS, attention_dropout=0.25)


def case_attn_dropout(tag, cfg, batch, samples, seed, ps, mask_seed):
    """As case_dropout, with attention_dropout > 0 as well.  HF's eager attention calls `nn.functional.dropout` on the
    softmax probabilities (HF:189-197) and nn.Dropout.forward calls it too, so `torch.nn.functional.dropout` itself is
    replaced by the call-order rule: the fixture pins the position of EVERY dropout of the training forward, the
    attention-probability one included (first call of each block)."""
    from transformers import Data2VecAudioForCTC
    import torch.nn.functional as TF
    sd = R.init_state_dict(cfg, seed=seed, dacs=False)
    hc = hf_config(cfg)
    for k, v in ps.items():
        setattr(hc, k, v)
    m = Data2VecAudioForCTC(hc)
    m.load_state_dict(sd, strict=True)
    m.freeze_feature_encoder()
    m.train()
    b = R.synth_batch(cfg, batch, samples, seed=1234 + seed, ragged=True, max_label=12)
    calls = {"k": 0, "ps": [], "shapes": []}
    orig = TF.dropout

    def patched(x, p=0.5, training=True, inplace=False):
        if not training:
            return x
        k = calls["k"]
        calls["k"] += 1
        calls["ps"].append(p)
        calls["shapes"].append(tuple(x.shape))
        return R.call_order_mask(x, p, mask_seed, k)

    TF.dropout = patched
    try:
        out = m(input_values=b["input_values"], attention_mask=b["attention_mask"], labels=b["labels"],
                output_hidden_states=True)
    finally:
        TF.dropout = orig
    out.loss.backward()
    assert calls["k"] == 3 + 4 * cfg.num_hidden_layers, calls
    g = {n: p.grad.detach().clone() for n, p in m.named_parameters() if p.grad is not None}
    keep = ["lm_head.weight", "data2vec_audio.feature_projection.projection.weight",
            "data2vec_audio.encoder.layers.0.attention.out_proj.bias",
            "data2vec_audio.encoder.layers.0.attention.v_proj.weight",
            "data2vec_audio.encoder.layers.0.attention.q_proj.weight",
            "data2vec_audio.encoder.layers.1.attention.k_proj.weight",
            "data2vec_audio.encoder.layers.1.feed_forward.output_dense.bias"]
    np.savez_compressed(
        os.path.join(OUT, f"{tag}.npz"),
        seed=seed, batch=batch, samples=samples, mask_seed=mask_seed, n_dropout_calls=calls["k"],
        dropout_call_ps=np.array(calls["ps"]), p_names=np.array(list(ps.keys())), p_values=np.array(list(ps.values())),
        loss=out.loss.detach().numpy(), logits=out.logits.detach().numpy(),
        hidden_last=out.hidden_states[-1].detach().numpy(),
        grad_names=np.array(keep), **{"grad:" + k: g[k].numpy() for k in keep},
        gnorm_names=np.array(list(g.keys())), gnorm_values=np.array([float(v.norm()) for v in g.values()]),
    )
    print(tag, "loss", float(out.loss.detach()), "dropout calls", calls["k"], calls["ps"], calls["shapes"][:6])


def case_variant_ctc(tag, cfg, batch, samples, seed, use_mask):
    """Config-5 variants (centralized/finetune_ASRs.py:154-172): HF Wav2Vec2ForCTC / HubertForCTC, eager attention.
    The oracle state_dict uses the transformers-4.17 weight_norm key names (weight_g / weight_v, the reference's
    pinned version); the installed transformers registers them as parametrizations.weight.original0/1."""
    import transformers as tf
    Config, Model = {"wav2vec2": (tf.Wav2Vec2Config, tf.Wav2Vec2ForCTC), "hubert": (tf.HubertConfig, tf.HubertForCTC)}[cfg.model_type]
    sd = R.init_state_dict(cfg, seed=seed, dacs=False)
    m = Model(Config(**cfg.to_hf_kwargs()))
    hf_sd = {}
    for k, v in sd.items():
        k = k.replace("conv.weight_g", "conv.parametrizations.weight.original0")
        k = k.replace("conv.weight_v", "conv.parametrizations.weight.original1")
        hf_sd[k] = v
    missing, unexpected = m.load_state_dict(hf_sd, strict=False)
    assert not unexpected, unexpected
    assert all("masked_spec_embed" in k for k in missing), missing
    m.freeze_feature_encoder()
    m.train()
    b = R.synth_batch(cfg, batch, samples, seed=1234 + seed, ragged=use_mask, max_label=12)
    out = m(input_values=b["input_values"], attention_mask=b["attention_mask"] if use_mask else None,
            labels=b["labels"], output_hidden_states=True)
    out.loss.backward()
    px = cfg.prefix
    back = lambda n: n.replace("conv.parametrizations.weight.original0", "conv.weight_g").replace(
        "conv.parametrizations.weight.original1", "conv.weight_v")
    keep = ["lm_head.weight", "lm_head.bias",
            px + "encoder.layers.0.attention.q_proj.weight", px + "encoder.layers.0.attention.out_proj.bias",
            px + "encoder.layers.0.layer_norm.weight", px + "encoder.layers.1.feed_forward.output_dense.bias",
            px + "encoder.layers.1.final_layer_norm.bias", px + "encoder.layer_norm.weight",
            px + "encoder.pos_conv_embed.conv.weight_g", px + "encoder.pos_conv_embed.conv.weight_v",
            px + "encoder.pos_conv_embed.conv.bias", px + "feature_projection.projection.weight"]
    g = {back(n): p.grad.detach().clone() for n, p in m.named_parameters() if p.grad is not None and back(n) in keep}
    gnorm = {back(n): float(p.grad.norm()) for n, p in m.named_parameters() if p.grad is not None}
    np.savez_compressed(
        os.path.join(OUT, f"{tag}.npz"),
        seed=seed, batch=batch, samples=samples, use_mask=int(use_mask),
        loss=out.loss.detach().numpy(), logits=out.logits.detach().numpy(), argmax=out.logits.argmax(-1).numpy(),
        hidden_last=out.hidden_states[-1].detach().numpy(), hidden_first=out.hidden_states[0].detach().numpy(),
        grad_names=np.array(list(g.keys())), **{"grad:" + k: v.numpy() for k, v in g.items()},
        gnorm_names=np.array(list(gnorm.keys())), gnorm_values=np.array(list(gnorm.values())),
    )
    print(tag, "loss", float(out.loss), "logits", tuple(out.logits.shape), "grads", len(gnorm))


VARIANTS = {
    # tag: (config factory kwargs, batch, samples, seed, use_mask)
    "w2v2_group_postln": ("tiny_wav2vec2", {}, 3, 8000, 30, True),           # wav2vec2-base-960h family
    "w2v2_layer_stable": ("tiny_wav2vec2", dict(feat_extract_norm="layer", conv_bias=True, do_stable_layer_norm=True), 2, 6000, 31, True),  # wav2vec2-large-960h-lv60
    "hubert_stable": ("tiny_hubert", {}, 3, 8000, 32, True),                # hubert-large / xlarge-ls960-ft family
    "hubert_base_nomask": ("tiny_hubert", dict(feat_extract_norm="group", conv_bias=False, do_stable_layer_norm=False,
                                               feat_proj_layer_norm=False), 2, 6000, 33, False),  # hubert-base-ls960
}


def variant_cfg(tag):
    fac, kw, *_ = VARIANTS[tag]
    return getattr(R.ArchConfig, fac)(**kw)


class _ReplayExponential:
    """Patch Tensor.exponential_ to replay queued Exp(1) samples (models.py:112)."""

    def __init__(self, queue):
        self.queue = list(queue)

    def __enter__(self):
        self.orig = torch.Tensor.exponential_
        q = self.queue

        def fake(t, *a, **k):
            src = q.pop(0)
            assert src.shape == t.shape, (src.shape, t.shape)
            return t.copy_(src)

        torch.Tensor.exponential_ = fake
        return self

    def __exit__(self, *a):
        torch.Tensor.exponential_ = self.orig


def case_dacs(tag, cfg, batch, samples, seed, stage, toggle_ratio=0, ad_loss="recall"):
    import models as refmodels  # /root/reference/federated/src/models.py
    sd = R.init_state_dict(cfg, seed=seed, dacs=True)
    args = SimpleNamespace(LAMBDA=0.5, TOGGLE_RATIO=toggle_ratio, GS_TAU=1, AD_loss=ad_loss, W_LOSS=None, STAGE=stage)
    hc = hf_config(cfg)
    m = refmodels.Data2VecAudioForCTC(hc, args)
    m.load_state_dict(sd, strict=True)
    m.config.ctc_zero_infinity = True
    m.train()
    b = R.synth_batch(cfg, batch, samples, seed=4321 + seed, ragged=True, max_label=12)
    T = int(R.conv_out_lengths(cfg, torch.tensor([samples]))[0])
    shape = (batch, T, cfg.hidden_size, 2)
    e_lm = torch.empty(shape).exponential_(generator=torch.Generator().manual_seed(1238 + seed))
    e_ad = torch.empty(shape).exponential_(generator=torch.Generator().manual_seed(2238 + seed))
    with _ReplayExponential([e_lm, e_ad]):
        out = m(input_values=b["input_values"], attention_mask=b["attention_mask"], labels=b["labels"],
                dementia_labels=b["dementia_labels"])
    out.loss.backward()
    g = {n: p.grad.detach().clone() for n, p in m.named_parameters()
         if p.grad is not None and (n.startswith(("arbitrator", "dementia_head", "lm_head")) or
                                    n.endswith("layers.1.final_layer_norm.weight"))}
    gnorm = {n: float(p.grad.norm()) for n, p in m.named_parameters() if p.grad is not None}
    np.savez_compressed(
        os.path.join(OUT, f"{tag}.npz"),
        seed=seed, batch=batch, samples=samples, stage=stage, toggle_ratio=toggle_ratio, ad_loss=ad_loss,
        exp_lm=e_lm.numpy(), exp_ad=e_ad.numpy(),
        loss=out.loss.detach().numpy(), logits=out.logits.detach().numpy(),
        grad_names=np.array(list(g.keys())), **{"grad:" + k: v.numpy() for k, v in g.items()},
        gnorm_names=np.array(list(gnorm.keys())), gnorm_values=np.array(list(gnorm.values())),
    )
    print(tag, "stage", stage, "loss", float(out.loss), "trainable-with-grad", len(gnorm))


def case_fedavg(tag):
    src = open(os.path.join(REF, "federated", "src", "utils.py"), encoding="utf-8").read().splitlines()
    body = "\n".join(src[239:249])  # lines 240-249
    ns = {"copy": copy, "torch": torch}
    exec(body, ns)
    average_weights = ns["average_weights"]
    out = {}
    for n_clients in (2, 3, 8):
        g = torch.Generator().manual_seed(77 + n_clients)
        ws = [{"a.weight": torch.randn(37, 19, generator=g) * (1 + i), "a.bias": torch.randn(19, generator=g),
               "b.weight": torch.randn(5, 3, 7, generator=g) * 1e-3} for i in range(n_clients)]
        avg = average_weights(copy.deepcopy(ws))
        for k, v in avg.items():
            out[f"n{n_clients}:{k}"] = v.numpy()
    np.savez_compressed(os.path.join(OUT, f"{tag}.npz"), **out)
    print(tag, "ok")


def main():
    cfg = R.ArchConfig.tiny()
    case_plain_ctc("plain_ctc_masked", cfg, batch=3, samples=8000, seed=0, use_mask=True)
    case_plain_ctc("plain_ctc_nomask", cfg, batch=2, samples=6000, seed=1, use_mask=False)
    cfgb = R.ArchConfig.tiny(conv_bias=True)
    case_plain_ctc("plain_ctc_convbias", cfgb, batch=2, samples=6000, seed=2, use_mask=True)
    for stage in (0, 1, 2):
        case_dacs(f"dacs_stage{stage}", cfg, batch=3, samples=8000, seed=10 + stage, stage=stage)
    case_dacs("dacs_stage2_toggle", cfg, batch=2, samples=6000, seed=20, stage=2, toggle_ratio=0.3)
    case_dacs("dacs_stage2_f1", cfg, batch=2, samples=6000, seed=21, stage=2, ad_loss="f1")
    case_fedavg("fedavg")
    case_layerdrop("plain_ctc_layerdrop", R.ArchConfig.tiny(num_hidden_layers=4), batch=2, samples=6000, seed=3,
                   layerdrop=0.5, coin_seed=5)
    case_dropout("plain_ctc_dropout", R.ArchConfig.tiny(), batch=2, samples=6000, seed=5, ps=DROPOUT_PS, mask_seed=900)
    case_attn_dropout("plain_ctc_attn_dropout", R.ArchConfig.tiny(), batch=2, samples=6000, seed=6, ps=ATTN_DROPOUT_PS,
                      mask_seed=901)
    for tag, (_, _, batch, samples, seed, use_mask) in VARIANTS.items():
        case_variant_ctc(tag, variant_cfg(tag), batch, samples, seed, use_mask)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "variants":  # only (re)generate the config-5 variant fixtures
        for tag, (_, _, batch, samples, seed, use_mask) in VARIANTS.items():
            case_variant_ctc(tag, variant_cfg(tag), batch, samples, seed, use_mask)
    elif len(sys.argv) > 1 and sys.argv[1] == "attn_dropout":
        case_attn_dropout("plain_ctc_attn_dropout", R.ArchConfig.tiny(), batch=2, samples=6000, seed=6, ps=ATTN_DROPOUT_PS,
                          mask_seed=901)
    elif len(sys.argv) > 1 and sys.argv[1] == "dropout":
        case_dropout("plain_ctc_dropout", R.ArchConfig.tiny(), batch=2, samples=6000, seed=5, ps=DROPOUT_PS, mask_seed=900)
    elif len(sys.argv) > 1 and sys.argv[1] == "layerdrop":
        case_layerdrop("plain_ctc_layerdrop", R.ArchConfig.tiny(num_hidden_layers=4), batch=2, samples=6000, seed=3,
                       layerdrop=0.5, coin_seed=5)
    else:
        main()
