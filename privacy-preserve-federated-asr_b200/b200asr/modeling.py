"""Drop-in nn.Modules for the reference's model classes (same sub-module names, state_dict keys, forward
signature and return type), computing through the b200asr C-ABI kernels.

  Data2VecAudioForCTC       <-> HF transformers Data2VecAudioForCTC as built at centralized/finetune_ASRs.py:164
  DacsData2VecAudioForCTC   <-> federated/src/models.py:262-631 (and centralized/trainer_data2vec_2st.py:159-505)

`load_state_dict(reference_model.state_dict())` works key for key.  Parameters stay fp32 (master weights);
`compute_dtype` selects the activation / GEMM operand type (bf16 = tcgen05 path, fp32 = parity mode).
"""
from __future__ import annotations

import contextlib
import math
import warnings
from dataclasses import dataclass
from typing import List, Optional, Tuple

import os as _os

import torch
import torch.nn as nn

from . import functional as Fn
from . import ops

try:  # the reference returns transformers' CausalLMOutput; use it when available so isinstance checks hold
    from transformers.modeling_outputs import CausalLMOutput
except Exception:  # pragma: no cover
    @dataclass
    class CausalLMOutput:  # type: ignore[no-redef]
        loss: Optional[torch.Tensor] = None
        logits: Optional[torch.Tensor] = None
        hidden_states: Optional[Tuple[torch.Tensor, ...]] = None
        attentions: Optional[Tuple[torch.Tensor, ...]] = None

        def __getitem__(self, i):
            return tuple(v for v in (self.loss, self.logits, self.hidden_states, self.attentions) if v is not None)[i]


def _cfg(config, name, default):
    return getattr(config, name, default)


class _Affine(nn.Module):
    """Parameter holder with the names nn.Linear / nn.Conv1d / nn.LayerNorm use (`weight`, `bias`)."""

    def __init__(self, wshape, bias=True, kind="linear", std=0.02):
        super().__init__()
        w = torch.empty(*wshape)
        if kind == "ln":
            w.fill_(1.0)
        elif kind == "conv":
            nn.init.kaiming_normal_(w)
        else:
            w.normal_(0.0, std)
        self.weight = nn.Parameter(w)
        if bias:
            self.bias = nn.Parameter(torch.zeros(wshape[0]))
        else:
            self.register_parameter("bias", None)


class _ConvLayer(nn.Module):
    def __init__(self, cin, cout, k, bias, norm=True):
        super().__init__()
        self.conv = _Affine((cout, cin, k), bias=bias, kind="conv")
        if norm:  # LayerNorm(C) ("layer") or GroupNorm(C groups, C) on layer 0 ("group"): both hold weight/bias [C]
            self.layer_norm = _Affine((cout,), kind="ln")


class _FeatureExtractor(nn.Module):
    def __init__(self, config):
        super().__init__()
        dims, ks = list(config.conv_dim), list(config.conv_kernel)
        group = _cfg(config, "feat_extract_norm", "layer") == "group"
        self.conv_layers = nn.ModuleList(
            [_ConvLayer(1 if i == 0 else dims[i - 1], dims[i], ks[i], bool(_cfg(config, "conv_bias", False)),
                        norm=(not group) or i == 0)
             for i in range(len(dims))])

    def _freeze_parameters(self):
        for p in self.parameters():
            p.requires_grad = False


class _FeatureProjection(nn.Module):
    def __init__(self, config, layer_norm=True):
        super().__init__()
        if layer_norm:
            self.layer_norm = _Affine((config.conv_dim[-1],), kind="ln")
        self.projection = _Affine((config.hidden_size, config.conv_dim[-1]))


class _PosConvLayer(nn.Module):
    def __init__(self, config):
        super().__init__()
        d, g, k = config.hidden_size, config.num_conv_pos_embedding_groups, config.conv_pos_kernel_size
        self.conv = _Affine((d, d // g, k), kind="conv")


class _PosConvEmbed(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.layers = nn.ModuleList([_PosConvLayer(config) for _ in range(config.num_conv_pos_embeddings)])


class _WeightNormConv(nn.Module):
    """nn.utils.weight_norm(nn.Conv1d(d, d, k, groups), dim=2): parameters `weight_g` [1,1,k], `weight_v` [d, d/g, k],
    `bias` [d] -- the transformers==4.17 (reference pin) key names; `parametrizations.weight.original0/1` keys of
    newer checkpoints are renamed on load."""

    def __init__(self, d, groups, k):
        super().__init__()
        v = torch.empty(d, d // groups, k).normal_(0.0, 2.0 * math.sqrt(1.0 / (k * d)))
        self.weight_g = nn.Parameter(v.norm(dim=(0, 1), keepdim=True))
        self.weight_v = nn.Parameter(v)
        self.bias = nn.Parameter(torch.zeros(d))
        self._register_load_state_dict_pre_hook(self._rename_parametrizations)

    @staticmethod
    def _rename_parametrizations(state_dict, prefix, *args):
        for new, old in (("parametrizations.weight.original0", "weight_g"), ("parametrizations.weight.original1", "weight_v")):
            if prefix + new in state_dict:
                state_dict[prefix + old] = state_dict.pop(prefix + new)


class _PosConvEmbedWeightNorm(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.conv = _WeightNormConv(config.hidden_size, config.num_conv_pos_embedding_groups,
                                    config.num_conv_pos_embeddings)


class _Attention(nn.Module):
    def __init__(self, d):
        super().__init__()
        # q, k, v registered back to back: in a FlatParams buffer the three weights (and the three biases) are adjacent,
        # so the fused [3d, d] projection, its bf16 shadow and its gradient are plain views (no concat / split kernels)
        self.q_proj, self.k_proj, self.v_proj, self.out_proj = (_Affine((d, d)) for _ in range(4))


class _FeedForward(nn.Module):
    def __init__(self, d, f):
        super().__init__()
        self.intermediate_dense = _Affine((f, d))
        self.output_dense = _Affine((d, f))


class _EncoderLayer(nn.Module):
    def __init__(self, config):
        super().__init__()
        d = config.hidden_size
        self.attention = _Attention(d)
        self.layer_norm = _Affine((d,), kind="ln")
        self.feed_forward = _FeedForward(d, config.intermediate_size)
        self.final_layer_norm = _Affine((d,), kind="ln")

    def params(self):
        a, f = self.attention, self.feed_forward
        return (a.q_proj.weight, a.q_proj.bias, a.k_proj.weight, a.k_proj.bias, a.v_proj.weight, a.v_proj.bias,
                a.out_proj.weight, a.out_proj.bias, self.layer_norm.weight, self.layer_norm.bias,
                f.intermediate_dense.weight, f.intermediate_dense.bias, f.output_dense.weight, f.output_dense.bias,
                self.final_layer_norm.weight, self.final_layer_norm.bias)


class _Encoder(nn.Module):
    def __init__(self, config, weight_norm_posconv=False):
        super().__init__()
        self.pos_conv_embed = _PosConvEmbedWeightNorm(config) if weight_norm_posconv else _PosConvEmbed(config)
        self.layer_norm = _Affine((config.hidden_size,), kind="ln")
        self.layers = nn.ModuleList([_EncoderLayer(config) for _ in range(config.num_hidden_layers)])


def conv_out_lengths(config, lengths: torch.Tensor) -> torch.Tensor:
    """_get_feat_extract_output_lengths (HF:507-526; centralized/Models.py:249-271 with the integer-floor
    semantics of the pinned torch 1.7.1): floor((len - k) / s) + 1 per conv layer."""
    for k, s in zip(config.conv_kernel, config.conv_stride):
        lengths = torch.div(lengths - k, s, rounding_mode="floor") + 1
    return lengths


def device_conv_out_lengths(config, attention_mask: torch.Tensor) -> torch.Tensor:
    """int32 frame lengths of a padded batch from its attention mask: conv_out_lengths(attention_mask.sum(-1)) with the
    whole per-layer chain in one launch (CUDA tensors; anything else takes the torch formula)."""
    n = attention_mask.sum(-1)
    if n.is_cuda and len(config.conv_kernel) <= 16:
        return ops.conv_out_lengths(n.to(torch.int64).contiguous(), list(config.conv_kernel), list(config.conv_stride))
    return conv_out_lengths(config, n).to(torch.int32).contiguous()


class Data2VecAudioModel(nn.Module):
    """`data2vec_audio` sub-module: feature extractor -> projection -> encoder (HF:744-800).  The wav2vec2 / hubert
    sub-classes below switch the positional conv (one weight-normed conv), the conv-stack norm ("group") and the
    LayerNorm ordering (do_stable_layer_norm) -- transformers wav2vec2/modeling_wav2vec2.py Wav2Vec2Model."""

    family = "data2vec"

    def __init__(self, config, compute_dtype=torch.bfloat16):
        super().__init__()
        self.config = config
        self.feature_extractor = _FeatureExtractor(config)
        self.feature_projection = _FeatureProjection(
            config, layer_norm=self.family != "hubert" or bool(_cfg(config, "feat_proj_layer_norm", True)))
        self.encoder = _Encoder(config, weight_norm_posconv=self.family != "data2vec")
        if _cfg(config, "mask_time_prob", 0.0) > 0.0 or _cfg(config, "mask_feature_prob", 0.0) > 0.0:
            self.masked_spec_embed = nn.Parameter(torch.empty(config.hidden_size).uniform_())
        self.cache = Fn.ShadowCache(compute_dtype)
        self._warned = False
        # LayerDrop coins (HF:392-404) are host RNG draws: "numpy" = np.random.uniform(0, 1) as in the pinned
        # transformers 4.17, "torch" = torch.rand([]) on the CPU generator as in transformers >= 4.3x.
        self.layerdrop_rng = "numpy"
        self.last_skipped_layers = ()
        # element dropout: masks are a pure function of (dropout_seed, site, dropout_counter, element index); the
        # counter lives on the device and is bumped once per training forward, so CUDA-graph replays draw fresh masks
        self.dropout_seed = None  # None: torch.initial_seed() at first use (follows torch.manual_seed)
        self.register_buffer("dropout_counter", torch.zeros(1, dtype=torch.int64), persistent=False)
        self._drop_ctr = None
        self._attn_keep = None            # keep bits of this forward's attention-probability dropout, one row per block
        self._act_keep = None             # (keep bits of the activation dropout, row pitch), one row per block
        # opt-in: every block draws its own attention keep bits under its QKV GEMM instead of one launch for all blocks under
        # the conv stack (A/B on one box, two runs each: 14.36 / 14.41 against 14.31 / 14.33 ms per step -- no gain)
        self.attn_keep_per_block = _os.environ.get("B200ASR_ATTN_KEEP_PER_BLOCK", "0") != "0"
        self.activation_keep_bits = False  # draw the activation dropout's decisions ahead as bits (measured slower: see below)
        self._segment_recorder = None     # set by graph.GraphedTrainStep while it captures an update with LayerDrop
        self._force_no_layerdrop = False

    @property
    def compute_dtype(self):
        return self.cache.dtype

    def set_compute_dtype(self, dtype):
        self.cache = Fn.ShadowCache(dtype)

    def _check_unsupported(self):
        c = self.config
        if self.training and _cfg(c, "mask_time_prob", 0.0) > 0.0 and _cfg(c, "apply_spec_augment", True):
            raise NotImplementedError("SpecAugment (mask_time_prob > 0) is not supported; every reference entry "
                                      "point sets mask_time_prob=0 (finetune_ASRs.py:162-163)")
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.feature_extractor.parameters()):
            raise NotImplementedError("the conv feature extractor is forward-only: call freeze_feature_encoder() "
                                      "as every reference entry point does (finetune_ASRs.py:200, models.py:302)")
        if _cfg(c, "feat_extract_norm", "layer") not in ("layer", "group"):
            raise ValueError(f"feat_extract_norm must be 'layer' or 'group', got {c.feat_extract_norm!r}")
        if self.family == "data2vec" and (_cfg(c, "feat_extract_norm", "layer") != "layer"
                                          or _cfg(c, "do_stable_layer_norm", False)):
            raise NotImplementedError("Data2VecAudio has no GroupNorm / stable-LayerNorm variant (HF:51-157, 359-418)")
        if _cfg(c, "conv_pos_batch_norm", False):
            raise NotImplementedError("hubert conv_pos_batch_norm is not supported")

    def _layerdrop_coins(self, n_layers):
        """HF:392-404: one uniform draw per block, every step, in block order; the block is skipped when
        training and the draw is below config.layerdrop.  Host RNG, no device sync.  Under stream capture the
        outcome would be frozen into the graph, so that combination is refused."""
        p = float(_cfg(self.config, "layerdrop", 0.0) or 0.0)
        if not self.training or p <= 0.0 or self._force_no_layerdrop:
            return ()
        if torch.cuda.is_current_stream_capturing():
            raise NotImplementedError("LayerDrop draws a fresh coin per block every step; a single captured CUDA graph "
                                      "would replay one outcome: use graph.GraphedTrainStep (one graph per block, coins "
                                      "drawn per replay), set config.layerdrop = 0, or run eagerly")
        if self.layerdrop_rng == "torch":
            coins = [float(torch.rand([])) for _ in range(n_layers)]
        elif self.layerdrop_rng == "numpy":
            import numpy as np
            coins = [float(np.random.uniform(0, 1)) for _ in range(n_layers)]
        else:
            raise ValueError(f"layerdrop_rng must be 'numpy' or 'torch', got {self.layerdrop_rng!r}")
        return tuple(l for l, u in enumerate(coins) if u < p)

    # dropout sites: 0 feat_proj_dropout, 1 encoder prologue (hidden_dropout), 2 final_dropout, 16 + 4 l + {0, 1, 2, 3}
    # = block l's {attention output, FFN activation, FFN output, attention probabilities}
    def _p_drop(self, name):
        return float(_cfg(self.config, name, 0.0) or 0.0) if self.training else 0.0

    def _begin_dropout_step(self):
        """One counter value per training forward; the clone is what this forward's autograd nodes keep, so a second
        forward before the backward (gradient accumulation) cannot shift the first one's masks."""
        self._drop_ctr = None
        self._attn_keep = None
        self._act_keep = None
        if self.training and any(self._p_drop(n) > 0 for n in ("hidden_dropout", "activation_dropout", "attention_dropout",
                                                               "feat_proj_dropout", "final_dropout")):
            if self.dropout_seed is None:
                self.dropout_seed = int(torch.initial_seed()) & (2 ** 63 - 1)
            self.dropout_counter.add_(1)
            self._drop_ctr = self.dropout_counter.clone()

    def _draw_attention_keep(self, B, T):
        """Keep bits of the attention-probability dropout (HF:189-197) for every block of this forward, drawn by ONE
        launch on the side stream while the conv stack runs (the decisions are data independent); the encoder loop
        joins the side stream before its first attention.  [n_layers, words] int32, held by the blocks' autograd nodes."""
        p = self._p_drop("attention_dropout")
        if p <= 0.0 or self._drop_ctr is None:
            return
        n = len(self.encoder.layers)
        H = int(self.config.num_attention_heads)
        keep = torch.empty(n, ops.attn_keep_words(B, T, H), dtype=torch.int32, device=self._drop_ctr.device)
        if not self.attn_keep_per_block:
            with ops.on_side_stream(keep, self._drop_ctr, manual_join=True):
                ops.attn_dropout_mask(n, B, T, H, p, self.dropout_seed, 16 + 3, 4, self._drop_ctr, out=keep)
        # (attn_keep_per_block: every block draws its own row at its start, on the side stream under its QKV GEMM)
        self._attn_keep = keep

    def _draw_activation_keep(self, rows):
        """Keep bits of the activation dropout (HF:335-339) of every block, drawn by ONE launch on the side stream while the
        conv stack runs; the W1 GEMM's epilogue reads them instead of running Philox itself (the epilogue bounds that GEMM).
        Same decisions as ops.dropout at site 16 + 4 l + 1.  bf16 path with 64-aligned intermediate size only."""
        self._act_keep = None
        p = self._p_drop("activation_dropout")
        inter = int(self.config.intermediate_size)
        # Opt-in (self.activation_keep_bits = True).  Measured at 8 x 10 s: the W1 GEMM drops from 48.9 us (Philox in the
        # epilogue) to 45.3 us, but drawing the bits is 243 us of GPU work per step that the conv stack does not hide (its
        # GEMMs fill the machine): 14.67 ms per step against 14.45 ms -- a loss, so the epilogue keeps its own Philox.
        if not self.activation_keep_bits:
            return
        if p <= 0.0 or self._drop_ctr is None or self.cache.dtype != torch.bfloat16 or inter % 64 != 0:
            return
        n = len(self.encoder.layers)
        mpad = ops.dropout_keep_mpad(rows)
        keep = torch.empty(n, ops.dropout_keep_words(inter, mpad), dtype=torch.int32, device=self._drop_ctr.device)
        with ops.on_side_stream(keep, self._drop_ctr, manual_join=True):
            ops.dropout_keep_bits(n, 16 + 1, 4, rows, inter, p, self.dropout_seed, self._drop_ctr, out=keep)
        self._act_keep = (keep, mpad)

    def apply_dropout(self, x, name, site):
        """nn.Dropout(config.<name>) at a stand-alone site (training mode only)."""
        p = self._p_drop(name)
        if p <= 0.0 or self._drop_ctr is None:
            return x
        return Fn.DropoutFn.apply(x, p, self.dropout_seed, site, self._drop_ctr)

    def forward(self, input_values, attention_mask=None, output_hidden_states=False):
        """Returns (hidden [B,T,d] in compute dtype, frame_lengths int32 [B] or None, hidden_states tuple|None)."""
        c = self.config
        self._check_unsupported()
        self._begin_dropout_step()
        cache = self.cache
        if self._drop_ctr is not None and self._p_drop("attention_dropout") > 0:
            L_in = int(input_values.shape[-1])
            self._draw_attention_keep(int(input_values.shape[0]), int(conv_out_lengths(c, torch.tensor(L_in))))
        if self._drop_ctr is not None and self._p_drop("activation_dropout") > 0:
            L_in = int(input_values.shape[-1])
            self._draw_activation_keep(int(input_values.shape[0]) * int(conv_out_lengths(c, torch.tensor(L_in))))
        fe = self.feature_extractor.conv_layers
        lns = [getattr(l, "layer_norm", None) for l in fe]
        feats = Fn.feature_extractor_fwd(
            input_values, [l.conv.weight for l in fe], [l.conv.bias for l in fe],
            [None if n is None else n.weight for n in lns], [None if n is None else n.bias for n in lns],
            list(c.conv_kernel), list(c.conv_stride), 1e-5, cache, norm=_cfg(c, "feat_extract_norm", "layer"))
        B, T, _ = feats.shape
        if self._segment_recorder is not None and self._segment_recorder.split_conv:
            self._segment_recorder.conv_mark()   # graph capture: the frozen conv stack becomes a graph of its own
        klen = None
        if attention_mask is not None:
            klen = device_conv_out_lengths(c, attention_mask)
        fp = self.feature_projection
        eps = float(_cfg(c, "layer_norm_eps", 1e-5))
        fln = getattr(fp, "layer_norm", None)
        h = Fn.FeatureProjectionFn.apply(feats, None if fln is None else fln.weight, None if fln is None else fln.bias,
                                         fp.projection.weight, fp.projection.bias, eps, cache, klen)
        h = self.apply_dropout(h, "feat_proj_dropout", 0)  # HF:171 (commutes with zeroing the padded frames)
        enc = self.encoder
        stable = bool(_cfg(c, "do_stable_layer_norm", False))
        if self.family == "data2vec":
            pcs = enc.pos_conv_embed.layers
            h = Fn.PosConvEmbedFn.apply(h, enc.layer_norm.weight, enc.layer_norm.bias, eps,
                                        int(c.num_conv_pos_embedding_groups), int(c.conv_pos_kernel_size), cache,
                                        *[l.conv.weight for l in pcs], *[l.conv.bias for l in pcs])
        else:
            pc = enc.pos_conv_embed.conv
            h = Fn.PosConvWeightNormFn.apply(h, None if stable else enc.layer_norm.weight,
                                             None if stable else enc.layer_norm.bias, eps,
                                             int(c.num_conv_pos_embedding_groups), int(c.num_conv_pos_embeddings), cache,
                                             pc.weight_g, pc.weight_v, pc.bias)
        h = self.apply_dropout(h, "hidden_dropout", 1)  # HF:380-381 (after the LayerNorm; pre-LN encoders: after the add)
        d = h.shape[-1]
        x = h.view(B * T, d)
        collected = [h] if output_hidden_states else None
        meta = (B, T, int(c.num_attention_heads), eps)
        block = Fn.EncoderLayerStableFn if stable else Fn.EncoderLayerFn
        ph, pa = self._p_drop("hidden_dropout"), self._p_drop("activation_dropout")
        pt = self._p_drop("attention_dropout") if self._attn_keep is not None else 0.0
        drop_block = None
        if self._drop_ctr is not None and (ph > 0 or pa > 0 or pt > 0):
            drop_block = Fn.EncoderLayerStableDropoutFn if stable else Fn.EncoderLayerDropoutFn
        if self._attn_keep is not None or self._act_keep is not None:
            ops.join_side_stream()   # the keep bits were drawn on the side stream while the conv stack ran
        skipped = self._layerdrop_coins(len(enc.layers))
        self.last_skipped_layers = skipped
        rec = self._segment_recorder
        for li, layer in enumerate(enc.layers):
            if rec is not None and rec.block_marks:   # graph capture with LayerDrop: one CUDA graph per block
                x = Fn.SegmentMarkFn.apply(x, li, rec)
            if li in skipped:  # a dropped block is the identity and its parameters get no gradient
                pass
            elif drop_block is not None:
                x = drop_block.apply(x, klen, meta, cache,
                                     (ph, pa, self.dropout_seed, 16 + 4 * li, self._drop_ctr,
                                      None if self._attn_keep is None else self._attn_keep[li], pt,
                                      None if self._act_keep is None else (self._act_keep[0][li], self._act_keep[1]),
                                      self.attn_keep_per_block),
                                     *layer.params())
            else:
                x = block.apply(x, klen, meta, cache, *layer.params())
            if collected is not None and not (stable and li + 1 == len(enc.layers)):
                collected.append(x.view(B, T, d))
        if rec is not None and rec.block_marks:
            x = Fn.SegmentMarkFn.apply(x, len(enc.layers), rec)
        if stable:  # Wav2Vec2EncoderStableLayerNorm: one LayerNorm after the last block
            x = Fn.LayerNormFn.apply(x, enc.layer_norm.weight, enc.layer_norm.bias, eps)
            if collected is not None:
                collected.append(x.view(B, T, d))
        return x.view(B, T, d), klen, (tuple(collected) if collected is not None else None)


class Wav2Vec2Model(Data2VecAudioModel):
    """`wav2vec2` sub-module of Wav2Vec2ForCTC (finetune_ASRs.py:154-158)."""
    family = "wav2vec2"


class HubertModel(Data2VecAudioModel):
    """`hubert` sub-module of HubertForCTC (finetune_ASRs.py:166-172)."""
    family = "hubert"


def _ctc_index_tensors(config, input_values, attention_mask, labels, klen):
    """models.py:486-496 / HF:884-895 without the host sync of `masked_select`: the padded label matrix is
    consumed in place (row b starts at b*Lmax; valid labels are the row prefix, as the collators build them)."""
    B, Lmax = labels.shape
    if klen is None:
        T = int(conv_out_lengths(config, torch.tensor(input_values.shape[-1])))
        in_len = torch.full((B,), T, dtype=torch.int32, device=labels.device)
    else:
        in_len = klen
    tgt_len = (labels >= 0).sum(-1).to(torch.int32)
    tgt_off = torch.arange(B, dtype=torch.int32, device=labels.device) * Lmax
    targets = labels.clamp(min=0).to(torch.int32).contiguous().view(-1)
    return targets, tgt_off, tgt_len, in_len, Lmax


class _CTCBase(nn.Module):
    validate_labels = True  # reference raises ValueError on labels >= vocab_size (costs one host sync)

    base_model_prefix = "data2vec_audio"

    @property
    def backbone(self):
        return getattr(self, self.base_model_prefix)

    def freeze_feature_encoder(self):
        self.backbone.feature_extractor._freeze_parameters()

    freeze_feature_extractor = freeze_feature_encoder

    def save_pretrained(self, save_directory, state_dict=None):
        """`pytorch_model.bin` + `config.json`, the files the reference's clients and server exchange (checkpoint.py)."""
        from .checkpoint import save_pretrained
        save_pretrained(self, save_directory, state_dict)

    @classmethod
    def from_pretrained(cls, directory, *model_args, **kw):
        """Open a directory written by HF `save_pretrained` (or by `save_pretrained` above); see checkpoint.from_pretrained."""
        from .checkpoint import from_pretrained
        return from_pretrained(cls, directory, *model_args, **kw)

    def _ctc(self, logits, idx):
        c = self.config
        targets, tgt_off, tgt_len, in_len, lmax = idx
        return Fn.CTCLossFn.apply(logits, targets, tgt_off, tgt_len, in_len, lmax, int(c.pad_token_id),
                                  bool(_cfg(c, "ctc_zero_infinity", False)), _cfg(c, "ctc_loss_reduction", "sum"))

    def _check_labels(self, labels):
        if self.validate_labels and not torch.cuda.is_current_stream_capturing():
            if int(labels.max()) >= self.config.vocab_size:
                raise ValueError(f"Label values must be <= vocab_size: {self.config.vocab_size}")

    @torch.no_grad()
    def greedy_decode(self, logits, klen=None):
        """argmax + repeat/blank collapse on device (finetune_ASRs.py:113-122); returns (ids, list of id lists)."""
        ids, col, clen = ops.greedy_decode(logits, klen, int(self.config.pad_token_id))
        col, clen = col.cpu(), clen.cpu()
        return ids.long(), [col[b, : int(clen[b])].tolist() for b in range(col.shape[0])]


class Data2VecAudioForCTC(_CTCBase):
    """HF Data2VecAudioForCTC (HF:803-916) as instantiated by centralized/finetune_ASRs.py:164."""

    backbone_cls = Data2VecAudioModel

    def __init__(self, config, compute_dtype=torch.bfloat16):
        super().__init__()
        if config.vocab_size is None:
            raise ValueError("config.vocab_size must be defined for the CTC head")
        self.config = config
        setattr(self, self.base_model_prefix, self.backbone_cls(config, compute_dtype))
        self.lm_head = _Affine((config.vocab_size, config.hidden_size))

    def forward(self, input_values, attention_mask=None, output_attentions=None, output_hidden_states=None,
                return_dict=None, labels=None):
        if output_attentions:
            raise NotImplementedError("output_attentions is not available from the fused attention path")
        if labels is not None:
            self._check_labels(labels)
        h, klen, hs = self.backbone(input_values, attention_mask, bool(output_hidden_states))
        h = self.backbone.apply_dropout(h, "final_dropout", 2)  # HF:877-878
        logits = Fn.LinearFn.apply(h, self.lm_head.weight, self.lm_head.bias, self.backbone.cache)
        loss = None
        if labels is not None:
            loss = self._ctc(logits, _ctc_index_tensors(self.config, input_values, attention_mask, labels, klen))
        if return_dict is False:
            out = (logits,) + ((hs,) if hs is not None else ())
            return ((loss,) + out) if loss is not None else out
        return CausalLMOutput(loss=loss, logits=logits, hidden_states=hs, attentions=None)


class Wav2Vec2ForCTC(Data2VecAudioForCTC):
    """HF Wav2Vec2ForCTC as instantiated by centralized/finetune_ASRs.py:154-158 (facebook/wav2vec2-base-960h:
    feat_extract_norm='group', post-LN) and its large / lv60 configurations (layer norm stack, do_stable_layer_norm).
    state_dict keys: `wav2vec2.*`, `lm_head.*`."""
    base_model_prefix = "wav2vec2"
    backbone_cls = Wav2Vec2Model


class HubertForCTC(Data2VecAudioForCTC):
    """HF HubertForCTC as instantiated by centralized/finetune_ASRs.py:166-172 (facebook/hubert-xlarge-ls960-ft:
    layer-norm conv stack with bias, do_stable_layer_norm).  state_dict keys: `hubert.*`, `lm_head.*`."""
    base_model_prefix = "hubert"
    backbone_cls = HubertModel


class _CosFaceHolder(nn.Module):
    """`criterion_similar` (AngularPenaltySMLoss, models.py:131-185): holds fc.weight [2, d] (no bias)."""

    def __init__(self, d):
        super().__init__()
        self.fc = _Affine((2, d), bias=False)
        self.s, self.m = 30.0, 0.4


class DacsData2VecAudioForCTC(_CTCBase):
    """federated/src/models.py:262-631 `Data2VecAudioForCTC(config, args)`: encoder + arbitrator (toggling
    network) + lm_head + dementia_head + CosFace; STAGE 0/1/2 freezing and loss selection (:304-323, :612-620).
    `args` carries LAMBDA, TOGGLE_RATIO, GS_TAU, AD_loss, W_LOSS, STAGE (federated/src/options.py)."""

    def __init__(self, config, args, compute_dtype=torch.bfloat16):
        super().__init__()
        if config.vocab_size is None:
            raise ValueError("config.vocab_size must be defined for the CTC head")
        self.config = config
        d = config.hidden_size
        self.data2vec_audio = Data2VecAudioModel(config, compute_dtype)
        self.alpha = float(args.LAMBDA)
        self.TOGGLE_RATIO = float(args.TOGGLE_RATIO)
        self.GS_TAU = float(args.GS_TAU)
        self.AD_loss = args.AD_loss
        self.W_LOSS = [0.1, 0.9] if args.W_LOSS is None else list(args.W_LOSS)
        self.STAGE = int(args.STAGE)
        self.arbitrator = _Affine((4 * d, d))
        self.lm_head = _Affine((config.vocab_size, d))
        self.dementia_head = _Affine((2, d))
        self.criterion_similar = _CosFaceHolder(d)
        self.freeze_feature_encoder()
        frozen = {0: ("dementia_head", "arbitrator", "criterion_similar"),
                  1: ("data2vec_audio", "lm_head", "arbitrator", "criterion_similar"),
                  2: ("data2vec_audio", "lm_head", "dementia_head", "criterion_similar")}[self.STAGE]
        for name in frozen:
            for p in getattr(self, name).parameters():
                p.requires_grad = False
        self.gumbel_seed = 0
        self._gumbel_calls = 0
        # Philox offset of the in-kernel Gumbel draws, kept on the device and advanced by a device-side add, so that a
        # forward captured in a CUDA graph draws fresh noise on every replay (the reference draws fresh noise per call)
        self.register_buffer("gumbel_counter", torch.zeros(1, dtype=torch.int64), persistent=False)

    def _ad_weights(self):
        if self.AD_loss == "recall":
            return self.W_LOSS
        if self.AD_loss == "prec":
            return [0.1, 0.9]
        return [0.5, 0.5]

    def _ad(self, mean_logits, dementia_labels):
        w = None if self.AD_loss == "cel" else self._ad_weights()
        return Fn.AdLossFn.apply(mean_logits, dementia_labels, self.AD_loss, w)

    def forward(self, input_values, attention_mask=None, output_attentions=None, output_hidden_states=None,
                return_dict=None, labels=None, dementia_labels=None, gumbel_noise=None, return_masks=False):
        """Same signature as models.py:375-384 plus two optional extras: `gumbel_noise=(g_lm, g_ad)` fp32
        [B,T,d,2] externally drawn Gumbel samples (parity runs), `return_masks` for the `_eval` twin's dict."""
        if output_attentions:
            raise NotImplementedError("output_attentions is not available from the fused attention path")
        c = self.config
        cache = self.data2vec_audio.cache
        if labels is not None:
            self._check_labels(labels)
        h, klen, hs = self.data2vec_audio(input_values, attention_mask, bool(output_hidden_states))
        h = self.data2vec_audio.apply_dropout(h, "final_dropout", 2)  # models.py:403
        lin = lambda x, m: Fn.LinearFn.apply(x, m.weight, m.bias, cache)
        stage = self.STAGE
        g_lm, g_ad = gumbel_noise if gumbel_noise is not None else (None, None)
        self._gumbel_calls += 1
        self.gumbel_counter.add_(1)
        # stages 0/1 only report the masked logits (their loss never touches the toggling branch): no graph needed
        with (contextlib.nullcontext() if stage == 2 else torch.no_grad()):
            score = lin(h, self.arbitrator)
            lm_masked, ad_masked, lm_mask, ad_mask = Fn.DacsMaskFn.apply(
                score, h, self.GS_TAU, self.TOGGLE_RATIO, g_lm, g_ad, self.gumbel_seed, self.gumbel_counter)
            logits = lin(lm_masked, self.lm_head)  # the returned logits are always the lm-masked ones (:629-631)
        final_loss = None
        if labels is not None:
            idx = _ctc_index_tensors(c, input_values, attention_mask, labels, klen)
            if stage == 0:
                final_loss = self._ctc(lin(h, self.lm_head), idx)
            elif stage == 1:
                final_loss = self._ad(Fn.MeanTimeFn.apply(lin(h, self.dementia_head)), dementia_labels)
            else:
                loss = self._ctc(logits, idx)
                ad_rev = self._ad(Fn.GradReverseFn.apply(Fn.MeanTimeFn.apply(lin(lm_masked, self.dementia_head)), self.alpha),
                                  dementia_labels)
                loss_r = self._ctc(Fn.GradReverseFn.apply(lin(ad_masked, self.lm_head), self.alpha), idx)
                ad = self._ad(Fn.MeanTimeFn.apply(lin(ad_masked, self.dementia_head)), dementia_labels)
                cf = self.criterion_similar
                score_loss = Fn.CosFaceFn.apply(lm_masked, ad_masked, cf.fc.weight, cf.s, cf.m)
                final_loss = loss + ad_rev + loss_r + ad + score_loss
        if return_masks:  # Data2VecAudioForCTC_eval (models.py:999-1004)
            logits = {"ASR logits": logits, "dementia logits": lin(lm_masked, self.dementia_head), "hidden_states": h,
                      "lm_mask": lm_mask, "dementia_mask": ad_mask}
        if return_dict is False:
            out = (logits,) + ((hs,) if hs is not None else ())
            return ((final_loss,) + out) if final_loss is not None else out
        return CausalLMOutput(loss=final_loss, logits=logits, hidden_states=hs, attentions=None)
