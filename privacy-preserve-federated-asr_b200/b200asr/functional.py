"""Autograd wiring of the hot path at block granularity (one autograd node per encoder layer, not per op).

Each Function's forward/backward is a hand-scheduled sequence of C-ABI kernels (ops.py); torch only owns the
tensors and the autograd graph.  Saved-for-backward tensors are chosen so that no activation is recomputed
except the attention probabilities (rebuilt from the saved log-sum-exp).
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import ops
from ._lib import ACT_GELU, ACT_GELU_SAVE_GRAD, ACT_NONE


# ----------------------------------------------------------------------------- weight shadows
class ShadowCache:
    """Compute-dtype copies of fp32 master parameters, refreshed when the parameter changes.

    fp32 mode returns the parameter storage itself.  bf16 mode keeps a bf16 shadow (what the tcgen05 GEMM
    reads); derived layouts (fused QKV, conv filters flattened for the implicit GEMM, packed pos-conv filters)
    are cached under the same versioning.  Parameters re-homed into a FlatParams buffer whose owner keeps a flat
    bf16 shadow current (the fused AdamW writes it in the same pass as the fp32 update) are served as views of that
    shadow: no per-step cast or concat kernels.
    """

    def __init__(self, dtype: torch.dtype):
        self.dtype = dtype
        self._cache = {}

    @staticmethod
    def invalidate_all():
        """Kept for callers that wrote parameters through raw pointers without a FlatParams owner."""
        ShadowCache._global_epoch += 1

    _global_epoch = 0

    @staticmethod
    def _ver(ps):
        return (ShadowCache._global_epoch,) + tuple(
            (p.data_ptr(), p._version, getattr(p, "_b200_flat").epoch if hasattr(p, "_b200_flat") else 0) for p in ps)

    def _get(self, key, ps, make):
        ver = self._ver(ps)
        hit = self._cache.get(key)
        if hit is not None and hit[0] == ver:
            return hit[1]
        with torch.no_grad():
            val = make()
        self._cache[key] = (ver, val, ps, make)
        return val

    def refresh_in_place(self) -> int:
        """Derived copies of parameters that no FlatParams owns (the frozen conv stack: bf16 / flattened filters) whose
        source was rewritten through torch (load_state_dict) are recomputed INTO the cached tensors, so that kernels
        captured in a CUDA graph -- which hold those pointers -- see the new values.  Returns the number refreshed.
        (Flat-owned parameters are refreshed through FlatParams.sync_params / the in-graph layout kernels.)"""
        n = 0
        for key, (ver, val, ps, make) in list(self._cache.items()):
            if any(hasattr(p, "_b200_flat") for p in ps) or self._ver(ps) == ver:
                continue
            with torch.no_grad():
                new = make()
            olds = val if isinstance(val, (tuple, list)) else (val,)
            news = new if isinstance(new, (tuple, list)) else (new,)
            for o, nw in zip(olds, news):
                if o.data_ptr() != nw.data_ptr():
                    o.copy_(nw)
            self._cache[key] = (self._ver(ps), val, ps, make)
            n += 1
        return n

    def _cast(self, t: torch.Tensor) -> torch.Tensor:
        t = t.detach()
        if not t.is_contiguous():
            t = t.contiguous()
        return t if self.dtype == torch.float32 else ops.cast_from_f32(t, self.dtype)

    def _flat_shadow(self, ps):
        """View of the owner's flat bf16 shadow covering the adjacent parameters ps, or None."""
        if self.dtype != torch.bfloat16:
            return None
        fp = getattr(ps[0], "_b200_flat", None)
        if fp is None or not fp.shadow_valid():
            return None
        off = ps[0]._b200_off
        nxt = off
        for p in ps:
            if getattr(p, "_b200_flat", None) is not fp or p._b200_off != nxt or p._version != p._b200_pver:
                return None
            nxt += p.numel()
        return fp.shadow[off:nxt]

    def weight(self, p: torch.Tensor) -> torch.Tensor:
        v = self._flat_shadow((p,))
        if v is not None:
            return v.view(p.shape)
        return self._get(("w", id(p)), (p,), lambda: self._cast(p))

    def qkv(self, q, k, v) -> torch.Tensor:
        """[3d, d] = rows of q_proj, k_proj, v_proj (the packed projection the attention kernel consumes)."""
        f = self._flat_shadow((q, k, v))
        if f is not None:
            return f.view(3 * q.shape[0], q.shape[1])
        return self._get(("qkv", id(q)), (q, k, v), lambda: self._cast(torch.cat([q.detach(), k.detach(), v.detach()], 0)))

    def qkv_bias(self, q, k, v) -> torch.Tensor:
        fp = getattr(q, "_b200_flat", None)
        if (fp is not None and getattr(k, "_b200_flat", None) is fp and getattr(v, "_b200_flat", None) is fp
                and k._b200_off == q._b200_off + q.numel() and v._b200_off == k._b200_off + k.numel()):
            return fp.flat[q._b200_off:q._b200_off + 3 * q.numel()]  # adjacent fp32 biases: a view, always current
        return self._get(("qkvb", id(q)), (q, k, v), lambda: torch.cat([q.detach(), k.detach(), v.detach()], 0).contiguous())

    def conv_flat(self, w) -> torch.Tensor:
        """Conv1d filter [Cout, Cin, k] -> [Cout, k*Cin] matching channels-last windows (row = [tap][channel])."""
        return self._get(("convflat", id(w)), (w,),
                         lambda: self._cast(w.detach().permute(0, 2, 1).reshape(w.shape[0], -1)))

    def posconv(self, w, groups):
        return self._get(("posconv", id(w)), (w,), lambda: ops.posconv_pack(self.weight(w), groups))

    def posconv_wn(self, g, v, groups):
        """weight_norm(dim=2) filter of the wav2vec2 / hubert positional conv: (w_fwd, w_bwd, norm[k]) rebuilt when
        weight_g / weight_v change (once per optimizer step)."""
        def make():
            w, norm = ops.weight_norm_fwd(v.detach(), g.detach())
            wf, wb = ops.posconv_pack(self._cast(w), groups)
            return wf, wb, norm
        return self._get(("posconv_wn", id(v)), (g, v), make)


# ----------------------------------------------------------------------------- gradient sinks
def grad_sink(ps):
    """Direct-write destination for the gradient of the adjacent parameters `ps` (a tuple): parameters owned by a
    FlatParams buffer expose their slice of the flat fp32 gradient; backward kernels write there and autograd
    receives None for them (no per-parameter accumulate / copy launches).  Returns (flat view, overwrite) or
    (None, True) when the parameters are not flat-owned, not adjacent, or disagree about being fresh."""
    fp = getattr(ps[0], "_b200_flat", None)
    if fp is None or not fp.direct_grads:
        return None, True
    off = ps[0]._b200_off
    nxt = off
    fresh = ps[0]._b200_fresh
    for p in ps:
        if getattr(p, "_b200_flat", None) is not fp or p._b200_off != nxt or p._b200_fresh != fresh:
            return None, True
        nxt += p.numel()
    for p in ps:
        p._b200_fresh = False
        p.grad = p._b200_grad
    return fp.grad[off:nxt], fresh


def _wgrad(ps, need, dy, x, side=True):
    """dW = dy^T x for one parameter or several adjacent ones stacked along the output dim.
    side=False: on the current stream (the last weight gradient of a block's backward pass, see TAIL_WGRAD_ON_MAIN)."""
    if not need:
        return None
    ps = ps if isinstance(ps, tuple) else (ps,)
    sink, fresh = grad_sink(ps)
    if sink is None:
        return ops.linear_wgrad(dy, x)
    if not side:
        if not fresh:
            ops.join_side_stream()   # the sink's first write may still be in flight on a side stream
        ops.linear_wgrad(dy, x, out=sink.view(dy.shape[1], x.shape[1]), accumulate=not fresh)
        return None
    with ops.on_side_stream(dy, x, ordered=not fresh):
        ops.linear_wgrad(dy, x, out=sink.view(dy.shape[1], x.shape[1]), accumulate=not fresh)
    return None


def _bgrad(ps, need, dy):
    if not need:
        return None
    ps = ps if isinstance(ps, tuple) else (ps,)
    sink, fresh = grad_sink(ps)
    if sink is None:
        return ops.colsum(dy)
    if fresh:
        with ops.on_side_stream(dy):
            ops.colsum(dy, out=sink)
    else:
        ops.join_side_stream()   # the sink's first write may still be in flight on a side stream
        ops.axpy_(sink, ops.colsum(dy))
    return None


import os as _os
# The QKV weight gradient is the last GEMM of a post-LN block's backward pass.  On the side stream it is still running when
# the block's segment graph ends (graph.GraphedTrainStep cuts there and has to join the side stream): it is therefore
# issued on the main stream, before the data gradient, and only the bias column sums stay on the side stream.
# A/B at 8 x 10 s with layerdrop 0.1 (two runs each): 14.204 / 14.209 ms per step against 14.334 / 14.340 ms.
TAIL_WGRAD_ON_MAIN = _os.environ.get("B200ASR_TAIL_WGRAD_MAIN", "1") != "0"
# A/B on one B200 (profiles/r01_dropout_bgrad_ab_v17.txt): the fused pass saves 48 launches per step but puts the column
# sums on the main stream's critical path (14.86-14.90 ms vs 14.79-14.80 ms with the sums on the side stream) -> opt-in.
FUSED_DROPOUT_BGRAD = _os.environ.get("B200ASR_FUSED_DROPOUT_BGRAD", "0") != "0"


def _dropout_bgrad(ps, need, dy, p, seed, site, ctr):
    """do = dropout(dy) (the mask the forward drew at `site`) and the bias gradient colsum(do) of the Linear whose
    output was dropped, in one pass over dy (b200asr_dropout_colsum).  Returns (do, dbias or None if delivered)."""
    if not need:
        return ops.dropout(dy, p, seed, site, ctr), None
    if not FUSED_DROPOUT_BGRAD:  # dropout on the main stream, column sums on the side stream
        do = ops.dropout(dy, p, seed, site, ctr)
        return do, _bgrad(ps, need, do)
    ps = ps if isinstance(ps, tuple) else (ps,)
    sink, fresh = grad_sink(ps)
    if sink is None:
        return ops.dropout_colsum(dy, p, seed, site, ctr)
    if fresh:
        do, _ = ops.dropout_colsum(dy, p, seed, site, ctr, out=sink)
        return do, None
    ops.join_side_stream()   # the sink's first write may still be in flight on a side stream
    do, cs = ops.dropout_colsum(dy, p, seed, site, ctr)
    ops.axpy_(sink, cs)
    return do, None


def _ln_bwd(dy, s, w, b, mean, rstd, need, in_bias=None, drop=None, **kw):
    """LayerNorm backward.  Affine gradients go straight to their sinks when fresh.  `in_bias` = the bias parameter
    of the Linear / conv whose output fed this LayerNorm: its gradient is the column sum of dx, which the kernel
    emits in the same pass.  Returns (dx, dw, db, d_in_bias) with None for everything already delivered to a sink.
    drop = (p, seed, site, counter): the Linear's output went through a hidden dropout before the residual sum; returns
    (dx, dw, db, d_in_bias, dx_drop) with dx_drop = dropout(dx) from the same pass and d_in_bias = colsum(dx_drop)."""
    cols = s.shape[-1]
    dxsum = bias_sink = None
    bias_fresh = True
    if in_bias is not None and need:
        bias_sink, bias_fresh = grad_sink((in_bias,))
        dxsum = bias_sink if (bias_sink is not None and bias_fresh) else torch.empty(cols, dtype=torch.float32, device=s.device)
    dropped = []

    def run(**k):
        out = ops.layernorm_bwd(dy, s, w, b, mean, rstd, dxsum_out=dxsum, drop=drop, **k, **kw)
        if drop is not None:
            dropped.append(out[3])
        return out[:3]

    dw = db = None
    if need and w is not None:
        sw, fw = grad_sink((w,))
        sb, fb = grad_sink((b,))
        if sw is not None and sb is not None and fw and fb:
            dx, _, _ = run(dw_out=sw, db_out=sb)
        else:
            dx, dw, db = run()
            if sw is not None:
                if fw: sw.copy_(dw)
                else: ops.axpy_(sw, dw)
                dw = None
            if sb is not None:
                if fb: sb.copy_(db)
                else: ops.axpy_(sb, db)
                db = None
    else:
        dx, _, _ = run(want_wgrad=False)
    d_in_bias = None
    if dxsum is not None and dxsum is not bias_sink:
        if bias_sink is not None:
            ops.axpy_(bias_sink, dxsum)
        else:
            d_in_bias = dxsum
    if drop is not None:
        return dx, dw, db, d_in_bias, dropped[0]
    return dx, dw, db, d_in_bias


# ----------------------------------------------------------------------------- frozen feature extractor
def feature_extractor_fwd(x: torch.Tensor, conv_w: List[torch.Tensor], conv_b: List[Optional[torch.Tensor]],
                          ln_w: List[torch.Tensor], ln_b: List[torch.Tensor], kernels, strides, eps: float,
                          cache: ShadowCache, norm: str = "layer") -> torch.Tensor:
    """HF:51-75,131-157 with the stack frozen (finetune_ASRs.py:200, models.py:302): forward only.
    x [B, L] fp32 -> [B, T, C] channels-last.  Layer 0 is the fused direct kernel, layers 1.. are GEMMs whose
    A operand is the overlapping-window view of the channels-last activations (row pitch = stride*C, row length
    k*C), followed by the fused LayerNorm+GELU pass."""
    cd = cache.dtype
    group = norm == "group"  # wav2vec2-base / hubert-base: GroupNorm after conv 0, plain conv + GELU afterwards
    with torch.no_grad():
        conv0 = ops.conv0_gn_gelu if group else ops.conv0_ln_gelu
        h = conv0(x.float(), conv_w[0].detach(), None if conv_b[0] is None else conv_b[0].detach(),
                  ln_w[0].detach(), ln_b[0].detach(), strides[0], eps, cd)
        for i in range(1, len(conv_w)):
            B, Tin, Cin = h.shape
            k, s = kernels[i], strides[i]
            Tout = (Tin - k) // s + 1
            Cout = conv_w[i].shape[0]
            wf = cache.conv_flat(conv_w[i])
            y = torch.empty(B, Tout, Cout, dtype=cd, device=h.device)
            ops.gemm_raw(h, wf, y, Tout, Cout, k * Cin, s * Cin, k * Cin, Cout, True, True,
                         bias=None if conv_b[i] is None else conv_b[i].detach(), act=ACT_GELU if group else ACT_NONE,
                         batch0=1, batch1=B, sa=(0, Tin * Cin), sb=(0, 0), sc=(0, Tout * Cout))
            if group:
                h = y  # GELU rode in the GEMM epilogue; no norm on layers 1..6
            else:
                h, _, _, _ = ops.layernorm_fwd(y, ln_w[i].detach(), ln_b[i].detach(), eps, act=ACT_GELU, want_stats=False)
    return h


# ----------------------------------------------------------------------------- feature projection
class FeatureProjectionFn(torch.autograd.Function):
    """HF:160-172: LayerNorm(C) -> Linear(C -> d).  The input comes from the frozen extractor (no dx)."""

    @staticmethod
    def forward(ctx, feats, ln_w, ln_b, w, b, eps, cache: ShadowCache, klen=None):
        M = feats.numel() // feats.shape[-1]
        x2 = feats.reshape(M, feats.shape[-1])
        if ln_w is None:  # HubertFeatureProjection with feat_proj_layer_norm=False
            n, mean, rstd = x2, None, None
        else:
            n, _, mean, rstd = ops.layernorm_fwd(x2, ln_w, ln_b, eps)
        wsh = cache.weight(w)
        y, _ = ops.linear_fwd(n, wsh, b)
        ctx.save_for_backward(x2, n, mean, rstd, wsh, klen)
        ctx.params = (ln_w, ln_b, w, b)
        ctx.out_shape = feats.shape[:-1] + (w.shape[0],)
        y = y.view(ctx.out_shape)
        if klen is not None:  # hidden_states[~attention_mask] = 0 (HF:369-371), fused here: y is our own buffer
            ops.mask_frames_(y, klen)
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, n, mean, rstd, wsh, klen = ctx.saved_tensors
        ln_w, ln_b, w, b = ctx.params
        need = ctx.needs_input_grad
        if klen is not None:
            dy = ops.mask_frames_(dy.clone(memory_format=torch.contiguous_format), klen)
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        db = _bgrad(b, need[4], dy2)
        dw = _wgrad(w, need[3], dy2, n)
        dlw = dlb = None
        if ln_w is not None and (need[1] or need[2]):
            dn = ops.linear_dgrad(dy2, wsh)
            _, dlw, dlb, _ = _ln_bwd(dn, x2, ln_w, ln_b, mean, rstd, True)
        return None, dlw, dlb, dw, db, None, None, None


# ----------------------------------------------------------------------------- positional conv embedding + LN
class PosConvEmbedFn(torch.autograd.Function):
    """HF:89-128 + HF:381-383: y = LN(h + posconv(h)); posconv = n x {grouped conv -> LN(no affine) -> GELU}.
    args: h [B,T,d], enc_ln_w, enc_ln_b, then conv weights (n), then conv biases (n)."""

    @staticmethod
    def forward(ctx, h, ln_w, ln_b, eps, groups, ksize, cache: ShadowCache, *wb):
        n = len(wb) // 2
        ws, bs = wb[:n], wb[n:]
        B, T, d = h.shape
        saved = []
        y = h
        packs = []
        for j in range(n):
            wf, wbk = cache.posconv(ws[j], groups)
            packs.append(wbk)
            c = ops.posconv_fwd(y, wf, bs[j], groups, ksize)
            y_next, _, mean, rstd = ops.layernorm_fwd(c.view(B * T, d), None, None, 1e-5, act=ACT_GELU)
            saved += [y, c, mean, rstd]
            y = y_next.view(B, T, d)
        out, s, mean, rstd = ops.layernorm_fwd(h.reshape(B * T, d), ln_w, ln_b, eps, res=y.reshape(B * T, d), want_sum=True)
        ctx.save_for_backward(s, mean, rstd, *saved, *packs)
        ctx.params = (ln_w, ln_b, ws, bs)
        ctx.n, ctx.groups, ctx.ksize, ctx.shape = n, groups, ksize, (B, T, d)
        return out.view(B, T, d)

    @staticmethod
    def backward(ctx, dout):
        n, groups, k = ctx.n, ctx.groups, ctx.ksize
        B, T, d = ctx.shape
        t = ctx.saved_tensors
        s, mean, rstd = t[:3]
        saved = t[3:3 + 4 * n]
        packs = t[3 + 4 * n:]
        ln_w, ln_b, ws, bs = ctx.params
        need = ctx.needs_input_grad
        dout2 = dout.reshape(B * T, d)
        ds, dlw, dlb, _ = _ln_bwd(dout2, s, ln_w, ln_b, mean, rstd, need[1] or need[2])
        dy = ds  # gradient into posconv output; ds is also the residual-path gradient
        dws, dbs = [None] * n, [None] * n
        for j in range(n - 1, -1, -1):
            y_in, c, cm, cr = saved[4 * j:4 * j + 4]
            dc, _, _, dbs[j] = _ln_bwd(dy, c.view(B * T, d), None, None, cm, cr, need[7 + n + j], in_bias=bs[j], act=ACT_GELU)
            need_dx = j > 0 or need[0]
            want_dw = need[7 + j]
            sink, fresh = grad_sink((ws[j],)) if want_dw else (None, True)
            if sink is not None and fresh:
                dx, _ = ops.posconv_bwd(y_in, packs[j], dc.view(B, T, d), groups, k, want_dx=need_dx,
                                        dw_out=sink.view(ws[j].shape))
            else:
                dx, dws[j] = ops.posconv_bwd(y_in, packs[j], dc.view(B, T, d), groups, k, want_dx=need_dx, want_dw=want_dw)
                if sink is not None:
                    ops.axpy_(sink, dws[j].view(-1))
                    dws[j] = None
            dy = dx.view(B * T, d) if dx is not None else None
        dh = None
        if ctx.needs_input_grad[0]:
            dh = ops.axpy_(dy, ds).view(B, T, d)
        return (dh, dlw, dlb, None, None, None, None, *dws, *dbs)


# ----------------------------------------------------------------------------- transformer encoder layer
class EncoderLayerFn(torch.autograd.Function):
    """HF:203-346, post-LN block (data2vec / wav2vec2-base ordering), dropout = 0:
         a = O(softmax(scale * Q K^T + keymask) V);  x1 = LN1(x + a);  y = LN2(x1 + W2 gelu(W1 x1)).
    Residual adds, biases, GELU and gelu' ride in GEMM epilogues; the QKV projection is one N = 3d GEMM.
    params order: q_w,q_b,k_w,k_b,v_w,v_b,o_w,o_b,ln1_w,ln1_b,w1,b1,w2,b2,ln2_w,ln2_b"""

    @staticmethod
    def forward(ctx, x, klen, meta, cache: ShadowCache, *P):
        (q_w, q_b, k_w, k_b, v_w, v_b, o_w, o_b, ln1_w, ln1_b, w1, b1, w2, b2, ln2_w, ln2_b) = P
        B, T, H, eps = meta
        M, d = x.shape
        dh = d // H
        scale = dh ** -0.5
        wqkv = cache.qkv(q_w, k_w, v_w)
        bqkv = cache.qkv_bias(q_b, k_b, v_b)
        wo, w1s, w2s = cache.weight(o_w), cache.weight(w1), cache.weight(w2)
        qkv, _ = ops.linear_fwd(x, wqkv, bqkv)
        attn, lse = ops.mha_fwd(qkv, klen, B, T, H, dh, scale)
        s1, _ = ops.linear_fwd(attn, wo, o_b, residual=x)
        x1, _, m1, r1 = ops.layernorm_fwd(s1, ln1_w, ln1_b, eps)
        g, u = ops.linear_fwd(x1, w1s, b1, act=ACT_GELU_SAVE_GRAD)  # u = gelu'(pre-activation), from the same erf evaluation
        s2, _ = ops.linear_fwd(g, w2s, b2, residual=x1)
        y, _, m2, r2 = ops.layernorm_fwd(s2, ln2_w, ln2_b, eps)
        ctx.save_for_backward(x, qkv, attn, lse, s1, m1, r1, x1, u, g, s2, m2, r2, wqkv, wo, w1s, w2s)
        ctx.params = P
        ctx.klen = klen
        ctx.meta = (B, T, H, dh, scale)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x, qkv, attn, lse, s1, m1, r1, x1, u, g, s2, m2, r2, wqkv, wo, w1s, w2s) = ctx.saved_tensors
        (q_w, q_b, k_w, k_b, v_w, v_b, o_w, o_b, ln1_w, ln1_b, w1, b1, w2, b2, ln2_w, ln2_b) = ctx.params
        B, T, H, dh, scale = ctx.meta
        d = x.shape[1]
        wg = any(ctx.needs_input_grad[4:])  # parameter gradients wanted?
        ds2, dln2w, dln2b, db2 = _ln_bwd(dy, s2, ln2_w, ln2_b, m2, r2, wg, in_bias=b2)
        dw2 = _wgrad(w2, wg, ds2, g)
        du = ops.linear_dgrad(ds2, w2s, gelu_grad=u)              # (ds2 W2) * gelu'(pre-activation)
        db1 = _bgrad(b1, wg, du)
        dw1 = _wgrad(w1, wg, du, x1)
        dx1 = ops.linear_dgrad(du, w1s, residual=ds2)            # + residual path of s2 = x1 + ...
        ds1, dln1w, dln1b, dbo = _ln_bwd(dx1, s1, ln1_w, ln1_b, m1, r1, wg, in_bias=o_b)
        dwo = _wgrad(o_w, wg, ds1, attn)
        dattn = ops.linear_dgrad(ds1, wo)
        dqkv = ops.mha_bwd(qkv, attn, dattn, lse, ctx.klen, B, T, H, dh, scale)
        dqw = dkw = dvw = dqb = dkb = dvb = None
        dbqkv = dwqkv = None
        if wg and TAIL_WGRAD_ON_MAIN:
            dbqkv = _bgrad((q_b, k_b, v_b), True, dqkv)
            dwqkv = _wgrad((q_w, k_w, v_w), True, dqkv, x, side=False)
        dx = ops.linear_dgrad(dqkv, wqkv, residual=ds1) if ctx.needs_input_grad[0] else None
        if wg:
            if not TAIL_WGRAD_ON_MAIN:
                dbqkv = _bgrad((q_b, k_b, v_b), True, dqkv)
                dwqkv = _wgrad((q_w, k_w, v_w), True, dqkv, x)
            if dwqkv is not None:
                dqw, dkw, dvw = dwqkv[:d], dwqkv[d:2 * d], dwqkv[2 * d:]
            if dbqkv is not None:
                dqb, dkb, dvb = dbqkv[:d], dbqkv[d:2 * d], dbqkv[2 * d:]
        return (dx, None, None, None, dqw, dqb, dkw, dkb, dvw, dvb, dwo, dbo, dln1w, dln1b, dw1, db1, dw2, db2,
                dln2w, dln2b)


class SegmentMarkFn(torch.autograd.Function):
    """Identity placed between encoder blocks while graph.GraphedTrainStep captures an update with LayerDrop: its
    forward and its backward tell the recorder that a block boundary has been reached (the recorder closes the running
    CUDA graph and opens the next one) and hand it the tensor / gradient that crosses the boundary, which a replay copies
    across a skipped block."""

    @staticmethod
    def forward(ctx, x, idx, recorder):
        ctx.idx, ctx.recorder = idx, recorder
        recorder.fwd_mark(idx, x)
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        if not g.is_contiguous():
            g = g.contiguous()
        ctx.recorder.bwd_mark(ctx.idx, g)
        return g, None, None


class DropoutFn(torch.autograd.Function):
    """nn.Dropout at a stand-alone site (after the encoder's first LayerNorm HF:380-381, feat_proj_dropout HF:171,
    final_dropout federated/src/models.py:403): y = keep(x) / (1 - p); backward re-draws the same mask."""

    @staticmethod
    def forward(ctx, x, p, seed, site, counter):
        ctx.key = (p, seed, site, counter)
        return ops.dropout(x, p, seed, site, counter)

    @staticmethod
    def backward(ctx, dy):
        p, seed, site, counter = ctx.key
        return ops.dropout(dy, p, seed, site, counter), None, None, None, None


class EncoderLayerDropoutFn(torch.autograd.Function):
    """EncoderLayerFn with the block's element dropouts applied (HF:300-346):
         a = O(attention(x));  x1 = LN1(x + drop_h(a));  f = drop_a(gelu(W1 x1));  y = LN2(x1 + drop_h(W2 f)).
    Each dropout is one b200asr_dropout pass fused with the residual add that follows it; masks are re-drawn from
    (seed, site, counter) in the backward pass, never stored.  Dropout on the attention probabilities
    (attention_dropout, HF:189-197) is applied inside the attention kernels from keep bits drawn ahead of time
    (site0 + 3, ops.attn_dropout_mask): `keep` int32 bits of this block or None, `p_attn` its probability.
    drop = (p_hidden, p_act, seed, site0, counter, keep, p_attn[, (act_keep_bits, mpad)]): sites site0 .. site0+3."""

    @staticmethod
    def forward(ctx, x, klen, meta, cache: ShadowCache, drop, *P):
        (q_w, q_b, k_w, k_b, v_w, v_b, o_w, o_b, ln1_w, ln1_b, w1, b1, w2, b2, ln2_w, ln2_b) = P
        B, T, H, eps = meta
        ph, pa, seed, site, ctr, keep, pt, akeep, lazy_keep = (tuple(drop) + (None, None))[:9]
        M, d = x.shape
        dh = d // H
        scale = dh ** -0.5
        wqkv = cache.qkv(q_w, k_w, v_w)
        bqkv = cache.qkv_bias(q_b, k_b, v_b)
        wo, w1s, w2s = cache.weight(o_w), cache.weight(w1), cache.weight(w2)
        if lazy_keep and keep is not None:   # this block's attention keep bits, drawn on the side stream under the QKV GEMM
            with ops.on_side_stream(keep, ctr, manual_join=True):
                ops.attn_dropout_mask(1, B, T, H, pt, seed, site + 3, 4, ctr, out=keep)
        qkv, _ = ops.linear_fwd(x, wqkv, bqkv)
        if lazy_keep and keep is not None:
            ops.join_side_stream()
        attn, lse = ops.mha_fwd(qkv, klen, B, T, H, dh, scale, keep=keep, p_drop=pt)
        o, _ = ops.linear_fwd(attn, wo, o_b)
        # s1 = x + drop_h(o) and LN1 in one pass (b200asr_layernorm_dropout_fwd)
        x1, s1, m1, r1 = ops.layernorm_fwd(o, ln1_w, ln1_b, eps, res=x, want_sum=True, drop=(ph, seed, site, ctr))
        # activation dropout inside the W1 GEMM epilogue: g and the saved gelu' both carry mask_a / (1 - p_a), so the
        # backward pass needs no dropout pass either (fp32 parity mode: separate passes, same mask)
        fused_a = pa > 0.0 and ops.gemm_dropout_fusable(x1, w1s)
        # (akeep: the decisions were drawn ahead of time for all blocks, modeling._draw_activation_keep; else Philox in the epilogue)
        adrop = None if not fused_a else ((pa, akeep[0], akeep[1]) if akeep is not None else (pa, seed, site + 1, ctr))
        g, u = ops.linear_fwd(x1, w1s, b1, act=ACT_GELU_SAVE_GRAD, drop=adrop)
        if not fused_a:
            g = ops.dropout(g, pa, seed, site + 1, ctr, out=g)
        ctx.fused_a = fused_a
        o2, _ = ops.linear_fwd(g, w2s, b2)
        y, s2, m2, r2 = ops.layernorm_fwd(o2, ln2_w, ln2_b, eps, res=x1, want_sum=True, drop=(ph, seed, site + 2, ctr))
        ctx.save_for_backward(x, qkv, attn, lse, s1, m1, r1, x1, u, g, s2, m2, r2, wqkv, wo, w1s, w2s)
        ctx.params = P
        ctx.klen = klen
        ctx.meta = (B, T, H, dh, scale)
        ctx.drop = drop
        return y

    @staticmethod
    def backward(ctx, dy):
        (x, qkv, attn, lse, s1, m1, r1, x1, u, g, s2, m2, r2, wqkv, wo, w1s, w2s) = ctx.saved_tensors
        (q_w, q_b, k_w, k_b, v_w, v_b, o_w, o_b, ln1_w, ln1_b, w1, b1, w2, b2, ln2_w, ln2_b) = ctx.params
        B, T, H, dh, scale = ctx.meta
        ph, pa, seed, site, ctr, keep, pt = ctx.drop[:7]
        d = x.shape[1]
        wg = any(ctx.needs_input_grad[5:])  # parameter gradients wanted?
        # LN2 backward emits ds2 (residual path), do2 = hidden-dropout backward of it (gradient of W2 f + b2) and b2's gradient
        ds2, dln2w, dln2b, db2, do2 = _ln_bwd(dy, s2, ln2_w, ln2_b, m2, r2, wg, in_bias=b2, drop=(ph, seed, site + 2, ctr))
        dw2 = _wgrad(w2, wg, do2, g)                              # g = the dropped activations W2 consumed
        du = ops.linear_dgrad(do2, w2s, gelu_grad=u)              # (do2 W2) * gelu'
        if not ctx.fused_a:
            du = ops.dropout(du, pa, seed, site + 1, ctr, out=du)     # ... * mask_a / (1 - p_a) (the factors commute)
        db1 = _bgrad(b1, wg, du)
        dw1 = _wgrad(w1, wg, du, x1)
        dx1 = ops.linear_dgrad(du, w1s, residual=ds2)
        ds1, dln1w, dln1b, dbo, do = _ln_bwd(dx1, s1, ln1_w, ln1_b, m1, r1, wg, in_bias=o_b, drop=(ph, seed, site, ctr))
        dwo = _wgrad(o_w, wg, do, attn)
        dattn = ops.linear_dgrad(do, wo)
        dqkv = ops.mha_bwd(qkv, attn, dattn, lse, ctx.klen, B, T, H, dh, scale, keep=keep, p_drop=pt)
        dqw = dkw = dvw = dqb = dkb = dvb = None
        dbqkv = dwqkv = None
        if wg and TAIL_WGRAD_ON_MAIN:
            dbqkv = _bgrad((q_b, k_b, v_b), True, dqkv)
            dwqkv = _wgrad((q_w, k_w, v_w), True, dqkv, x, side=False)
        dx = ops.linear_dgrad(dqkv, wqkv, residual=ds1) if ctx.needs_input_grad[0] else None
        if wg:
            if not TAIL_WGRAD_ON_MAIN:
                dbqkv = _bgrad((q_b, k_b, v_b), True, dqkv)
                dwqkv = _wgrad((q_w, k_w, v_w), True, dqkv, x)
            if dwqkv is not None:
                dqw, dkw, dvw = dwqkv[:d], dwqkv[d:2 * d], dwqkv[2 * d:]
            if dbqkv is not None:
                dqb, dkb, dvb = dbqkv[:d], dbqkv[d:2 * d], dbqkv[2 * d:]
        return (dx, None, None, None, None, dqw, dqb, dkw, dkb, dvw, dvb, dwo, dbo, dln1w, dln1b, dw1, db1, dw2, db2,
                dln2w, dln2b)


def _deliver(p, grad):
    """Hand a finished fp32 gradient to p's flat-gradient sink (copy when fresh, add otherwise); returns what autograd
    should receive (None when delivered)."""
    sink, fresh = grad_sink((p,))
    if sink is None:
        return grad
    if fresh:
        sink.copy_(grad.reshape(-1))
    else:
        ops.axpy_(sink, grad.reshape(-1).contiguous())
    return None


class PosConvWeightNormFn(torch.autograd.Function):
    """Wav2Vec2PositionalConvEmbedding / HubertPositionalConvEmbedding + the encoder's residual add (and LayerNorm for
    the post-LN encoder): s = h + gelu(conv_{g * v / ||v||}(h)),  out = LN(s) or s (do_stable_layer_norm).
    transformers wav2vec2/modeling_wav2vec2.py: PositionalConvEmbedding, Wav2Vec2Encoder.forward,
    Wav2Vec2EncoderStableLayerNorm.forward."""

    @staticmethod
    def forward(ctx, h, ln_w, ln_b, eps, groups, ksize, cache: ShadowCache, g, v, bias):
        B, T, d = h.shape
        wf, wb, norm = cache.posconv_wn(g, v, groups)
        c = ops.posconv_fwd(h, wf, bias, groups, ksize)
        s = ops.gelu_add(c, h)
        mean = rstd = None
        out = s
        if ln_w is not None:
            out, _, mean, rstd = ops.layernorm_fwd(s.view(B * T, d), ln_w, ln_b, eps)
        ctx.save_for_backward(h, c, s if ln_w is not None else None, mean, rstd, wb, norm)
        ctx.params = (ln_w, ln_b, g, v, bias)
        ctx.cfg = (groups, ksize, (B, T, d))
        return out.view(B, T, d)

    @staticmethod
    def backward(ctx, dout):
        h, c, s, mean, rstd, wb, norm = ctx.saved_tensors
        ln_w, ln_b, g, v, bias = ctx.params
        groups, k, (B, T, d) = ctx.cfg
        need = ctx.needs_input_grad
        dout2 = dout.reshape(B * T, d)
        if not dout2.is_contiguous():
            dout2 = dout2.contiguous()
        dlw = dlb = None
        ds = dout2
        if ln_w is not None:
            ds, dlw, dlb, _ = _ln_bwd(dout2, s.view(B * T, d), ln_w, ln_b, mean, rstd, need[1] or need[2])
        dc = ops.gelu_bwd(ds, c.view(B * T, d))
        dbias = _bgrad(bias, need[9], dc)
        want_dw = need[7] or need[8]
        dx, dW = ops.posconv_bwd(h, wb, dc.view(B, T, d), groups, k, want_dx=need[0], want_dw=want_dw)
        dg = dv = None
        if want_dw:
            dv, dg = ops.weight_norm_bwd(dW, v.detach(), g.detach(), norm)
            dv, dg = _deliver(v, dv), _deliver(g, dg)
        dh = None
        if need[0]:
            dh = ops.axpy_(dx.view(B * T, d), ds).view(B, T, d)
        return (dh, dlw, dlb, None, None, None, None, dg, dv, dbias)


class LayerNormFn(torch.autograd.Function):
    """A stand-alone affine LayerNorm over the last dim (the final encoder.layer_norm of the pre-LN encoder)."""

    @staticmethod
    def forward(ctx, x, w, b, eps):
        x2 = x.reshape(-1, x.shape[-1])
        y, _, mean, rstd = ops.layernorm_fwd(x2, w, b, eps)
        ctx.save_for_backward(x2, mean, rstd)
        ctx.params = (w, b)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, mean, rstd = ctx.saved_tensors
        w, b = ctx.params
        need = ctx.needs_input_grad
        dy2 = dy.reshape(x2.shape)
        dx, dw, db, _ = _ln_bwd(dy2, x2, w, b, mean, rstd, need[1] or need[2])
        return dx.view(dy.shape), dw, db, None


class EncoderLayerStableFn(torch.autograd.Function):
    """Pre-LN block (wav2vec2/modeling_wav2vec2.py Wav2Vec2EncoderLayerStableLayerNorm; wav2vec2-large-lv60,
    hubert-large/xlarge), dropout = 0:
         s1 = x + O(attn(LN1(x)));   y = s1 + W2 gelu(W1 LN2(s1)).
    Same kernels as the post-LN block; the residual adds ride in the O / W2 GEMM epilogues and in the LayerNorm
    backward (`dres_add`).  params order as EncoderLayerFn (ln1 = layer_norm, ln2 = final_layer_norm)."""

    @staticmethod
    def forward(ctx, x, klen, meta, cache: ShadowCache, *P):
        (q_w, q_b, k_w, k_b, v_w, v_b, o_w, o_b, ln1_w, ln1_b, w1, b1, w2, b2, ln2_w, ln2_b) = P
        B, T, H, eps = meta
        M, d = x.shape
        dh = d // H
        scale = dh ** -0.5
        wqkv = cache.qkv(q_w, k_w, v_w)
        bqkv = cache.qkv_bias(q_b, k_b, v_b)
        wo, w1s, w2s = cache.weight(o_w), cache.weight(w1), cache.weight(w2)
        n1, _, m1, r1 = ops.layernorm_fwd(x, ln1_w, ln1_b, eps)
        qkv, _ = ops.linear_fwd(n1, wqkv, bqkv)
        attn, lse = ops.mha_fwd(qkv, klen, B, T, H, dh, scale)
        s1, _ = ops.linear_fwd(attn, wo, o_b, residual=x)
        n2, _, m2, r2 = ops.layernorm_fwd(s1, ln2_w, ln2_b, eps)
        g, u = ops.linear_fwd(n2, w1s, b1, act=ACT_GELU_SAVE_GRAD)
        y, _ = ops.linear_fwd(g, w2s, b2, residual=s1)
        ctx.save_for_backward(x, m1, r1, n1, qkv, attn, lse, s1, m2, r2, n2, u, g, wqkv, wo, w1s, w2s)
        ctx.params = P
        ctx.klen = klen
        ctx.meta = (B, T, H, dh, scale)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x, m1, r1, n1, qkv, attn, lse, s1, m2, r2, n2, u, g, wqkv, wo, w1s, w2s) = ctx.saved_tensors
        (q_w, q_b, k_w, k_b, v_w, v_b, o_w, o_b, ln1_w, ln1_b, w1, b1, w2, b2, ln2_w, ln2_b) = ctx.params
        B, T, H, dh, scale = ctx.meta
        d = x.shape[1]
        wg = any(ctx.needs_input_grad[4:])
        if not dy.is_contiguous():
            dy = dy.contiguous()
        db2 = _bgrad(b2, wg, dy)
        dw2 = _wgrad(w2, wg, dy, g)
        du = ops.linear_dgrad(dy, w2s, gelu_grad=u)
        db1 = _bgrad(b1, wg, du)
        dw1 = _wgrad(w1, wg, du, n2)
        dn2 = ops.linear_dgrad(du, w1s)
        ds1, dln2w, dln2b, _ = _ln_bwd(dn2, s1, ln2_w, ln2_b, m2, r2, wg, dres_add=dy)   # + residual path of y
        dbo = _bgrad(o_b, wg, ds1)
        dwo = _wgrad(o_w, wg, ds1, attn)
        dattn = ops.linear_dgrad(ds1, wo)
        dqkv = ops.mha_bwd(qkv, attn, dattn, lse, ctx.klen, B, T, H, dh, scale)
        dqw = dkw = dvw = dqb = dkb = dvb = None
        if wg:
            dbqkv = _bgrad((q_b, k_b, v_b), True, dqkv)
            dwqkv = _wgrad((q_w, k_w, v_w), True, dqkv, n1)
            if dwqkv is not None:
                dqw, dkw, dvw = dwqkv[:d], dwqkv[d:2 * d], dwqkv[2 * d:]
            if dbqkv is not None:
                dqb, dkb, dvb = dbqkv[:d], dbqkv[d:2 * d], dbqkv[2 * d:]
        dx = dln1w = dln1b = None
        if ctx.needs_input_grad[0] or wg:
            dn1 = ops.linear_dgrad(dqkv, wqkv)
            dx, dln1w, dln1b, _ = _ln_bwd(dn1, x, ln1_w, ln1_b, m1, r1, wg, dres_add=ds1)  # + residual path of s1
            if not ctx.needs_input_grad[0]:
                dx = None
        return (dx, None, None, None, dqw, dqb, dkw, dkb, dvw, dvb, dwo, dbo, dln1w, dln1b, dw1, db1, dw2, db2,
                dln2w, dln2b)


class EncoderLayerStableDropoutFn(torch.autograd.Function):
    """EncoderLayerStableFn with the block's element dropouts (Wav2Vec2EncoderLayerStableLayerNorm.forward):
         s1 = x + drop_h(O(attn(LN1(x))));   y = s1 + drop_h(W2 drop_a(gelu(W1 LN2(s1)))).
    drop = (p_hidden, p_act, seed, site0, counter, keep, p_attn) as in EncoderLayerDropoutFn."""

    @staticmethod
    def forward(ctx, x, klen, meta, cache: ShadowCache, drop, *P):
        (q_w, q_b, k_w, k_b, v_w, v_b, o_w, o_b, ln1_w, ln1_b, w1, b1, w2, b2, ln2_w, ln2_b) = P
        B, T, H, eps = meta
        ph, pa, seed, site, ctr, keep, pt, akeep, lazy_keep = (tuple(drop) + (None, None))[:9]
        M, d = x.shape
        dh = d // H
        scale = dh ** -0.5
        wqkv = cache.qkv(q_w, k_w, v_w)
        bqkv = cache.qkv_bias(q_b, k_b, v_b)
        wo, w1s, w2s = cache.weight(o_w), cache.weight(w1), cache.weight(w2)
        n1, _, m1, r1 = ops.layernorm_fwd(x, ln1_w, ln1_b, eps)
        if lazy_keep and keep is not None:
            with ops.on_side_stream(keep, ctr, manual_join=True):
                ops.attn_dropout_mask(1, B, T, H, pt, seed, site + 3, 4, ctr, out=keep)
        qkv, _ = ops.linear_fwd(n1, wqkv, bqkv)
        if lazy_keep and keep is not None:
            ops.join_side_stream()
        attn, lse = ops.mha_fwd(qkv, klen, B, T, H, dh, scale, keep=keep, p_drop=pt)
        o, _ = ops.linear_fwd(attn, wo, o_b)
        # s1 = x + drop_h(o) and LN2 in one pass (b200asr_layernorm_dropout_fwd)
        n2, s1, m2, r2 = ops.layernorm_fwd(o, ln2_w, ln2_b, eps, res=x, want_sum=True, drop=(ph, seed, site, ctr))
        # activation dropout inside the W1 GEMM epilogue: g and the saved gelu' both carry mask_a / (1 - p_a), so the
        # backward pass needs no dropout pass either (fp32 parity mode: separate passes, same mask)
        fused_a = pa > 0.0 and ops.gemm_dropout_fusable(n2, w1s)
        # (akeep: the decisions were drawn ahead of time for all blocks, modeling._draw_activation_keep; else Philox in the epilogue)
        adrop = None if not fused_a else ((pa, akeep[0], akeep[1]) if akeep is not None else (pa, seed, site + 1, ctr))
        g, u = ops.linear_fwd(n2, w1s, b1, act=ACT_GELU_SAVE_GRAD, drop=adrop)
        if not fused_a:
            g = ops.dropout(g, pa, seed, site + 1, ctr, out=g)
        ctx.fused_a = fused_a
        o2, _ = ops.linear_fwd(g, w2s, b2)
        y = ops.dropout(o2, ph, seed, site + 2, ctr, residual=s1, out=o2)
        ctx.save_for_backward(x, m1, r1, n1, qkv, attn, lse, s1, m2, r2, n2, u, g, wqkv, wo, w1s, w2s)
        ctx.params = P
        ctx.klen = klen
        ctx.meta = (B, T, H, dh, scale)
        ctx.drop = drop
        return y

    @staticmethod
    def backward(ctx, dy):
        (x, m1, r1, n1, qkv, attn, lse, s1, m2, r2, n2, u, g, wqkv, wo, w1s, w2s) = ctx.saved_tensors
        (q_w, q_b, k_w, k_b, v_w, v_b, o_w, o_b, ln1_w, ln1_b, w1, b1, w2, b2, ln2_w, ln2_b) = ctx.params
        B, T, H, dh, scale = ctx.meta
        ph, pa, seed, site, ctr, keep, pt = ctx.drop[:7]
        d = x.shape[1]
        wg = any(ctx.needs_input_grad[5:])
        if not dy.is_contiguous():
            dy = dy.contiguous()
        do2, db2 = _dropout_bgrad(b2, wg, dy, ph, seed, site + 2, ctr)
        dw2 = _wgrad(w2, wg, do2, g)
        du = ops.linear_dgrad(do2, w2s, gelu_grad=u)
        if not ctx.fused_a:
            du = ops.dropout(du, pa, seed, site + 1, ctr, out=du)
        db1 = _bgrad(b1, wg, du)
        dw1 = _wgrad(w1, wg, du, n2)
        dn2 = ops.linear_dgrad(du, w1s)
        # LN2 backward (+ the residual path of y) emits ds1, do = hidden-dropout backward of it (gradient of O attn + o_b) and
        # o_b's gradient in one pass
        ds1, dln2w, dln2b, dbo, do = _ln_bwd(dn2, s1, ln2_w, ln2_b, m2, r2, wg, in_bias=o_b, drop=(ph, seed, site, ctr),
                                             dres_add=dy)
        dwo = _wgrad(o_w, wg, do, attn)
        dattn = ops.linear_dgrad(do, wo)
        dqkv = ops.mha_bwd(qkv, attn, dattn, lse, ctx.klen, B, T, H, dh, scale, keep=keep, p_drop=pt)
        dqw = dkw = dvw = dqb = dkb = dvb = None
        if wg:
            dbqkv = _bgrad((q_b, k_b, v_b), True, dqkv)
            dwqkv = _wgrad((q_w, k_w, v_w), True, dqkv, n1)
            if dwqkv is not None:
                dqw, dkw, dvw = dwqkv[:d], dwqkv[d:2 * d], dwqkv[2 * d:]
            if dbqkv is not None:
                dqb, dkb, dvb = dbqkv[:d], dbqkv[d:2 * d], dbqkv[2 * d:]
        dx = dln1w = dln1b = None
        if ctx.needs_input_grad[0] or wg:
            dn1 = ops.linear_dgrad(dqkv, wqkv)
            dx, dln1w, dln1b, _ = _ln_bwd(dn1, x, ln1_w, ln1_b, m1, r1, wg, dres_add=ds1)  # + residual path of s1
            if not ctx.needs_input_grad[0]:
                dx = None
        return (dx, None, None, None, None, dqw, dqb, dkw, dkb, dvw, dvb, dwo, dbo, dln1w, dln1b, dw1, db1, dw2, db2,
                dln2w, dln2b)


# ----------------------------------------------------------------------------- heads
class LinearFn(torch.autograd.Function):
    """nn.Linear on [.., K] activations (lm_head, dementia_head, arbitrator: models.py:292-296)."""

    @staticmethod
    def forward(ctx, x, w, b, cache: ShadowCache):
        K = x.shape[-1]
        x2 = x.reshape(-1, K)
        wsh = cache.weight(w)
        y, _ = ops.linear_fwd(x2, wsh, b)
        ctx.save_for_backward(x2, wsh)
        ctx.params = (w, b)
        return y.view(x.shape[:-1] + (w.shape[0],))

    @staticmethod
    def backward(ctx, dy):
        x2, wsh = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = ops.linear_dgrad(dy2, wsh).view(dy.shape[:-1] + (wsh.shape[1],))
        w, b = ctx.params
        dw = _wgrad(w, ctx.needs_input_grad[1], dy2, x2)
        db = _bgrad(b, ctx.needs_input_grad[2], dy2) if b is not None else None
        return dx, dw, db, None


class GradReverseFn(torch.autograd.Function):
    """ReverseLayerF (models.py:61-73): identity forward, grad * (-lambda) backward."""

    @staticmethod
    def forward(ctx, x, lam):
        ctx.lam = float(lam)
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return ops.scale_(g.clone(memory_format=torch.contiguous_format), -ctx.lam), None


class CTCLossFn(torch.autograd.Function):
    """log_softmax(fp32) + F.ctc_loss (models.py:498-533; HF:897-908), loss and d loss / d logits in one launch.
    Returns the reduced loss (0-dim fp32): reduction 'sum' or 'mean' (per-utterance / target length, batch mean)."""

    @staticmethod
    def forward(ctx, logits, targets, tgt_off, tgt_len, in_len, max_tgt, blank, zero_infinity, reduction):
        B = logits.shape[0]
        gs = None
        if reduction == "mean":
            gs = (1.0 / (tgt_len.clamp(min=1).float() * B)).view(1, B).contiguous()
        loss, grad = ops.ctc_fwd_bwd(logits.unsqueeze(0), targets, tgt_off, tgt_len, in_len, max_tgt, blank,
                                     zero_infinity, grad_scale=gs, want_grad=ctx.needs_input_grad[0])
        ctx.save_for_backward(grad)
        ctx.shape = logits.shape
        per = loss[0]
        if reduction == "mean":
            return (per * gs[0]).sum()
        return per.sum()

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        if grad is None:
            return (None,) * 9
        out = ops.scale_(grad.clone(), 1.0, g.reshape(1).float().contiguous())
        return (out.view(ctx.shape),) + (None,) * 8


class DacsMaskFn(torch.autograd.Function):
    """Fused gumbel_softmax(hard) masks + masked hidden states (models.py:428-455, :75-129)."""

    @staticmethod
    def forward(ctx, scores, h, tau, ratio, noise_lm, noise_ad, seed, offset):
        """offset: host int, or a device int64 counter tensor (graph-replayable draws)."""
        d = h.shape[-1]
        s2, h2 = scores.reshape(-1, 4 * d), h.reshape(-1, d)
        dev_off = offset if isinstance(offset, torch.Tensor) else None
        lm, ad, sl, sa, ml, ma = ops.dacs_mask_fwd(s2, h2, tau, ratio, noise_lm, noise_ad, seed,
                                                   0 if dev_off is not None else offset, dev_offset=dev_off)
        ctx.save_for_backward(h2, sl, sa, ml, ma)
        ctx.cfg = (tau, ratio, scores.shape, h.shape)
        ctx.mark_non_differentiable(ml, ma)
        return lm.view(h.shape), ad.view(h.shape), ml.view(h.shape), ma.view(h.shape)

    @staticmethod
    def backward(ctx, d_lm, d_ad, _dml, _dma):
        h2, sl, sa, ml, ma = ctx.saved_tensors
        tau, ratio, sshape, hshape = ctx.cfg
        d = h2.shape[1]
        if d_lm is None:
            d_lm = torch.zeros_like(h2)
        if d_ad is None:
            d_ad = torch.zeros_like(h2)
        ds, dh = ops.dacs_mask_bwd(d_lm.reshape(-1, d), d_ad.reshape(-1, d), h2, sl, sa, ml, ma, tau, ratio,
                                   want_dh=ctx.needs_input_grad[1])
        return ds.view(sshape), (dh.view(hshape) if dh is not None else None), None, None, None, None, None, None


class MeanTimeFn(torch.autograd.Function):
    """torch.mean(x, dim=1) of per-frame AD logits (models.py:468-476) -> fp32 [B, C]."""

    @staticmethod
    def forward(ctx, x):
        ctx.T, ctx.dtype = x.shape[1], x.dtype
        return ops.mean_time_fwd(x)

    @staticmethod
    def backward(ctx, g):
        return ops.mean_time_bwd(g.float(), ctx.T, ctx.dtype)


class AdLossFn(torch.autograd.Function):
    """RecallLoss / CrossEntropy on [B, C] mean AD logits (models.py:187-260, :535-541)."""

    @staticmethod
    def forward(ctx, logits, target, kind, weights):
        loss, dl = ops.ad_loss(logits.float(), target, kind, weights, want_grad=ctx.needs_input_grad[0])
        ctx.save_for_backward(dl)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return ops.scale_(dl.clone(), 1.0, g.reshape(1).float().contiguous()), None, None, None


class CosFaceFn(torch.autograd.Function):
    """AngularPenaltySMLoss('cosface') on cat(lm_masked, AD_masked) with labels 0/1 (models.py:131-185, :592-607)."""

    @staticmethod
    def forward(ctx, x1, x2, fc_w, s, m):
        d = x1.shape[-1]
        need = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        loss, dx1, dx2 = ops.cosface_fwd_bwd(x1.reshape(-1, d), x2.reshape(-1, d), fc_w.detach().float().contiguous(), s, m,
                                             want_grad=need)
        ctx.save_for_backward(dx1, dx2)
        ctx.shape = x1.shape
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        dx1, dx2 = ctx.saved_tensors
        if dx1 is None:
            return None, None, None, None, None
        gs = g.reshape(1).float().contiguous()
        return (ops.scale_(dx1, 1.0, gs).view(ctx.shape), ops.scale_(dx2, 1.0, gs).view(ctx.shape), None, None, None)
