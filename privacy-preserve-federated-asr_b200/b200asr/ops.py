"""Raw (non-autograd) operators: torch tensors in, C-ABI calls on the current CUDA stream, torch tensors out.

torch is used for device memory and streams only.  Every function requires CUDA tensors and raises if the
library reports an error -- there is no eager fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import ACT_GELU, ACT_GELU_SAVE_GRAD, ACT_NONE, BF16, F32, GemmArgs, check, lib

_DT = {torch.float32: F32, torch.bfloat16: BF16}


def dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"b200asr supports float32 and bfloat16 tensors, got {t.dtype}") from None


def _p(t: Optional[torch.Tensor]):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("b200asr operators need CUDA tensors (no CPU fallback)")
    return t.data_ptr()


def _st():
    return torch.cuda.current_stream().cuda_stream


def _c(t: torch.Tensor) -> torch.Tensor:
    return t if t.is_contiguous() else t.contiguous()


_WS = {}

# bench.py's roofline leg: when set to a list, every GEMM / attention call is bracketed by CUDA events on the
# launching stream and (kind, start, end, flops) is appended.  None (the default) adds no work.
PROFILE = None


def _prof_begin():
    if PROFILE is None:
        return None
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def _prof_end(kind, e0, flops):
    if e0 is not None:
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record()
        PROFILE.append((kind, e0, e1, flops))


def workspace(nbytes: int, device) -> torch.Tensor:
    """Scratch buffer shared by the library calls of one stream (they are serialised on it)."""
    key = (torch.device(device).index, torch.cuda.current_stream().cuda_stream, torch.cuda.is_current_stream_capturing())
    buf = _WS.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _WS[key] = buf
    return buf


# ----------------------------------------------------------------------------- side stream (weight gradients)
# The weight / bias gradient kernels of a layer depend only on tensors that already exist when the data-gradient
# chain moves on, so they run on a second stream: their CTAs fill the SMs the persistent GEMM's last wave leaves idle.
import os as _os

SIDE_STREAM_ENABLED = _os.environ.get("B200ASR_SIDE_STREAM", "1") != "0"
# B200ASR_SIDE_STREAMS > 1 rotates several side streams (independent weight-gradient GEMMs that each fill only part of the machine
# could then overlap); measured 1 vs 2 vs 3 streams: 13.86 / 13.85 / 13.91 ms per step, so one stream stays the default.
N_SIDE_STREAMS = max(1, int(_os.environ.get("B200ASR_SIDE_STREAMS", "1")))
_SIDE = {}
_SIDE_NEXT = [0]
_SIDE_DIRTY = set()
_SIDE_CB_QUEUED = [False]


def _side_stream():
    dev = torch.cuda.current_device()
    sts = _SIDE.get(dev)
    if sts is None:
        sts = _SIDE[dev] = [torch.cuda.Stream(device=dev) for _ in range(N_SIDE_STREAMS)]
    i = _SIDE_NEXT[0] % len(sts)
    _SIDE_NEXT[0] = i + 1
    return sts[i]


def join_side_stream():
    """Make the current stream wait for everything queued on the side streams (end of backward, optimizer step)."""
    _SIDE_CB_QUEUED[0] = False
    dev = torch.cuda.current_device()
    if dev in _SIDE_DIRTY:
        for st in _SIDE[dev]:
            torch.cuda.current_stream().wait_stream(st)
        _SIDE_DIRTY.discard(dev)


class on_side_stream:
    """Context manager: fork the side stream from the current one and run the enclosed launches there.  `tensors` are
    the operands read inside (their memory must not be recycled by the caching allocator until the side stream is
    done with them).  Outputs must be long-lived buffers (the flat gradient)."""

    def __init__(self, *tensors, ordered=False, manual_join=False):
        self.tensors = tensors
        self.ctx = None
        self.ordered = ordered   # the launches update a buffer an earlier side-stream launch wrote: order after all of them
        self.manual_join = manual_join   # the caller joins (join_side_stream) before it consumes the results

    def __enter__(self):
        if not SIDE_STREAM_ENABLED:
            return self
        side = _side_stream()
        side.wait_stream(torch.cuda.current_stream())
        if self.ordered:
            for other in _SIDE[torch.cuda.current_device()]:
                if other is not side:
                    side.wait_stream(other)
        for t in self.tensors:
            if t is not None:
                t.record_stream(side)
        self.ctx = torch.cuda.stream(side)
        self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.ctx is None:
            return False
        self.ctx.__exit__(*exc)
        _SIDE_DIRTY.add(torch.cuda.current_device())
        if self.manual_join:
            return False
        if not _SIDE_CB_QUEUED[0]:
            try:  # joined automatically when the running backward pass ends
                torch.autograd.Variable._execution_engine.queue_callback(join_side_stream)
                _SIDE_CB_QUEUED[0] = True
            except RuntimeError:  # not inside a backward pass: join right away
                join_side_stream()
        return False


# ----------------------------------------------------------------------------- norm / elementwise
def layernorm_fwd(x, w, b, eps, res=None, act=ACT_NONE, want_sum=False, want_stats=True, drop=None):
    """y = act(LN(x (+res)) * w + b) over the last dim.  Returns (y, s_or_None, mean, rstd).
    drop = (p, seed, site, counter): x goes through an inverted dropout first (keys as ops.dropout), i.e.
    s = res + dropout(x), in the same pass."""
    x = _c(x)
    cols = x.shape[-1]
    rows = x.numel() // cols
    y = torch.empty_like(x)
    s = torch.empty_like(x) if (want_sum and res is not None) else None
    mean = torch.empty(rows, dtype=torch.float32, device=x.device) if want_stats else None
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device) if want_stats else None
    if res is not None:
        res = _c(res)
    if drop is not None:
        p, seed, site, counter = drop
        dev = counter if isinstance(counter, torch.Tensor) else None
        host = 0 if dev is not None or counter is None else int(counter)
        check(lib.b200asr_layernorm_dropout_fwd(_p(x), _p(res), _p(w), _p(b), _p(y), _p(s), _p(mean), _p(rstd), rows, cols,
                                                float(eps), act, float(p), int(seed) & (2 ** 64 - 1), int(site), host,
                                                _p(dev), dt(x), _st()), "layernorm_dropout_fwd")
        return y, s, mean, rstd
    check(lib.b200asr_layernorm_fwd(_p(x), _p(res), _p(w), _p(b), _p(y), _p(s), _p(mean), _p(rstd), rows, cols,
                                    float(eps), act, dt(x), _st()), "layernorm_fwd")
    return y, s, mean, rstd


def layernorm_bwd(dy, s, w, b, mean, rstd, act=ACT_NONE, dres_add=None, want_wgrad=True, dw_out=None, db_out=None,
                  dxsum_out=None, drop=None):
    """Returns (dx, dw, db); dw/db are None when the LN has no affine or want_wgrad is False.
    dw_out/db_out: optional preallocated fp32 [cols] destinations (overwritten).
    dxsum_out: optional fp32 [cols] receiving the column sums of dx (bias gradient of the producing Linear).
    drop = (p, seed, site, counter): returns (dx, dw, db, dx_drop) with dx_drop = ops.dropout(dx, ...) written by the same
    pass (b200asr_layernorm_bwd_dropout); dxsum_out then receives the column sums of dx_drop."""
    dy = _c(dy)
    s = _c(s)
    cols = s.shape[-1]
    rows = s.numel() // cols
    dx = torch.empty_like(s)
    dw = db = ws = None
    nws = 0
    if w is not None and want_wgrad:
        dw = dw_out if dw_out is not None else torch.empty(cols, dtype=torch.float32, device=s.device)
        db = db_out if db_out is not None else torch.empty(cols, dtype=torch.float32, device=s.device)
    if dw is not None or dxsum_out is not None:
        nws = lib.b200asr_layernorm_bwd_workspace_bytes(cols)
        ws = workspace(nws, s.device)
    if drop is not None:
        pd, seed, site, counter = drop
        dev = counter if isinstance(counter, torch.Tensor) else None
        host = 0 if dev is not None or counter is None else int(counter)
        dxd = torch.empty_like(s)
        check(lib.b200asr_layernorm_bwd_dropout(_p(dy), _p(s), _p(w), _p(b), _p(mean), _p(rstd), _p(dres_add), _p(dx), _p(dxd),
                                                _p(dw), _p(db), _p(dxsum_out), _p(ws), nws, rows, cols, act, float(pd),
                                                int(seed) & (2 ** 64 - 1), int(site), host, _p(dev), dt(s), _st()),
              "layernorm_bwd_dropout")
        return dx, dw, db, dxd
    check(lib.b200asr_layernorm_bwd(_p(dy), _p(s), _p(w), _p(b), _p(mean), _p(rstd), _p(dres_add), _p(dx), _p(dw),
                                    _p(db), _p(dxsum_out), _p(ws), nws, rows, cols, act, dt(s), _st()), "layernorm_bwd")
    return dx, dw, db


def colsum(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 column sums of x viewed as [rows, cols] (bias gradient); `out` fp32 [cols] is overwritten if given."""
    x = _c(x)
    cols = x.shape[-1]
    rows = x.numel() // cols
    if out is None:
        out = torch.empty(cols, dtype=torch.float32, device=x.device)
    nws = lib.b200asr_colsum_workspace_bytes(cols)
    ws = workspace(nws, x.device)
    check(lib.b200asr_colsum(_p(x), _p(out), _p(ws), nws, rows, cols, cols, dt(x), _st()), "colsum")
    return out


def dropout_colsum(x: torch.Tensor, p: float, seed: int, site: int, counter=None, out: Optional[torch.Tensor] = None):
    """(y, colsum): y = dropout(x) (keys as ops.dropout) and the fp32 column sums of y, one pass over x [rows, cols]."""
    x = _c(x)
    cols = x.shape[-1]
    rows = x.numel() // cols
    y = torch.empty_like(x)
    if out is None:
        out = torch.empty(cols, dtype=torch.float32, device=x.device)
    dev = counter if isinstance(counter, torch.Tensor) else None
    host = 0 if dev is not None or counter is None else int(counter)
    nws = lib.b200asr_colsum_workspace_bytes(cols)
    ws = workspace(nws, x.device)
    check(lib.b200asr_dropout_colsum(_p(x), _p(y), _p(out), _p(ws), nws, rows, cols, float(p), int(seed) & (2 ** 64 - 1),
                                     int(site), host, _p(dev), dt(x), _st()), "dropout_colsum")
    return y, out


def conv_out_lengths(lengths: torch.Tensor, kernels, strides) -> torch.Tensor:
    """int32 frame lengths after the conv feature extractor for int64 sample counts `lengths` [n] on the device
    (b200asr_conv_out_lengths: the whole floor((len - k) / s) + 1 chain in one launch)."""
    assert lengths.is_cuda and lengths.dtype == torch.int64 and lengths.is_contiguous() and len(kernels) == len(strides)
    out = torch.empty(lengths.numel(), dtype=torch.int32, device=lengths.device)
    nl = len(kernels)
    ks = (C.c_int * nl)(*[int(k) for k in kernels])
    ss = (C.c_int * nl)(*[int(v) for v in strides])
    check(lib.b200asr_conv_out_lengths(_p(lengths), _p(out), lengths.numel(), ks, ss, nl, _st()), "conv_out_lengths")
    return out.view(lengths.shape)


def mask_frames_(x: torch.Tensor, lengths: torch.Tensor) -> torch.Tensor:
    """x [B, T, C]: zero frames t >= lengths[b] in place."""
    B, T, Cc = x.shape
    assert x.is_contiguous() and lengths.dtype == torch.int32
    check(lib.b200asr_mask_frames(_p(x), _p(lengths), B, T, Cc, dt(x), _st()), "mask_frames")
    return x


def cast_from_f32(src: torch.Tensor, dtype: torch.dtype, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    assert src.dtype == torch.float32
    src = _c(src)
    if out is None:
        out = torch.empty(src.shape, dtype=dtype, device=src.device)
    check(lib.b200asr_cast_f32(_p(src), _p(out), src.numel(), _DT[dtype], _st()), "cast_f32")
    return out


def cast_to_f32(src: torch.Tensor) -> torch.Tensor:
    src = _c(src)
    out = torch.empty(src.shape, dtype=torch.float32, device=src.device)
    check(lib.b200asr_cast_to_f32(_p(src), _p(out), src.numel(), dt(src), _st()), "cast_to_f32")
    return out


def scale_(x: torch.Tensor, host_scale: float = 1.0, dev_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    assert x.is_contiguous()
    if dev_scale is not None:
        assert dev_scale.dtype == torch.float32 and dev_scale.numel() == 1
    check(lib.b200asr_scale(_p(x), x.numel(), float(host_scale), _p(dev_scale), dt(x), _st()), "scale")
    return x


def axpy_(y: torch.Tensor, x: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
    assert y.is_contiguous() and x.is_contiguous() and y.dtype == x.dtype and y.numel() == x.numel()
    check(lib.b200asr_axpy(_p(y), _p(x), float(alpha), y.numel(), dt(y), _st()), "axpy")
    return y


# ----------------------------------------------------------------------------- feature extractor layer 0
def conv0_ln_gelu(x, w, bias, ln_w, ln_b, stride: int, eps: float, out_dtype: torch.dtype) -> torch.Tensor:
    """x [B, L] fp32 -> [B, Tout, C] channels-last (HF:51-75, layer 0)."""
    x = _c(x)
    B, L = x.shape
    Cc, cin, k = w.shape
    assert cin == 1 and x.dtype == torch.float32
    Tout = (L - k) // stride + 1
    y = torch.empty(B, Tout, Cc, dtype=out_dtype, device=x.device)
    check(lib.b200asr_conv0_ln_gelu_fwd(_p(x), _p(_c(w)), _p(bias), _p(ln_w), _p(ln_b), _p(y), B, L, Cc, k, stride,
                                        float(eps), _DT[out_dtype], _st()), "conv0_ln_gelu_fwd")
    return y


def conv0_gn_gelu(x, w, bias, gn_w, gn_b, stride: int, eps: float, out_dtype: torch.dtype) -> torch.Tensor:
    """x [B, L] fp32 -> [B, Tout, C] channels-last: Conv1d(1->C) + GroupNorm(C groups) + GELU (wav2vec2
    GroupNormConvLayer, feat_extract_norm='group')."""
    x = _c(x)
    B, L = x.shape
    Cc, cin, k = w.shape
    assert cin == 1 and x.dtype == torch.float32
    Tout = (L - k) // stride + 1
    y = torch.empty(B, Tout, Cc, dtype=out_dtype, device=x.device)
    nws = lib.b200asr_conv0_gn_gelu_workspace_bytes(B, L, Cc, k, stride)
    ws = workspace(nws, x.device)
    check(lib.b200asr_conv0_gn_gelu_fwd(_p(x), _p(_c(w)), _p(bias), _p(gn_w), _p(gn_b), _p(y), _p(ws), nws, B, L, Cc, k,
                                        stride, float(eps), _DT[out_dtype], _st()), "conv0_gn_gelu_fwd")
    return y


def weight_norm_fwd(v: torch.Tensor, g: torch.Tensor):
    """weight_norm(dim=2): v fp32 [C, Cg, k], g fp32 [1, 1, k] -> (w fp32 [C, Cg, k], norm fp32 [k])."""
    v, g = _c(v), _c(g)
    assert v.dtype == torch.float32 and g.dtype == torch.float32
    k = v.shape[-1]
    rows = v.numel() // k
    w = torch.empty_like(v)
    norm = torch.empty(k, dtype=torch.float32, device=v.device)
    nws = lib.b200asr_weight_norm_workspace_bytes(k)
    ws = workspace(nws, v.device)
    check(lib.b200asr_weight_norm_fwd(_p(v), _p(g), _p(w), _p(norm), _p(ws), nws, rows, k, _st()), "weight_norm_fwd")
    return w, norm


def weight_norm_bwd(dw, v, g, norm, dv_out=None, dg_out=None):
    """(dv [C, Cg, k], dg [1, 1, k]) from dw fp32 [C, Cg, k]; optional preallocated fp32 destinations."""
    dw, v, g = _c(dw), _c(v), _c(g)
    k = v.shape[-1]
    rows = v.numel() // k
    dv = dv_out if dv_out is not None else torch.empty_like(v)
    dg = dg_out if dg_out is not None else torch.empty(g.shape, dtype=torch.float32, device=v.device)
    nws = lib.b200asr_weight_norm_workspace_bytes(k)
    ws = workspace(nws, v.device)
    check(lib.b200asr_weight_norm_bwd(_p(dw), _p(v), _p(g), _p(norm), _p(dv), _p(dg), _p(ws), nws, rows, k, _st()),
          "weight_norm_bwd")
    return dv, dg


def gelu_add(c: torch.Tensor, res: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = res + gelu(c)."""
    c = _c(c)
    if res is not None:
        res = _c(res)
        assert res.dtype == c.dtype and res.numel() == c.numel()
    y = torch.empty_like(c)
    check(lib.b200asr_gelu_add_fwd(_p(c), _p(res), _p(y), c.numel(), dt(c), _st()), "gelu_add_fwd")
    return y


def gelu_bwd(dy: torch.Tensor, c: torch.Tensor) -> torch.Tensor:
    """dc = dy * gelu'(c)."""
    dy, c = _c(dy), _c(c)
    dc = torch.empty_like(c)
    check(lib.b200asr_gelu_bwd(_p(dy), _p(c), _p(dc), c.numel(), dt(c), _st()), "gelu_bwd")
    return dc


def dropout(x: torch.Tensor, p: float, seed: int, site: int, counter=None, residual: Optional[torch.Tensor] = None,
            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = residual + keep(x) / (1 - p) (torch.nn.functional.dropout semantics, b200asr_dropout).  The mask is a pure
    function of (seed, site, counter, element index): the backward pass calls this again on dy with the same keys.
    counter: host int, or a device int64 tensor [1] (graph-replayable draws)."""
    x = _c(x)
    if residual is not None:
        residual = _c(residual)
        assert residual.shape == x.shape and residual.dtype == x.dtype
    y = torch.empty_like(x) if out is None else out
    dev = counter if isinstance(counter, torch.Tensor) else None
    host = 0 if dev is not None or counter is None else int(counter)
    check(lib.b200asr_dropout(_p(x), _p(residual), _p(y), x.numel(), float(p), int(seed) & (2 ** 64 - 1), int(site),
                              host, _p(dev), dt(x), _st()), "dropout")
    return y


# ----------------------------------------------------------------------------- GEMM
def gemm_raw(A, B, Cout, M, N, K, lda, ldb, ldc, a_kmajor=True, b_kmajor=True, bias=None, residual=None,
             aux_out=None, aux_in=None, act=ACT_NONE, mul_gelu_grad=False, alpha=1.0, batch0=1, batch1=1,
             sa=(0, 0), sb=(0, 0), sc=(0, 0), a_off=0, b_off=0, c_off=0, backend: Optional[int] = None, drop=None):
    """C[m,n] = epi(sum_k A(m,k) B(n,k)); offsets in elements.  See include/b200asr.h b200asr_gemm."""
    g = GemmArgs()
    esa, esc = A.element_size(), Cout.element_size()
    g.A = A.data_ptr() + a_off * esa
    g.B = B.data_ptr() + b_off * esa
    g.C = Cout.data_ptr() + c_off * esc
    g.bias = _p(bias)
    g.residual = (residual.data_ptr() + c_off * esc) if residual is not None else None
    g.aux_out = (aux_out.data_ptr() + c_off * esc) if aux_out is not None else None
    g.aux_in = (aux_in.data_ptr() + c_off * esc) if aux_in is not None else None
    g.M, g.N, g.K = M, N, K
    g.lda, g.ldb, g.ldc = lda, ldb, ldc
    g.a_kmajor, g.b_kmajor = int(a_kmajor), int(b_kmajor)
    g.act, g.mul_gelu_grad, g.alpha = act, int(mul_gelu_grad), float(alpha)  # mul_gelu_grad: 0 | 1 (gelu'(aux_in)) | 2 (aux_in)
    g.batch0, g.batch1 = batch0, batch1
    g.sa0, g.sa1 = sa
    g.sb0, g.sb1 = sb
    g.sc0, g.sc1 = sc
    assert A.dtype == B.dtype
    g.dtype, g.c_dtype = dt(A), dt(Cout)
    if bias is not None:
        assert bias.dtype == torch.float32
    e0 = _prof_begin()
    if drop is not None and len(drop) == 3:   # (p, keep_bits, mpad): decisions drawn ahead of time (dropout_keep_bits)
        pd, bits, mpad = drop
        assert bits.dtype == torch.int32 and bits.is_contiguous() and bits.numel() >= dropout_keep_words(N, mpad)
        check(lib.b200asr_gemm_dropout_bits(C.byref(g), _p(bits), int(mpad), float(pd), _st()), "gemm_dropout_bits")
    elif drop is not None:   # (p, seed, site, counter): element dropout on C and aux_out inside the GELU_SAVE_GRAD epilogue
        pd, seed, site, counter = drop
        dev = counter if isinstance(counter, torch.Tensor) else None
        host = 0 if dev is not None or counter is None else int(counter)
        check(lib.b200asr_gemm_dropout(C.byref(g), float(pd), int(seed) & (2 ** 64 - 1), int(site), host, _p(dev), _st()),
              "gemm_dropout")
    elif backend is None:
        check(lib.b200asr_gemm(C.byref(g), _st()), "gemm")
    else:
        check(lib.b200asr_gemm_backend(C.byref(g), backend, _st()), "gemm_backend")
    _prof_end("gemm", e0, 2.0 * M * N * K * batch0 * batch1)
    return Cout


def dropout_keep_mpad(rows: int) -> int:
    """Row pitch of the keep-bit words for a GEMM with `rows` output rows (covers the largest tile's row overhang)."""
    return (int(rows) + 511) // 512 * 512


def dropout_keep_words(cols: int, mpad: int) -> int:
    return int(lib.b200asr_dropout_keep_words(int(cols), int(mpad)))


def dropout_keep_bits(nsites: int, site0: int, site_stride: int, rows: int, cols: int, p: float, seed: int, counter=None,
                      out: Optional[torch.Tensor] = None, device=None) -> torch.Tensor:
    """Keep bits of `nsites` dropout sites over a [rows, cols] tensor each (b200asr_dropout_keep_bits): int32
    [nsites, cols / 32 * mpad], the decisions of ops.dropout(p, seed, site0 + i * site_stride, counter)."""
    mpad = dropout_keep_mpad(rows)
    words = dropout_keep_words(cols, mpad)
    dev = counter if isinstance(counter, torch.Tensor) else None
    host = 0 if dev is not None or counter is None else int(counter)
    if out is None:
        out = torch.empty(nsites, words, dtype=torch.int32, device=device if device is not None else dev.device)
    assert out.dtype == torch.int32 and out.is_contiguous() and out.numel() >= nsites * words
    check(lib.b200asr_dropout_keep_bits(_p(out), int(nsites), int(site0), int(site_stride), int(rows), int(cols), mpad, float(p),
                                        int(seed) & (2 ** 64 - 1), host, _p(dev), _st()), "dropout_keep_bits")
    return out


def gemm_dropout_fusable(x, w, out_dtype=None) -> bool:
    """True when linear_fwd(..., act=ACT_GELU_SAVE_GRAD, drop=...) can apply the dropout in the GEMM epilogue (tcgen05
    back end, bf16 in and out, aligned rows); otherwise the caller runs ops.dropout on the outputs."""
    return (x.dtype == torch.bfloat16 and (out_dtype or x.dtype) == torch.bfloat16 and x.shape[-1] % 8 == 0
            and w.shape[0] % 8 == 0 and x.data_ptr() % 16 == 0 and w.data_ptr() % 16 == 0)


def linear_fwd(x, w, bias=None, act=ACT_NONE, residual=None, want_preact=False, out_dtype=None, drop=None):
    """y = act(x @ w.T + bias) (+ residual).  x [M, K], w [N, K] (same dtype).  Returns (y, preact_or_None);
    with act=ACT_GELU_SAVE_GRAD the second tensor holds gelu'(pre-activation) (feed it to linear_dgrad(gelu_grad=)).
    drop = (p, seed, site, counter) with ACT_GELU_SAVE_GRAD: y and the second tensor both come out multiplied by the
    dropout mask / (1 - p) of ops.dropout's keys (b200asr_gemm_dropout; check gemm_dropout_fusable first)."""
    if act == ACT_GELU_SAVE_GRAD:
        want_preact = True
    x = _c(x)
    M, K = x.shape
    N = w.shape[0]
    y = torch.empty(M, N, dtype=out_dtype or x.dtype, device=x.device)
    pre = torch.empty_like(y) if want_preact else None
    gemm_raw(x, w, y, M, N, K, K, K, N, True, True, bias=bias, residual=residual, aux_out=pre, act=act, drop=drop)
    return y, pre


def linear_dgrad(dy, w, residual=None, gelu_preact=None, gelu_grad=None):
    """dx = dy @ w  (* gelu'(preact) | * gelu_grad) (+ residual).  dy [M, N], w [N, K] -> [M, K]."""
    dy = _c(dy)
    M, N = dy.shape
    K = w.shape[1]
    dx = torch.empty(M, K, dtype=dy.dtype, device=dy.device)
    aux = gelu_grad if gelu_grad is not None else gelu_preact
    gemm_raw(dy, w, dx, M, K, N, N, K, K, True, False, residual=residual, aux_in=aux,
             mul_gelu_grad=2 if gelu_grad is not None else int(gelu_preact is not None))
    return dx


def linear_wgrad(dy, x, out=None, accumulate=False):
    """dw = dy.T @ x in fp32.  dy [M, N], x [M, K] -> [N, K].  `out` (fp32, contiguous [N, K]) receives the result
    (added to its contents when `accumulate`: the epilogue's residual operand aliases the output)."""
    dy = _c(dy)
    x = _c(x)
    M, N = dy.shape
    K = x.shape[1]
    dw = out if out is not None else torch.empty(N, K, dtype=torch.float32, device=dy.device)
    gemm_raw(dy, x, dw, N, K, M, N, K, K, False, False, residual=dw if (accumulate and out is not None) else None)
    return dw


# ----------------------------------------------------------------------------- attention
def attn_keep_words(B, T, H) -> int:
    return int(lib.b200asr_attn_keep_words(B, T, H))


def attn_dropout_mask(n_sites, B, T, H, p, seed, site0, site_stride, counter=None, out=None) -> torch.Tensor:
    """Keep bits of the attention-probability dropout for `n_sites` sites (encoder blocks) in one launch:
    int32 [n_sites, words] (b200asr_attn_dropout_mask layout).  counter: host int or device int64 [1] step counter."""
    words = attn_keep_words(B, T, H)
    if out is None:
        out = torch.empty(n_sites, words, dtype=torch.int32, device="cuda")
    assert out.dtype == torch.int32 and out.is_contiguous() and out.numel() >= n_sites * words
    dev = counter if isinstance(counter, torch.Tensor) else None
    host = 0 if dev is not None or counter is None else int(counter)
    check(lib.b200asr_attn_dropout_mask(_p(out), n_sites, B, T, H, float(p), int(seed) & (2 ** 64 - 1), int(site0),
                                        int(site_stride), host, _p(dev), _st()), "attn_dropout_mask")
    return out


def attn_keep_dense(keep: torch.Tensor, B, T, H) -> torch.Tensor:
    """Test helper: the keep bits of one site as a bool tensor [B, H, T(query), T(key)]."""
    KW, TQ = (T + 31) // 32, (T + 127) // 128 * 128
    w = keep.view(B, H, KW, TQ)[..., :T].to(torch.int64) & 0xffffffff       # [B, H, KW, Tq]
    bits = (w.unsqueeze(-1) >> torch.arange(32, device=keep.device)) & 1    # [B, H, KW, Tq, 32]
    return bits.permute(0, 1, 3, 2, 4).reshape(B, H, T, KW * 32)[..., :T].bool()


def mha_fwd(qkv, klen, B, T, H, dh, scale, keep=None, p_drop=0.0):
    """qkv [B*T, 3*H*dh] packed; returns (ctx [B*T, H*dh], lse [B, H, T]).  keep: int32 keep bits of this site
    (attn_dropout_mask) -> dropout with probability p_drop on the softmax probabilities."""
    assert qkv.is_contiguous()
    out = torch.empty(B * T, H * dh, dtype=qkv.dtype, device=qkv.device)
    lse = torch.empty(B, H, T, dtype=torch.float32, device=qkv.device)
    nws = lib.b200asr_mha_fwd_workspace_bytes(B, T, H, dh, dt(qkv))
    ws = workspace(nws, qkv.device)
    e0 = _prof_begin()
    check(lib.b200asr_mha_fwd_dropout(_p(qkv), _p(out), _p(lse), _p(klen), _p(keep), float(p_drop), _p(ws), nws, B, T, H, dh,
                                      float(scale), dt(qkv), _st()), "mha_fwd")
    _prof_end("mha_fwd", e0, 4.0 * B * H * T * T * dh)
    return out, lse


def mha_bwd(qkv, out, dout, lse, klen, B, T, H, dh, scale, keep=None, p_drop=0.0):
    dout = _c(dout)
    dqkv = torch.empty_like(qkv)
    nws = lib.b200asr_mha_bwd_workspace_bytes(B, T, H, dh, dt(qkv))
    ws = workspace(nws, qkv.device)
    e0 = _prof_begin()
    check(lib.b200asr_mha_bwd_dropout(_p(qkv), _p(out), _p(dout), _p(lse), _p(klen), _p(keep), float(p_drop), _p(dqkv),
                                      _p(ws), nws, B, T, H, dh, float(scale), dt(qkv), _st()), "mha_bwd")
    _prof_end("mha_bwd", e0, 10.0 * B * H * T * T * dh)
    return dqkv


# ----------------------------------------------------------------------------- positional conv
def posconv_pack(w: torch.Tensor, groups: int):
    """HF weight [C, Cg, k] (compute dtype) -> (w_fwd, w_bwd) packed [groups, k, Cg, Cg]."""
    w = _c(w)
    Cc, Cg, k = w.shape
    wf = torch.empty(groups, k, Cg, Cg, dtype=w.dtype, device=w.device)
    wb = torch.empty_like(wf)
    check(lib.b200asr_posconv_pack_weight(_p(w), _p(wf), _p(wb), Cc, groups, k, dt(w), _st()), "posconv_pack_weight")
    return wf, wb


def posconv_fwd(x, w_fwd, bias, groups, k):
    x = _c(x)
    B, T, Cc = x.shape
    y = torch.empty_like(x)
    check(lib.b200asr_posconv_fwd(_p(x), _p(w_fwd), _p(bias), _p(y), B, T, Cc, groups, k, dt(x), _st()), "posconv_fwd")
    return y


def posconv_bwd(x, w_bwd, dy, groups, k, want_dx=True, want_dw=True, dw_out=None):
    x, dy = _c(x), _c(dy)
    B, T, Cc = x.shape
    dx = torch.empty_like(x) if want_dx else None
    dw = None
    if want_dw:
        dw = dw_out if dw_out is not None else torch.empty(Cc, Cc // groups, k, dtype=torch.float32, device=x.device)
    nws = lib.b200asr_posconv_bwd_workspace_bytes(Cc, groups, k) if want_dw else 0
    ws = workspace(nws, x.device) if want_dw else None
    check(lib.b200asr_posconv_bwd(_p(x), _p(w_bwd), _p(dy), _p(dx), _p(dw), _p(ws), nws, B, T, Cc, groups, k, dt(x),
                                  _st()), "posconv_bwd")
    return dx, dw


# ----------------------------------------------------------------------------- CTC / greedy decode
def ctc_fwd_bwd(logits, targets, tgt_off, tgt_len, in_len, max_tgt_len, blank, zero_infinity, grad_scale=None,
                want_grad=True):
    """logits [n_heads, B, T, V]; int32 index tensors.  Returns (loss [n_heads, B] fp32, grad or None)."""
    logits = _c(logits)
    nh, B, T, V = logits.shape
    loss = torch.empty(nh, B, dtype=torch.float32, device=logits.device)
    grad = torch.empty_like(logits) if want_grad else None
    nws = lib.b200asr_ctc_workspace_bytes(nh, B, T, max(1, max_tgt_len))
    ws = workspace(nws, logits.device)
    check(lib.b200asr_ctc_fwd_bwd(_p(logits), _p(targets), _p(tgt_off), _p(tgt_len), _p(in_len), _p(grad_scale),
                                  _p(loss), _p(grad), _p(ws), nws, nh, B, T, V, max_tgt_len, blank,
                                  int(bool(zero_infinity)), dt(logits), _st()), "ctc_fwd_bwd")
    return loss, grad


def greedy_decode(logits, in_len=None, blank=0):
    """logits [B, T, V] -> (ids [B,T] int32, collapsed [B,T] int32 padded -1, collapsed_len [B] int32)."""
    logits = _c(logits)
    B, T, V = logits.shape
    ids = torch.empty(B, T, dtype=torch.int32, device=logits.device)
    col = torch.empty(B, T, dtype=torch.int32, device=logits.device)
    clen = torch.empty(B, dtype=torch.int32, device=logits.device)
    check(lib.b200asr_greedy_decode(_p(logits), _p(in_len), _p(ids), _p(col), _p(clen), B, T, V, blank, dt(logits),
                                    _st()), "greedy_decode")
    return ids, col, clen


# ----------------------------------------------------------------------------- DACS
def dacs_mask_fwd(scores, h, tau, toggle_ratio, noise_lm=None, noise_ad=None, seed=0, offset=0, want_masks=True,
                  dev_offset=None):
    """scores [M, 4d], h [M, d] -> (lm_masked, ad_masked, soft_lm, soft_ad, lm_mask, ad_mask)."""
    scores, h = _c(scores), _c(h)
    M, d = h.shape
    outs = [torch.empty_like(h) for _ in range(6 if want_masks else 4)]
    lm_mask, ad_mask = (outs[4], outs[5]) if want_masks else (None, None)
    if noise_lm is not None:
        assert noise_lm.dtype == torch.float32 and noise_ad.dtype == torch.float32
        noise_lm, noise_ad = _c(noise_lm), _c(noise_ad)
    if dev_offset is not None:
        assert dev_offset.dtype == torch.int64 and dev_offset.numel() == 1 and dev_offset.is_cuda
    check(lib.b200asr_dacs_mask_fwd_ctr(_p(scores), _p(h), _p(noise_lm), _p(noise_ad), _p(outs[0]), _p(outs[1]), _p(outs[2]),
                                        _p(outs[3]), _p(lm_mask), _p(ad_mask), M, d, float(tau), float(toggle_ratio),
                                        int(seed), int(offset), _p(dev_offset), dt(h), _st()), "dacs_mask_fwd")
    return outs[0], outs[1], outs[2], outs[3], lm_mask, ad_mask


def dacs_mask_bwd(d_lm, d_ad, h, soft_lm, soft_ad, lm_mask, ad_mask, tau, toggle_ratio, want_dh=False):
    d_lm, d_ad = _c(d_lm), _c(d_ad)
    M, d = h.shape
    d_scores = torch.empty(M, 4 * d, dtype=h.dtype, device=h.device)
    dh = torch.empty_like(h) if want_dh else None
    check(lib.b200asr_dacs_mask_bwd(_p(d_lm), _p(d_ad), _p(h), _p(soft_lm), _p(soft_ad), _p(lm_mask), _p(ad_mask),
                                    _p(d_scores), _p(dh), M, d, float(tau), float(toggle_ratio), dt(h), _st()),
          "dacs_mask_bwd")
    return d_scores, dh


def mean_time_fwd(x):
    x = _c(x)
    B, T, Cc = x.shape
    out = torch.empty(B, Cc, dtype=torch.float32, device=x.device)
    check(lib.b200asr_mean_time_fwd(_p(x), _p(out), B, T, Cc, dt(x), _st()), "mean_time_fwd")
    return out


def mean_time_bwd(dout, T, dtype, scale=1.0):
    dout = _c(dout)
    B, Cc = dout.shape
    dx = torch.empty(B, T, Cc, dtype=dtype, device=dout.device)
    check(lib.b200asr_mean_time_bwd(_p(dout), _p(dx), B, T, Cc, float(scale), _DT[dtype], _st()), "mean_time_bwd")
    return dx


AD_LOSS_KINDS = {"recall": 0, "prec": 1, "f1": 2, "recall_ori": 3, "prec_ori": 4, "cel": 5}


def ad_loss(logits, target, kind: str, weights: Optional[Sequence[float]], want_grad=True):
    """logits fp32 [B, C], target int64 [B] -> (loss [1] fp32, dlogits [B, C] or None)."""
    logits = _c(logits)
    B, Cc = logits.shape
    assert logits.dtype == torch.float32 and target.dtype == torch.int64
    loss = torch.empty(1, dtype=torch.float32, device=logits.device)
    dl = torch.empty_like(logits) if want_grad else None
    w = (C.c_float * Cc)(*[float(v) for v in weights]) if weights is not None else None
    check(lib.b200asr_ad_loss(_p(logits), _p(target), _p(loss), _p(dl), B, Cc, AD_LOSS_KINDS[kind], w, _st()), "ad_loss")
    return loss, dl


def cosface_fwd_bwd(x1, x2, fc_w, s=30.0, m=0.4, want_grad=True):
    x1, x2 = _c(x1), _c(x2)
    n, d = x1.shape
    loss = torch.empty(1, dtype=torch.float32, device=x1.device)
    dx1 = torch.empty_like(x1) if want_grad else None
    dx2 = torch.empty_like(x2) if want_grad else None
    nws = 2 * n * 4
    ws = workspace(nws, x1.device)
    check(lib.b200asr_cosface_fwd_bwd(_p(x1), _p(x2), _p(fc_w), _p(loss), _p(dx1), _p(dx2), _p(ws), nws, n, d, float(s),
                                      float(m), dt(x1), _st()), "cosface_fwd_bwd")
    return loss, dx1, dx2


# ----------------------------------------------------------------------------- FedAvg / optimizer
def fedavg_scale_accum(inputs: Sequence[torch.Tensor], weights: Optional[Sequence[float]], out: torch.Tensor):
    K = len(inputs)
    n = out.numel()
    for t in inputs:
        assert t.dtype == torch.float32 and t.is_contiguous() and t.numel() == n and t.is_cuda
    ptrs = (C.c_void_p * K)(*[t.data_ptr() for t in inputs])
    w = (C.c_float * K)(*[float(v) for v in weights]) if weights is not None else None
    check(lib.b200asr_fedavg_scale_accum(ptrs, w, _p(out), K, n, _st()), "fedavg_scale_accum")
    return out


def fedavg_prescale_(buf: torch.Tensor, w: float):
    assert buf.dtype == torch.float32 and buf.is_contiguous()
    check(lib.b200asr_fedavg_prescale(_p(buf), float(w), buf.numel(), _st()), "fedavg_prescale")
    return buf


def fedavg_peer_reduce(peer_ptrs: Sequence[int], weights: Optional[Sequence[float]], rank: int, n: int,
                       abort_flag: Optional[torch.Tensor] = None):
    """abort_flag: device int32 [1] (the barrier's timeout flag): when set, the kernel leaves every buffer untouched."""
    world = len(peer_ptrs)
    ptrs = (C.c_void_p * world)(*[int(p) for p in peer_ptrs])
    w = (C.c_float * world)(*[float(v) for v in weights]) if weights is not None else None
    check(lib.b200asr_fedavg_peer_reduce_guarded(ptrs, w, world, rank, n, _p(abort_flag), _st()), "fedavg_peer_reduce")


def adamw_step_(p, grad, m, v, shadow, lr, beta1, beta2, eps, weight_decay, step, gnorm_sq=None, max_norm=1.0):
    for t in (p, grad, m, v):
        assert t.dtype == torch.float32 and t.is_contiguous()
    check(lib.b200asr_adamw_step(_p(p), _p(grad), _p(m), _p(v), _p(shadow), p.numel(), float(lr), float(beta1),
                                 float(beta2), float(eps), float(weight_decay), int(step), _p(gnorm_sq),
                                 float(max_norm), _st()), "adamw_step")


def adamw_step_dev_(p, grad, m, v, shadow, hyper, beta1, beta2, eps, weight_decay, gnorm_sq=None, max_norm=1.0):
    """AdamW with {lr, bias-corrected step size} read from the device tensor `hyper` (graph-replayable launch)."""
    for t in (p, grad, m, v, hyper):
        assert t.dtype == torch.float32 and t.is_contiguous()
    check(lib.b200asr_adamw_step_dev(_p(p), _p(grad), _p(m), _p(v), _p(shadow), p.numel(), _p(hyper), float(beta1),
                                     float(beta2), float(eps), float(weight_decay), _p(gnorm_sq), float(max_norm), _st()),
          "adamw_step_dev")


def adamw_step_chunks_(p, grad, m, v, shadow, chunks, hyper, beta1, beta2, eps, weight_decay, gnorm_sq=None, max_norm=1.0):
    """AdamW with per-parameter {lr, step size} rows in the device table `hyper` [n_params, 2] (a negative step size =
    the parameter is idle in this step) over the chunk table `chunks` int32 [n_chunks, 4] (b200asr_adamw_step_chunks)."""
    for t in (p, grad, m, v, hyper):
        assert t.dtype == torch.float32 and t.is_contiguous()
    assert chunks.dtype == torch.int32 and chunks.is_contiguous() and chunks.shape[1] == 4
    check(lib.b200asr_adamw_step_chunks(_p(p), _p(grad), _p(m), _p(v), _p(shadow), _p(chunks), chunks.shape[0], _p(hyper),
                                        float(beta1), float(beta2), float(eps), float(weight_decay), _p(gnorm_sq),
                                        float(max_norm), _st()), "adamw_step_chunks")


def sumsq_(x: torch.Tensor, out: torch.Tensor, ws: Optional[torch.Tensor] = None):
    """out[0] += sum(x^2); `ws` = caller-owned zero-initialised workspace (uint8, >= b200asr_sumsq_workspace_bytes())."""
    assert x.dtype == torch.float32 and x.is_contiguous() and out.dtype == torch.float32
    if ws is None:
        check(lib.b200asr_sumsq(_p(x), x.numel(), _p(out), _st()), "sumsq")
    else:
        check(lib.b200asr_sumsq_ws(_p(x), x.numel(), _p(out), _p(ws), ws.numel() * ws.element_size(), _st()), "sumsq_ws")
    return out
