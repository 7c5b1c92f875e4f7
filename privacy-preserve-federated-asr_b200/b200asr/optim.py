"""Fused AdamW + global-norm clipping on flat parameter buffers.

Semantics of the reference's optimizer = HF Trainer "adamw_hf" (transformers.AdamW, correct_bias=True) with
max_grad_norm clipping, weight decay on everything except biases and LayerNorm weights, linear warm-up then
linear decay (centralized/finetune_ASRs.py:202-220, federated/src/update.py:651-678).

transformers.AdamW keeps `state["step"]` PER PARAMETER and skips parameters whose `grad is None` (a block dropped by
LayerDrop, an unused head, `masked_spec_embed` without SpecAugment).  Here that is a device table: one {lr, bias-corrected
step size} row per parameter (negative step size = idle in this step) consumed by `b200asr_adamw_step_chunks` -- the same
launch serves every pattern of idle parameters, so it can be captured in a CUDA graph; the eager step() and the captured
step_captured() go through the same arithmetic and agree bit for bit.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np
import torch

from . import ops
from ._lib import lib
from .fedavg import FlatParams

CHUNK_VECS = 2048  # float4 per chunk of the AdamW chunk table (8192 parameters: one 256-thread block, 8 float4 per thread)


def _no_decay(name: str) -> bool:
    return name.endswith(".bias") or "layer_norm" in name


class FusedAdamW:
    def __init__(self, named_params: Iterable, lr=1e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.005,
                 max_grad_norm: Optional[float] = 1.0, warmup_steps: int = 0, total_steps: Optional[int] = None,
                 bf16_shadow: bool = True, peer: bool = False):
        """bf16_shadow: keep a flat bf16 copy of the decayed group (the GEMM weights) current inside the update
        kernel, which the bf16 compute path then reads directly (functional.ShadowCache)."""
        wd_shadow = bf16_shadow
        named = [(n, p) for n, p in named_params if p.requires_grad]
        dec = [(n, p) for n, p in named if not _no_decay(n)]
        nod = [(n, p) for n, p in named if _no_decay(n)]
        self.groups = []
        for items, wd in ((dec, weight_decay), (nod, 0.0)):
            if items:
                fp = FlatParams(items, with_shadow=(wd_shadow and items is dec), peer=peer)
                self.groups.append(dict(fp=fp, wd=wd, m=torch.zeros_like(fp.flat), v=torch.zeros_like(fp.flat)))
        self.lr, self.betas, self.eps = lr, betas, eps
        self.max_grad_norm = max_grad_norm
        self.warmup_steps, self.total_steps = warmup_steps, total_steps
        self.step_count = 0
        dev = self.groups[0]["fp"].flat.device
        self.gnorm_sq = torch.zeros(1, dtype=torch.float32, device=dev)
        # caller-owned partials + arrival counter of the deterministic sum of squares: two optimizers (two clients on one
        # GPU, a side stream) never share reduction state
        self.sumsq_ws = torch.zeros(int(lib.b200asr_sumsq_workspace_bytes()), dtype=torch.uint8, device=dev)

    def current_lr(self) -> float:
        s = self.step_count
        if self.warmup_steps and s < self.warmup_steps:
            return self.lr * s / max(1, self.warmup_steps)
        if self.total_steps:
            return self.lr * max(0.0, (self.total_steps - s) / max(1, self.total_steps - self.warmup_steps))
        return self.lr

    def zero_grad(self):
        """No memset: every parameter is marked fresh, the backward kernels overwrite their slice of the flat
        gradient (functional.grad_sink); parameters that end up with no gradient are zeroed in step()."""
        for g in self.groups:
            g["fp"].sync_params()   # weights replaced through torch since the last step (load_state_dict): refresh the shadow
            for p in g["fp"].params:
                p.grad = None
                p._b200_fresh = True

    def _gather_stray_grads(self):
        """Returns, per group, the indices of the parameters that received no gradient in this step (a block
        dropped by LayerDrop, an unused head): their slice of the flat gradient is zeroed so the global norm is
        unaffected, and step() leaves them alone as transformers.AdamW does (`if p.grad is None: continue`)."""
        untouched = []
        for g in self.groups:
            fp = g["fp"]
            idle = []
            for i, p in enumerate(fp.params):
                if p.grad is None:
                    if p._b200_fresh:  # untouched this step
                        p._b200_grad.zero_()
                        p._b200_fresh = False
                        idle.append(i)
                    p.grad = p._b200_grad
                elif p.grad.data_ptr() != p._b200_grad.data_ptr():  # autograd delivered its own tensor
                    if p._b200_fresh:
                        p._b200_grad.copy_(p.grad)
                        p._b200_fresh = False
                    else:
                        p._b200_grad.add_(p.grad)
                    p.grad = p._b200_grad
            untouched.append(idle)
        return untouched

    # ---- per-parameter step table
    def _table(self, g):
        """(chunks int32 [n, 4] on the device, hyper fp32 [n_params, 2] on the device) of one group, built once."""
        if "chunks" not in g:
            fp = g["fp"]
            rows = []
            for i, (p, o) in enumerate(zip(fp.params, fp.offsets)):
                v0, nv = o // 4, (p.numel() + 3) // 4
                for c in range(0, nv, CHUNK_VECS):
                    rows.append((v0 + c, min(CHUNK_VECS, nv - c), i, 0))
            g["chunks"] = torch.tensor(rows, dtype=torch.int32, device=fp.flat.device)
            # first chunk-table row of every parameter (+ the end): a run of parameters is a run of rows
            first, r = [], 0
            for p in fp.params:
                first.append(r)
                r += ((p.numel() + 3) // 4 + CHUNK_VECS - 1) // CHUNK_VECS
            g["param_row"] = first + [r]
            g["hyper_dev"] = torch.zeros(len(fp.params), 2, dtype=torch.float32, device=fp.flat.device)
        return g["chunks"], g["hyper_dev"]

    def _upload_hyper(self, g, idle: Sequence[int], lr: float):
        """Advance the step count of every non-idle parameter of the group and upload its {lr, step size} rows
        (same arithmetic as b200asr_adamw_step: lr and betas rounded to fp32 first, bias correction in double)."""
        fp = g["fp"]
        _, hyper_dev = self._table(g)
        n = len(fp.params)
        steps = g.get("param_steps")
        if steps is None:
            steps = g["param_steps"] = np.full(n, self.step_count - 1, dtype=np.int64)
        active = np.ones(n, dtype=bool)
        if len(idle):
            active[np.asarray(list(idle), dtype=np.int64)] = False
        steps[active] += 1
        lr32 = float(np.float32(lr))
        b1, b2 = float(np.float32(self.betas[0])), float(np.float32(self.betas[1]))
        rows = np.empty((n, 2), dtype=np.float32)
        rows[:, 0] = lr32
        cache: Dict[int, float] = {}
        for i in range(n):   # (C pow through Python floats: bit-identical to b200asr_adamw_step's double arithmetic)
            if not active[i]:
                rows[i, 1] = -1.0
                continue
            t = int(max(steps[i], 1))
            if t not in cache:
                cache[t] = lr32 * (1.0 - b2 ** t) ** 0.5 / (1.0 - b1 ** t)
            rows[i, 1] = cache[t]
        # a fresh pageable tensor per step: the copy is staged before the call returns, so the next step's values can
        # never overtake this one in the stream
        hyper_dev.copy_(torch.from_numpy(rows))

    def _clip_norm(self):
        if self.max_grad_norm is None:
            return None
        self.gnorm_sq.zero_()
        for g in self.groups:
            ops.sumsq_(g["fp"].grad, self.gnorm_sq, self.sumsq_ws)
        return self.gnorm_sq

    def step(self):
        ops.join_side_stream()
        untouched = self._gather_stray_grads()
        lr = self.current_lr()
        self.step_count += 1
        gn = self._clip_norm()
        for g, idle in zip(self.groups, untouched):
            fp = g["fp"]
            if not idle and "param_steps" not in g:  # every parameter has always been updated: one scalar launch per group
                ops.adamw_step_(fp.flat, fp.grad, g["m"], g["v"], fp.shadow, lr, self.betas[0], self.betas[1], self.eps,
                                g["wd"], self.step_count, gn, self.max_grad_norm or 0.0)
            else:
                self._upload_hyper(g, idle, lr)
                chunks, hyper = self._table(g)
                ops.adamw_step_chunks_(fp.flat, fp.grad, g["m"], g["v"], fp.shadow, chunks, hyper, self.betas[0],
                                       self.betas[1], self.eps, g["wd"], gn, self.max_grad_norm or 0.0)
            fp.mark_updated(shadow_written=fp.shadow is not None)

    # ---- CUDA-graph friendly split: per-step scalars live in device memory (graph.GraphedTrainStep)
    def prepare_hyper(self, idle: Optional[Sequence[Sequence[int]]] = None):
        """Host side of the next captured update: advances the step counters and uploads the per-parameter
        {lr, lr*sqrt(1-b2^t)/(1-b1^t)} table the captured AdamW launch reads.  Call once before every replay (outside the
        graph).  `idle`: per group, the indices of the parameters without a gradient in THIS step (blocks LayerDrop
        skipped) in addition to the ones that never get one (`static_idle`, found when the update was captured)."""
        lr = self.current_lr()
        self.step_count += 1
        for gi, g in enumerate(self.groups):
            skip = set(g.get("static_idle", ()))
            if idle is not None:
                skip |= set(idle[gi])
            self._upload_hyper(g, sorted(skip), lr)

    def step_captured(self, update: bool = True):
        """Device side of the update, with no host-dependent scalar: what a CUDA graph captures.
        update=False: gradient bookkeeping and the global norm only; the AdamW launches follow outside the graph
        (`adamw_params`), e.g. chunk by chunk, interleaved with the exchange of a pipelined FedAvg."""
        ops.join_side_stream()
        untouched = self._gather_stray_grads()
        gn = self._clip_norm()
        for g, idle in zip(self.groups, untouched):
            fp = g["fp"]
            # parameters that get no gradient from this (fixed) launch sequence never will in a replay either: from now
            # on prepare_hyper marks them idle, as the eager step() does
            g["static_idle"] = tuple(idle)
            chunks, hyper = self._table(g)
            if not update:
                continue
            ops.adamw_step_chunks_(fp.flat, fp.grad, g["m"], g["v"], fp.shadow, chunks, hyper, self.betas[0], self.betas[1],
                                   self.eps, g["wd"], gn, self.max_grad_norm or 0.0)
            fp.mark_updated(shadow_written=fp.shadow is not None)

    def adamw_params(self, gi: int, p_lo: int = 0, p_hi: Optional[int] = None):
        """The AdamW launch for parameters [p_lo, p_hi) of group gi (default: the whole group) on the current stream, from
        the per-parameter table prepare_hyper uploaded and the clip norm step_captured(update=False) left behind."""
        g = self.groups[gi]
        fp = g["fp"]
        chunks, hyper = self._table(g)
        p_hi = len(fp.params) if p_hi is None else p_hi
        r0, r1 = g["param_row"][p_lo], g["param_row"][p_hi]
        if r1 > r0:
            ops.adamw_step_chunks_(fp.flat, fp.grad, g["m"], g["v"], fp.shadow, chunks[r0:r1], hyper, self.betas[0], self.betas[1],
                                   self.eps, g["wd"], self.gnorm_sq if self.max_grad_norm is not None else None,
                                   self.max_grad_norm or 0.0)
        # (also for a partial range: the epoch only has to differ from the one the derived layouts -- packed positional
        # conv filters, fp32 QKV concat -- were cached at; a capture that hit those caches would replay stale weights)
        fp.mark_updated(shadow_written=fp.shadow is not None)

    def param_indices(self, predicate) -> List[List[int]]:
        """Per group, the indices of the parameters whose name satisfies `predicate` (e.g. one encoder block)."""
        return [[i for i, n in enumerate(g["fp"].names) if predicate(n)] for g in self.groups]

    def state_tensors(self):
        out = []
        for g in self.groups:
            out += [g["fp"].flat, g["m"], g["v"]] + ([g["fp"].shadow] if g["fp"].shadow is not None else [])
        return out

    def host_state(self) -> Dict:
        """Host-side counters (restored together with state_tensors() after a graph capture's warm-up steps)."""
        return dict(step_count=self.step_count,
                    param_steps=[None if g.get("param_steps") is None else g["param_steps"].copy() for g in self.groups])

    def load_host_state(self, st: Dict):
        self.step_count = st["step_count"]
        for g, ps in zip(self.groups, st["param_steps"]):
            if ps is None:
                g.pop("param_steps", None)
            else:
                g["param_steps"] = ps.copy()
