"""Fused AdamW + global-norm clipping on flat parameter buffers.

Semantics of the reference's optimizer = HF Trainer "adamw_hf" (transformers.AdamW, correct_bias=True) with
max_grad_norm clipping, weight decay on everything except biases and LayerNorm weights, linear warm-up then
linear decay (centralized/finetune_ASRs.py:202-220, federated/src/update.py:651-678).
"""
from __future__ import annotations

from typing import Iterable, Optional

import torch

from . import ops
from .fedavg import FlatParams


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

    def _segments(self, g, idle):
        """Per-parameter step counts (transformers.AdamW keeps `state["step"]` per parameter and only advances it
        when the parameter has a gradient).  Yields (start, end, step) over the flat buffer: maximal runs of
        consecutive parameters that are updated in this step with the same step number."""
        fp = g["fp"]
        steps = g.setdefault("param_steps", [self.step_count - 1] * len(fp.params))
        idle = set(idle)
        run = None
        for i, (p, o) in enumerate(zip(fp.params, fp.offsets)):
            if i in idle:
                if run:
                    yield run
                run = None
                continue
            steps[i] += 1
            end = o + (p.numel() + 3) // 4 * 4
            if run and run[2] == steps[i]:
                run = (run[0], end, run[2])
            else:
                if run:
                    yield run
                run = (o, end, steps[i])
        if run:
            yield run

    def _clip_norm(self):
        if self.max_grad_norm is None:
            return None
        self.gnorm_sq.zero_()
        for g in self.groups:
            ops.sumsq_(g["fp"].grad, self.gnorm_sq)
        return self.gnorm_sq

    def step(self):
        ops.join_side_stream()
        untouched = self._gather_stray_grads()
        lr = self.current_lr()
        self.step_count += 1
        gn = self._clip_norm()
        for g, idle in zip(self.groups, untouched):
            fp = g["fp"]
            if not idle and "param_steps" not in g:  # every parameter has always been updated: one launch per group
                ops.adamw_step_(fp.flat, fp.grad, g["m"], g["v"], fp.shadow, lr, self.betas[0], self.betas[1], self.eps,
                                g["wd"], self.step_count, gn, self.max_grad_norm or 0.0)
            else:
                for a, b, t in self._segments(g, idle):
                    ops.adamw_step_(fp.flat[a:b], fp.grad[a:b], g["m"][a:b], g["v"][a:b],
                                    None if fp.shadow is None else fp.shadow[a:b], lr, self.betas[0], self.betas[1],
                                    self.eps, g["wd"], t, gn, self.max_grad_norm or 0.0)
            fp.mark_updated(shadow_written=fp.shadow is not None)

    # ---- CUDA-graph friendly split: per-step scalars live in device memory (graph.GraphedTrainStep)
    def prepare_hyper(self):
        """Host side of the next update: advances the step counter and uploads {lr, lr*sqrt(1-b2^t)/(1-b1^t)} to the
        device tensor the captured AdamW launch reads.  Call once before every replay (outside the graph)."""
        if not hasattr(self, "hyper_dev"):
            dev = self.groups[0]["fp"].flat.device
            self.hyper_dev = torch.zeros(2, dtype=torch.float32, device=dev)
        lr = self.current_lr()
        self.step_count += 1
        t = self.step_count
        # a fresh pageable tensor per step: the copy is staged before the call returns, so the next step's values
        # can never overtake this one in the stream
        # (same arithmetic as b200asr_adamw_step: lr rounded to fp32 first, bias correction in double -> bit-identical
        # to the eager path)
        lr32 = float(torch.tensor(lr, dtype=torch.float32))
        b1, b2 = float(torch.tensor(self.betas[0], dtype=torch.float32)), float(torch.tensor(self.betas[1], dtype=torch.float32))
        self.hyper_dev.copy_(torch.tensor([lr32, lr32 * (1.0 - b2 ** t) ** 0.5 / (1.0 - b1 ** t)], dtype=torch.float32))

    def step_captured(self):
        """Device side of the update, with no host-dependent scalar: what a CUDA graph captures."""
        ops.join_side_stream()
        self._gather_stray_grads()
        gn = self._clip_norm()
        for g in self.groups:
            fp = g["fp"]
            ops.adamw_step_dev_(fp.flat, fp.grad, g["m"], g["v"], fp.shadow, self.hyper_dev, self.betas[0], self.betas[1],
                                self.eps, g["wd"], gn, self.max_grad_norm or 0.0)
            fp.mark_updated(shadow_written=fp.shadow is not None)

    def state_tensors(self):
        out = []
        for g in self.groups:
            out += [g["fp"].flat, g["m"], g["v"]] + ([g["fp"].shadow] if g["fp"].shadow is not None else [])
        return out
