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
                 max_grad_norm: Optional[float] = 1.0, warmup_steps: int = 0, total_steps: Optional[int] = None):
        named = [(n, p) for n, p in named_params if p.requires_grad]
        dec = [(n, p) for n, p in named if not _no_decay(n)]
        nod = [(n, p) for n, p in named if _no_decay(n)]
        self.groups = []
        for items, wd in ((dec, weight_decay), (nod, 0.0)):
            if items:
                fp = FlatParams(items)
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
        for g in self.groups:
            fp = g["fp"]
            fp.grad.zero_()
            for p, o in zip(fp.params, fp.offsets):  # autograd may have replaced .grad; re-home it into the flat buffer
                if p.grad is None or p.grad.data_ptr() != fp.grad.data_ptr() + 4 * o:
                    p.grad = fp.grad[o:o + p.numel()].view(p.shape)

    def _gather_stray_grads(self):
        for g in self.groups:
            fp = g["fp"]
            for p, o in zip(fp.params, fp.offsets):
                if p.grad is not None and p.grad.data_ptr() != fp.grad.data_ptr() + 4 * o:
                    fp.grad[o:o + p.numel()].view(p.shape).copy_(p.grad)
                    p.grad = fp.grad[o:o + p.numel()].view(p.shape)

    def step(self):
        self._gather_stray_grads()
        lr = self.current_lr()
        self.step_count += 1
        gn = None
        if self.max_grad_norm is not None:
            self.gnorm_sq.zero_()
            for g in self.groups:
                ops.sumsq_(g["fp"].grad, self.gnorm_sq)
            gn = self.gnorm_sq
        for g in self.groups:
            fp = g["fp"]
            ops.adamw_step_(fp.flat, fp.grad, g["m"], g["v"], None, lr, self.betas[0], self.betas[1], self.eps, g["wd"],
                            self.step_count, gn, self.max_grad_norm or 0.0)
            fp.mark_updated()
