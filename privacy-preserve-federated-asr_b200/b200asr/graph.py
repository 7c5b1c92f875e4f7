"""Whole client-local update as ONE CUDA graph (forward -> CTC -> backward -> clip -> fused AdamW).

The update of a fixed-shape batch is a fixed sequence of ~700 kernel launches on two streams; issuing them from
Python costs more host time than the B200 needs to execute them.  GraphedTrainStep captures the sequence once
(torch.cuda.graph: stream capture of our C-ABI launches, including the side-stream fork/join of the weight-gradient
kernels and the cluster launches of the CTA-pair GEMM) and replays it per step.  Per-step host values (learning
rate, Adam bias correction) are read by the captured kernels from a small device tensor refreshed before each replay.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch


class GraphedTrainStep:
    """step = GraphedTrainStep(model, opt, example_batch); loss = step(batch)

    `example_batch`: dict of DEVICE tensors with the shapes every later batch will have (`input_values` [B, L],
    optional `attention_mask` [B, L], `labels` [B, <= max_label_len] padded with -100, optional `dementia_labels`).
    Construction runs `warmup` eager updates and one capture; model, optimizer state and step counter are restored
    afterwards, so building the graph has no training side effect.  The returned loss is a device tensor that the
    next call overwrites."""

    def __init__(self, model, opt, example_batch: Dict[str, torch.Tensor], max_label_len: Optional[int] = None,
                 warmup: int = 3):
        self.model, self.opt = model, opt
        if model.training and float(getattr(getattr(model, "config", None), "layerdrop", 0.0) or 0.0) > 0.0:
            raise NotImplementedError("GraphedTrainStep: LayerDrop changes the launch sequence from step to step "
                                      "(HF:392-404); set config.layerdrop = 0 or run the update eagerly")
        self.max_label_len = int(max_label_len or example_batch["labels"].shape[1])
        if hasattr(model, "validate_labels"):
            model.validate_labels = False  # the reference's host-side range check would synchronise (models.py:483)
        self.static = {k: self._fit(k, v).clone() for k, v in example_batch.items()}
        saved = [t.clone() for t in opt.state_tensors()]
        step_count = opt.step_count
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                opt.prepare_hyper()
                self._body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        from . import _lib
        self.graph = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count()
        with torch.cuda.graph(self.graph):
            self.loss = self._body()
        self.launches_per_replay = _lib.launch_count() - n0  # library kernels captured = launched by every replay
        for t, s in zip(opt.state_tensors(), saved):
            t.copy_(s)
        opt.step_count = step_count
        torch.cuda.synchronize()

    def _fit(self, key, v):
        if key == "labels" and v.shape[1] != self.max_label_len:
            out = torch.full((v.shape[0], self.max_label_len), -100, dtype=v.dtype, device=v.device)
            out[:, : v.shape[1]] = v[:, : self.max_label_len]
            return out
        return v

    def _body(self):
        self.opt.zero_grad()
        out = self.model(**self.static)
        out.loss.backward()
        self.opt.step_captured()
        return out.loss.detach()

    def __call__(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        for k, dst in self.static.items():
            src = batch[k]
            if k == "labels" and src.shape[1] != self.max_label_len:
                dst.fill_(-100)
                dst[:, : src.shape[1]].copy_(src, non_blocking=True)
            else:
                dst.copy_(src, non_blocking=True)
        self.opt.prepare_hyper()
        self.graph.replay()
        return self.loss
