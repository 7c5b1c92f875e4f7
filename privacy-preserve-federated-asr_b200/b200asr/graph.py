"""Whole client-local update as CUDA graphs (forward -> CTC -> backward -> clip -> fused AdamW).

The update of a fixed-shape batch is a fixed sequence of ~700 kernel launches on two streams; issuing them from
Python costs more host time than the B200 needs to execute them.  GraphedTrainStep captures the sequence once
(stream capture of our C-ABI launches, including the side-stream fork/join of the weight-gradient kernels and the
cluster launches of the CTA-pair GEMM) and replays it per step.  Per-step host values (learning rate, per-parameter
Adam bias correction, which parameters are idle) are read by the captured kernels from a small device table
refreshed before each replay (optim.FusedAdamW.prepare_hyper).

LayerDrop (HF:392-404, `config.layerdrop`, 0.1 in every reference training run) changes the launch sequence from step
to step: every block is skipped with probability p by a host coin.  With layerdrop > 0 the update is therefore captured
as 2 L + 3 graphs that share one memory pool -- prologue, one per block forward, head + loss + head backward, one per
block backward, epilogue + optimizer -- cut at identity autograd nodes placed between the blocks (functional.SegmentMarkFn).
A replay draws the coins on the host exactly as the eager path does, launches the graphs of the surviving blocks and
replaces a skipped block by what it is: a copy of its input to its output (forward), of its output gradient to its input
gradient (backward), zeroed parameter gradients and idle rows in the AdamW table (transformers.AdamW skips parameters
without a gradient and does not advance their step count).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import ops


def _layerdrop_p(model) -> float:
    return float(getattr(getattr(model, "config", None), "layerdrop", 0.0) or 0.0) if model.training else 0.0


class _SegmentRecorder:
    """Cuts one captured update into graphs at the block boundaries.  `boundary()` is called from the forward pass (main
    thread) and from the backward pass (autograd's device thread) with the capture stream current: it joins the side
    streams, ends the running capture and begins the next one in the same memory pool (relaxed capture mode: the captures
    begin and end on different threads)."""

    def __init__(self, pool, n_layers: int, block_marks: bool = True, split_conv: bool = False):
        self.pool = pool
        self.L = n_layers
        self.block_marks = block_marks   # cut before every block and after the last one (LayerDrop)
        self.split_conv = split_conv     # cut after the frozen conv stack (its graph may run while FedAvg is in flight)
        self.graphs: List[torch.cuda.CUDAGraph] = []
        self.fwd_x: Dict[int, torch.Tensor] = {}   # tensor at mark l (block l's input = block l-1's output)
        self.bwd_g: Dict[int, torch.Tensor] = {}   # gradient seen at mark l (d loss / d block l's input)
        self.cur: Optional[torch.cuda.CUDAGraph] = None
        self.launches: List[int] = []              # library kernels captured per graph
        self._n0 = 0

    def begin(self):
        from . import _lib
        self._n0 = _lib.launch_count()
        self.cur = torch.cuda.CUDAGraph()
        self.cur.capture_begin(pool=self.pool, capture_error_mode="relaxed")

    def end(self):
        ops.join_side_stream()
        self.cur.capture_end()
        from . import _lib
        self.launches.append(_lib.launch_count() - self._n0)
        self.graphs.append(self.cur)
        self.cur = None

    def boundary(self):
        self.end()
        self.begin()

    def conv_mark(self):
        self.boundary()

    def fwd_mark(self, idx: int, x: torch.Tensor):
        self.fwd_x[idx] = x
        self.boundary()

    def bwd_mark(self, idx: int, g: torch.Tensor):
        self.bwd_g[idx] = g
        self.boundary()


class GraphedTrainStep:
    """step = GraphedTrainStep(model, opt, example_batch); loss = step(batch)

    `example_batch`: dict of DEVICE tensors with the shapes every later batch will have (`input_values` [B, L],
    optional `attention_mask` [B, L], `labels` [B, <= max_label_len] padded with -100, optional `dementia_labels`).
    Construction runs `warmup` eager updates and one capture; model, optimizer state and step counters are restored
    afterwards, so building the graph has no training side effect.  The returned loss is a device tensor that the
    next call overwrites.  `last_skipped_layers` holds the blocks LayerDrop skipped in the latest call."""

    def __init__(self, model, opt, example_batch: Dict[str, torch.Tensor], max_label_len: Optional[int] = None,
                 warmup: int = 3, split_conv: bool = False, defer_adamw: bool = False, per_block_graphs: bool = False):
        """split_conv: capture the frozen conv feature extractor as a graph of its own, replayed by `begin(batch)`; the
        rest of the update is replayed by `finish()`.  The conv stack does not depend on the trainable weights, so a
        federated client can run `begin(next batch)` while the FedAvg of the previous round is still in flight."""
        self.model, self.opt = model, opt
        # defer_adamw: the graphs end with the global gradient norm; the AdamW launches are issued after the replay --
        # by finish() (two launches) or, with finish(adamw=False), by the caller chunk by chunk (fedavg.PipelinedFedAvg)
        self.defer_adamw = bool(defer_adamw)
        self.max_label_len = int(max_label_len or example_batch["labels"].shape[1])
        if hasattr(model, "validate_labels"):
            model.validate_labels = False  # the reference's host-side range check would synchronise (models.py:483)
        self.backbone = getattr(model, "backbone", None)
        self.layerdrop = _layerdrop_p(model)
        # per_block_graphs: cut the update at the block boundaries even without LayerDrop (the layer-wise pipelined FedAvg
        # makes every block's graph wait only for the average of its own parameters)
        self.segmented = self.layerdrop > 0.0 or bool(per_block_graphs)
        self.split_conv = bool(split_conv)
        if (self.segmented or self.split_conv) and self.backbone is None:
            raise ValueError("GraphedTrainStep: LayerDrop needs a model with a `backbone` (Data2VecAudioModel)")
        self.static = {k: self._fit(k, v).clone() for k, v in example_batch.items()}
        saved = [t.clone() for t in opt.state_tensors()]
        host = opt.host_state()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                opt.prepare_hyper()
                self._body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        from . import _lib
        n0 = _lib.launch_count()
        self.last_skipped_layers = ()
        self.conv_graph, self.conv_launches = None, 0
        if self.segmented or self.split_conv:
            self._capture_segmented()
        else:
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.loss = self._body()
        self.launches_per_replay = _lib.launch_count() - n0  # library kernels captured = launched by a replay of everything
        for t, s in zip(opt.state_tensors(), saved):
            t.copy_(s)
        opt.load_host_state(host)
        for g in opt.groups:
            g["fp"].sync_params()
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
        self.opt.step_captured(update=not self.defer_adamw)
        if self.defer_adamw and not torch.cuda.is_current_stream_capturing():
            for gi in range(len(self.opt.groups)):   # (eager warm-up steps: complete the update)
                self.opt.adamw_params(gi)
        return out.loss.detach()

    # ------------------------------------------------------------------ LayerDrop: one graph per block and direction
    def _capture_segmented(self):
        bb = self.backbone
        L = len(bb.encoder.layers)
        torch.cuda.synchronize()
        pool = torch.cuda.graph_pool_handle()
        rec = _SegmentRecorder(pool, L, block_marks=self.segmented, split_conv=self.split_conv)
        stream = torch.cuda.Stream()
        stream.wait_stream(torch.cuda.current_stream())
        bb._segment_recorder = rec      # modeling.forward places a SegmentMarkFn before every block and after the last
        bb._force_no_layerdrop = True   # the capture holds every block; the coins are drawn per replay
        try:
            with torch.cuda.stream(stream):
                rec.begin()
                self.loss = self._body()
                rec.end()
        finally:
            bb._segment_recorder = None
            bb._force_no_layerdrop = False
        torch.cuda.current_stream().wait_stream(stream)
        c0 = 1 if self.split_conv else 0
        want = c0 + (2 * L + 3 if self.segmented else 1)
        nm = L + 1 if self.segmented else 0
        if len(rec.graphs) != want or len(rec.fwd_x) != nm or len(rec.bwd_g) != nm:
            raise RuntimeError(f"GraphedTrainStep: expected {want} segments, captured {len(rec.graphs)} "
                               f"({len(rec.fwd_x)} forward / {len(rec.bwd_g)} backward marks)")
        self.L = L
        self.conv_graph = rec.graphs[0] if self.split_conv else None
        self.conv_launches = rec.launches[0] if self.split_conv else 0
        self.graphs = rec.graphs[c0:]
        rec.launches = rec.launches[c0:]
        self.fwd_x, self.bwd_g = rec.fwd_x, rec.bwd_g
        self.graph_launches = rec.launches
        self.launches_last_replay = self.conv_launches + sum(rec.launches)
        # per block: the parameters it owns (idle rows of the AdamW table when skipped) and their slices of the flat gradients
        self.layer_idle, self.layer_grad_slices = [], []
        for l in range(L):
            tag = f"encoder.layers.{l}."
            idx = self.opt.param_indices(lambda n, tag=tag: tag in n)
            self.layer_idle.append(idx)
            slices = []
            for g, ii in zip(self.opt.groups, idx):
                fp = g["fp"]
                if ii:
                    if ii != list(range(ii[0], ii[-1] + 1)):
                        raise RuntimeError("GraphedTrainStep: a block's parameters are not adjacent in the flat buffer")
                    a = fp.offsets[ii[0]]
                    b = fp.offsets[ii[-1]] + (fp.params[ii[-1]].numel() + 3) // 4 * 4
                    slices.append(fp.grad[a:b])
            self.layer_grad_slices.append(slices)

    def fedavg_chunk_params(self):
        """fedavg_chunks() in parameter indices: per chunk a list of (group index, first parameter, end parameter)."""
        chunks = [[] for _ in range(self.L + 1)]
        for gi, g in enumerate(self.opt.groups):
            pos = 0
            for l in range(self.L):
                ii = self.layer_idle[l][gi]
                if not ii:
                    continue
                if ii[0] > pos:
                    chunks[0].append((gi, pos, ii[0]))
                chunks[1 + l].append((gi, ii[0], ii[-1] + 1))
                pos = ii[-1] + 1
            if len(g["fp"].params) > pos:
                chunks[0].append((gi, pos, len(g["fp"].params)))
        return chunks

    def fedavg_chunks(self):
        """The flat-buffer ranges in the order the forward pass reads them, for fedavg.PipelinedFedAvg: chunk 0 = every
        parameter outside the encoder blocks (projection, positional conv, encoder LayerNorm, heads), chunk 1 + l = block l.
        Each chunk is a list of (group index, start, end) in elements."""
        if not self.segmented:
            raise RuntimeError("fedavg_chunks: needs the per-block graphs (config.layerdrop > 0 or per_block_graphs=True)")
        chunks = [[] for _ in range(self.L + 1)]
        for gi, g in enumerate(self.opt.groups):
            fp = g["fp"]
            pos = 0
            for l in range(self.L):
                ii = self.layer_idle[l][gi]
                if not ii:
                    continue
                a = fp.offsets[ii[0]]
                b = fp.offsets[ii[-1]] + (fp.params[ii[-1]].numel() + 3) // 4 * 4
                if a > pos:
                    chunks[0].append((gi, pos, a))
                chunks[1 + l].append((gi, a, b))
                pos = b
            if fp.numel > pos:
                chunks[0].append((gi, pos, fp.numel))
        return chunks

    def _replay_segmented(self, wait_events=None):
        L = self.L
        main = torch.cuda.current_stream()
        if wait_events is not None and not self.segmented:
            main.wait_event(wait_events[-1])
        if not self.segmented:   # only the conv stack was split off
            self.opt.prepare_hyper()
            self.launches_last_replay = self.conv_launches + self.graph_launches[0]
            self.graphs[0].replay()
            return
        skipped = set(self.backbone._layerdrop_coins(L))
        self.last_skipped_layers = tuple(sorted(skipped))
        self.backbone.last_skipped_layers = self.last_skipped_layers
        idle = [[i for l in skipped for i in self.layer_idle[l][gi]] for gi in range(len(self.opt.groups))]
        self.opt.prepare_hyper(idle)
        g = self.graphs
        gl = self.graph_launches
        self.launches_last_replay = self.conv_launches + sum(gl) - sum(gl[1 + l] + gl[2 + L + (L - 1 - l)] for l in skipped)
        if wait_events is not None:
            main.wait_event(wait_events[0])                 # (pipelined FedAvg: the non-block parameters have been averaged)
        g[0].replay()                                       # conv stack, projection, positional conv
        for l in range(L):
            if l in skipped:
                self.fwd_x[l + 1].copy_(self.fwd_x[l])      # a dropped block is the identity
            else:
                if wait_events is not None:
                    main.wait_event(wait_events[1 + l])     # block l's parameters have been averaged
                g[1 + l].replay()
        if wait_events is not None:
            main.wait_event(wait_events[-1])                # everything (skipped blocks included) before backward / AdamW
        g[1 + L].replay()                                   # head, CTC, head backward
        for l in range(L - 1, -1, -1):
            if l in skipped:
                self.bwd_g[l].copy_(self.bwd_g[l + 1])
                for s in self.layer_grad_slices[l]:         # no gradient: zero for the global norm, idle in AdamW
                    s.zero_()
            else:
                g[2 + L + (L - 1 - l)].replay()
        g[2 + 2 * L].replay()                               # positional conv / projection backward, clip, AdamW

    def __call__(self, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
        self.begin(batch)
        return self.finish()

    def begin(self, batch: Dict[str, torch.Tensor]):
        """Stage the batch and (split_conv) run the frozen conv feature extractor -- everything that does not depend on
        the trainable weights."""
        for k, dst in self.static.items():
            src = batch[k]
            if k == "labels" and src.shape[1] != self.max_label_len:
                dst.fill_(-100)
                dst[:, : src.shape[1]].copy_(src, non_blocking=True)
            else:
                dst.copy_(src, non_blocking=True)
        if self.backbone is not None:
            self.backbone.cache.refresh_in_place()   # cached copies of the frozen conv stack's filters (load_state_dict)
        if self.conv_graph is not None:
            self.conv_graph.replay()

    def finish(self, wait_events=None, adamw: bool = True) -> torch.Tensor:
        """The rest of the update (reads the trainable weights: run after the aggregation of the previous round).
        wait_events: the per-chunk events of fedavg.PipelinedFedAvg.launch() (chunks = self.fedavg_chunks()): each block's
        graph then waits only for ITS parameters, the aggregation of the later blocks overlaps the forward pass."""
        for g in self.opt.groups:   # weights replaced through torch since the last step (load_state_dict of the global
            g["fp"].sync_params()   # model at the start of a federated round): the captured kernels read the bf16 shadow
        if self.segmented or self.split_conv:
            self._replay_segmented(wait_events)
        else:
            if wait_events is not None:
                torch.cuda.current_stream().wait_event(wait_events[-1])
            self.opt.prepare_hyper()
            self.graph.replay()
        if self.defer_adamw and adamw:
            for gi in range(len(self.opt.groups)):
                self.opt.adamw_params(gi)
        return self.loss


class GraphedInference:
    """Greedy CTC inference (finetune_ASRs.py:113-122: forward, argmax, repeat / blank collapse) of a fixed-shape batch as
    ONE CUDA graph: infer = GraphedInference(model, input_values[, attention_mask]); logits, ids, seqs = infer(input_values).

    At batch 1 x 5 s the ~330 launches of the forward pass execute faster than Python can issue them; the replay costs
    one launch.  The model must be in eval mode (no dropout, no LayerDrop); weights replaced through torch
    (`load_state_dict`) are picked up on the next call as in GraphedTrainStep."""

    def __init__(self, model, input_values: torch.Tensor, attention_mask: Optional[torch.Tensor] = None, warmup: int = 2):
        if model.training:
            raise ValueError("GraphedInference: call model.eval() first")
        self.model = model
        self.backbone = getattr(model, "backbone", None)
        self.x = input_values.clone()
        self.am = None if attention_mask is None else attention_mask.clone()
        self.blank = int(model.config.pad_token_id)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):
                self._body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.logits, self.ids, self.collapsed, self.clen = self._body()
        self._params = list(model.parameters())
        self._params_version = sum(p._version for p in self._params)
        self._caches = [m.cache for m in model.modules() if hasattr(getattr(m, "cache", None), "refresh_in_place")]
        self._host = (torch.empty(self.collapsed.shape, dtype=self.collapsed.dtype, pin_memory=True),
                      torch.empty(self.clen.shape, dtype=self.clen.dtype, pin_memory=True))

    def _body(self):
        out = self.model(self.x, self.am)
        klen = None
        if self.am is not None:   # frames beyond an utterance's length are not decoded (HF pads the logits there)
            from .modeling import device_conv_out_lengths
            klen = device_conv_out_lengths(self.model.config, self.am)
        ids, col, clen = ops.greedy_decode(out.logits, klen, self.blank)
        ops.join_side_stream()
        return out.logits, ids.long(), col, clen

    def __call__(self, input_values: torch.Tensor, attention_mask: Optional[torch.Tensor] = None):
        self.x.copy_(input_values, non_blocking=True)
        if self.am is not None:
            self.am.copy_(attention_mask, non_blocking=True)
        ver = sum(p._version for p in self._params)
        if ver != self._params_version:   # weights replaced through torch since the last call: recompute the cached copies
            for cache in self._caches:    # INTO the tensors the captured kernels read (one integer sum per call otherwise)
                cache.refresh_in_place()
            self._params_version = ver
        self.graph.replay()
        col, clen = self._host
        col.copy_(self.collapsed, non_blocking=True)
        clen.copy_(self.clen, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.logits, self.ids, [col[b, : int(clen[b])].tolist() for b in range(col.shape[0])]
