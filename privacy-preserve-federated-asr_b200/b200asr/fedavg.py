"""FedAvg on the device (federated/src/utils.py:240-249 `average_weights`; callers federated_main.py:138-141).

Three layers, all ending in the same fp32 kernels of libb200asr.so:
  * average_weights(list_of_state_dicts)  -- drop-in signature; every key is averaged by the scale-accumulate kernel
  * FlatParams                            -- one flat fp32 buffer per sub-network (the selection made by
                                             get_model_weight, federated/src/update.py:268-283), parameters become views
  * fedavg_allreduce / PeerFedAvg         -- one federated client per rank/GPU: NCCL all-reduce over NVLink, or the
                                             fused peer-memory kernel (scale-accumulate straight from peer HBM)
"""
from __future__ import annotations

import copy
from typing import Dict, Iterable, List, Optional, Sequence

import torch

from . import ops


def average_weights(w: List[Dict[str, torch.Tensor]], weights: Optional[Sequence[float]] = None) -> Dict[str, torch.Tensor]:
    """Returns the average of the weights (uniform when `weights` is None: sequential sum then / N, bit-identical
    to the reference).  Tensors must be fp32; CPU tensors are staged through pinned memory to the current device
    and the result is returned on the device of w[0]'s tensors."""
    if len(w) == 0:
        raise ValueError("average_weights: empty client list")
    keys = list(w[0].keys())
    dev = torch.device("cuda", torch.cuda.current_device())
    out: Dict[str, torch.Tensor] = {}
    for k in keys:
        src = [wi[k] for wi in w]
        t0 = src[0]
        if not t0.dtype.is_floating_point:  # integer buffers: the reference's `+=` then torch.div; none exist in this model
            raise TypeError(f"average_weights: non-float tensor {k} ({t0.dtype}) is not supported")
        ins = [s.detach().to(device=dev, dtype=torch.float32, non_blocking=True).contiguous().view(-1) for s in src]
        ins = [_aligned(t) for t in ins]
        res = torch.empty_like(ins[0])
        ops.fedavg_scale_accum(ins, weights, res)
        out[k] = res.view(t0.shape).to(device=t0.device, dtype=t0.dtype)
    return out


def _aligned(t: torch.Tensor) -> torch.Tensor:
    return t if t.data_ptr() % 16 == 0 else t.clone()


class FlatParams:
    """Re-homes the given parameters as views into ONE flat fp32 buffer (16-byte aligned segments), so FedAvg,
    the optimizer and the gradient norm each touch the sub-network with a single streaming kernel.

    Also the hub of two per-step shortcuts of the host layer (functional.py):
      * `grad` is the flat fp32 gradient; with `direct_grads` the backward kernels write each parameter's slice in
        place (functional.grad_sink) instead of handing tensors to autograd's per-parameter accumulation;
      * `shadow` (optional) is a flat bf16 copy of `flat` that whoever updates `flat` keeps current (the fused AdamW
        kernel writes it in the same pass): the tensor-core GEMMs read weight views of it, so there is no per-step
        cast or QKV concat.  `epoch` counts raw-pointer updates of `flat`; the shadow is valid iff it was refreshed
        at the current epoch."""

    def __init__(self, named_params: Iterable, device=None, with_shadow: bool = False, direct_grads: bool = True):
        named = [(n, p) for n, p in named_params]
        if not named:
            raise ValueError("FlatParams: no parameters")
        device = device or named[0][1].device
        self.names = [n for n, _ in named]
        self.params = [p for _, p in named]
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + 3) // 4 * 4
        self.numel = off
        self.flat = torch.zeros(off, dtype=torch.float32, device=device)
        self.grad = torch.zeros(off, dtype=torch.float32, device=device)
        self.shadow = torch.zeros(off, dtype=torch.bfloat16, device=device) if with_shadow else None
        self.epoch = 0
        self.shadow_epoch = -1
        self.direct_grads = direct_grads
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                v = self.flat[o:o + p.numel()].view(p.shape)
                v.copy_(p.data)
                p.data = v
                p.grad = self.grad[o:o + p.numel()].view(p.shape)
                p._b200_flat, p._b200_off, p._b200_grad, p._b200_fresh = self, o, p.grad, False
                p._b200_pver = p._version
        if with_shadow:
            self.refresh_shadow()

    def manifest(self):
        return [(n, o, tuple(p.shape)) for n, o, p in zip(self.names, self.offsets, self.params)]

    def shadow_valid(self) -> bool:
        return self.shadow is not None and self.shadow_epoch == self.epoch

    def refresh_shadow(self):
        """One streaming cast of the whole flat buffer (after FedAvg, load_state_dict, ...)."""
        if self.shadow is not None:
            ops.cast_from_f32(self.flat, torch.bfloat16, out=self.shadow)
            self.shadow_epoch = self.epoch
            for p in self.params:
                p._b200_pver = p._version

    def mark_updated(self, shadow_written: bool = False):
        """A kernel wrote the flat buffer behind autograd's back: bump the epoch so cached weight shadows and derived
        layouts of THESE parameters are rebuilt (frozen parameters elsewhere keep theirs)."""
        self.epoch += 1
        if self.shadow is not None:
            if shadow_written:
                self.shadow_epoch = self.epoch
            else:
                self.refresh_shadow()


def fedavg_local(flats: Sequence[torch.Tensor], weights: Optional[Sequence[float]] = None,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Single-GPU N-way average of flat client buffers (read N*4P + write 4P bytes)."""
    out = out if out is not None else torch.empty_like(flats[0])
    return ops.fedavg_scale_accum(list(flats), weights, out)


def fedavg_allreduce(flat: torch.Tensor, weight: Optional[float] = None, group=None) -> torch.Tensor:
    """One client per rank: flat <- sum_r w_r * flat_r, in place on every rank (no server hop, no re-broadcast).
    The scale runs as our streaming kernel; the sum is an NCCL all-reduce over NVLink/NVSwitch (gloo on CPU tests
    is handled by the caller: this function requires CUDA tensors)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    w = (1.0 / world) if weight is None else float(weight)
    ops.fedavg_prescale_(flat, w)
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def fedavg_shard_plan(n: int, world: int):
    """Slice [lo, hi) of the flat buffer owned by each rank in the fused peer kernel (multiples of 4 elements;
    the last rank takes the remainder).  Mirrors b200asr_fedavg_peer_reduce."""
    per = (n // world) // 4 * 4
    return [(per * r, n if r == world - 1 else per * (r + 1)) for r in range(world)]
