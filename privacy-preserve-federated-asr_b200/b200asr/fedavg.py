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

    def __init__(self, named_params: Iterable, device=None, with_shadow: bool = False, direct_grads: bool = True,
                 peer: bool = False):
        """peer=True places `flat` in a peer-visible allocation (PeerBuffer) so that PeerFedAvg can reduce it straight
        out of the other clients' HBM."""
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
        # peer=True: plain cudaMalloc exported through CUDA IPC (PeerFedAvg); peer="symm": torch symmetric memory, whose
        # allocation can be bound to an NVSwitch multicast object as well (SymmFedAvg: NVLS in-switch reduction)
        self.peer_buffer = PeerBuffer(off * 4, device) if peer is True else None
        self.symm = None
        if peer == "symm":
            import torch.distributed._symmetric_memory as symm_mem
            self.symm = symm_mem.empty(off, dtype=torch.float32, device=device)
            self.symm.zero_()
            self.flat = self.symm
        elif peer:
            self.flat = self.peer_buffer.as_tensor(torch.float32, off)
        else:
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

    def sync_params(self) -> bool:
        """Parameters were written through torch (`load_state_dict`, `p.copy_`, `p.data.mul_` ... bump `p._version`)
        since the shadow was last refreshed -- the per-round "load the global weights" step of a federated client
        (federated/src/update.py:228).  Refreshes the bf16 shadow and bumps the epoch so derived layouts are rebuilt.
        Cheap when nothing changed (one integer compare per parameter); called by FusedAdamW.zero_grad and before every
        CUDA-graph replay, whose captured kernels read the shadow through fixed pointers."""
        if all(p._version == p._b200_pver for p in self.params):
            return False
        self.epoch += 1
        if self.shadow is not None:
            self.refresh_shadow()
        else:
            for p in self.params:
                p._b200_pver = p._version
        return True

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


class PeerBuffer:
    """A zero-filled device allocation made by the library (plain cudaMalloc, so CUDA IPC can export it) and viewed as
    a torch tensor.  `export()` gives the 64-byte IPC handle another process opens with `PeerBuffer.open`."""

    def __init__(self, nbytes: int, device=None, _ptr=None, _opened=False):
        import ctypes as C
        from ._lib import check, lib
        self.nbytes = int(nbytes)
        self.device = torch.device(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
        self.opened = _opened
        if _ptr is None:
            with torch.cuda.device(self.device):
                p = C.c_void_p()
                check(lib.b200asr_peer_alloc(C.byref(p), self.nbytes), "peer_alloc")
                _ptr = p.value
        self.ptr = int(_ptr)

    def as_tensor(self, dtype=torch.float32, numel=None) -> torch.Tensor:
        item = torch.empty(0, dtype=dtype).element_size()
        n = self.nbytes // item if numel is None else int(numel)
        typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
        holder = type("_PeerView", (), {})()
        holder.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (self.ptr, False), "version": 2}
        holder._owner = self  # keeps the allocation alive as long as the tensor
        return torch.as_tensor(holder, device=self.device)

    def export(self) -> bytes:
        import ctypes as C
        from ._lib import check, lib
        buf = C.create_string_buffer(64)
        check(lib.b200asr_ipc_export(self.ptr, buf), "ipc_export")
        return buf.raw

    @staticmethod
    def open(handle: bytes, nbytes: int, device=None) -> "PeerBuffer":
        import ctypes as C
        from ._lib import check, lib
        p = C.c_void_p()
        check(lib.b200asr_ipc_open(C.create_string_buffer(handle, 64), C.byref(p)), "ipc_open")
        return PeerBuffer(nbytes, device, _ptr=p.value, _opened=True)

    def close(self):
        from ._lib import lib
        if self.ptr:
            (lib.b200asr_ipc_close if self.opened else lib.b200asr_peer_free)(self.ptr)
            self.ptr = 0


class PeerFedAvg:
    """The fused multi-GPU FedAvg (one federated client per rank / GPU, federated_main.py:138-141): every rank reduces
    its slice of each flat buffer straight out of the peers' HBM over NVLink and writes the average back into every
    peer -- scale, reduce-scatter and all-gather in one kernel (`b200asr_fedavg_peer_reduce`), bracketed by the
    peer-memory barrier kernel.  No NCCL on the data path; torch.distributed only carries the IPC handles once.

      agg = PeerFedAvg([g["fp"] for g in opt.groups])      # FlatParams built with peer=True
      agg()                                                # after the local step, on every rank
    """

    def __init__(self, flats: Sequence["FlatParams"], group=None):
        import torch.distributed as dist
        self.flats = list(flats)
        for fp in self.flats:
            if fp.peer_buffer is None:
                raise ValueError("PeerFedAvg needs FlatParams(..., peer=True)")
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        dev = self.flats[0].flat.device
        self.sync = PeerBuffer(4096, dev)
        self.timeout = torch.zeros(1, dtype=torch.int32, device=dev)
        mine = {"flats": [(fp.peer_buffer.export(), fp.peer_buffer.nbytes, fp.numel) for fp in self.flats],
                "sync": self.sync.export()}
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=group)
        self._opened = []
        self.flag_ptrs, self.flat_ptrs = [], [[] for _ in self.flats]
        for r, info in enumerate(everyone):
            if [n for _, _, n in info["flats"]] != [fp.numel for fp in self.flats]:
                raise ValueError(f"PeerFedAvg: rank {r} holds differently sized flat buffers")
            if r == self.rank:
                self.flag_ptrs.append(self.sync.ptr)
                for i, fp in enumerate(self.flats):
                    self.flat_ptrs[i].append(fp.peer_buffer.ptr)
                continue
            sb = PeerBuffer.open(info["sync"], 4096, dev)
            self._opened.append(sb)
            self.flag_ptrs.append(sb.ptr)
            for i, (h, nbytes, _) in enumerate(info["flats"]):
                pb = PeerBuffer.open(h, nbytes, dev)
                self._opened.append(pb)
                self.flat_ptrs[i].append(pb.ptr)
        self.epoch = 0
        dist.barrier(group=group)  # every rank has mapped every peer before the first kernel touches them

    def _barrier(self):
        import ctypes as C
        from ._lib import check, lib
        self.epoch += 1
        ptrs = (C.c_void_p * self.world)(*self.flag_ptrs)
        check(lib.b200asr_fedavg_peer_barrier(ptrs, self.world, self.rank, self.epoch, self.timeout.data_ptr(),
                                              torch.cuda.current_stream().cuda_stream), "fedavg_peer_barrier")

    def __call__(self, weights: Optional[Sequence[float]] = None):
        """weights: per-rank FedAvg weights (None = the reference's uniform 1/N, sequential sum then divide)."""
        self._barrier()   # every client's local update is complete and visible
        for fp, ptrs in zip(self.flats, self.flat_ptrs):   # (a no-op on every rank that saw the barrier time out)
            ops.fedavg_peer_reduce(ptrs, weights, self.rank, fp.numel, abort_flag=self.timeout)
        self._barrier()   # every slice has been written back everywhere
        for fp in self.flats:
            fp.mark_updated()

    def barrier(self):
        self._barrier()

    def reduce_range(self, gi: int, a: int, b: int, impl: str = "peer"):
        """The exchange of elements [a, b) of flat buffer `gi` only (a, b multiples of 4), no barriers (PipelinedFedAvg)."""
        ops.fedavg_peer_reduce([p + 4 * a for p in self.flat_ptrs[gi]], None, self.rank, b - a, abort_flag=self.timeout)

    def check(self):
        """Host-synchronising: raises if a barrier ever timed out (a peer never arrived)."""
        if int(self.timeout.item()) != 0:
            raise RuntimeError("PeerFedAvg: cross-rank barrier timed out")

    def close(self):
        for pb in self._opened:
            pb.close()
        self._opened = []


class SymmFedAvg:
    """Multi-GPU FedAvg over torch symmetric memory (FlatParams(..., peer="symm")): the same flat buffers are reachable
    both through peer pointers (the fused peer-memory kernel, bit-exact left-to-right sum) and through ONE NVSwitch
    multicast address (`b200asr_fedavg_nvls_reduce`: the switch sums the ranks' slices and multicasts the average back --
    4P/N bytes each way per GPU at any N).  torch.distributed only does the rendezvous; every byte of the exchange is moved
    by our kernels, bracketed by our peer-memory barrier kernel.

      agg = SymmFedAvg([g["fp"] for g in opt.groups]); agg()            # impl="nvls" (default) | "peer"
    """

    def __init__(self, flats: Sequence["FlatParams"], group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        self.flats = list(flats)
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        dev = self.flats[0].flat.device
        self.mc_ptrs, self.flat_ptrs, self._handles = [], [], []
        for fp in self.flats:
            if fp.symm is None:
                raise ValueError('SymmFedAvg needs FlatParams(..., peer="symm")')
            h = symm_mem.rendezvous(fp.symm, group)
            self._handles.append(h)
            off = fp.symm.data_ptr() - int(h.buffer_ptrs[self.rank])
            self.flat_ptrs.append([int(p) + off for p in h.buffer_ptrs])
            self.mc_ptrs.append(int(h.multicast_ptr) + off if int(h.multicast_ptr) else 0)
        self._flags = symm_mem.empty(1024, dtype=torch.int32, device=dev)
        self._flags.zero_()
        fh = symm_mem.rendezvous(self._flags, group)
        self._handles.append(fh)
        foff = self._flags.data_ptr() - int(fh.buffer_ptrs[self.rank])
        self.flag_ptrs = [int(p) + foff for p in fh.buffer_ptrs]
        self.timeout = torch.zeros(1, dtype=torch.int32, device=dev)
        self.has_multicast = all(self.mc_ptrs)
        self.epoch = 0
        torch.cuda.synchronize()
        dist.barrier(group=group)

    def _barrier(self):
        import ctypes as C
        from ._lib import check, lib
        self.epoch += 1
        ptrs = (C.c_void_p * self.world)(*self.flag_ptrs)
        check(lib.b200asr_fedavg_peer_barrier(ptrs, self.world, self.rank, self.epoch, self.timeout.data_ptr(),
                                              torch.cuda.current_stream().cuda_stream), "fedavg_peer_barrier")

    def __call__(self, weights: Optional[Sequence[float]] = None, impl: str = "nvls", max_blocks: int = 0, mark: bool = True):
        from ._lib import check, lib
        if impl == "nvls" and (weights is not None or not self.has_multicast):
            impl = "peer"   # per-client weights need the scale-accumulate form; no multicast object on this fabric
        self._barrier()   # every client's local update is complete and visible
        for fp, ptrs, mc in zip(self.flats, self.flat_ptrs, self.mc_ptrs):
            if impl == "nvls":
                check(lib.b200asr_fedavg_nvls_reduce(mc, self.world, self.rank, fp.numel, self.timeout.data_ptr(), int(max_blocks),
                                                     torch.cuda.current_stream().cuda_stream), "fedavg_nvls_reduce")
            else:
                ops.fedavg_peer_reduce(ptrs, weights, self.rank, fp.numel, abort_flag=self.timeout)
        self._barrier()   # every slice has been written back everywhere
        if mark:
            for fp in self.flats:
                fp.mark_updated()

    def barrier(self):
        self._barrier()

    def reduce_range(self, gi: int, a: int, b: int, impl: str = "nvls"):
        """The exchange of elements [a, b) of flat buffer `gi` only (a, b multiples of 4), no barriers (PipelinedFedAvg)."""
        from ._lib import check, lib
        if impl == "nvls" and self.has_multicast:
            check(lib.b200asr_fedavg_nvls_reduce(self.mc_ptrs[gi] + 4 * a, self.world, self.rank, b - a, self.timeout.data_ptr(), 0,
                                                 torch.cuda.current_stream().cuda_stream), "fedavg_nvls_reduce")
        else:
            ops.fedavg_peer_reduce([p + 4 * a for p in self.flat_ptrs[gi]], None, self.rank, b - a, abort_flag=self.timeout)

    def check(self):
        """Host-synchronising: raises if a barrier ever timed out (a peer never arrived)."""
        if int(self.timeout.item()) != 0:
            raise RuntimeError("SymmFedAvg: cross-rank barrier timed out")


class PipelinedFedAvg:
    """Layer-wise FedAvg that hides behind the next round's forward pass.

    The flat buffers are cut into chunks in the order the forward pass consumes them (the non-block parameters, then block
    0, 1, ...).  `launch()` enqueues, on a communication stream, for every chunk: exchange of its ranges -> cross-rank
    barrier -> refresh of its slice of the bf16 weight shadow -> event.  The caller's forward pass waits for chunk c's
    event right before the first kernel that reads chunk c (graph.GraphedTrainStep.finish(wait_events=...): one CUDA graph
    per block), so the aggregation of block l+1.. overlaps the compute of block l.  The exchange kernels run with a capped
    grid (`max_blocks`): they are NVLink bound with a fraction of the SMs and leave the rest to the forward pass.

    `agg`: PeerFedAvg | SymmFedAvg | None (impl "nccl": scale kernel + NCCL all-reduce per range); uniform weights.
    `chunks`: list of lists of (flat-buffer index, start, end) in elements."""

    def __init__(self, flats: Sequence["FlatParams"], chunks, agg=None, max_blocks: int = 32):
        self.flats, self.chunks, self.agg = list(flats), chunks, agg
        self.max_blocks = int(max_blocks)
        self.events = [torch.cuda.Event() for _ in chunks]

    def launch(self, comm_stream, impl: str = "peer", update_chunk=None):
        """Returns the per-chunk events (valid until the next launch).
        update_chunk(c): optional -- launches the optimizer update of chunk c on the CURRENT stream (graph.GraphedTrainStep(
        defer_adamw=True) + FusedAdamW.adamw_params): the exchange of chunk c then starts as soon as ITS update is done, while
        the update of chunk c + 1 is still running (HBM bound) -- the aggregation starts a whole AdamW pass earlier."""
        from ._lib import lib
        main = torch.cuda.current_stream()
        ready = None
        if update_chunk is None:
            comm_stream.wait_stream(main)   # the local update is complete
        else:
            ready = []
            for c in range(len(self.chunks)):
                update_chunk(c)
                e = torch.cuda.Event()
                e.record(main)
                ready.append(e)
        old = lib.b200asr_fedavg_set_max_blocks(self.max_blocks)
        try:
            with torch.cuda.stream(comm_stream):
                if impl != "nccl" and ready is None:
                    self.agg.barrier()   # every client's local update is complete and visible
                def landed(c):   # chunk c's average is in place on every rank: refresh its bf16 shadow slice, signal it
                    for gi, a, b in self.chunks[c]:
                        fp = self.flats[gi]
                        if fp.shadow is not None:
                            ops.cast_from_f32(fp.flat[a:b], torch.bfloat16, out=fp.shadow[a:b])
                    self.events[c].record(comm_stream)

                merged = ready is not None and impl != "nccl"
                for c, ranges in enumerate(self.chunks):
                    if ready is not None:
                        comm_stream.wait_event(ready[c])
                        if impl != "nccl":
                            # ONE barrier between two exchanges: every client has updated chunk c AND every rank's slice of
                            # chunk c - 1 has landed everywhere
                            self.agg.barrier()
                            if c > 0:
                                landed(c - 1)
                    for gi, a, b in ranges:
                        if impl == "nccl":
                            fedavg_allreduce(self.flats[gi].flat[a:b])
                        else:
                            self.agg.reduce_range(gi, a, b, impl)
                    if not merged:
                        if impl != "nccl":
                            self.agg.barrier()   # every rank's slice of this chunk has landed everywhere
                        landed(c)
                if merged:
                    self.agg.barrier()
                    landed(len(self.chunks) - 1)
        finally:
            lib.b200asr_fedavg_set_max_blocks(old)
        for fp in self.flats:
            fp.mark_updated(shadow_written=True)   # (host bookkeeping: the shadow slices were refreshed chunk by chunk)
        return self.events


def fedavg_shard_plan(n: int, world: int):
    """Slice [lo, hi) of the flat buffer owned by each rank in the fused peer kernel (multiples of 4 elements;
    the last rank takes the remainder).  Mirrors b200asr_fedavg_peer_reduce."""
    per = (n // world) // 4 * 4
    return [(per * r, n if r == world - 1 else per * (r + 1)) for r in range(world)]
