"""The steps either side of the model, on the device.

  collate_on_device  <-> Wav2Vec2Processor normalisation + DataCollatorCTCWithPadding
                         (centralized/finetune_ASRs.py:59-96; with AD labels centralized/Models.py:166-195)
  error_rate         <-> compute_metrics' jiwer.wer on the decoded strings (federated/src/update.py:36-50), computed on
                         token ids: the greedy transcript never leaves the GPU until the final two integers
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence

import torch

from . import ops
from ._lib import check, lib


def collate_on_device(waves: Sequence[torch.Tensor], labels: Optional[Sequence[Sequence[int]]] = None,
                      conv_kernel: Sequence[int] = (10, 3, 3, 3, 3, 2, 2), conv_stride: Sequence[int] = (5, 2, 2, 2, 2, 2, 2),
                      device=None, dementia_labels: Optional[Sequence[int]] = None) -> Dict[str, torch.Tensor]:
    """Raw 16 kHz utterances (1-D float tensors / arrays, any lengths) -> the batch dict the model takes:
    `input_values` [B, Lmax] fp32 (each utterance normalised to zero mean / unit variance over its own samples, zero
    padded), `attention_mask` [B, Lmax] int32, `frame_lengths` [B] int32, `labels` [B, Tmax] int64 (-100 padded).
    One pinned staging buffer, one host->device copy, two kernels."""
    device = torch.device(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
    B = len(waves)
    if B == 0:
        raise ValueError("collate_on_device: empty batch")
    lens = [int(w.numel() if isinstance(w, torch.Tensor) else len(w)) for w in waves]
    Lmax = max(lens)
    if Lmax == 0:
        raise ValueError("collate_on_device: all utterances are empty")
    offs = [0]
    for n in lens:
        offs.append(offs[-1] + n)
    stage = torch.empty(offs[-1], dtype=torch.float32, pin_memory=True)
    for w, o, n in zip(waves, offs, lens):
        stage[o:o + n].copy_(torch.as_tensor(w, dtype=torch.float32).reshape(-1))
    packed = stage.to(device, non_blocking=True)
    offsets = torch.tensor(offs, dtype=torch.int64).to(device, non_blocking=True)
    ck = torch.tensor(list(conv_kernel), dtype=torch.int32).to(device, non_blocking=True)
    cs = torch.tensor(list(conv_stride), dtype=torch.int32).to(device, non_blocking=True)
    x = torch.empty(B, Lmax, dtype=torch.float32, device=device)
    mask = torch.empty(B, Lmax, dtype=torch.int32, device=device)
    frames = torch.empty(B, dtype=torch.int32, device=device)
    nws = lib.b200asr_normalize_pad_workspace_bytes(B, Lmax)
    ws = ops.workspace(nws, device)
    check(lib.b200asr_normalize_pad(packed.data_ptr(), offsets.data_ptr(), x.data_ptr(), mask.data_ptr(), frames.data_ptr(),
                                    ws.data_ptr(), nws, B, Lmax, ck.data_ptr(), cs.data_ptr(), len(conv_kernel),
                                    torch.cuda.current_stream().cuda_stream), "normalize_pad")
    packed.record_stream(torch.cuda.current_stream())
    out = {"input_values": x, "attention_mask": mask, "frame_lengths": frames}
    if labels is not None:
        Tmax = max(1, max(len(l) for l in labels))
        lab = torch.full((B, Tmax), -100, dtype=torch.int64)
        for b, l in enumerate(labels):
            if len(l):
                lab[b, : len(l)] = torch.as_tensor(list(l), dtype=torch.int64)
        out["labels"] = lab.to(device, non_blocking=True)
    if dementia_labels is not None:
        out["dementia_labels"] = torch.as_tensor(list(dementia_labels), dtype=torch.int64).to(device, non_blocking=True)
    return out


def _tokens(ids: torch.Tensor, lengths: Optional[torch.Tensor], delim: int):
    ids = ids.to(torch.int32).contiguous()
    B, ld = ids.shape
    words = torch.empty(B, ld, dtype=torch.int64, device=ids.device)  # u64 hashes, stored in int64 bit patterns
    nwords = torch.empty(B, dtype=torch.int32, device=ids.device)
    lp = None if lengths is None else lengths.to(torch.int32).contiguous()
    check(lib.b200asr_words_from_ids(ids.data_ptr(), None if lp is None else lp.data_ptr(), B, ld, int(delim), words.data_ptr(),
                                     nwords.data_ptr(), ld, torch.cuda.current_stream().cuda_stream), "words_from_ids")
    return words, nwords


def error_rate(hyp_ids: torch.Tensor, hyp_len: Optional[torch.Tensor], ref_ids: torch.Tensor, delim: int = 4,
               unit: str = "word"):
    """hyp_ids int [B, T]: collapsed greedy ids (padding < 0, as `ops.greedy_decode` returns them) with `hyp_len` [B];
    ref_ids int [B, L]: label ids padded with negative values (-100).  `delim` = id of the word delimiter `|`.
    unit 'word' -> WER terms, 'char' -> CER terms.  Returns (edit_distances [B] int32, ref_token_counts [B] int32) on the
    device; the corpus rate jiwer reports is dist.sum() / count.sum()."""
    if unit not in ("word", "char"):
        raise ValueError("unit must be 'word' or 'char'")
    d = int(delim) if unit == "word" else -1
    hw, hn = _tokens(hyp_ids, hyp_len, d)
    rw, rn = _tokens(ref_ids, None, d)
    B = hw.shape[0]
    dist = torch.empty(B, dtype=torch.int32, device=hw.device)
    check(lib.b200asr_edit_distance(hw.data_ptr(), hn.data_ptr(), rw.data_ptr(), rn.data_ptr(), dist.data_ptr(), B, hw.shape[1],
                                    rw.shape[1], torch.cuda.current_stream().cuda_stream), "edit_distance")
    return dist, rn


def wer(hyp_ids, hyp_len, ref_ids, delim: int = 4) -> float:
    """Corpus word error rate (one host read of two integers)."""
    dist, n = error_rate(hyp_ids, hyp_len, ref_ids, delim, "word")
    tot = torch.stack([dist.sum(), n.sum()]).cpu()
    return float(tot[0]) / max(1, int(tot[1]))
