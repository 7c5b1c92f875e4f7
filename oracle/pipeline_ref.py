"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the steps either side of the model.

  normalize_and_pad : Wav2Vec2FeatureExtractor.zero_mean_unit_var_norm as `processor(audio, sampling_rate=16000)` applies it
                      per utterance at centralized/finetune_ASRs.py:88-96 ((x - x.mean()) / sqrt(x.var() + 1e-7), float32 numpy)
                      + DataCollatorCTCWithPadding (finetune_ASRs.py:59-86): zero padding, attention mask, -100 label padding
  word_error_rate   : what jiwer.wer(label_str, pred_str) computes at federated/src/update.py:36-50 (jiwer is not in this
                      image; its published definition: Levenshtein distance over whitespace-split words, summed over the
                      sentences, divided by the total number of reference words), restated on token ids with `delim` = '|'
Parity unpinned by reference fixtures: the reference holds no golden vectors for these steps; the numpy
normalisation IS the dependency's own two-line formula.
"""
from typing import List, Sequence

import numpy as np


def normalize_and_pad(waves: Sequence[np.ndarray], conv_kernel=(10, 3, 3, 3, 3, 2, 2), conv_stride=(5, 2, 2, 2, 2, 2, 2)):
    lens = [len(w) for w in waves]
    Lmax = max(lens)
    x = np.zeros((len(waves), Lmax), dtype=np.float32)
    mask = np.zeros((len(waves), Lmax), dtype=np.int32)
    frames = []
    for b, w in enumerate(waves):
        w = np.asarray(w, dtype=np.float32)
        x[b, : len(w)] = (w - w.mean()) / np.sqrt(w.var() + 1e-7)
        mask[b, : len(w)] = 1
        l = len(w)
        for k, s in zip(conv_kernel, conv_stride):
            l = (l - k) // s + 1 if l >= k else 0
        frames.append(l)
    return x, mask, np.asarray(frames, dtype=np.int32)


def split_words(ids: Sequence[int], delim: int) -> List[tuple]:
    words, cur = [], []
    for v in ids:
        if v < 0:
            continue
        if v == delim:
            if cur:
                words.append(tuple(cur))
                cur = []
        else:
            cur.append(int(v))
    if cur:
        words.append(tuple(cur))
    return words


def levenshtein(a: Sequence, b: Sequence) -> int:
    prev = list(range(len(b) + 1))
    for i in range(1, len(a) + 1):
        cur = [i] + [0] * len(b)
        for j in range(1, len(b) + 1):
            cur[j] = min(prev[j - 1] + (a[i - 1] != b[j - 1]), prev[j] + 1, cur[j - 1] + 1)
        prev = cur
    return prev[len(b)]


def word_error_terms(hyp: Sequence[Sequence[int]], ref: Sequence[Sequence[int]], delim: int = 4, unit: str = "word"):
    """Per-sentence (edit distance, reference token count)."""
    out = []
    for h, r in zip(hyp, ref):
        if unit == "word":
            hw, rw = split_words(h, delim), split_words(r, delim)
        else:
            hw, rw = [v for v in h if v >= 0], [v for v in r if v >= 0]
        out.append((levenshtein(hw, rw), len(rw)))
    return out
