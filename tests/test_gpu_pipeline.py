"""GPU parity of the device-side input pipeline and WER against oracle/pipeline_ref.py."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from b200asr import data, ops  # noqa: E402
from oracle import pipeline_ref as P  # noqa: E402


@pytest.mark.parametrize("lens", [[16000], [80000, 51234, 79999, 400], [480000, 123457], [5, 11, 10]])
def test_collate_on_device_matches_processor_and_collator(lens):
    rng = np.random.default_rng(sum(lens))
    waves = [(rng.standard_normal(n) * rng.uniform(0.01, 3.0) + rng.uniform(-0.5, 0.5)).astype(np.float32) for n in lens]
    labels = [list(rng.integers(1, 32, size=rng.integers(0, 9))) for _ in lens]
    x_ref, m_ref, f_ref = P.normalize_and_pad(waves)
    out = data.collate_on_device([torch.from_numpy(w) for w in waves], labels, dementia_labels=[i % 2 for i in range(len(lens))])
    torch.cuda.synchronize()
    x = out["input_values"].cpu().numpy()
    assert x.shape == x_ref.shape and x.dtype == np.float32
    # fp32 tolerance 1e-5 absolute on unit-variance data (numpy's float32 mean/var vs the kernel's double accumulation)
    np.testing.assert_allclose(x, x_ref, rtol=0, atol=2e-5)
    assert (out["attention_mask"].cpu().numpy() == m_ref).all()
    assert (out["frame_lengths"].cpu().numpy() == f_ref).all()
    lab = out["labels"].cpu().numpy()
    for b, l in enumerate(labels):
        assert list(lab[b, : len(l)]) == [int(v) for v in l] and (lab[b, len(l):] == -100).all()
    assert out["dementia_labels"].tolist() == [i % 2 for i in range(len(lens))]


@pytest.mark.parametrize("unit", ["word", "char"])
def test_error_rate_matches_levenshtein_oracle(unit):
    rng = np.random.default_rng(7)
    B, T, L = 37, 150, 120
    hyp = np.full((B, T), -1, dtype=np.int32)
    ref = np.full((B, L), -100, dtype=np.int64)
    hlen = np.zeros(B, dtype=np.int32)
    hyps, refs = [], []
    for b in range(B):
        n = int(rng.integers(0, L + 1))
        r = rng.integers(4, 12, size=n)  # small alphabet with many delimiters (id 4): plenty of words and of matches
        h = list(r)
        for _ in range(int(rng.integers(0, 12))):  # random edits of the reference
            if h and rng.random() < 0.6:
                h[int(rng.integers(0, len(h)))] = int(rng.integers(4, 12))
            elif h and rng.random() < 0.5:
                del h[int(rng.integers(0, len(h)))]
            else:
                h.insert(int(rng.integers(0, len(h) + 1)), int(rng.integers(4, 12)))
        h = h[:T]
        if b == 0:
            h, r = [], r           # empty hypothesis
        if b == 1:
            r = np.zeros(0, dtype=np.int64)  # empty reference
        hyp[b, : len(h)] = h
        hlen[b] = len(h)
        ref[b, : len(r)] = r
        hyps.append(list(h)); refs.append([int(v) for v in r])
    terms = P.word_error_terms(hyps, refs, delim=4, unit=unit)
    dist, cnt = data.error_rate(torch.from_numpy(hyp).cuda(), torch.from_numpy(hlen).cuda(), torch.from_numpy(ref).cuda(), 4, unit)
    assert dist.cpu().tolist() == [t[0] for t in terms]   # integers: exact
    assert cnt.cpu().tolist() == [t[1] for t in terms]
    if unit == "word":
        want = sum(t[0] for t in terms) / max(1, sum(t[1] for t in terms))
        got = data.wer(torch.from_numpy(hyp).cuda(), torch.from_numpy(hlen).cuda(), torch.from_numpy(ref).cuda(), 4)
        assert abs(got - want) < 1e-12


def test_greedy_decode_to_wer_end_to_end():
    """logits -> greedy ids -> collapsed ids -> WER without leaving the device."""
    torch.manual_seed(0)
    B, T, V = 4, 60, 32
    ref = torch.randint(4, 10, (B, 20))
    logits = torch.full((B, T, V), -5.0)
    for b in range(B):
        for i, v in enumerate(ref[b].tolist()):   # each label held for 2 frames with a blank in between
            logits[b, 3 * i, v] = 5.0
            logits[b, 3 * i + 1, v] = 5.0
            logits[b, 3 * i + 2, 0] = 5.0
    ids, col, clen = ops.greedy_decode(logits.cuda(), None, 0)
    assert data.wer(col, clen, ref.cuda(), 4) == 0.0
    bad = ref.clone()
    bad[:, 0] = 11  # not a delimiter: the first word of every reference changes
    dist, cnt = data.error_rate(col, clen, bad.cuda(), 4, "word")
    assert (dist.cpu() >= 1).all()


def test_conv_out_lengths_kernel_matches_the_torch_chain():
    """b200asr_conv_out_lengths: the per-layer floor((len - k) / s) + 1 chain of _get_feat_extract_output_lengths in one
    launch equals modeling.conv_out_lengths (torch.div(rounding_mode="floor")), too-short inputs (negative numerators)
    included; wired into the model through modeling.device_conv_out_lengths."""
    from b200asr import modeling as M2
    from b200asr import ops as O2
    from oracle import data2vec_ref as R
    cfg = R.ArchConfig()
    lens = torch.tensor([160000, 159999, 80000, 400, 399, 25, 10, 9, 1, 0, 480000, 123457], dtype=torch.int64, device="cuda")
    want = M2.conv_out_lengths(cfg, lens).to(torch.int32)
    got = O2.conv_out_lengths(lens, list(cfg.conv_kernel), list(cfg.conv_stride))
    assert got.dtype == torch.int32 and torch.equal(got, want)
    mask = (torch.arange(2000, device="cuda")[None] < torch.tensor([2000, 1234, 400, 7], device="cuda")[:, None]).long()
    assert torch.equal(M2.device_conv_out_lengths(cfg, mask), M2.conv_out_lengths(cfg, mask.sum(-1)).to(torch.int32))
