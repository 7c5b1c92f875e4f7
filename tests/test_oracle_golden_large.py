"""CPU: the oracle restatement at the FULL-SIZE configurations of BASELINE.json reproduces what the reference's own
classes produced there (tests/golden/large_*.npz from tests/golden/make_golden_large.py): data2vec-audio-large greedy
inference (configs[0]), the B=8 x 10 s CTC training step (configs[1]), DACS stage 2 (configs[2]), and the
wav2vec2-large-lv60 / hubert-xlarge members of configs[4].  About 2.5 minutes on 8 cores."""
import numpy as np
import pytest
import torch

from oracle import dacs_ref as D
from oracle import data2vec_ref as R

import large_cases as LC

torch.set_num_threads(max(4, torch.get_num_threads()))


def test_large_c1_greedy_inference_matches_hf(golden_dir):
    g, cfg, sd, b = LC.setup(golden_dir, "large_c1_infer")
    with torch.no_grad():
        col = []
        loss, logits = R.ctc_forward(sd, cfg, b["input_values"], None, b["labels"], collect=col)
    assert LC.relinf(logits, g["logits"]) < 1e-4
    np.testing.assert_allclose(float(loss), float(g["loss"]), rtol=1e-5)
    ids, seqs = R.greedy_ids(logits)
    assert (ids.numpy() == g["argmax"]).all()
    assert seqs[0] == g["collapsed"].tolist() and len(seqs[0]) > 50   # a non-degenerate transcript
    fs = int(g["frame_stride"])
    L = cfg.num_hidden_layers
    for tag, i in (("first", 0), ("mid", L // 2), ("last", L)):
        assert LC.relinf(col[i][:, ::fs], g["hidden_" + tag]) < 1e-4, tag
    rms = torch.stack([h.pow(2).mean((1, 2)).sqrt() for h in col]).numpy()
    np.testing.assert_allclose(rms, g["hidden_rms"], rtol=1e-4)


@pytest.mark.parametrize("name", ["large_c2_train", "large_c5_w2v2_lv60", "large_c5_hubert_xlarge"])
def test_large_training_step_matches_hf(golden_dir, name):
    g, cfg, sd, b = LC.setup(golden_dir, name)
    fe = cfg.prefix + "feature_extractor"
    sd = {k: v.requires_grad_(not k.startswith(fe)) for k, v in sd.items()}
    col = []
    loss, logits = R.ctc_forward(sd, cfg, b["input_values"], b["attention_mask"], b["labels"], collect=col)
    loss.backward()
    assert LC.relinf(logits, g["logits"]) < 2e-4
    np.testing.assert_allclose(float(loss), float(g["loss"]), rtol=1e-5)
    assert (logits.argmax(-1).numpy() == g["argmax"]).mean() > 0.9995   # (fp32 summation order: a near-tie may flip)
    fs, blk = int(g["frame_stride"]), int(g["grad_block"])
    L = cfg.num_hidden_layers
    for tag, i in (("first", 0), ("mid", L // 2), ("last", L)):
        assert LC.relinf(col[i][:, ::fs], g["hidden_" + tag]) < 2e-4, tag
    # gradients: two fp32 implementations (different summation orders in conv / matmul) drift apart through 24-48 layers
    # of backward; measured up to 1.6e-3 on the earliest parameters of the pre-LN stacks with the thread count the fixture
    # was written with and 7.4e-3 with another (torch's CPU reductions change their summation order with the number of
    # threads; these posteriors are peaked: tools/grad_diag.py) -- the bound is the GPU test's fp32 gradient tolerance
    for n in g["grad_names"]:
        assert LC.relinf(LC.grad_block(sd[str(n)].grad, blk), g["grad:" + str(n)]) < 1.5e-2, n
    for n, v in zip(g["gnorm_names"], g["gnorm_values"]):
        if str(n).endswith("k_proj.bias"):
            continue  # exactly zero in exact arithmetic (softmax shift invariance): rounding noise on both sides
        assert abs(float(sd[str(n)].grad.norm()) - v) <= 1.5e-2 * max(v, 1e-3), n


def test_large_c3_dacs_stage2_matches_reference(golden_dir):
    """The 2 x 8 x 499 x 1024 hard Gumbel decisions sit behind a comparison: the oracle's encoder differs from the
    reference's by fp32 rounding (~1e-6), which flips the handful of decisions that are tied to that precision.  The
    test states the flip count, compares loss / logits / gradients with the oracle's own decisions, and the logits
    element-wise on every frame without a flipped decision."""
    g, cfg, sd, b = LC.setup(golden_dir, "large_c3_dacs2")
    gl, ga = LC.dacs_noise(g, cfg)
    ref_lm, ref_ad = LC.dacs_ref_masks(g)
    train = ("arbitrator.weight", "arbitrator.bias")
    sd = {k: v.requires_grad_(k in train) for k, v in sd.items()}
    with torch.no_grad():
        hidden, _ = R.base_model(sd, cfg, b["input_values"], b["attention_mask"])
    out = D.dacs_forward(sd, cfg, D.default_args(2), b["input_values"], b["attention_mask"], b["labels"],
                         b["dementia_labels"], gl, ga, hidden=hidden)
    out["loss"].backward()
    assert {str(n) for n in g["gnorm_names"]} == set(train)
    flips_lm = (out["lm_mask"].detach() > 0.5) != ref_lm
    flips_ad = (out["ad_mask"].detach() > 0.5) != ref_ad
    n_flip = int(flips_lm.sum()) + int(flips_ad.sum())
    print(f"[c3 oracle] flipped hard decisions: {n_flip} of {2 * ref_lm.numel()}")
    assert n_flip <= 2e-5 * 2 * ref_lm.numel()
    clean = ~flips_lm.any(-1)   # frames whose lm mask agrees everywhere
    assert clean.float().mean() > 0.99
    ref_logits = torch.from_numpy(g["logits"])
    assert LC.relinf(out["logits"].detach()[clean], ref_logits[clean]) < 2e-4
    np.testing.assert_allclose(float(out["loss"]), float(g["loss"]), rtol=1e-3)
    blk = int(g["grad_block"])
    for n in g["grad_names"]:
        assert LC.relinf(LC.grad_block(sd[str(n)].grad, blk), g["grad:" + str(n)]) < 1e-2, n
    for n, v in zip(g["gnorm_names"], g["gnorm_values"]):
        assert abs(float(sd[str(n)].grad.norm()) - v) <= 3e-3 * v, n
