"""Shared by tests/test_oracle_golden_large.py (CPU) and tests/test_gpu_large.py (B200): the full-size configurations of
BASELINE.json whose reference outputs live in tests/golden/large_*.npz (written by tests/golden/make_golden_large.py from
HF Data2VecAudioForCTC / Wav2Vec2ForCTC / HubertForCTC and federated/src/models.py::Data2VecAudioForCTC)."""
import os

import numpy as np
import torch

from oracle import data2vec_ref as R

SR = 16000

CASES = {
    # name: (config factory, kind)
    "large_c1_infer": (lambda: R.ArchConfig(), "infer"),
    "large_c2_train": (lambda: R.ArchConfig(), "train"),
    "large_c3_dacs2": (lambda: R.ArchConfig(), "dacs2"),
    "large_c5_w2v2_lv60": (lambda: R.ArchConfig(model_type="wav2vec2", feat_extract_norm="layer", conv_bias=True,
                                                do_stable_layer_norm=True, conv_pos_kernel_size=128), "train"),
    "large_c5_hubert_xlarge": (lambda: R.ArchConfig(model_type="hubert", feat_extract_norm="layer", conv_bias=True,
                                                    do_stable_layer_norm=True, conv_pos_kernel_size=128, hidden_size=1280,
                                                    num_hidden_layers=48, num_attention_heads=16, intermediate_size=5120),
                               "train"),
}


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


def setup(golden_dir, name):
    """(fixture, cfg, state_dict, batch) of one large case, rebuilt from the seeds the fixture records."""
    g = load(golden_dir, name)
    cfg = CASES[name][0]()
    kind = CASES[name][1]
    seed = int(g["seed"])
    sd = R.init_state_dict(cfg, seed=seed, dacs=(kind == "dacs2"), branch_gain=float(g["branch_gain"]),
                           qk_gain=float(g["qk_gain"]))
    batch, samples = int(g["batch"]), int(g["samples"])
    b = R.synth_batch(cfg, batch, samples, seed=(4321 if kind == "dacs2" else 1234) + seed, ragged=(kind != "infer"),
                      max_label=120)
    return g, cfg, sd, b


def dacs_noise(g, cfg):
    """The Exp(1) draws the reference consumed (models.py:112), regenerated from the recorded seeds and checked against
    the fixture's checksum -- 2 x 33 MB that need not be committed."""
    batch, samples = int(g["batch"]), int(g["samples"])
    T = int(R.conv_out_lengths(cfg, torch.tensor([samples]))[0])
    shape = (batch, T, cfg.hidden_size, 2)
    e_lm = torch.empty(shape).exponential_(generator=torch.Generator().manual_seed(int(g["noise_seed_lm"])))
    e_ad = torch.empty(shape).exponential_(generator=torch.Generator().manual_seed(int(g["noise_seed_ad"])))
    assert np.array_equal(e_lm.flatten()[:16].numpy(), g["noise_head_lm"]), "CPU generator stream differs from the fixture's"
    assert np.array_equal(e_ad.flatten()[:16].numpy(), g["noise_head_ad"])
    np.testing.assert_allclose([float(e_lm.double().sum()), float(e_ad.double().sum())], g["noise_check"], rtol=1e-12)
    return -e_lm.log(), -e_ad.log()


def dacs_ref_masks(g):
    """The hard masks the reference drew, (lm_mask, ad_mask) bool [B, T, d]."""
    shape = tuple(int(v) for v in g["mask_shape"])
    n = int(np.prod(shape))
    return (torch.from_numpy(np.unpackbits(g["lm_mask_bits"])[:n].reshape(shape).astype(bool)),
            torch.from_numpy(np.unpackbits(g["ad_mask_bits"])[:n].reshape(shape).astype(bool)))


def grad_block(t, block):
    t = t.detach()
    return t if t.dim() == 1 else t[:block, :block]


def relinf(a, b):
    a = torch.as_tensor(np.asarray(a) if not torch.is_tensor(a) else a).float().cpu()
    b = torch.as_tensor(np.asarray(b) if not torch.is_tensor(b) else b).float().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-12))


def rel2(a, b):
    """relative L2 (Frobenius) error ||a - b|| / ||b||."""
    a = torch.as_tensor(np.asarray(a) if not torch.is_tensor(a) else a).double().cpu()
    b = torch.as_tensor(np.asarray(b) if not torch.is_tensor(b) else b).double().cpu()
    return float((a - b).norm() / max(float(b.norm()), 1e-30))
