"""the weight exchange files of the federated loop (`pytorch_model.bin` + `config.json`, written by
`trainer.save_model` and re-opened with `from_pretrained`, federated/src/federated_main.py:69-145) round-trip between the
b200asr modules and HF transformers on the host -- no GPU involved (the modules are only constructed and (de)serialised)."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "privacy-preserve-federated-asr_b200")]

from b200asr import checkpoint as ck  # noqa: E402
from b200asr import modeling as M  # noqa: E402
from oracle.data2vec_ref import ArchConfig  # noqa: E402


def hf_config(kind, cfg):
    tr = pytest.importorskip("transformers")
    common = dict(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, num_hidden_layers=cfg.num_hidden_layers,
                  num_attention_heads=cfg.num_attention_heads, intermediate_size=cfg.intermediate_size,
                  conv_dim=list(cfg.conv_dim), conv_kernel=list(cfg.conv_kernel), conv_stride=list(cfg.conv_stride),
                  conv_bias=cfg.conv_bias, num_conv_pos_embeddings=cfg.num_conv_pos_embeddings,
                  num_conv_pos_embedding_groups=cfg.num_conv_pos_embedding_groups, layer_norm_eps=cfg.layer_norm_eps,
                  pad_token_id=0, ctc_loss_reduction="sum", ctc_zero_infinity=True, mask_time_prob=0.0,
                  feat_extract_norm=cfg.feat_extract_norm, num_feat_extract_layers=len(cfg.conv_dim))
    if kind == "data2vec":
        return tr.Data2VecAudioConfig(conv_pos_kernel_size=cfg.conv_pos_kernel_size, **common), tr.Data2VecAudioForCTC
    common["do_stable_layer_norm"] = cfg.do_stable_layer_norm
    if kind == "wav2vec2":
        return tr.Wav2Vec2Config(**common), tr.Wav2Vec2ForCTC
    return tr.HubertConfig(feat_proj_layer_norm=cfg.feat_proj_layer_norm, **common), tr.HubertForCTC


CASES = [("data2vec", ArchConfig.tiny, M.Data2VecAudioForCTC), ("wav2vec2", ArchConfig.tiny_wav2vec2, M.Wav2Vec2ForCTC),
         ("hubert", ArchConfig.tiny_hubert, M.HubertForCTC)]


@pytest.mark.parametrize("kind,mk,cls", CASES, ids=[c[0] for c in CASES])
def test_hf_directory_opens_in_b200asr_and_back(tmp_path, kind, mk, cls):
    cfg = mk()
    hcfg, hcls = hf_config(kind, cfg)
    torch.manual_seed(0)
    hf = hcls(hcfg).eval()
    d1 = str(tmp_path / "hf")
    os.makedirs(d1)
    hf.config.save_pretrained(d1)   # the two files the reference's transformers version writes
    torch.save(hf.state_dict(), os.path.join(d1, ck.WEIGHTS_NAME))
    ours = cls.from_pretrained(d1, compute_dtype=torch.float32)
    d0 = str(tmp_path / "hf_current")   # what the installed transformers writes (model.safetensors) opens as well
    hf.save_pretrained(d0)
    ours0 = cls.from_pretrained(d0, compute_dtype=torch.float32)
    for (k, a), (_, b) in zip(ours.state_dict().items(), ours0.state_dict().items()):
        assert torch.equal(a, b), k
    ref = hf.state_dict()
    got = ours.state_dict()
    assert len(got) > 0
    for k, v in got.items():
        if k in ref:
            assert torch.equal(v.cpu(), ref[k]), k
    # and back: the directory we write opens in HF without missing / unexpected keys, tensors identical
    d2 = str(tmp_path / "ours")
    ours.save_pretrained(d2)
    hf2, info = hcls.from_pretrained(d2, output_loading_info=True)
    assert not info["missing_keys"] and not info["unexpected_keys"], info
    for k, v in hf2.state_dict().items():
        assert torch.equal(v, ref[k]), k
    # our own round trip is exact too
    ours2 = cls.from_pretrained(d2, compute_dtype=torch.float32)
    for (k, a), (_, b) in zip(ours.state_dict().items(), ours2.state_dict().items()):
        assert torch.equal(a, b), k


def test_config_json_fields(tmp_path):
    cfg = ArchConfig.tiny()
    hcfg, hcls = hf_config("data2vec", cfg)
    hf = hcls(hcfg)
    d = str(tmp_path / "m")
    hf.save_pretrained(d)
    c = ck.load_config(d)
    assert c.model_type == "data2vec-audio" and c.hidden_size == cfg.hidden_size and list(c.conv_kernel) == list(cfg.conv_kernel)
    with pytest.raises(FileNotFoundError):
        ck.load_state_dict_file(str(tmp_path / "nothing"))


def test_plain_ctc_checkpoint_opens_in_the_dacs_class(tmp_path):
    """The reference's main flow: `Data2VecAudioForCTC.from_pretrained(pretrain_name, config=config, args=args)` on a
    plain HF CTC checkpoint (federated/src/update.py:228) -- HF initialises the heads the checkpoint lacks (arbitrator,
    dementia_head, criterion_similar) and reports them; strict=True is the opt-in that raises."""
    from types import SimpleNamespace
    cfg = ArchConfig.tiny()
    hcfg, hcls = hf_config("data2vec", cfg)
    torch.manual_seed(1)
    hf = hcls(hcfg).eval()
    d = str(tmp_path / "ctc")
    os.makedirs(d)
    hf.config.save_pretrained(d)
    torch.save(hf.state_dict(), os.path.join(d, ck.WEIGHTS_NAME))
    args = SimpleNamespace(LAMBDA=0.5, TOGGLE_RATIO=0, GS_TAU=1, AD_loss="recall", W_LOSS=None, STAGE=2)
    with pytest.warns(UserWarning, match="newly initialised"):
        m = M.DacsData2VecAudioForCTC.from_pretrained(d, args, compute_dtype=torch.float32)
    ref = hf.state_dict()
    for k, v in m.state_dict().items():
        if k in ref:
            assert torch.equal(v, ref[k]), k
    assert m.arbitrator.weight.shape == (4 * cfg.hidden_size, cfg.hidden_size)
    with pytest.raises(RuntimeError, match="missing keys"):
        M.DacsData2VecAudioForCTC.from_pretrained(d, args, compute_dtype=torch.float32, strict=True)
