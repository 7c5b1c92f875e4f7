"""The weight exchange format of the reference's federated loop.

Every client round ends with `trainer.save_model(dir)` (HF `save_pretrained`: `pytorch_model.bin` = torch.save of the
state_dict under HF key names + `config.json`), the server re-opens the directories with
`Data2VecAudioForCTC.from_pretrained(dir)`, averages the state_dicts key by key on the CPU
(federated/src/utils.py:240-249, round loop federated/src/federated_main.py:69-145) and saves the global model the same
way.  This module reads and writes exactly those two files for the b200asr modules, so directories written by either side
open on the other, and performs the aggregation of saved client checkpoints on the device:

  save_pretrained(model, dir)                  <-> model.save_pretrained(dir)
  from_pretrained(cls, dir, ...)               <-> cls.from_pretrained(dir)
  average_checkpoints(dirs, out_dir, weights)  <-> average_weights([m.state_dict() for m in ...]) + save

`average_checkpoints` packs each client's tensors into one flat fp32 buffer in HBM and runs the N-way FedAvg kernel
(`b200asr_fedavg`) once over the whole model instead of one CPU add per key; integer buffers, if any, are taken from the
first client (the reference's `torch.div` would turn them into floats; the models of this path have none).
"""
from __future__ import annotations

import json
import os
from collections import OrderedDict
from types import SimpleNamespace
from typing import Dict, Optional, Sequence

import torch

WEIGHTS_NAME = "pytorch_model.bin"
CONFIG_NAME = "config.json"

# HF `model_type` / `architectures` per backbone prefix (transformers configuration_*.py)
_FAMILY = {
    "data2vec_audio": ("data2vec-audio", "Data2VecAudioForCTC"),
    "wav2vec2": ("wav2vec2", "Wav2Vec2ForCTC"),
    "hubert": ("hubert", "HubertForCTC"),
}


def config_to_dict(config) -> Dict:
    """JSON-able dict of a model config: a transformers PretrainedConfig (`to_dict`) or any attribute bag."""
    if hasattr(config, "to_dict"):
        d = dict(config.to_dict())
    else:
        d = {k: v for k, v in vars(config).items() if not k.startswith("_")}
    out = {}
    for k, v in d.items():
        if isinstance(v, tuple):
            v = list(v)
        if isinstance(v, torch.dtype):
            v = str(v).replace("torch.", "")
        try:
            json.dumps(v)
        except TypeError:
            continue
        out[k] = v
    return out


def config_from_dict(d: Dict):
    """Attribute bag with the fields the b200asr modules read (lists stay lists: they are only iterated / indexed)."""
    return SimpleNamespace(**d)


def cpu_state_dict(model) -> "OrderedDict[str, torch.Tensor]":
    """state_dict with every tensor on the host, detached, contiguous, under the reference's key names (one device->host
    copy per flat parameter group would be faster still; checkpoints are written once per round)."""
    sd = OrderedDict()
    for k, v in model.state_dict().items():
        sd[k] = v.detach().to("cpu").contiguous().clone()
    return sd


def save_pretrained(model, save_directory: str, state_dict: Optional[Dict[str, torch.Tensor]] = None) -> None:
    """Write `pytorch_model.bin` + `config.json` as HF's save_pretrained does for the reference's models."""
    os.makedirs(save_directory, exist_ok=True)
    cfg = config_to_dict(model.config)
    prefix = getattr(model, "base_model_prefix", "data2vec_audio")
    mtype, arch = _FAMILY.get(prefix, _FAMILY["data2vec_audio"])
    cfg.setdefault("model_type", mtype)
    cfg.setdefault("architectures", [arch])
    with open(os.path.join(save_directory, CONFIG_NAME), "w") as f:
        json.dump(cfg, f, indent=2, sort_keys=True)
    sd = state_dict if state_dict is not None else cpu_state_dict(model)
    torch.save(OrderedDict((k, v.detach().cpu()) for k, v in sd.items()), os.path.join(save_directory, WEIGHTS_NAME))


def load_state_dict_file(directory_or_file: str) -> Dict[str, torch.Tensor]:
    """`pytorch_model.bin` (what the reference's transformers version writes); `model.safetensors` (what current
    transformers writes) is read too when that is all the directory holds."""
    path = directory_or_file
    if os.path.isdir(path):
        st = os.path.join(path, "model.safetensors")
        path = os.path.join(path, WEIGHTS_NAME)
        if not os.path.isfile(path) and os.path.isfile(st):
            from safetensors.torch import load_file
            return load_file(st)
    if not os.path.isfile(path):
        raise FileNotFoundError(f"no {WEIGHTS_NAME} at {directory_or_file}")
    sd = torch.load(path, map_location="cpu", weights_only=True)
    if not isinstance(sd, dict):
        raise ValueError(f"{path}: expected a state_dict")
    return sd


def load_config(directory: str):
    with open(os.path.join(directory, CONFIG_NAME)) as f:
        return config_from_dict(json.load(f))


def from_pretrained(cls, directory: str, *model_args, config=None, strict: bool = False, device=None, **model_kwargs):
    """cls(config, *model_args, **model_kwargs) with the weights of `directory` (the reference's
    `Data2VecAudioForCTC.from_pretrained(path)`; for DacsData2VecAudioForCTC pass `args` as the reference does).
    Keys of the reference's weight-normed positional conv (`weight_g` / `weight_v` or `parametrizations.*`) are
    accepted either way (modeling._WeightNormConv).
    strict=False (default) is HF's behaviour, which the reference's main flow relies on: `Data2VecAudioForCTC.from_pretrained(
    pretrain_name, config, args)` opens a plain CTC checkpoint and leaves arbitrator / dementia_head / criterion_similar
    freshly initialised (federated/src/update.py:228); missing and unexpected keys are reported by a warning.  strict=True
    raises instead."""
    config = config if config is not None else load_config(directory)
    model = cls(config, *model_args, **model_kwargs)
    sd = load_state_dict_file(directory)
    missing, unexpected = model.load_state_dict(sd, strict=False)
    if strict and (missing or unexpected):
        raise RuntimeError(f"from_pretrained({directory}): missing keys {list(missing)[:5]}..., unexpected {list(unexpected)[:5]}...")
    if missing or unexpected:
        import warnings
        warnings.warn(f"from_pretrained({directory}): newly initialised (not in the checkpoint): {sorted(missing)}; "
                      f"unused checkpoint keys: {sorted(unexpected)}")
    if device is not None:
        model.to(device)
    return model


def average_state_dicts(state_dicts: Sequence[Dict[str, torch.Tensor]], weights: Optional[Sequence[float]] = None,
                        device=None) -> "OrderedDict[str, torch.Tensor]":
    """FedAvg of host state_dicts on the device: one flat fp32 buffer per client, ONE N-way kernel, one copy back.
    weights=None is the reference's uniform mean (sum in client order, then divide by N: bit-identical)."""
    from .fedavg import fedavg_local
    if not state_dicts:
        raise ValueError("average_state_dicts: no clients")
    device = torch.device(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
    ref = state_dicts[0]
    fkeys = [k for k, v in ref.items() if v.is_floating_point()]
    for sd in state_dicts[1:]:
        if list(sd.keys()) != list(ref.keys()):
            raise ValueError("average_state_dicts: clients disagree on the state_dict keys")
    offs, off = [], 0
    for k in fkeys:
        offs.append(off)
        off += (ref[k].numel() + 3) // 4 * 4
    flats = []
    for sd in state_dicts:
        host = torch.zeros(off, dtype=torch.float32, pin_memory=torch.cuda.is_available())
        for k, o in zip(fkeys, offs):
            if sd[k].shape != ref[k].shape:
                raise ValueError(f"average_state_dicts: shape mismatch at {k}")
            host[o:o + sd[k].numel()].copy_(sd[k].reshape(-1).to(torch.float32))
        flats.append(host.to(device, non_blocking=True))
    avg = fedavg_local(flats, weights).cpu()
    out = OrderedDict()
    for k, v in ref.items():
        out[k] = v.clone()
    for k, o in zip(fkeys, offs):
        out[k] = avg[o:o + ref[k].numel()].view(ref[k].shape).to(ref[k].dtype).clone()
    return out


def average_checkpoints(directories: Sequence[str], out_directory: Optional[str] = None,
                        weights: Optional[Sequence[float]] = None, device=None):
    """The server side of one round over saved client directories; writes the global checkpoint when `out_directory` is
    given (config.json copied from the first client) and returns the averaged state_dict."""
    sds = [load_state_dict_file(d) for d in directories]
    avg = average_state_dicts(sds, weights, device)
    if out_directory is not None:
        os.makedirs(out_directory, exist_ok=True)
        with open(os.path.join(directories[0], CONFIG_NAME)) as f:
            cfg = f.read()
        with open(os.path.join(out_directory, CONFIG_NAME), "w") as f:
            f.write(cfg)
        torch.save(avg, os.path.join(out_directory, WEIGHTS_NAME))
    return avg
