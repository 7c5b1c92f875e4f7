"""SURVEY 8f-3 on the device: the server side of one federated round over saved client directories
(`average_weights`, federated/src/utils.py:240-249, applied to the `pytorch_model.bin` files of the clients) must match
the reference's key-by-key CPU mean bit for bit in the uniform case."""
import copy
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "privacy-preserve-federated-asr_b200")]

from b200asr import checkpoint as ck  # noqa: E402
from b200asr import modeling as M  # noqa: E402
from oracle.data2vec_ref import ArchConfig  # noqa: E402


def reference_average_weights(w):   # federated/src/utils.py:240-249, restated
    w_avg = copy.deepcopy(w[0])
    for key in w_avg.keys():
        for i in range(1, len(w)):
            w_avg[key] += w[i][key]
        w_avg[key] = torch.div(w_avg[key], len(w))
    return w_avg


@pytest.mark.parametrize("nclients", [2, 5])
def test_average_checkpoints_matches_average_weights(tmp_path, nclients):
    cfg = ArchConfig.tiny()
    dirs, sds = [], []
    for c in range(nclients):
        torch.manual_seed(100 + c)
        m = M.Data2VecAudioForCTC(cfg.to_namespace() if hasattr(cfg, "to_namespace") else cfg, compute_dtype=torch.float32)
        with torch.no_grad():
            for p in m.parameters():
                p.normal_(0.0, 0.5 + 0.1 * c)
        d = str(tmp_path / f"client{c}")
        m.save_pretrained(d)
        dirs.append(d)
        sds.append(ck.load_state_dict_file(d))
    out = str(tmp_path / "global")
    avg = ck.average_checkpoints(dirs, out)
    ref = reference_average_weights(sds)
    assert list(avg.keys()) == list(ref.keys())
    for k in ref:
        assert torch.equal(avg[k], ref[k]), k
    # the written global checkpoint opens again, and carries the same tensors
    g = M.Data2VecAudioForCTC.from_pretrained(out, compute_dtype=torch.float32)
    for k, v in g.state_dict().items():
        assert torch.equal(v.cpu(), ref[k]), k
    # weighted variant (sample-count weights): within 1e-6 of the fp32 formula
    w = [1.0 + c for c in range(nclients)]
    tot = sum(w)
    w = [x / tot for x in w]
    wavg = ck.average_state_dicts(sds, w)
    for k in ref:
        r = sum(wi * sd[k] for wi, sd in zip(w, sds))
        assert (wavg[k] - r).abs().max() <= 1e-6 * max(1.0, float(r.abs().max())), k
