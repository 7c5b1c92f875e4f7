"""CPU-only checks: the C-ABI library loads and exports every symbol include/b200asr.h declares, the Python
binding lists them all, the drop-in modules expose the reference's state_dict keys, host-side length / shard
math, and the world_size-2 FedAvg plumbing over gloo.  No compute call is made (there is no GPU here)."""
import os
import re
import subprocess
import sys
from types import SimpleNamespace

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "b200asr.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200asr_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from b200asr import _lib
    syms = header_symbols()
    assert len(syms) >= 35
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (b200asr_[a-z0-9_]+)", out))
    missing = [s for s in syms if s not in exported]
    assert not missing, missing
    assert sorted(_lib.SIGNATURES) == syms  # the ctypes table binds exactly the header's API
    assert _lib.lib.b200asr_abi_version() == 2
    assert _lib.lib.b200asr_layernorm_bwd_workspace_bytes(1024) == 148 * 3 * 1024 * 4


def test_library_is_sm100a_tcgen05_tma():
    from b200asr import _lib
    r = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in r.stdout
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):
        assert mnemonic in r.stdout, mnemonic


def test_bad_arguments_fail_loudly_without_a_gpu():
    from b200asr import _lib
    rc = _lib.lib.b200asr_layernorm_fwd(None, None, None, None, None, None, None, None, 4, 8, 1e-5, 0, 0, None)
    assert rc == -1 and "null" in _lib.last_error()
    with pytest.raises(ValueError):
        _lib.check(rc, "layernorm_fwd")
    from b200asr import ops
    with pytest.raises(RuntimeError):
        ops.colsum(torch.zeros(4, 8))  # CPU tensor: no fallback


def _ns(cfg):
    return SimpleNamespace(**cfg.__dict__)


def test_state_dict_keys_match_reference_models():
    from b200asr import modeling as M
    from oracle import data2vec_ref as R, dacs_ref as D
    for conv_bias in (False, True):
        cfg = R.ArchConfig.tiny(conv_bias=conv_bias)
        ref = R.init_state_dict(cfg, seed=0, dacs=False)  # loads strict=True into HF Data2VecAudioForCTC (make_golden.py)
        m = M.Data2VecAudioForCTC(_ns(cfg), compute_dtype=torch.float32)
        sd = m.state_dict()
        assert list(sorted(sd)) == list(sorted(ref))
        assert all(sd[k].shape == ref[k].shape for k in ref)
    cfg = R.ArchConfig.tiny()
    ref = R.init_state_dict(cfg, seed=0, dacs=True)  # strict-loads into federated/src/models.py::Data2VecAudioForCTC
    for stage, trainable_prefix in ((0, ("data2vec_audio.", "lm_head.")), (1, ("dementia_head.",)), (2, ("arbitrator.",))):
        m = M.DacsData2VecAudioForCTC(_ns(cfg), D.default_args(stage), compute_dtype=torch.float32)
        assert sorted(m.state_dict()) == sorted(ref)
        for n, p in m.named_parameters():
            want = n.startswith(trainable_prefix) and not n.startswith("data2vec_audio.feature_extractor")
            assert p.requires_grad == want, (stage, n)


def test_full_size_parameter_count():
    from b200asr import modeling as M
    from oracle import data2vec_ref as R
    with torch.device("meta"):
        m = M.Data2VecAudioForCTC(_ns(R.ArchConfig(conv_bias=True)), compute_dtype=torch.bfloat16)
    assert sum(p.numel() for p in m.parameters()) == 313_311_776  # SURVEY 0.3 (conv_bias=True)


def test_conv_out_lengths_integer_floor():
    from b200asr.modeling import conv_out_lengths
    from oracle import data2vec_ref as R
    cfg = R.ArchConfig()
    L = torch.tensor([160000, 80000, 480000, 400, 96001, 123457])
    assert conv_out_lengths(cfg, L).tolist() == R.conv_out_lengths(cfg, L).tolist()
    assert conv_out_lengths(cfg, L).tolist()[:3] == [499, 249, 1499]


def test_fedavg_shard_plan_covers_buffer_once():
    from b200asr.fedavg import fedavg_shard_plan
    for n in (1, 3, 16, 1003, 313_311_776):
        for world in (1, 2, 4, 8):
            plan = fedavg_shard_plan(n, world)
            assert plan[0][0] == 0 and plan[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(plan, plan[1:]))
            assert all(lo % 4 == 0 for lo, _ in plan)


_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "privacy-preserve-federated-asr_b200"))
from b200asr.fedavg import fedavg_shard_plan
from oracle import dacs_ref as D
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:" + sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank, world, n = dist.get_rank(), 2, 1003
clients = [torch.randn(n, generator=torch.Generator().manual_seed(100 + r)) for r in range(world)]
ref = D.average_weights([{"w": c.clone()} for c in clients])["w"]
# path 1: prescale by 1/N then sum all-reduce (fedavg_allreduce's schedule, CPU stand-in for the scale kernel)
flat = clients[rank].clone() * (1.0 / world)
dist.all_reduce(flat)
assert torch.allclose(flat, ref, atol=1e-6), float((flat - ref).abs().max())
# path 2: the peer kernel's schedule: rank r reduces slice r of every client and publishes it to all
lo, hi = fedavg_shard_plan(n, world)[rank]
gathered = [torch.empty(n) for _ in range(world)]
dist.all_gather(gathered, clients[rank])
mine = torch.zeros(n)
acc = gathered[0][lo:hi].clone()
for j in range(1, world):
    acc += gathered[j][lo:hi]
mine[lo:hi] = acc / world
dist.all_reduce(mine)  # disjoint slices: the sum is the concatenation
assert torch.equal(mine, ref)
dist.destroy_process_group()
print("ok", rank)
"""


def test_fedavg_two_rank_schedule_over_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
