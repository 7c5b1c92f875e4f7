"""GPU: the CUDA-graph update with LayerDrop (graph.GraphedTrainStep: one graph per block and direction, host coins per
replay), the per-parameter AdamW table, and the weight-shadow refresh after load_state_dict."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from b200asr import modeling as M  # noqa: E402
from b200asr import ops  # noqa: E402
from b200asr.graph import GraphedTrainStep  # noqa: E402
from b200asr.optim import FusedAdamW  # noqa: E402
from oracle import data2vec_ref as R  # noqa: E402

DEV = "cuda"


def ns(cfg, **kw):
    c = SimpleNamespace(**{k: (tuple(v) if isinstance(v, (list, tuple)) else v) for k, v in cfg.__dict__.items()})
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def dev(b):
    return {k: v.to(DEV) for k, v in b.items() if k in ("input_values", "attention_mask", "labels")}


def relinf(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-6))


def make(cfg, sd, dtype=torch.bfloat16, **cfg_kw):
    m = M.Data2VecAudioForCTC(ns(cfg, **cfg_kw), compute_dtype=dtype)
    m.load_state_dict(sd)
    m.freeze_feature_encoder()
    m.to(DEV).train()
    m.validate_labels = False
    return m, FusedAdamW(m.named_parameters(), lr=1e-3, weight_decay=0.01, warmup_steps=2)


@pytest.mark.parametrize("mode", ["plain", "split_conv", "deferred_adamw", "split_conv+deferred_adamw"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_layerdrop_graph_matches_eager_with_the_same_coins(dtype, mode):
    """Six updates with layerdrop 0.4 on a 4-block model: the segmented CUDA-graph replay and the eager path, driven by
    the same host coin stream (np.random.uniform, as the pinned transformers 4.17), skip the same blocks, report the same
    losses and end with the same parameters, Adam moments and per-parameter step counts."""
    cfg = R.ArchConfig.tiny(num_hidden_layers=4)
    sd = R.init_state_dict(cfg, seed=21, dacs=False)
    batches = [R.synth_batch(cfg, 2, 6000, seed=170 + i, ragged=True, max_label=12) for i in range(6)]
    m1, o1 = make(cfg, sd, dtype, layerdrop=0.4)
    np.random.seed(5)
    eager, eager_skips = [], []
    for b in batches:
        o1.zero_grad()
        out = m1(**dev(b))
        out.loss.backward()
        o1.step()
        eager.append(float(out.loss))
        eager_skips.append(tuple(m1.backbone.last_skipped_layers))
    assert any(eager_skips) and len(set(eager_skips)) > 1, eager_skips
    m2, o2 = make(cfg, sd, dtype, layerdrop=0.4)
    before = [t.clone() for t in o2.state_tensors()]
    split, defer = "split_conv" in mode, "deferred_adamw" in mode   # conv stack as its own graph; AdamW after the graphs
    step = GraphedTrainStep(m2, o2, dev(batches[0]), max_label_len=16, split_conv=split, defer_adamw=defer)
    assert step.segmented and len(step.graphs) == 2 * 4 + 3 and (step.conv_graph is not None) == split
    if split:   # the chunk lists of the pipelined FedAvg cover every parameter exactly once
        ch, cp = step.fedavg_chunks(), step.fedavg_chunk_params()
        assert len(ch) == len(cp) == 4 + 1
        for gi, g in enumerate(o2.groups):
            covered = sorted((a, b) for c in ch for (i, a, b) in c if i == gi)
            assert covered[0][0] == 0 and covered[-1][1] == g["fp"].numel
            assert all(covered[k][1] == covered[k + 1][0] for k in range(len(covered) - 1))
            assert sorted(p for c in cp for (i, lo, hi) in c if i == gi for p in range(lo, hi)) == list(range(len(g["fp"].params)))
    for t, b in zip(o2.state_tensors(), before):
        assert torch.equal(t, b)   # building the graphs left weights / moments alone
    assert o2.step_count == 0
    np.random.seed(5)
    graphed, graph_skips = [], []
    for b in batches:
        graphed.append(float(step(dev(b))))
        graph_skips.append(tuple(step.last_skipped_layers))
    assert graph_skips == eager_skips
    for a, b in zip(eager, graphed):
        assert abs(a - b) <= 1e-4 * abs(a), (eager, graphed)
    for (n, p1), (_, p2) in zip(m1.named_parameters(), m2.named_parameters()):
        assert relinf(p2.detach(), p1.detach()) < 1e-4, n
    for g1, g2 in zip(o1.groups, o2.groups):
        assert np.array_equal(g1["param_steps"], g2["param_steps"])
        assert relinf(g2["m"], g1["m"]) < 1e-4 and relinf(g2["v"], g1["v"]) < 1e-4
    # a skipped block's parameters really stayed put in that step
    l = next(iter(next(s for s in graph_skips if s)))
    assert min(g["param_steps"].min() for g in o2.groups) < o2.step_count


def test_layerdrop_graph_with_all_dropouts_runs_and_trains():
    """The reference's real training defaults (every dropout 0.1, layerdrop 0.1, attention dropout included) under the
    segmented graph: finite losses, fresh masks and coins per replay, loss decreasing on a repeated batch."""
    cfg = R.ArchConfig.tiny(num_hidden_layers=6)
    sd = R.init_state_dict(cfg, seed=22, dacs=False)
    m, o = make(cfg, sd, torch.bfloat16, layerdrop=0.3, hidden_dropout=0.1, activation_dropout=0.1, attention_dropout=0.1,
                final_dropout=0.1, feat_proj_dropout=0.0)
    b = dev(R.synth_batch(cfg, 2, 6000, seed=180, ragged=True, max_label=12))
    step = GraphedTrainStep(m, o, b, max_label_len=16)
    np.random.seed(1)
    losses, skips = [], set()
    for _ in range(30):
        losses.append(float(step(b)))
        skips.add(step.last_skipped_layers)
    assert all(np.isfinite(losses)) and len(skips) > 3
    assert np.mean(losses[-5:]) < 0.8 * np.mean(losses[:5]), losses


def test_graph_with_idle_parameter_matches_eager():
    """A trainable parameter that never gets a gradient (an unused head registered on the model) is left alone by the
    captured update exactly as by the eager one (transformers.AdamW: `if p.grad is None: continue`) -- no weight decay,
    no moment decay, no step count."""
    cfg = R.ArchConfig.tiny()
    sd = R.init_state_dict(cfg, seed=23, dacs=False)
    batches = [dev(R.synth_batch(cfg, 2, 6000, seed=190 + i, ragged=True, max_label=12)) for i in range(3)]

    def build():
        m = M.Data2VecAudioForCTC(ns(cfg), compute_dtype=torch.bfloat16)
        m.load_state_dict(sd)
        m.spare_head = torch.nn.Linear(cfg.hidden_size, 3)   # never used by forward
        with torch.no_grad():
            m.spare_head.weight.fill_(0.5)
            m.spare_head.bias.fill_(0.25)
        m.freeze_feature_encoder()
        m.to(DEV).train()
        m.validate_labels = False
        return m, FusedAdamW(m.named_parameters(), lr=1e-2, weight_decay=0.1)

    m1, o1 = build()
    for b in batches:
        o1.zero_grad()
        m1(**b).loss.backward()
        o1.step()
    m2, o2 = build()
    step = GraphedTrainStep(m2, o2, batches[0], max_label_len=16)
    for b in batches:
        step(b)
    torch.cuda.synchronize()
    for m in (m1, m2):
        assert float((m.spare_head.weight - 0.5).abs().max()) == 0.0 and float((m.spare_head.bias - 0.25).abs().max()) == 0.0
    for (n, p1), (_, p2) in zip(m1.named_parameters(), m2.named_parameters()):
        assert relinf(p2.detach(), p1.detach()) < 1e-4, n


def test_load_state_dict_after_graph_build_is_seen_by_the_replay():
    """The per-round "load the global weights" step of a federated client (update.py:228): load_state_dict AFTER the
    optimizer and the graph exist must reach the bf16 shadow the captured kernels read."""
    cfg = R.ArchConfig.tiny()
    sd_a = R.init_state_dict(cfg, seed=24, dacs=False)
    sd_b = R.init_state_dict(cfg, seed=25, dacs=False)
    b = dev(R.synth_batch(cfg, 2, 6000, seed=200, ragged=True, max_label=12))
    m1, o1 = make(cfg, sd_b)               # reference: built on B from the start, eager
    o1.zero_grad()
    l_ref = m1(**b).loss
    l_ref.backward()
    o1.step()
    m2, o2 = make(cfg, sd_a)               # graph built on A, then the global model B arrives
    step = GraphedTrainStep(m2, o2, b, max_label_len=16)
    m2.load_state_dict(sd_b)
    l_graph = float(step(b))
    assert abs(l_graph - float(l_ref)) <= 1e-4 * abs(float(l_ref)), (l_graph, float(l_ref))
    for (n, p1), (_, p2) in zip(m1.named_parameters(), m2.named_parameters()):
        assert relinf(p2.detach(), p1.detach()) < 1e-4, n
    m3, o3 = make(cfg, sd_a)               # the same on the eager path
    o3.zero_grad(); m3(**b).loss.backward(); o3.step()
    m3.load_state_dict(sd_b)
    o3.zero_grad()
    l3 = m3(**b).loss
    assert abs(float(l3) - float(l_ref)) <= 1e-4 * abs(float(l_ref))


def test_adamw_chunk_table_equals_the_scalar_kernel():
    """b200asr_adamw_step_chunks with one {lr, step size} row per parameter == b200asr_adamw_step on each parameter's slice
    (bit for bit), and idle rows leave p / m / v / shadow untouched."""
    torch.manual_seed(3)
    shapes = [(1000, 33), (7,), (4096,), (513, 17), (5,)]
    params = [(f"p{i}.weight", torch.nn.Parameter(torch.randn(*s, device=DEV))) for i, s in enumerate(shapes)]
    opt = FusedAdamW(params, lr=1e-2, weight_decay=0.05, max_grad_norm=None)
    g = opt.groups[0]
    fp = g["fp"]
    fp.grad.copy_(torch.randn_like(fp.grad))
    g["m"].copy_(torch.rand_like(fp.flat) * 0.1)
    g["v"].copy_(torch.rand_like(fp.flat) * 0.01)
    ref_p, ref_m, ref_v = fp.flat.clone(), g["m"].clone(), g["v"].clone()
    ref_s = fp.shadow.clone()
    steps = [3, 1, 7, 2, 5]
    idle = [3]
    for i, (p, o) in enumerate(zip(fp.params, fp.offsets)):
        if i in idle:
            continue
        a, b = o, o + (p.numel() + 3) // 4 * 4
        ops.adamw_step_(ref_p[a:b], fp.grad[a:b], ref_m[a:b], ref_v[a:b], ref_s[a:b], 1e-2, 0.9, 0.999, 1e-8, 0.05, steps[i])
    opt.step_count = 0
    g["param_steps"] = np.array([s - 1 for s in steps], dtype=np.int64)
    g["param_steps"][3] = 2
    opt.step_count = 1
    opt._upload_hyper(g, idle, 1e-2)
    chunks, hyper = opt._table(g)
    ops.adamw_step_chunks_(fp.flat, fp.grad, g["m"], g["v"], fp.shadow, chunks, hyper, 0.9, 0.999, 1e-8, 0.05)
    torch.cuda.synchronize()
    assert torch.equal(fp.flat, ref_p) and torch.equal(g["m"], ref_m) and torch.equal(g["v"], ref_v)
    assert torch.equal(fp.shadow, ref_s)
    assert list(g["param_steps"]) == [3, 1, 7, 2, 5]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("ragged", [False, True])
def test_graphed_inference_matches_eager(dtype, ragged):
    """graph.GraphedInference (forward + greedy decode as one CUDA graph) returns the eager path's logits, ids and collapsed
    sequences -- for new inputs of the captured shape, and for new weights loaded after the capture."""
    from b200asr.graph import GraphedInference
    cfg = R.ArchConfig.tiny()
    sd = R.init_state_dict(cfg, seed=31, dacs=False)
    m = M.Data2VecAudioForCTC(ns(cfg), compute_dtype=dtype)
    m.load_state_dict(sd)
    m.to(DEV).eval()
    b0 = dev(R.synth_batch(cfg, 3, 7000, seed=300, ragged=ragged, max_label=8))
    am0 = b0["attention_mask"] if ragged else None
    infer = GraphedInference(m, b0["input_values"], am0)

    def eager(b):
        with torch.no_grad():
            am = b["attention_mask"] if ragged else None
            o = m(b["input_values"], am)
            klen = None if am is None else M.conv_out_lengths(m.config, am.sum(-1)).to(torch.int32)
            return (o.logits,) + tuple(m.greedy_decode(o.logits, klen))

    for seed in (301, 302):
        b = dev(R.synth_batch(cfg, 3, 7000, seed=seed, ragged=ragged, max_label=8))
        lg, ids, seqs = infer(b["input_values"], b["attention_mask"] if ragged else None)
        lg_e, ids_e, seqs_e = eager(b)
        assert torch.equal(lg, lg_e) and torch.equal(ids, ids_e) and seqs == seqs_e
    with pytest.raises(ValueError):
        GraphedInference(m.train(), b0["input_values"], am0)
    m.eval()
    m.load_state_dict(R.init_state_dict(cfg, seed=32, dacs=False))   # e.g. the next round's global model
    lg, ids, seqs = infer(b["input_values"], b["attention_mask"] if ragged else None)
    lg_e, ids_e, seqs_e = eager(b)
    assert torch.equal(lg, lg_e) and torch.equal(ids, ids_e) and seqs == seqs_e


def test_per_block_graphs_without_layerdrop_match_eager():
    """per_block_graphs=True cuts the captured update at the block boundaries although no block is ever dropped (the
    layer-wise pipelined FedAvg wants one graph per block): same losses and parameters as the eager path."""
    cfg = R.ArchConfig.tiny(num_hidden_layers=3)
    sd = R.init_state_dict(cfg, seed=33, dacs=False)
    batches = [R.synth_batch(cfg, 2, 6000, seed=190 + i, ragged=True, max_label=12) for i in range(4)]
    m1, o1 = make(cfg, sd, torch.float32)
    eager = []
    for b in batches:
        o1.zero_grad()
        out = m1(**dev(b))
        out.loss.backward()
        o1.step()
        eager.append(float(out.loss))
    m2, o2 = make(cfg, sd, torch.float32)
    step = GraphedTrainStep(m2, o2, dev(batches[0]), max_label_len=16, split_conv=True, defer_adamw=True, per_block_graphs=True)
    assert step.segmented and len(step.graphs) == 2 * 3 + 3 and len(step.fedavg_chunks()) == 3 + 1
    graphed = [float(step(dev(b))) for b in batches]
    assert step.last_skipped_layers == ()
    for a, b in zip(eager, graphed):
        assert abs(a - b) <= 1e-4 * abs(a), (eager, graphed)
    for (n, p1), (_, p2) in zip(m1.named_parameters(), m2.named_parameters()):
        assert relinf(p2.detach(), p1.detach()) < 1e-4, n
