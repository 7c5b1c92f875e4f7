"""GPU: element dropout on the product path (b200asr_dropout + the dropout variants of the encoder blocks).

torch's dropout RNG stream cannot be reproduced by another kernel, so parity is defined GIVEN the masks: the masks the
library drew in a training step are read back (the same entry point applied to a tensor of ones with the same keys),
handed to the CPU oracle through its `drop(site, x)` hook -- whose dropout placement is pinned against HF by
tests/golden/plain_ctc_dropout.npz (tests/test_oracle_golden.py) -- and loss / logits / every gradient are compared."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from b200asr import modeling as M  # noqa: E402
from b200asr import ops  # noqa: E402
from oracle import data2vec_ref as R  # noqa: E402

DEV = "cuda"
TOL = {torch.float32: 1e-3, torch.bfloat16: 2e-2}  # BASELINE.json: loss / logits relative tolerance
PS = dict(hidden_dropout=0.1, activation_dropout=0.15, feat_proj_dropout=0.05, final_dropout=0.2)


def relinf(a, b):
    a = torch.as_tensor(a).float().cpu()
    b = torch.as_tensor(b).float().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-6))


@pytest.mark.parametrize("p", [0.1, 0.5])
def test_dropout_kernel_mask_properties(p):
    n = 1 << 20
    ones32 = torch.ones(n, device=DEV)
    y = ops.dropout(ones32, p, seed=1234, site=7, counter=3)
    vals = torch.unique(y)
    assert vals.numel() == 2 and float(vals[0]) == 0.0 and abs(float(vals[1]) - 1.0 / (1.0 - p)) < 1e-6
    keep = float((y != 0).float().mean())
    p_q = round(p * 65536) / 65536  # the drop probability is p rounded to 1/65536
    assert abs(keep - (1.0 - p_q)) < 5 * (p_q * (1 - p_q) / n) ** 0.5
    # deterministic in its keys; dtype-independent; device counter == host counter; other keys -> other masks
    assert torch.equal(y, ops.dropout(ones32, p, seed=1234, site=7, counter=3))
    yb = ops.dropout(torch.ones(n, device=DEV, dtype=torch.bfloat16), p, seed=1234, site=7, counter=3)
    assert torch.equal(yb != 0, y != 0)
    ctr = torch.tensor([3], dtype=torch.int64, device=DEV)
    assert torch.equal(y, ops.dropout(ones32, p, seed=1234, site=7, counter=ctr))
    for other in (dict(seed=1235, site=7, counter=3), dict(seed=1234, site=8, counter=3), dict(seed=1234, site=7, counter=4)):
        z = ops.dropout(ones32, p, **other)
        agree = float(((z != 0) == (y != 0)).float().mean())
        assert abs(agree - (p_q ** 2 + (1 - p_q) ** 2)) < 0.01, other  # independent masks
    # residual add and in-place use
    x = torch.randn(n, device=DEV)
    r = torch.randn(n, device=DEV)
    ref = r + x * y
    got = ops.dropout(x, p, seed=1234, site=7, counter=3, residual=r)
    assert torch.allclose(got, ref, rtol=1e-6, atol=1e-6)
    x2 = x.clone()
    assert ops.dropout(x2, p, seed=1234, site=7, counter=3, residual=r, out=x2).data_ptr() == x2.data_ptr()
    assert torch.equal(x2, got)
    # no long runs: neighbouring elements are not correlated
    k = (y != 0).float()
    corr = float(((k[1:] - k.mean()) * (k[:-1] - k.mean())).mean() / k.var())
    assert abs(corr) < 0.01


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,cols", [(3992, 1024), (37, 128), (8, 2048)])
def test_fused_dropout_entry_points_equal_the_composition(rows, cols, dtype):
    """b200asr_layernorm_dropout_fwd == dropout(+residual) then layernorm; b200asr_dropout_colsum == dropout then colsum
    (bit-exact: same mask keys, same rounding points)."""
    g = torch.Generator().manual_seed(rows + cols)
    x = torch.randn(rows, cols, generator=g).to(DEV).to(dtype)
    r = torch.randn(rows, cols, generator=g).to(DEV).to(dtype)
    w = (1 + 0.1 * torch.randn(cols, generator=g)).to(DEV)
    b = (0.1 * torch.randn(cols, generator=g)).to(DEV)
    ctr = torch.tensor([5], dtype=torch.int64, device=DEV)
    keys = (0.2, 99, 21, ctr)
    s_ref = ops.dropout(x, *keys, residual=r)
    y_ref, _, m_ref, r_ref = ops.layernorm_fwd(s_ref, w, b, 1e-5)
    y, s, m, rs = ops.layernorm_fwd(x, w, b, 1e-5, res=r, want_sum=True, drop=keys)
    assert torch.equal(s, s_ref) and torch.equal(y, y_ref) and torch.equal(m, m_ref) and torch.equal(rs, r_ref)
    do_ref = ops.dropout(x, *keys)
    cs_ref = ops.colsum(do_ref)
    do, cs = ops.dropout_colsum(x, *keys)
    assert torch.equal(do, do_ref) and torch.equal(cs, cs_ref)
    ref64 = do_ref.double().sum(0)
    assert float((cs.double() - ref64).abs().max()) <= 1e-3 * max(1.0, float(ref64.abs().max()))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,cols", [(3992, 1024), (37, 128), (8, 2048)])
@pytest.mark.parametrize("with_res", [False, True])
def test_layernorm_bwd_dropout_equals_the_composition(rows, cols, dtype, with_res):
    """b200asr_layernorm_bwd_dropout: dx, dw, db as b200asr_layernorm_bwd; dx_drop == b200asr_dropout(dx) bit for bit;
    the column sums are those of dx_drop (the dropped Linear's bias gradient)."""
    g = torch.Generator().manual_seed(rows * 3 + cols)
    s_in = torch.randn(rows, cols, generator=g).to(DEV).to(dtype)
    dy = torch.randn(rows, cols, generator=g).to(DEV).to(dtype)
    res = torch.randn(rows, cols, generator=g).to(DEV).to(dtype) if with_res else None
    w = (1 + 0.1 * torch.randn(cols, generator=g)).to(DEV)
    b = (0.1 * torch.randn(cols, generator=g)).to(DEV)
    _, _, mean, rstd = ops.layernorm_fwd(s_in, w, b, 1e-5)
    for ctr in (7, torch.tensor([7], dtype=torch.int64, device=DEV)):
        keys = (0.15, 1234, 18, ctr)
        dx0, dw0, db0 = ops.layernorm_bwd(dy, s_in, w, b, mean, rstd, dres_add=res)
        cs = torch.empty(cols, device=DEV)
        dx, dw, db, dxd = ops.layernorm_bwd(dy, s_in, w, b, mean, rstd, dres_add=res, dxsum_out=cs, drop=keys)
        assert torch.equal(dx, dx0) and torch.equal(dw, dw0) and torch.equal(db, db0)
        want = ops.dropout(dx0, *keys)
        assert torch.equal(dxd, want)
        ref64 = want.double().sum(0)
        assert float((cs.double() - ref64).abs().max()) <= 1e-3 * max(1.0, float(ref64.abs().max()))
        # without the affine gradients / column sums (frozen parameters)
        dx1, _, _, dxd1 = ops.layernorm_bwd(dy, s_in, w, b, mean, rstd, dres_add=res, want_wgrad=False, drop=keys)
        assert torch.equal(dx1, dx0) and torch.equal(dxd1, want)


def _masks_from_library(backbone, cfg, B, T, ps):
    """Read the masks of the last training forward back: dropout(ones) with that forward's keys."""
    seed, ctr = backbone.dropout_seed, backbone._drop_ctr
    d, inter = cfg.hidden_size, cfg.intermediate_size

    def mask(site, cols, p):
        if p <= 0:
            return None
        return ops.dropout(torch.ones(B * T * cols, device=DEV), p, seed, site, ctr).view(B, T, cols).cpu()

    masks = {0: mask(0, d, ps["feat_proj_dropout"]), 1: mask(1, d, ps["hidden_dropout"]), 2: mask(2, d, ps["final_dropout"])}
    for l in range(cfg.num_hidden_layers):
        masks[16 + 4 * l] = mask(16 + 4 * l, d, ps["hidden_dropout"])
        masks[16 + 4 * l + 1] = mask(16 + 4 * l + 1, inter, ps["activation_dropout"])
        masks[16 + 4 * l + 2] = mask(16 + 4 * l + 2, d, ps["hidden_dropout"])
    return masks


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("family", ["data2vec_postln", "hubert_stable"])
def test_dropout_training_step_vs_oracle_given_masks(family, dtype):
    if family == "data2vec_postln":
        cfg, cls = R.ArchConfig.tiny(), M.Data2VecAudioForCTC
        c = SimpleNamespace(**{k: (tuple(v) if isinstance(v, (list, tuple)) else v) for k, v in cfg.__dict__.items()})
    else:
        cfg, cls = R.ArchConfig.tiny_hubert(), M.HubertForCTC
        c = SimpleNamespace(**cfg.to_hf_kwargs())
    for k, v in PS.items():
        setattr(c, k, v)
    c.attention_dropout = 0.0
    sd = R.init_state_dict(cfg, seed=6, dacs=False)
    model = cls(c, compute_dtype=dtype)
    model.load_state_dict(sd, strict=True)
    model.freeze_feature_encoder()
    model.to(DEV).train()
    model.backbone.dropout_seed = 4242
    b = R.synth_batch(cfg, 2, 6000, seed=77, ragged=True, max_label=12)
    out = model(input_values=b["input_values"].to(DEV), attention_mask=b["attention_mask"].to(DEV), labels=b["labels"].to(DEV))
    out.loss.backward()
    B, T = out.logits.shape[:2]
    masks = _masks_from_library(model.backbone, cfg, B, T, PS)
    assert all(m is not None and 0 < float((m != 0).float().mean()) < 1 for m in masks.values())
    trainable = {n for n, p in model.named_parameters() if p.requires_grad}
    leaf = {k: v.clone().requires_grad_(k in trainable) for k, v in sd.items()}
    loss_ref, logits_ref = R.ctc_forward(leaf, cfg, b["input_values"], b["attention_mask"], b["labels"],
                                         drop=lambda site, x: x * masks[site] if site in masks else x)  # (attention site: p = 0 here)
    loss_ref.backward()
    tol = TOL[dtype]
    assert abs(float(out.loss) - float(loss_ref)) / abs(float(loss_ref)) < tol
    assert relinf(out.logits, logits_ref.detach()) < tol
    gt = 2e-3 if dtype == torch.float32 else 6e-2
    params = dict(model.named_parameters())
    for n in sorted(trainable):
        if n.endswith("k_proj.bias"):
            continue  # zero in exact arithmetic (softmax shift invariance)
        ref = leaf[n].grad
        got = params[n].grad
        assert got is not None, n
        assert abs(float(got.float().norm()) - float(ref.norm())) <= gt * max(float(ref.norm()), 1e-3), n
        if dtype == torch.float32:
            assert relinf(got, ref) < gt, n
    # a second forward draws other masks; eval mode applies none (deterministic, equals the dropout-free oracle)
    out2 = model(input_values=b["input_values"].to(DEV), attention_mask=b["attention_mask"].to(DEV), labels=b["labels"].to(DEV))
    assert float(out2.loss) != float(out.loss)
    model.eval()
    with torch.no_grad():
        e1 = model(input_values=b["input_values"].to(DEV), attention_mask=b["attention_mask"].to(DEV))
    _, logits_eval = R.ctc_forward(sd, cfg, b["input_values"], b["attention_mask"])
    assert relinf(e1.logits, logits_eval) < tol


def test_dropout_graphed_step_draws_fresh_masks():
    """The device counter is bumped inside the captured step: every replay of the CUDA graph draws new masks."""
    from b200asr.graph import GraphedTrainStep
    from b200asr.optim import FusedAdamW
    cfg = R.ArchConfig.tiny()
    c = SimpleNamespace(**{k: (tuple(v) if isinstance(v, (list, tuple)) else v) for k, v in cfg.__dict__.items()})
    c.hidden_dropout, c.activation_dropout, c.final_dropout = 0.1, 0.1, 0.1
    model = M.Data2VecAudioForCTC(c, compute_dtype=torch.bfloat16)
    model.load_state_dict(R.init_state_dict(cfg, seed=8, dacs=False))
    model.freeze_feature_encoder()
    model.to(DEV).train()
    opt = FusedAdamW(model.named_parameters(), lr=0.0, weight_decay=0.0)  # lr 0: the weights never move
    b = {k: v.to(DEV) for k, v in R.synth_batch(cfg, 2, 6000, seed=78, ragged=True, max_label=12).items()
         if k in ("input_values", "attention_mask", "labels")}
    step = GraphedTrainStep(model, opt, b)
    c0 = int(model.backbone.dropout_counter)
    losses = [float(step(b)) for _ in range(4)]
    assert int(model.backbone.dropout_counter) == c0 + 4
    assert len(set(losses)) == 4, losses  # same weights, same batch, different masks
    assert max(losses) / min(losses) < 1.5


@pytest.mark.parametrize("cfg", [0, 1064, 1128, 1256, 2128, 2256])
@pytest.mark.parametrize("shape", [(300, 256, 64), (1000, 512, 128), (3992, 4096, 1024)])
def test_gemm_epilogue_dropout_equals_gemm_then_dropout(shape, cfg):
    """b200asr_gemm_dropout (activation dropout inside the W1 GEMM's GELU_SAVE_GRAD epilogue): same keep decisions as
    b200asr_dropout for the same keys, applied to BOTH outputs -- gelu(v) and the saved gelu'(v) -- on every tile
    configuration (staged TMA-store epilogue and the direct one), host and device step counters."""
    from b200asr._lib import ACT_GELU_SAVE_GRAD, lib
    Mr, N, K = shape
    g = torch.Generator(device="cuda").manual_seed(5)
    x = (torch.randn(Mr, K, device="cuda", generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * (K ** -0.5)).bfloat16()
    b = torch.randn(N, device="cuda", generator=g) * 0.1
    p, seed, site = 0.1, 99, 17
    old = lib.b200asr_gemm_force_config(cfg)
    try:
        y0, u0 = ops.linear_fwd(x, w, b, act=ACT_GELU_SAVE_GRAD)
        assert ops.gemm_dropout_fusable(x, w)
        for counter in (3, torch.tensor([3], device="cuda", dtype=torch.int64)):
            y1, u1 = ops.linear_fwd(x, w, b, act=ACT_GELU_SAVE_GRAD, drop=(p, seed, site, counter))
            # the same decisions read from bits drawn ahead of the GEMM: identical outputs
            bits = ops.dropout_keep_bits(2, site - 4, 4, Mr, N, p, seed, counter, device="cuda")
            y2, u2 = ops.linear_fwd(x, w, b, act=ACT_GELU_SAVE_GRAD, drop=(p, bits[1], ops.dropout_keep_mpad(Mr)))
            assert torch.equal(y2, y1) and torch.equal(u2, u1)
            mask = ops.dropout(torch.ones(Mr, N, device="cuda", dtype=torch.bfloat16), p, seed, site, counter).float()
            assert 0.85 < float((mask > 0).float().mean()) < 0.95
            for got, plain in ((y1, y0), (u1, u0)):
                want = plain.float() * mask          # (the separate pass rounds gelu to bf16 before scaling: one more rounding)
                assert bool(((got.float() == 0) == (want == 0))[plain != 0].all())
                err = (got.float() - want).abs().max() / want.abs().max()
                assert float(err) < 1e-2, float(err)
    finally:
        lib.b200asr_gemm_force_config(old)


def test_gemm_epilogue_dropout_rejects_what_it_cannot_fuse():
    from b200asr._lib import ACT_GELU_SAVE_GRAD
    x = torch.randn(64, 64, device="cuda")
    w = torch.randn(64, 64, device="cuda")
    assert not ops.gemm_dropout_fusable(x, w)
    with pytest.raises(RuntimeError, match="gemm_dropout"):
        ops.linear_fwd(x, w, None, act=ACT_GELU_SAVE_GRAD, drop=(0.1, 1, 2, 3))


@pytest.mark.parametrize("rows,cols", [(300, 256), (3992, 4096), (17, 64)])
def test_dropout_keep_bits_are_the_dropout_kernels_decisions(rows, cols):
    """b200asr_dropout_keep_bits: bit (n & 31) of word [site][n / 32][m] is set exactly where b200asr_dropout keeps element
    (m, n) for the same keys -- for every site of the launch, host and device step counters; pad rows hold zeros."""
    p, seed, site0, stride = 0.25, 4321, 17, 4
    mpad = ops.dropout_keep_mpad(rows)
    for counter in (9, torch.tensor([9], device="cuda", dtype=torch.int64)):
        bits = ops.dropout_keep_bits(3, site0, stride, rows, cols, p, seed, counter, device="cuda")
        assert bits.shape == (3, cols // 32 * mpad)
        w = bits.view(3, cols // 32, mpad).cpu().numpy().astype(np.uint32)
        for i in range(3):
            mask = ops.dropout(torch.ones(rows, cols, device="cuda"), p, seed, site0 + i * stride, counter).cpu().numpy() > 0
            got = ((w[i][:, :rows, None] >> np.arange(32, dtype=np.uint32)[None, None, :]) & 1).astype(bool)   # [cg, m, bit]
            got = got.transpose(1, 0, 2).reshape(rows, cols)
            assert np.array_equal(got, mask)
            assert not w[i][:, rows:].any()
        assert 0.70 < float(mask.mean()) < 0.80
