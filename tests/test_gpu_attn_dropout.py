"""GPU: attention-probability dropout (HF:189-197 `nn.functional.dropout(attn_weights, p=config.attention_dropout)`) and
the dQ accumulation inside the fused attention backward.

torch's dropout RNG stream cannot be reproduced by another kernel, so parity is defined GIVEN the masks: the keep bits the
library drew are read back (ops.attn_keep_dense) and handed to an fp32 eager reference / to the CPU oracle, whose placement
of the attention dropout is pinned against HF by tests/golden/plain_ctc_attn_dropout.npz (tests/test_oracle_golden.py)."""
from types import SimpleNamespace

import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from b200asr import modeling as M  # noqa: E402
from b200asr import ops  # noqa: E402
from b200asr._lib import lib  # noqa: E402
from oracle import data2vec_ref as R  # noqa: E402

DEV = "cuda"
TOL = {torch.float32: 1e-3, torch.bfloat16: 2e-2}


def rel(a, b):
    a = torch.as_tensor(a).float().cpu()
    b = torch.as_tensor(b).float().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-6))


def _mha_ref(qkv, dout, klen, scale, keep=None, p=0.0):
    """fp32 eager attention with an explicit keep mask [B,H,T,T] on the probabilities."""
    B, T, _, H, dh = qkv.shape
    x = qkv.float().requires_grad_(True)
    q, k, v = (x[:, :, i].transpose(1, 2) for i in range(3))
    s = (q * scale) @ k.transpose(-1, -2)
    if klen is not None:
        km = torch.arange(T)[None, :] < klen[:, None]
        s = s.masked_fill(~km[:, None, None, :], torch.finfo(torch.float32).min)
    pr = torch.softmax(s, -1)
    if keep is not None:
        pr = pr * keep.float() / (1.0 - p)
    ref = (pr @ v).transpose(1, 2).reshape(B, T, H * dh)
    ref.backward(dout.float())
    return ref.detach(), x.grad


@pytest.mark.parametrize("p", [0.1, 0.5])
def test_keep_mask_kernel_properties(p):
    B, T, H, n_sites = 2, 200, 3, 4
    ctr = torch.tensor([7], dtype=torch.int64, device=DEV)
    keep = ops.attn_dropout_mask(n_sites, B, T, H, p, seed=99, site0=19, site_stride=4, counter=ctr)
    assert keep.shape == (n_sites, ops.attn_keep_words(B, T, H))
    dense = torch.stack([ops.attn_keep_dense(keep[s], B, T, H) for s in range(n_sites)])   # [S,B,H,T,T]
    p_q = round(p * 65536) / 65536
    n = dense.numel()
    assert abs(float(dense.float().mean()) - (1 - p_q)) < 5 * (p_q * (1 - p_q) / n) ** 0.5
    # deterministic in its keys; host counter == device counter; one launch for all sites == one launch per site
    assert torch.equal(keep, ops.attn_dropout_mask(n_sites, B, T, H, p, 99, 19, 4, counter=7))
    one = ops.attn_dropout_mask(1, B, T, H, p, 99, 19 + 4 * 2, 4, counter=7)
    assert torch.equal(one[0], keep[2])
    # other site / step / seed -> independent masks
    a = dense[0]
    for other in (dense[1], ops.attn_keep_dense(ops.attn_dropout_mask(1, B, T, H, p, 99, 19, 4, counter=8)[0], B, T, H),
                  ops.attn_keep_dense(ops.attn_dropout_mask(1, B, T, H, p, 100, 19, 4, counter=7)[0], B, T, H)):
        agree = float((a == other).float().mean())
        assert abs(agree - (p_q ** 2 + (1 - p_q) ** 2)) < 0.01
    # no structure along keys or queries
    k = a.float()
    for shifted, base in ((k[..., 1:], k[..., :-1]), (k[..., 1:, :], k[..., :-1, :])):
        corr = float(((shifted - k.mean()) * (base - k.mean())).mean() / k.var())
        assert abs(corr) < 0.01


@pytest.mark.parametrize("dq_fused", [1, 0])
@pytest.mark.parametrize("fused", [1, 0])
@pytest.mark.parametrize("B,T,H,klens,p", [
    (2, 499, 3, None, 0.1),            # BASELINE configs[1] frame count
    (3, 499, 2, [499, 300, 64], 0.1),  # ragged: key blocks partly / fully masked
    (2, 128, 2, [128, 1], 0.25),       # exact tile, single valid key
    (1, 1499, 2, [1311], 0.1),         # 30 s utterances
    (2, 65, 1, [65, 64], 0.5),
    (2, 8, 16, None, 0.1),
])
def test_mha_dropout_vs_reference_given_masks(fused, dq_fused, B, T, H, klens, p):
    """bf16, head dim 64: fused tcgen05 kernels (fused=1; dQ in-kernel or via dS^T + GEMM) and the composed path (fused=0)
    with keep bits, against fp32 eager attention with the same mask."""
    if not fused and not dq_fused:
        pytest.skip("dq mode only matters on the fused path")
    g = torch.Generator().manual_seed(140 + T)
    dh, D_ = 64, H * 64
    qkv = torch.randn(B, T, 3, H, dh, generator=g).to(torch.bfloat16)
    dout = torch.randn(B, T, D_, generator=g).to(torch.bfloat16)
    klen = torch.tensor(klens, dtype=torch.int32) if klens is not None else None
    scale = dh ** -0.5
    keep = ops.attn_dropout_mask(1, B, T, H, p, seed=5, site0=19, site_stride=4, counter=3)[0]
    dense = ops.attn_keep_dense(keep, B, T, H).cpu()
    ref, gref = _mha_ref(qkv, dout, klen, scale, dense, p)
    ref0, _ = _mha_ref(qkv, dout, klen, scale)
    old = lib.b200asr_mha_set_fused(fused)
    oldq = lib.b200asr_mha_set_dq_fused(dq_fused)
    try:
        qd = qkv.reshape(B * T, 3 * D_).to(DEV)
        kl = klen.to(DEV) if klen is not None else None
        out, lse = ops.mha_fwd(qd, kl, B, T, H, dh, scale, keep=keep, p_drop=p)
        dqkv = ops.mha_bwd(qd, out, dout.reshape(B * T, D_).to(DEV), lse, kl, B, T, H, dh, scale, keep=keep, p_drop=p)
        out0, lse0 = ops.mha_fwd(qd, kl, B, T, H, dh, scale)
        torch.cuda.synchronize()
    finally:
        lib.b200asr_mha_set_fused(old)
        lib.b200asr_mha_set_dq_fused(oldq)
    assert rel(out.view(B, T, D_), ref) < 2e-2
    assert rel(out0.view(B, T, D_), ref0) < 2e-2
    assert torch.equal(lse, lse0)   # the saved log-sum-exp is that of the undropped softmax
    got = dqkv.view(B, T, 3, H, dh).float().cpu()
    # (an utterance with ONE valid key has dS = dP_drop - D = 0 in exact arithmetic; in bf16 D is formed from the rounded
    # output, which leaves rounding noise of the size of the other utterance's true gradient)
    gt = 8e-2 if (klens is not None and min(klens) == 1) else 3e-2
    for i, name in enumerate("qkv"):
        assert rel(got[:, :, i], gref[:, :, i]) < gt, name


@pytest.mark.parametrize("B,T,H,klens", [(2, 499, 4, None), (3, 499, 2, [499, 300, 64]), (1, 1499, 2, [1311]), (2, 130, 16, [130, 7]),
                                         (8, 499, 16, None)])
def test_fused_dq_matches_the_gemm_path(B, T, H, klens):
    """dQ accumulated inside the backward kernel (fp32 bulk reduce-add over the key blocks) vs the deterministic
    dS^T + batched GEMM path: same dK / dV bit for bit, dQ equal to bf16 rounding; the deterministic path repeats exactly."""
    g = torch.Generator().manual_seed(240 + T + B)
    dh, D_ = 64, H * 64
    qd = torch.randn(B * T, 3 * D_, generator=g).to(torch.bfloat16).to(DEV)
    dout = torch.randn(B * T, D_, generator=g).to(torch.bfloat16).to(DEV)
    kl = torch.tensor(klens, dtype=torch.int32, device=DEV) if klens is not None else None
    scale = dh ** -0.5
    out, lse = ops.mha_fwd(qd, kl, B, T, H, dh, scale)
    res = {}
    old = lib.b200asr_mha_set_dq_fused(1)
    try:
        for mode in (1, 0, 0, 1):
            lib.b200asr_mha_set_dq_fused(mode)
            res.setdefault(mode, []).append(ops.mha_bwd(qd, out, dout, lse, kl, B, T, H, dh, scale).view(B, T, 3, H, dh).clone())
        torch.cuda.synchronize()
    finally:
        lib.b200asr_mha_set_dq_fused(old)
    assert torch.equal(res[0][0], res[0][1])                       # deterministic path: bit-reproducible
    assert torch.equal(res[1][0][:, :, 1:], res[0][0][:, :, 1:])   # dK, dV do not depend on the dQ mode
    assert torch.equal(res[1][1][:, :, 1:], res[0][0][:, :, 1:])
    for r in res[1]:
        assert rel(r[:, :, 0], res[0][0][:, :, 0]) < 1.2e-2        # bf16 rounding of an fp32 sum in another order
    assert float((res[1][0][:, :, 0].float() - res[1][1][:, :, 0].float()).abs().max()) <= 2 ** -7 * float(res[0][0][:, :, 0].float().abs().max())


@pytest.mark.parametrize("B,T,H,klens", [(2, 499, 4, None), (3, 200, 2, [200, 77, 1]), (1, 1499, 2, [1016])])
@pytest.mark.parametrize("share", [4, 2])
@pytest.mark.parametrize("p_drop", [0.0, 0.1])
def test_polynomial_exp2_share_matches_the_mufu_kernels(B, T, H, klens, share, p_drop):
    """b200asr_mha_set_exp2_poly: a quarter / half of the softmax exponentials on the FMA pipe (degree-3 polynomial, 1e-4
    relative) -- forward output, log-sum-exp and all three gradients equal the all-MUFU kernels' to bf16 rounding, with and
    without attention dropout, ragged key lengths included."""
    g = torch.Generator().manual_seed(7 * T + B)
    dh, D_ = 64, H * 64
    qd = (torch.randn(B * T, 3 * D_, generator=g) * 1.5).to(torch.bfloat16).to(DEV)
    dout = torch.randn(B * T, D_, generator=g).to(torch.bfloat16).to(DEV)
    kl = torch.tensor(klens, dtype=torch.int32, device=DEV) if klens is not None else None
    keep = ops.attn_dropout_mask(1, B, T, H, p_drop, 77, 19, 4, 3)[0] if p_drop > 0 else None
    scale = dh ** -0.5
    res = {}
    old = lib.b200asr_mha_set_exp2_poly(0, 0)
    try:
        for mode in (0, share):
            lib.b200asr_mha_set_exp2_poly(mode, mode)
            out, lse = ops.mha_fwd(qd, kl, B, T, H, dh, scale, keep=keep, p_drop=p_drop)
            dqkv = ops.mha_bwd(qd, out, dout, lse, kl, B, T, H, dh, scale, keep=keep, p_drop=p_drop)
            res[mode] = (out.float().clone(), lse.clone(), dqkv.float().clone())
        torch.cuda.synchronize()
    finally:
        lib.b200asr_mha_set_exp2_poly(old >> 4, old & 15)
    (o0, l0, g0), (o1, l1, g1) = res[0], res[share]
    assert torch.isfinite(o1).all() and torch.isfinite(g1).all() and torch.isfinite(l1).all()
    assert rel(o1, o0) < 1.2e-2
    assert float((l1 - l0).abs().max()) < 5e-4          # log-sum-exp: 1e-4 relative error of the row sum
    g0, g1 = g0.view(B, T, 3, H, dh), g1.view(B, T, 3, H, dh)
    for i in range(3):
        assert rel(g1[:, :, i], g0[:, :, i]) < 1.5e-2, i
    if kl is not None:   # keys beyond an utterance's length get no gradient in either mode
        masked = torch.arange(T, device=DEV)[None] >= kl[:, None]
        assert float(g1[:, :, 1:][masked].abs().max()) == 0.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, "bf16+activation_keep_bits"])
@pytest.mark.parametrize("family", ["data2vec_postln", "hubert_stable"])
def test_all_dropouts_training_step_vs_oracle_given_masks(family, dtype):
    """hidden / activation / feat_proj / final AND attention dropout in one training step: every mask the library drew is
    read back and replayed by the CPU oracle; loss, logits and all gradients are compared.  (Third variant: the activation
    dropout's decisions drawn ahead as keep bits, Data2VecAudioModel.activation_keep_bits -- same masks by construction.)"""
    act_bits = dtype == "bf16+activation_keep_bits"
    dtype = torch.bfloat16 if act_bits else dtype
    PS = dict(hidden_dropout=0.1, activation_dropout=0.15, feat_proj_dropout=0.05, final_dropout=0.2, attention_dropout=0.2)
    if family == "data2vec_postln":
        cfg, cls = R.ArchConfig.tiny(), M.Data2VecAudioForCTC
        c = SimpleNamespace(**{k: (tuple(v) if isinstance(v, (list, tuple)) else v) for k, v in cfg.__dict__.items()})
    else:
        cfg, cls = R.ArchConfig.tiny_hubert(), M.HubertForCTC
        c = SimpleNamespace(**cfg.to_hf_kwargs())
    for k, v in PS.items():
        setattr(c, k, v)
    sd = R.init_state_dict(cfg, seed=16, dacs=False)
    model = cls(c, compute_dtype=dtype)
    model.load_state_dict(sd, strict=True)
    model.freeze_feature_encoder()
    model.to(DEV).train()
    bb = model.backbone
    bb.dropout_seed = 777
    bb.activation_keep_bits = act_bits
    b = R.synth_batch(cfg, 3, 8000, seed=177, ragged=True, max_label=12)
    out = model(input_values=b["input_values"].to(DEV), attention_mask=b["attention_mask"].to(DEV), labels=b["labels"].to(DEV))
    out.loss.backward()
    assert (bb._act_keep is not None) == act_bits
    B, T = out.logits.shape[:2]
    H, d, inter = cfg.num_attention_heads, cfg.hidden_size, cfg.intermediate_size
    seed, ctr = bb.dropout_seed, bb._drop_ctr

    def emask(site, cols, p):
        return ops.dropout(torch.ones(B * T * cols, device=DEV), p, seed, site, ctr).view(B, T, cols).cpu()

    masks = {0: emask(0, d, PS["feat_proj_dropout"]), 1: emask(1, d, PS["hidden_dropout"]), 2: emask(2, d, PS["final_dropout"])}
    assert bb._attn_keep is not None and bb._attn_keep.shape[0] == cfg.num_hidden_layers
    for l in range(cfg.num_hidden_layers):
        masks[16 + 4 * l] = emask(16 + 4 * l, d, PS["hidden_dropout"])
        masks[16 + 4 * l + 1] = emask(16 + 4 * l + 1, inter, PS["activation_dropout"])
        masks[16 + 4 * l + 2] = emask(16 + 4 * l + 2, d, PS["hidden_dropout"])
        masks[16 + 4 * l + 3] = (ops.attn_keep_dense(bb._attn_keep[l], B, T, H).float() / (1 - PS["attention_dropout"])).cpu()
        frac = float((masks[16 + 4 * l + 3] != 0).float().mean())
        assert abs(frac - 0.8) < 0.02
    trainable = {n for n, p in model.named_parameters() if p.requires_grad}
    leaf = {k: v.clone().requires_grad_(k in trainable) for k, v in sd.items()}
    loss_ref, logits_ref = R.ctc_forward(leaf, cfg, b["input_values"], b["attention_mask"], b["labels"],
                                         drop=lambda site, x: x * masks[site])
    loss_ref.backward()
    tol = TOL[dtype]
    assert abs(float(out.loss) - float(loss_ref)) / abs(float(loss_ref)) < tol
    assert rel(out.logits, logits_ref.detach()) < tol
    gt = 2e-3 if dtype == torch.float32 else 6e-2
    params = dict(model.named_parameters())
    for n in sorted(trainable):
        if n.endswith("k_proj.bias"):
            continue  # zero in exact arithmetic (softmax shift invariance)
        ref, got = leaf[n].grad, params[n].grad
        assert got is not None, n
        assert abs(float(got.float().norm()) - float(ref.norm())) <= gt * max(float(ref.norm()), 1e-3), n
        if dtype == torch.float32:
            assert rel(got, ref) < gt, n
    # the oracle WITHOUT the attention mask must disagree: the test is sensitive to the attention dropout
    with torch.no_grad():
        _, logits_no = R.ctc_forward(sd, cfg, b["input_values"], b["attention_mask"], None,
                                     drop=lambda site, x: x if (site >= 16 and (site - 16) % 4 == 3) else x * masks[site])
    if dtype == torch.float32:
        assert rel(out.logits, logits_no) > 10 * tol
    # a second forward draws other masks
    out2 = model(input_values=b["input_values"].to(DEV), attention_mask=b["attention_mask"].to(DEV), labels=b["labels"].to(DEV))
    assert float(out2.loss) != float(out.loss)
