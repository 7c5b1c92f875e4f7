"""GPU parity, part 2: attention, positional conv, CTC, greedy decode, DACS kernels, FedAvg, AdamW."""
import copy
import math
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from b200asr import fedavg as FA  # noqa: E402
from b200asr import ops  # noqa: E402
from oracle import dacs_ref as D  # noqa: E402
from oracle import data2vec_ref as R  # noqa: E402

DEV = "cuda"


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-6))


# ----------------------------------------------------------------------------- attention
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,T,H,dh,masked", [(2, 37, 2, 64, True), (3, 130, 4, 64, True), (1, 249, 16, 64, False)])
def test_mha_fwd_bwd(dtype, B, T, H, dh, masked):
    g = torch.Generator().manual_seed(4)
    D_ = H * dh
    qkv = (torch.randn(B, T, 3, H, dh, generator=g)).to(dtype)
    dout = torch.randn(B, T, D_, generator=g).to(dtype)
    klen = torch.tensor([T] + [max(1, T - 7 * (i + 1)) for i in range(B - 1)], dtype=torch.int32) if masked else None
    scale = dh ** -0.5
    x = qkv.float().requires_grad_(True)
    q, k, v = (x[:, :, i].transpose(1, 2) for i in range(3))  # [B,H,T,dh]
    s = (q * scale) @ k.transpose(-1, -2)
    if masked:
        km = torch.arange(T)[None, :] < klen[:, None]
        s = s.masked_fill(~km[:, None, None, :], torch.finfo(torch.float32).min)
    ref = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B, T, D_)
    ref.backward(dout.float())
    qd = qkv.reshape(B * T, 3 * D_).to(DEV)
    kl = klen.to(DEV) if masked else None
    out, lse = ops.mha_fwd(qd, kl, B, T, H, dh, scale)
    t = 3e-5 if dtype == torch.float32 else 2e-2
    assert rel(out.view(B, T, D_), ref.detach()) < t
    dqkv = ops.mha_bwd(qd, out, dout.reshape(B * T, D_).to(DEV), lse, kl, B, T, H, dh, scale)
    assert rel(dqkv.view(B, T, 3, H, dh), x.grad) < (1e-4 if dtype == torch.float32 else 3e-2)


def _mha_ref(qkv, dout, klen, scale):
    B, T, _, H, dh = qkv.shape
    x = qkv.float().requires_grad_(True)
    q, k, v = (x[:, :, i].transpose(1, 2) for i in range(3))
    s = (q * scale) @ k.transpose(-1, -2)
    if klen is not None:
        km = torch.arange(T)[None, :] < klen[:, None]
        s = s.masked_fill(~km[:, None, None, :], torch.finfo(torch.float32).min)
    ref = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B, T, H * dh)
    ref.backward(dout.float())
    return ref.detach(), x.grad


@pytest.mark.parametrize("fused", [1, 0])
@pytest.mark.parametrize("B,T,H,klens", [
    (2, 499, 3, None),                 # BASELINE configs[1] frame count, all keys valid
    (3, 499, 2, [499, 300, 64]),       # ragged: key blocks partly / fully masked
    (2, 128, 2, [128, 1]),             # exact tile, single valid key
    (1, 1499, 2, [1311]),              # 30 s utterances (configs[4]): 12 query tiles x 21 key blocks
    (2, 65, 1, [65, 64]),              # one frame past a block boundary
    (2, 8, 16, None),                  # tiny T
])
def test_mha_bf16_fused_vs_composed(fused, B, T, H, klens):
    """bf16, head_dim 64: the fused tcgen05 flash kernels (fused=1) and the composed GEMM+softmax path (fused=0)
    against the fp32 eager reference of HF:189-197."""
    from b200asr._lib import lib
    g = torch.Generator().manual_seed(40 + T)
    dh, D_ = 64, H * 64
    qkv = torch.randn(B, T, 3, H, dh, generator=g).to(torch.bfloat16)
    dout = torch.randn(B, T, D_, generator=g).to(torch.bfloat16)
    klen = torch.tensor(klens, dtype=torch.int32) if klens is not None else None
    scale = dh ** -0.5
    ref, gref = _mha_ref(qkv, dout, klen, scale)
    old = lib.b200asr_mha_set_fused(fused)
    try:
        qd = qkv.reshape(B * T, 3 * D_).to(DEV)
        kl = klen.to(DEV) if klen is not None else None
        out, lse = ops.mha_fwd(qd, kl, B, T, H, dh, scale)
        dqkv = ops.mha_bwd(qd, out, dout.reshape(B * T, D_).to(DEV), lse, kl, B, T, H, dh, scale)
        torch.cuda.synchronize()
    finally:
        lib.b200asr_mha_set_fused(old)
    assert rel(out.view(B, T, D_), ref) < 2e-2
    # log-sum-exp of the scaled logits (saved for backward) against fp32
    x = qkv.float()
    s = (x[:, :, 0].transpose(1, 2) * scale) @ x[:, :, 1].transpose(1, 2).transpose(-1, -2)
    if klen is not None:
        km = torch.arange(T)[None, :] < klen[:, None]
        s = s.masked_fill(~km[:, None, None, :], float("-inf"))
    assert float((lse.cpu() - torch.logsumexp(s, -1)).abs().max()) < 2e-2
    got = dqkv.view(B, T, 3, H, dh).float().cpu()
    for i, name in enumerate("qkv"):
        assert rel(got[:, :, i], gref[:, :, i]) < 3e-2, name
    if klen is not None:  # masked keys receive exactly zero gradient
        for b in range(B):
            assert float(got[b, int(klen[b]):, 1:].abs().max() if int(klen[b]) < T else 0.0) == 0.0


# ----------------------------------------------------------------------------- positional conv
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("C,groups,k,T", [(128, 2, 19, 50), (1024, 16, 19, 77), (128, 2, 128, 150), (1024, 16, 19, 499), (64, 2, 19, 70),
                                       (1024, 16, 128, 499), (1280, 16, 128, 200)])  # wav2vec2/hubert-large, hubert-xlarge (Cg=80)
def test_posconv_fwd_bwd(dtype, C, groups, k, T):
    g = torch.Generator().manual_seed(5)
    B = 2
    Cg = C // groups
    x = torch.randn(B, T, C, generator=g).to(dtype)
    w = (torch.randn(C, Cg, k, generator=g) * (1.0 / math.sqrt(Cg * k))).to(dtype)
    bias = torch.randn(C, generator=g) * 0.1
    dy = torch.randn(B, T, C, generator=g).to(dtype)
    xr, wr = x.float().requires_grad_(True), w.float().requires_grad_(True)
    ref = F.conv1d(xr.transpose(1, 2), wr, bias, padding=k // 2, groups=groups)
    if k % 2 == 0:
        ref = ref[:, :, :-1]
    ref = ref.transpose(1, 2)
    ref.backward(dy.float())
    wf, wb = ops.posconv_pack(w.to(DEV), groups)
    y = ops.posconv_fwd(x.to(DEV), wf, bias.to(DEV), groups, k)
    t = 3e-5 if dtype == torch.float32 else 2e-2
    assert rel(y, ref.detach()) < t
    dx, dw = ops.posconv_bwd(x.to(DEV), wb, dy.to(DEV), groups, k)
    assert rel(dx, xr.grad) < t
    assert rel(dw, wr.grad) < (1e-4 if dtype == torch.float32 else 2e-2)


# ----------------------------------------------------------------------------- CTC
def _ctc_case(dtype, B, T, V, lens, tlens, seed, zero_inf=True, repeat=False):
    g = torch.Generator().manual_seed(seed)
    logits = (torch.randn(B, T, V, generator=g) * 2).to(dtype)
    Lmax = max(max(tlens), 1)
    labels = torch.full((B, Lmax), -100, dtype=torch.long)
    for b, n in enumerate(tlens):
        lab = torch.randint(1, V, (n,), generator=g)
        if repeat and n > 3:
            lab[1] = lab[0]
            lab[3] = lab[2]
        labels[b, :n] = lab
    il = torch.tensor(lens, dtype=torch.long)
    tl = torch.tensor(tlens, dtype=torch.long)
    lr = logits.float().requires_grad_(True)
    lp = F.log_softmax(lr, -1).transpose(0, 1)
    ref = F.ctc_loss(lp, labels.masked_select(labels >= 0), il, tl, blank=0, reduction="none", zero_infinity=zero_inf)
    ref.sum().backward()
    targets = labels.clamp(min=0).to(torch.int32).view(-1).to(DEV)
    off = (torch.arange(B, dtype=torch.int32) * Lmax).to(DEV)
    loss, grad = ops.ctc_fwd_bwd(logits.to(DEV).unsqueeze(0), targets, off, tl.to(torch.int32).to(DEV),
                                 il.to(torch.int32).to(DEV), Lmax, 0, zero_inf)
    t = 2e-5 if dtype == torch.float32 else 1e-3
    np.testing.assert_allclose(loss[0].cpu().numpy(), ref.detach().numpy(), rtol=t, atol=1e-4)
    # fp32 log-space recursions over T steps: alpha/beta ~ -1e3 carry ~1e-4 absolute rounding in either implementation
    assert rel(grad[0], lr.grad) < (2e-3 if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_ctc_matches_torch(dtype):
    _ctc_case(dtype, 3, 50, 32, [50, 41, 30], [12, 7, 1], 0)
    _ctc_case(dtype, 2, 60, 32, [60, 33], [20, 15], 1, repeat=True)
    _ctc_case(dtype, 4, 499, 32, [499, 450, 300, 499], [120, 60, 20, 100], 2)  # config-2 shape: S up to 241
    _ctc_case(dtype, 2, 40, 8, [40, 10], [5, 9], 3)                          # utterance 1 infeasible -> zero_infinity
    _ctc_case(dtype, 2, 30, 40, [30, 30], [3, 0], 4)                         # empty target, V > 32
    _ctc_case(dtype, 2, 25, 150, [25, 20], [4, 6], 5)                        # V > 128: the re-reading softmax path


@pytest.mark.parametrize("gain", [1.0, 8.0, 20.0])
def test_ctc_peaked_posteriors_ragged_vs_torch(gain):
    """Full-size shapes (8 utterances of 327..499 frames, 34..120 labels) with PEAKED posteriors (logit range up to +-80):
    emission ratios of 2^-100 per frame drive neighbouring lattice states hundreds of binades apart -- the case the
    per-lane shared exponent of round 1 got wrong (gradients off by 50-100 % on some utterances at BASELINE configs[1]).
    fp64 torch CTC on the CPU is the judge."""
    g = torch.Generator().manual_seed(33)
    B, T, V = 8, 499, 32
    logits = torch.randn(B, T, V, generator=g) * gain
    in_len = torch.tensor([499, 356, 428, 439, 372, 327, 413, 395], dtype=torch.int32)
    tl = torch.tensor([34, 67, 88, 78, 44, 120, 43, 39], dtype=torch.int32)
    Lmax = int(tl.max())
    labels = torch.full((B, Lmax), -100, dtype=torch.long)
    for b in range(B):
        labels[b, : int(tl[b])] = torch.randint(1, V, (int(tl[b]),), generator=g)
    x = logits.double().requires_grad_(True)
    lp = F.log_softmax(x, -1).transpose(0, 1)
    ref = F.ctc_loss(lp, labels.clamp(min=0), in_len.long(), tl.long(), blank=0, reduction="none", zero_infinity=True)
    ref.sum().backward()
    tg = labels.clamp(min=0).to(torch.int32).contiguous().view(-1).to(DEV)
    off = (torch.arange(B, dtype=torch.int32) * Lmax).to(DEV)
    loss, grad = ops.ctc_fwd_bwd(logits.unsqueeze(0).to(DEV), tg, off, tl.to(DEV), in_len.to(DEV), Lmax, 0, True)
    assert float((loss[0].cpu().double() - ref.detach()).abs().max() / ref.detach().abs().max()) < 2e-5
    for b in range(B):
        gb, rb = grad[0, b].cpu().double(), x.grad[b]
        assert float((gb - rb).abs().max()) < 2e-4, (b, float((gb - rb).abs().max()))   # |d loss / d logit| <= 1
        assert float(gb[int(in_len[b]):].abs().max() if int(in_len[b]) < T else 0.0) == 0.0


def test_kernel_traces_record_and_do_not_change_results():
    """b200asr_gemm_set_trace / b200asr_attn_set_trace (diagnostics): stamps are written for the CTAs of the launch, the
    results are bit-identical with and without them, NULL stops the recording."""
    from b200asr._lib import check, lib
    g = torch.Generator().manual_seed(3)
    M, N, K = 1000, 512, 512
    x = (torch.randn(M, K, generator=g) * 0.2).bfloat16().to(DEV)
    w = (torch.randn(N, K, generator=g) * 0.2).bfloat16().to(DEV)
    b = torch.randn(N, generator=g).to(DEV)
    y0 = ops.linear_fwd(x, w, b)
    y0 = y0[0] if isinstance(y0, tuple) else y0
    tr = torch.zeros(148 * 20, dtype=torch.int64, device=DEV)
    check(lib.b200asr_gemm_set_trace(tr.data_ptr()), "set_trace")
    y1 = ops.linear_fwd(x, w, b)
    check(lib.b200asr_gemm_set_trace(None), "set_trace")
    y1 = y1[0] if isinstance(y1, tuple) else y1
    torch.cuda.synchronize()
    assert torch.equal(y0, y1)
    t = tr.view(148, 20).cpu()
    used = t[:, 1] != 0
    assert int(used.sum()) >= 1 and bool((t[used, 8] > t[used, 1]).all())   # exit after entry (SM clock)
    tr.zero_()
    ops.linear_fwd(x, w, b)
    torch.cuda.synchronize()
    assert int(tr.abs().sum()) == 0
    B, T, H = 2, 150, 2
    qkv = (torch.randn(B * T, 3 * H * 64, generator=g) * 0.5).bfloat16().to(DEV)
    do = (torch.randn(B * T, H * 64, generator=g) * 0.1).bfloat16().to(DEV)
    out, lse = ops.mha_fwd(qkv, None, B, T, H, 64, 0.125)
    d0 = ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
    nc = ((T + 127) // 128) * H * B
    tra = torch.zeros(nc * 24, dtype=torch.int64, device=DEV)
    check(lib.b200asr_attn_set_trace(tra.data_ptr()), "set_trace")
    d1 = ops.mha_bwd(qkv, out, do, lse, None, B, T, H, 64, 0.125)
    check(lib.b200asr_attn_set_trace(None), "set_trace")
    torch.cuda.synchronize()
    assert torch.equal(d0, d1)
    ta = tra.view(nc, 24).cpu()
    assert bool((ta[:, 7] > ta[:, 0]).all())


def test_ctc_third_opinion_fp64():
    g = torch.Generator().manual_seed(11)
    T, V = 40, 12
    logits = torch.randn(1, T, V, generator=g)
    tgt = [3, 3, 5, 1, 1, 7]
    nll, grad_lp = R.ctc_alpha_beta_numpy(F.log_softmax(logits[0], -1).numpy(), tgt)
    loss, grad = ops.ctc_fwd_bwd(logits.to(DEV).unsqueeze(0), torch.tensor(tgt, dtype=torch.int32, device=DEV),
                                 torch.zeros(1, dtype=torch.int32, device=DEV),
                                 torch.tensor([len(tgt)], dtype=torch.int32, device=DEV),
                                 torch.tensor([T], dtype=torch.int32, device=DEV), len(tgt), 0, True)
    assert abs(float(loss[0, 0]) - nll) < 1e-4
    p = torch.softmax(logits[0], -1).numpy()
    g_logits = grad_lp - p * grad_lp.sum(-1, keepdims=True)
    np.testing.assert_allclose(grad[0, 0].cpu().numpy(), g_logits, atol=3e-5)


def test_greedy_decode():
    g = torch.Generator().manual_seed(6)
    logits = torch.randn(3, 64, 32, generator=g)
    logits[0, 5] = logits[0, 4]  # force a repeat
    ids_ref, seqs_ref = R.greedy_ids(logits)
    ids, col, clen = ops.greedy_decode(logits.to(DEV))
    assert torch.equal(ids.cpu().long(), ids_ref)
    for b in range(3):
        assert col[b, : int(clen[b])].cpu().tolist() == seqs_ref[b]
    lens = torch.tensor([64, 10, 0], dtype=torch.int32, device=DEV)
    _, col, clen = ops.greedy_decode(logits.to(DEV), lens)
    _, s10 = R.greedy_ids(logits[1:2, :10])
    assert col[1, : int(clen[1])].cpu().tolist() == s10[0] and int(clen[2]) == 0


# ----------------------------------------------------------------------------- DACS
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("tau,ratio", [(1.0, 0.0), (0.5, 0.3)])
def test_dacs_mask_fwd_bwd(dtype, tau, ratio):
    g = torch.Generator().manual_seed(7)
    M, d = 77, 128
    scores = torch.randn(M, 4 * d, generator=g).to(dtype)
    h = torch.randn(M, d, generator=g).to(dtype)
    g_lm, g_ad = D.gen_gumbels((M, d, 2), 1), D.gen_gumbels((M, d, 2), 2)
    dl, da = torch.randn(M, d, generator=g).to(dtype), torch.randn(M, d, generator=g).to(dtype)
    sr, hr = scores.float().requires_grad_(True), h.float().requires_grad_(True)
    lm_s = torch.stack((sr[:, :d], sr[:, d:2 * d]), -1)
    ad_s = torch.stack((sr[:, 2 * d:3 * d], sr[:, 3 * d:]), -1)
    if ratio:
        lm_s = torch.stack(((lm_s[..., 1] - lm_s[..., 0]) * ratio + lm_s[..., 0], lm_s[..., 1]), -1)
        ad_s = torch.stack(((ad_s[..., 1] - ad_s[..., 0]) * ratio + ad_s[..., 0], ad_s[..., 1]), -1)
    lm_mask = D.gumbel_hard_mask(lm_s, g_lm, tau)
    ad_mask = D.gumbel_hard_mask(ad_s, g_ad, tau)
    ((lm_mask * hr) * dl.float()).sum().add(((ad_mask * hr) * da.float()).sum()).backward()
    cu = lambda t: t.to(DEV)
    lm, ad, sl, sa, ml, ma = ops.dacs_mask_fwd(cu(scores), cu(h), tau, ratio, cu(g_lm), cu(g_ad))
    assert torch.equal(ml.float().cpu().round(), lm_mask.detach().round())
    assert torch.equal(ma.float().cpu().round(), ad_mask.detach().round())
    assert rel(lm, (lm_mask * hr).detach()) < (1e-6 if dtype == torch.float32 else 1e-2)
    ds, dh = ops.dacs_mask_bwd(cu(dl), cu(da), cu(h), sl, sa, ml, ma, tau, ratio, want_dh=True)
    t = 2e-5 if dtype == torch.float32 else 2e-2
    assert rel(ds, sr.grad) < t
    assert rel(dh, hr.grad) < t


def test_dacs_mask_internal_rng_statistics():
    M, d = 512, 256
    scores = torch.zeros(M, 4 * d, device=DEV)
    h = torch.ones(M, d, device=DEV)
    lm, ad, sl, sa, ml, ma = ops.dacs_mask_fwd(scores, h, 1.0, 0.0, seed=123, offset=1)
    assert abs(float(ml.mean()) - 0.5) < 0.01 and abs(float(ma.mean()) - 0.5) < 0.01
    lm2 = ops.dacs_mask_fwd(scores, h, 1.0, 0.0, seed=123, offset=1)[0]
    assert torch.equal(lm, lm2)
    lm3 = ops.dacs_mask_fwd(scores, h, 1.0, 0.0, seed=123, offset=2)[0]
    assert not torch.equal(lm, lm3)
    assert abs(float((ml * ma).mean()) - 0.25) < 0.01  # lm / AD masks independent


@pytest.mark.parametrize("kind", ["recall", "prec", "f1", "recall_ori", "prec_ori", "cel"])
def test_ad_loss(kind):
    g = torch.Generator().manual_seed(8)
    x = torch.randn(6, 2, generator=g)
    y = torch.tensor([0, 1, 1, 0, 1, 0])
    xr = x.clone().requires_grad_(True)
    w = [0.1, 0.9] if kind in ("recall", "prec") else [0.5, 0.5]
    ref = F.cross_entropy(xr, y) if kind == "cel" else D.recall_loss(xr, y, w, kind)
    ref.backward()
    loss, dl = ops.ad_loss(x.to(DEV), y.to(DEV), kind, None if kind == "cel" else w)
    assert abs(float(loss) - float(ref)) < 1e-5 * max(1, abs(float(ref)))
    # 'prec' variants: d precision / d pt cancels exactly for the target class, what is left is rounding noise
    assert rel(dl, xr.grad) < (3e-3 if kind.startswith('prec') else 1e-4)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_mean_time_and_cosface(dtype):
    g = torch.Generator().manual_seed(9)
    x = torch.randn(3, 41, 2, generator=g).to(dtype)
    assert rel(ops.mean_time_fwd(x.to(DEV)), x.float().mean(1)) < 1e-5
    dout = torch.randn(3, 2, generator=g)
    assert rel(ops.mean_time_bwd(dout.to(DEV), 41, dtype), (dout / 41)[:, None, :].expand(3, 41, 2)) < 1e-2
    n, d = 50, 128
    x1, x2 = torch.randn(n, d, generator=g).to(dtype), torch.randn(n, d, generator=g).to(dtype)
    fc = torch.randn(2, d, generator=g) * 0.2
    a, b = x1.float().requires_grad_(True), x2.float().requires_grad_(True)
    labels = torch.cat((torch.zeros(n, dtype=torch.long), torch.ones(n, dtype=torch.long)))
    ref = D.cosface_loss(torch.cat((a, b), 0), labels, fc)
    ref.backward()
    loss, d1, d2 = ops.cosface_fwd_bwd(x1.to(DEV), x2.to(DEV), fc.to(DEV))
    t = 1e-5 if dtype == torch.float32 else 1e-3
    assert abs(float(loss) - float(ref)) < t * max(1.0, abs(float(ref)))
    assert rel(d1, a.grad) < (1e-4 if dtype == torch.float32 else 2e-2)
    assert rel(d2, b.grad) < (1e-4 if dtype == torch.float32 else 2e-2)


# ----------------------------------------------------------------------------- FedAvg / AdamW
def test_fedavg_bit_exact_vs_reference_golden(golden_dir):
    gold = np.load(os.path.join(golden_dir, "fedavg.npz"))
    for n_clients in (2, 3, 8):
        gen = torch.Generator().manual_seed(77 + n_clients)
        ws = [{"a.weight": torch.randn(37, 19, generator=gen) * (1 + i), "a.bias": torch.randn(19, generator=gen),
               "b.weight": torch.randn(5, 3, 7, generator=gen) * 1e-3} for i in range(n_clients)]
        avg = FA.average_weights([{k: v.to(DEV) for k, v in w.items()} for w in ws])
        for k, v in avg.items():
            assert np.array_equal(v.cpu().numpy(), gold[f"n{n_clients}:{k}"]), (n_clients, k)
        avg_cpu = FA.average_weights(ws)  # CPU state_dicts in, CPU out (the reference's calling convention)
        for k, v in avg_cpu.items():
            assert v.device.type == "cpu" and np.array_equal(v.numpy(), gold[f"n{n_clients}:{k}"])


def test_fedavg_flat_weighted_and_large():
    g = torch.Generator().manual_seed(10)
    n = (1 << 22) + 3
    xs = [torch.randn(n, generator=g) for _ in range(4)]
    ref = D.average_weights([{"w": x} for x in xs])["w"]
    out = FA.fedavg_local([x.to(DEV) for x in xs])
    assert torch.equal(out.cpu(), ref)
    w = [0.1, 0.2, 0.3, 0.4]
    outw = FA.fedavg_local([x.to(DEV) for x in xs], w)
    refw = sum(wi * x for wi, x in zip(w, xs))
    assert rel(outw, refw) < 1e-6
    # the fused peer kernel, emulated on one GPU: every "rank" reduces its slice of all buffers in place
    bufs = [x.to(DEV).clone() for x in xs]
    for r in range(4):
        ops.fedavg_peer_reduce([b.data_ptr() for b in bufs], None, r, n)
    torch.cuda.synchronize()
    for b in bufs:
        assert torch.equal(b.cpu(), ref)


def test_adamw_matches_transformers_formula():
    g = torch.Generator().manual_seed(12)
    n = 10007
    p = torch.randn(n, generator=g)
    m, v = torch.zeros(n), torch.zeros(n)
    pd, md, vd = p.to(DEV).clone(), m.to(DEV), v.to(DEV)
    pad = lambda t: t  # noqa
    lr, b1, b2, eps, wd = 1e-3, 0.9, 0.999, 1e-8, 0.005
    for step in range(1, 4):
        grad = torch.randn(n, generator=g) * 3
        gn = grad.norm()
        clip = min(1.0, 1.0 / (float(gn) + 1e-6))
        gg = grad * clip
        m = b1 * m + (1 - b1) * gg
        v = b2 * v + (1 - b2) * gg * gg
        step_size = lr * math.sqrt(1 - b2 ** step) / (1 - b1 ** step)
        p = p - step_size * m / (v.sqrt() + eps)
        p = p - lr * wd * p
        gd = grad.to(DEV)
        gsq = torch.zeros(1, device=DEV)
        ops.sumsq_(gd, gsq)
        assert abs(float(gsq.sqrt()) - float(gn)) < 1e-3 * float(gn)
        shadow = torch.empty(n, dtype=torch.bfloat16, device=DEV)
        ops.adamw_step_(pd, gd, md, vd, shadow, lr, b1, b2, eps, wd, step, gsq, 1.0)
        assert rel(pd, p) < 1e-5
        assert torch.equal(shadow.cpu(), pd.cpu().to(torch.bfloat16))
