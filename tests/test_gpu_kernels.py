"""GPU parity of every C-ABI kernel against plain torch fp32 / the CPU oracle (run with -m gpu on the B200)."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from b200asr import ops  # noqa: E402
from b200asr._lib import ACT_GELU, ACT_NONE  # noqa: E402
from oracle import dacs_ref as D  # noqa: E402
from oracle import data2vec_ref as R  # noqa: E402

DEV = "cuda"


def rel(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-6))


def tol(dtype, f32=2e-5, bf16=2e-2):
    return f32 if dtype == torch.float32 else bf16


# ----------------------------------------------------------------------------- layernorm / colsum
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("cols,affine,act,res", [(1024, True, ACT_NONE, False), (512, True, ACT_GELU, False),
                                                  (1024, False, ACT_GELU, False), (128, True, ACT_NONE, True),
                                                  (64, True, ACT_GELU, True)])
def test_layernorm_fwd_bwd(dtype, cols, affine, act, res):
    g = torch.Generator().manual_seed(1)
    rows = 333
    x = torch.randn(rows, cols, generator=g).to(dtype)
    r = torch.randn(rows, cols, generator=g).to(dtype) if res else None
    w = (1 + 0.1 * torch.randn(cols, generator=g)) if affine else None
    b = (0.1 * torch.randn(cols, generator=g)) if affine else None
    dy = torch.randn(rows, cols, generator=g).to(dtype)
    dres = torch.randn(rows, cols, generator=g).to(dtype)
    # reference in fp32 from the same (rounded) inputs
    s_ref = (x.float() + (r.float() if res else 0)).to(dtype).float().requires_grad_(True)
    wr = w.clone().requires_grad_(True) if affine else None
    br = b.clone().requires_grad_(True) if affine else None
    y_ref = F.layer_norm(s_ref, (cols,), wr, br, 1e-5)
    if act == ACT_GELU:
        y_ref = F.gelu(y_ref)
    y_ref.backward(dy.float())
    cu = lambda t: None if t is None else t.to(DEV)
    y, s, mean, rstd = ops.layernorm_fwd(cu(x), cu(w), cu(b), 1e-5, res=cu(r), act=act, want_sum=True)
    assert rel(y, y_ref.detach()) < tol(dtype)
    s_in = s if res else cu(x)
    dxsum = torch.empty(cols, dtype=torch.float32, device=DEV)
    dx, dw, db = ops.layernorm_bwd(cu(dy), s_in, cu(w), cu(b), mean, rstd, act=act, dres_add=cu(dres), dxsum_out=dxsum)
    assert rel(dx, s_ref.grad + dres.float()) < tol(dtype)
    assert rel(dxsum, dx.float().sum(0)) < 1e-5  # column sums of the stored dx (bias gradient of the producing Linear)
    if affine:
        assert rel(dw, wr.grad) < tol(dtype, 1e-4, 3e-2)
        assert rel(db, br.grad) < tol(dtype, 1e-4, 3e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_colsum_mask_scale_axpy(dtype):
    g = torch.Generator().manual_seed(2)
    x = torch.randn(1000, 96, generator=g).to(dtype)
    assert rel(ops.colsum(x.to(DEV)), x.float().sum(0)) < tol(dtype, 1e-5, 1e-5)
    h = torch.randn(3, 17, 32, generator=g).to(dtype)
    lens = torch.tensor([17, 5, 0], dtype=torch.int32)
    m = ops.mask_frames_(h.to(DEV).clone(), lens.to(DEV)).cpu()
    ref = h.clone()
    ref[1, 5:] = 0
    ref[2] = 0
    assert torch.equal(m, ref)
    y = torch.randn(1001, generator=g).to(dtype)
    z = ops.axpy_(y.to(DEV).clone(), x.reshape(-1)[:1001].contiguous().to(DEV), 0.5)
    assert rel(z, y.float() + 0.5 * x.reshape(-1)[:1001].float()) < tol(dtype, 1e-6, 1e-2)
    sc = ops.scale_(y.to(DEV).clone(), -0.5, torch.tensor([3.0], device=DEV))
    assert rel(sc, y.float() * -1.5) < tol(dtype, 1e-6, 1e-2)
    f = torch.randn(4099, generator=g)
    assert torch.equal(ops.cast_from_f32(f.to(DEV), torch.bfloat16).cpu(), f.to(torch.bfloat16))
    assert torch.equal(ops.cast_to_f32(f.to(torch.bfloat16).to(DEV)).cpu(), f.to(torch.bfloat16).float())


# ----------------------------------------------------------------------------- conv0
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("C,bias", [(512, False), (64, True)])
def test_conv0_ln_gelu(dtype, C, bias):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(3, 4001, generator=g)
    w = torch.randn(C, 1, 10, generator=g) * 0.4
    b = torch.randn(C, generator=g) * 0.1 if bias else None
    lw, lb = 1 + 0.1 * torch.randn(C, generator=g), 0.1 * torch.randn(C, generator=g)
    ref = F.conv1d(x[:, None], w, b, stride=5).transpose(1, 2)
    ref = F.gelu(F.layer_norm(ref, (C,), lw, lb, 1e-5))
    cu = lambda t: None if t is None else t.to(DEV)
    y = ops.conv0_ln_gelu(cu(x), cu(w), cu(b), cu(lw), cu(lb), 5, 1e-5, dtype)
    assert y.shape == ref.shape
    assert rel(y, ref) < tol(dtype, 2e-5, 1e-2)


# ----------------------------------------------------------------------------- GEMM
def _gemm_case(backend, dtype, M, N, K, a_k, b_k, epi, c_dtype=None, batch=(1, 1), seed=0):
    g = torch.Generator().manual_seed(seed)
    b0, b1 = batch
    nb = b0 * b1
    c_dtype = c_dtype or dtype
    A = (torch.randn(nb, M, K, generator=g) * 0.5).to(dtype)
    Bm = (torch.randn(nb, N, K, generator=g) * 0.5).to(dtype)
    bias = torch.randn(N, generator=g) if epi in ("bias", "gelu", "all") else None
    resid = torch.randn(nb, M, N, generator=g).to(c_dtype) if epi in ("res", "all") else None
    aux_in = torch.randn(nb, M, N, generator=g).to(c_dtype) if epi == "gelugrad" else None
    alpha = 0.37 if epi == "alpha" else 1.0
    ref = torch.einsum("bmk,bnk->bmn", A.float(), Bm.float()) * alpha
    if bias is not None:
        ref = ref + bias
    pre = ref.clone()
    if epi in ("gelu", "all"):
        ref = F.gelu(ref)
    if aux_in is not None:
        xa = aux_in.float()
        ref = ref * (0.5 * (1 + torch.erf(xa / math.sqrt(2))) + xa * torch.exp(-0.5 * xa * xa) / math.sqrt(2 * math.pi))
    if resid is not None:
        ref = ref + resid.float()
    Mp, Np = (M + 7) // 8 * 8, (N + 7) // 8 * 8  # padded pitch of MN-major operands (TMA needs 16-byte pitches)

    def mn_major(X, R, Rp):
        out = torch.full((nb, K, Rp), 7.0, dtype=dtype)  # poison in the padding: must never be read as data
        out[:, :, :R] = X.transpose(1, 2)
        return out.to(DEV)

    Ad = A.to(DEV) if a_k else mn_major(A, M, Mp)
    Bd = Bm.to(DEV) if b_k else mn_major(Bm, N, Np)
    Cd = torch.zeros(nb, M, N, dtype=c_dtype, device=DEV)
    aux_out = torch.zeros_like(Cd) if epi == "all" else None
    cu = lambda t: None if t is None else t.to(DEV)
    ops.gemm_raw(Ad, Bd, Cd, M, N, K, K if a_k else Mp, K if b_k else Np, N, a_k, b_k, bias=cu(bias), residual=cu(resid),
                 aux_out=aux_out, aux_in=cu(aux_in), act=ACT_GELU if epi in ("gelu", "all") else ACT_NONE,
                 mul_gelu_grad=aux_in is not None, alpha=alpha, batch0=b0, batch1=b1,
                 sa=((M if a_k else Mp) * K, b0 * (M if a_k else Mp) * K),
                 sb=((N if b_k else Np) * K, b0 * (N if b_k else Np) * K), sc=(M * N, b0 * M * N), backend=backend)
    torch.cuda.synchronize()
    # fp32: 1e-4 (BASELINE asks for 1e-3); the CPU reference's own summation order varies between host CPUs by a few 1e-5
    t = 1e-4 if (dtype == torch.float32 and c_dtype == torch.float32) else (1e-2 if c_dtype == torch.bfloat16 else 2e-5)
    assert rel(Cd, ref) < t, (rel(Cd, ref), M, N, K, a_k, b_k, epi)
    if aux_out is not None:
        assert rel(aux_out, pre) < t


@pytest.mark.parametrize("a_k,b_k", [(True, True), (True, False), (False, False), (False, True)])
@pytest.mark.parametrize("epi", ["none", "all", "gelugrad", "alpha"])
def test_gemm_simt_fp32(a_k, b_k, epi):
    _gemm_case(0, torch.float32, 131, 70, 45, a_k, b_k, epi)
    _gemm_case(0, torch.float32, 64, 33, 100, a_k, b_k, epi, batch=(3, 2), seed=1)


def test_gemm_simt_bf16_skinny():
    _gemm_case(0, torch.bfloat16, 200, 2, 128, True, True, "bias")
    _gemm_case(0, torch.bfloat16, 200, 128, 2, True, False, "none")


@pytest.mark.parametrize("a_k,b_k", [(True, True), (True, False), (False, False), (False, True)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 128, 256), (200, 72, 136), (499, 1024, 1024),
                                    (1000, 4096, 512), (130, 64, 64), (3992, 32, 1024), (77, 40, 72)])
def test_gemm_tcgen05_shapes(a_k, b_k, M, N, K):
    _gemm_case(1, torch.bfloat16, M, N, K, a_k, b_k, "none")


@pytest.mark.parametrize("epi", ["bias", "gelu", "res", "all", "gelugrad", "alpha"])
@pytest.mark.parametrize("c_dtype", [torch.bfloat16, torch.float32])
def test_gemm_tcgen05_epilogues(epi, c_dtype):
    _gemm_case(1, torch.bfloat16, 300, 264, 192, True, True, epi, c_dtype=c_dtype)


def test_gemm_tcgen05_batched_and_persistent():
    _gemm_case(1, torch.bfloat16, 250, 250, 64, True, True, "alpha", batch=(4, 3))
    _gemm_case(1, torch.bfloat16, 250, 64, 250 + 6, True, False, "none", batch=(4, 3), seed=5)
    _gemm_case(1, torch.bfloat16, 250, 64, 256, False, False, "none", batch=(2, 2), seed=6)
    _gemm_case(1, torch.bfloat16, 128 * 40, 256 * 5, 128, True, True, "bias", seed=7)  # > 148 tiles: several per CTA


@pytest.mark.parametrize("cfg", [4256, 2256, 2128, 1256, 1128, 1064])
@pytest.mark.parametrize("a_k,b_k", [(True, True), (True, False), (False, False), (False, True)])
def test_gemm_tcgen05_forced_tile_configs(cfg, a_k, b_k):
    """Every tile configuration of the tcgen05 kernel (ncta*1000 + bn: single CTA 128 x {64,128,256}, CTA pair =
    cluster of 2 with cta_group::2 MMA 256 x {128,256}) on every operand-major combination, ragged edges, several
    tiles per worker, batches, and the full epilogue."""
    from b200asr._lib import lib
    old = lib.b200asr_gemm_force_config(cfg)
    try:
        _gemm_case(1, torch.bfloat16, 300, 264, 192, a_k, b_k, "all")
        _gemm_case(1, torch.bfloat16, 3992, 1024, 320, a_k, b_k, "bias", seed=3)   # model shape: 16 pair rows, tail rows
        _gemm_case(1, torch.bfloat16, 700, 520, 136, a_k, b_k, "gelugrad", batch=(3, 2), seed=4)
        _gemm_case(1, torch.bfloat16, 1024, 1024, 1000, a_k, b_k, "res", c_dtype=torch.float32, seed=5)
        _gemm_case(1, torch.bfloat16, 128 * 45, 256 * 4, 128, a_k, b_k, "none", seed=6)  # > 74 pair tiles
    finally:
        lib.b200asr_gemm_force_config(old)


def test_gemm_tcgen05_overlapping_rows_conv_view():
    """Conv layers 1-6: A rows are overlapping windows of a channels-last tensor (row pitch = stride*C)."""
    g = torch.Generator().manual_seed(9)
    B, Tin, Cc, k, s, Co = 2, 203, 64, 3, 2, 64
    h = torch.randn(B, Tin, Cc, generator=g).to(torch.bfloat16)
    w = (torch.randn(Co, Cc, k, generator=g) * 0.1).to(torch.bfloat16)
    ref = F.conv1d(h.float().transpose(1, 2), w.float(), stride=s).transpose(1, 2)
    Tout = (Tin - k) // s + 1
    wf = w.permute(0, 2, 1).reshape(Co, k * Cc).contiguous().to(DEV)
    y = torch.empty(B, Tout, Co, dtype=torch.bfloat16, device=DEV)
    ops.gemm_raw(h.to(DEV), wf, y, Tout, Co, k * Cc, s * Cc, k * Cc, Co, True, True, batch0=1, batch1=B,
                 sa=(0, Tin * Cc), sb=(0, 0), sc=(0, Tout * Co), backend=1)
    assert rel(y, ref) < 1e-2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,N,K", [(300, 264, 192), (3992, 4096, 256), (77, 40, 64)])
def test_gemm_gelu_save_grad_modes(dtype, M, N, K):
    """act = GELU_SAVE_GRAD writes gelu'(pre-activation) next to gelu(pre-activation); mul_gelu_grad = 2 multiplies the
    data-gradient GEMM by that saved derivative (HF:330-343 backward without transcendental math in the epilogue)."""
    from b200asr._lib import ACT_GELU_SAVE_GRAD
    g = torch.Generator().manual_seed(M + N)
    x = torch.randn(M, K, generator=g).to(dtype)
    w = (torch.randn(N, K, generator=g) / math.sqrt(K)).to(dtype)
    b = torch.randn(N, generator=g) * 0.3
    pre = (x.float() @ w.float().t() + b).requires_grad_()
    ref = F.gelu(pre)
    (gp_ref,) = torch.autograd.grad(ref.sum(), pre)
    y, gp = ops.linear_fwd(x.to(DEV), w.to(DEV), b.to(DEV), act=ACT_GELU_SAVE_GRAD)
    t = 2e-5 if dtype == torch.float32 else 1e-2
    assert rel(y, ref.detach()) < t
    assert rel(gp, gp_ref) < t
    dy = torch.randn(M, N, generator=g).to(dtype)
    w2 = (torch.randn(N, K, generator=g) / math.sqrt(N)).to(dtype)       # [N, K]: dx[M, K] = dy @ w2
    gsave = torch.rand(M, K, generator=g).to(dtype)
    dref = (dy.float() @ w2.float()) * gsave.float()
    dx = ops.linear_dgrad(dy.to(DEV), w2.to(DEV), gelu_grad=gsave.to(DEV))
    assert rel(dx, dref) < t
