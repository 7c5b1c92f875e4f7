"""GPU: b200asr_colsum_static -- the column-sum entry point that allocates nothing (its scratch is a device array built into
the library).  Called through the C ABI with ctypes, compared with b200asr_colsum (caller-provided workspace) on two calls
in a row: the second call finds the first one's partial sums in the shared scratch and must not be affected by them.
Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from b200asr import _lib  # noqa: E402

lib = _lib.lib
DEV = "cuda"
DT = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16}


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _colsum_ws(x, out, ws):
    rows, cols = x.shape
    rc = lib.b200asr_colsum(_ptr(x), _ptr(out), _ptr(ws), ws.numel() * ws.element_size(), rows, cols, cols,
                            DT[x.dtype], _stream())
    assert rc == 0, _lib.last_error()


def _colsum_static(x, out):
    rows, cols = x.shape
    return lib.b200asr_colsum_static(_ptr(x), _ptr(out), rows, cols, cols, DT[x.dtype], _stream())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_colsum_static_matches_colsum_on_two_calls_in_a_row(dtype):
    max_cols = lib.b200asr_colsum_static_max_cols()
    assert max_cols == 8192
    g = torch.Generator().manual_seed(11)
    # two different problems back to back: the second is narrower and shorter than the first, so stale partial sums of the
    # first call sit in every part of the scratch the second one uses -- and beyond
    shapes = [(3992, 4096), (517, 1024)]
    xs = [torch.randn(r, c, generator=g).to(DEV).to(dtype) for r, c in shapes]
    outs, refs = [], []
    for x in xs:                     # references first (their own workspace), everything allocated before the calls under test
        rows, cols = x.shape
        ws = torch.empty(lib.b200asr_colsum_workspace_bytes(cols), dtype=torch.uint8, device=DEV)
        ref = torch.empty(cols, dtype=torch.float32, device=DEV)
        _colsum_ws(x, ref, ws)
        refs.append(ref)
        outs.append(torch.full((cols,), float("nan"), dtype=torch.float32, device=DEV))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for x, out in zip(xs, outs):     # the two calls in a row, same stream
        assert _colsum_static(x, out) == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for x, out, ref in zip(xs, outs, refs):
        assert torch.equal(out, ref)                       # same kernels, same partition: bit-identical
        ref64 = x.double().sum(0)
        tol = 1e-3 * max(1.0, float(ref64.abs().max()))   # the tolerance of the other column-sum tests
        assert float((out.double() - ref64).abs().max()) <= tol
    # and once more in the opposite order (the wide problem after the narrow one)
    for x, out in zip(xs[::-1], outs[::-1]):
        out.fill_(float("nan"))
        assert _colsum_static(x, out) == 0, _lib.last_error()
    torch.cuda.synchronize()
    for out, ref in zip(outs, refs):
        assert torch.equal(out, ref)


def test_colsum_static_rejects_sizes_beyond_its_scratch():
    max_cols = lib.b200asr_colsum_static_max_cols()
    x = torch.zeros(4, max_cols + 8, device=DEV)
    out = torch.zeros(max_cols + 8, device=DEV)
    rc = _colsum_static(x, out)
    assert rc != 0
    msg = _lib.last_error()
    assert "exceeds the built-in scratch" in msg and str(max_cols) in msg
    x = torch.ones(4, max_cols, device=DEV)      # the maximum itself is served
    out = torch.zeros(max_cols, device=DEV)
    assert _colsum_static(x, out) == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert torch.equal(out, torch.full_like(out, 4.0))
