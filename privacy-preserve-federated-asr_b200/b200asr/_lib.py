"""ctypes binding of libb200asr.so (the C ABI declared in include/b200asr.h).

The product path has no CPU or eager-PyTorch fallback: if the library is missing this module raises at
import, and every entry point raises if the library call fails.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200asr.so")

F32, BF16 = 0, 1
ACT_NONE, ACT_GELU, ACT_GELU_SAVE_GRAD = 0, 1, 2

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `python privacy-preserve-federated-asr_b200/build.py` "
        "(or __graft_entry__.build()); b200asr has no fallback path")

lib = C.CDLL(LIB_PATH)

vp, i32, i64, f32, u64, sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong, C.c_size_t


class GemmArgs(C.Structure):
    _fields_ = [
        ("A", vp), ("B", vp), ("C", vp), ("bias", vp), ("residual", vp), ("aux_out", vp), ("aux_in", vp),
        ("M", i32), ("N", i32), ("K", i32), ("lda", i32), ("ldb", i32), ("ldc", i32),
        ("a_kmajor", i32), ("b_kmajor", i32), ("act", i32), ("mul_gelu_grad", i32), ("alpha", f32),
        ("batch0", i32), ("batch1", i32),
        ("sa0", i64), ("sa1", i64), ("sb0", i64), ("sb1", i64), ("sc0", i64), ("sc1", i64),
        ("dtype", i32), ("c_dtype", i32),
    ]


# name -> (restype, argtypes); must list every symbol include/b200asr.h declares (tests check this)
SIGNATURES = {
    "b200asr_last_error": (C.c_char_p, []),
    "b200asr_abi_version": (i32, []),
    "b200asr_launch_count": (i64, []),
    "b200asr_layernorm_fwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, f32, i32, i32, vp]),
    "b200asr_layernorm_bwd_workspace_bytes": (sz, [i32]),
    "b200asr_layernorm_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, i32, i32, i32, i32, vp]),
    "b200asr_layernorm_bwd_dropout": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, i32, i32, i32, f32, u64, u64,
                                            u64, vp, i32, vp]),
    "b200asr_colsum_workspace_bytes": (sz, [i32]),
    "b200asr_colsum": (i32, [vp, vp, vp, sz, i32, i32, i32, i32, vp]),
    "b200asr_colsum_static_max_cols": (i32, []),
    "b200asr_colsum_static": (i32, [vp, vp, i32, i32, i32, i32, vp]),
    "b200asr_mask_frames": (i32, [vp, vp, i32, i32, i32, i32, vp]),
    "b200asr_cast_f32": (i32, [vp, vp, i64, i32, vp]),
    "b200asr_cast_to_f32": (i32, [vp, vp, i64, i32, vp]),
    "b200asr_scale": (i32, [vp, i64, f32, vp, i32, vp]),
    "b200asr_axpy": (i32, [vp, vp, f32, i64, i32, vp]),
    "b200asr_conv0_ln_gelu_fwd": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, i32, vp]),
    "b200asr_gemm": (i32, [C.POINTER(GemmArgs), vp]),
    "b200asr_gemm_dropout": (i32, [C.POINTER(GemmArgs), f32, u64, u64, u64, vp, vp]),
    "b200asr_gemm_dropout_bits": (i32, [C.POINTER(GemmArgs), vp, i32, f32, vp]),
    "b200asr_dropout_keep_words": (sz, [i32, i32]),
    "b200asr_dropout_keep_bits": (i32, [vp, i32, u64, u64, i32, i32, i32, f32, u64, u64, vp, vp]),
    "b200asr_gemm_force_config": (i32, [i32]),
    "b200asr_gemm_set_trace": (i32, [vp]),
    "b200asr_attn_set_trace": (i32, [vp]),
    "b200asr_ln_set_trace": (i32, [vp]),
    "b200asr_gemm_backend": (i32, [C.POINTER(GemmArgs), i32, vp]),
    "b200asr_softmax_rows": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "b200asr_softmax_rows_bwd": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "b200asr_mha_fwd_workspace_bytes": (sz, [i32, i32, i32, i32, i32]),
    "b200asr_mha_fwd": (i32, [vp, vp, vp, vp, vp, sz, i32, i32, i32, i32, f32, i32, vp]),
    "b200asr_mha_set_fused": (i32, [i32]),
    "b200asr_mha_set_dq_fused": (i32, [i32]),
    "b200asr_mha_set_exp2_poly": (i32, [i32, i32]),
    "b200asr_attn_keep_words": (sz, [i32, i32, i32]),
    "b200asr_attn_dropout_mask": (i32, [vp, i32, i32, i32, i32, f32, u64, u64, u64, u64, vp, vp]),
    "b200asr_mha_fwd_dropout": (i32, [vp, vp, vp, vp, vp, f32, vp, sz, i32, i32, i32, i32, f32, i32, vp]),
    "b200asr_mha_bwd_dropout": (i32, [vp, vp, vp, vp, vp, vp, f32, vp, vp, sz, i32, i32, i32, i32, f32, i32, vp]),
    "b200asr_mha_bwd_workspace_bytes": (sz, [i32, i32, i32, i32, i32]),
    "b200asr_mha_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, sz, i32, i32, i32, i32, f32, i32, vp]),
    "b200asr_posconv_pack_weight": (i32, [vp, vp, vp, i32, i32, i32, i32, vp]),
    "b200asr_posconv_fwd": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "b200asr_posconv_bwd_workspace_bytes": (sz, [i32, i32, i32]),
    "b200asr_posconv_bwd": (i32, [vp, vp, vp, vp, vp, vp, sz, i32, i32, i32, i32, i32, i32, vp]),
    "b200asr_conv0_gn_gelu_workspace_bytes": (sz, [i32, i32, i32, i32, i32]),
    "b200asr_conv0_gn_gelu_fwd": (i32, [vp, vp, vp, vp, vp, vp, vp, sz, i32, i32, i32, i32, i32, f32, i32, vp]),
    "b200asr_weight_norm_workspace_bytes": (sz, [i32]),
    "b200asr_weight_norm_fwd": (i32, [vp, vp, vp, vp, vp, sz, i32, i32, vp]),
    "b200asr_weight_norm_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, sz, i32, i32, vp]),
    "b200asr_gelu_add_fwd": (i32, [vp, vp, vp, i64, i32, vp]),
    "b200asr_gelu_bwd": (i32, [vp, vp, vp, i64, i32, vp]),
    "b200asr_layernorm_dropout_fwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, f32, i32, f32, u64, u64, u64, vp, i32, vp]),
    "b200asr_dropout_colsum": (i32, [vp, vp, vp, vp, sz, i32, i32, f32, u64, u64, u64, vp, i32, vp]),
    "b200asr_dropout": (i32, [vp, vp, vp, i64, f32, u64, u64, u64, vp, i32, vp]),
    "b200asr_ctc_workspace_bytes": (sz, [i32, i32, i32, i32]),
    "b200asr_ctc_fwd_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "b200asr_greedy_decode": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]),
    "b200asr_dacs_mask_fwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, u64, u64, i32, vp]),
    "b200asr_dacs_mask_fwd_ctr": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, u64, u64, vp, i32, vp]),
    "b200asr_dacs_mask_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, i32, vp]),
    "b200asr_mean_time_fwd": (i32, [vp, vp, i32, i32, i32, i32, vp]),
    "b200asr_mean_time_bwd": (i32, [vp, vp, i32, i32, i32, f32, i32, vp]),
    "b200asr_ad_loss": (i32, [vp, vp, vp, vp, i32, i32, i32, C.POINTER(f32), vp]),
    "b200asr_cosface_fwd_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, sz, i64, i32, f32, f32, i32, vp]),
    "b200asr_fedavg_scale_accum": (i32, [C.POINTER(vp), C.POINTER(f32), vp, i32, i64, vp]),
    "b200asr_fedavg_prescale": (i32, [vp, f32, i64, vp]),
    "b200asr_fedavg_peer_reduce": (i32, [C.POINTER(vp), C.POINTER(f32), i32, i32, i64, vp]),
    "b200asr_fedavg_peer_reduce_guarded": (i32, [C.POINTER(vp), C.POINTER(f32), i32, i32, i64, vp, vp]),
    "b200asr_fedavg_nvls_reduce": (i32, [vp, i32, i32, i64, vp, i32, vp]),
    "b200asr_fedavg_set_max_blocks": (i32, [i32]),
    "b200asr_fedavg_peer_barrier": (i32, [C.POINTER(vp), i32, i32, C.c_uint, vp, vp]),
    "b200asr_fedavg_peer_barrier_set_timeout": (i32, [C.c_double]),
    "b200asr_peer_alloc": (i32, [C.POINTER(vp), sz]),
    "b200asr_peer_free": (i32, [vp]),
    "b200asr_ipc_export": (i32, [vp, C.c_char_p]),
    "b200asr_ipc_open": (i32, [C.c_char_p, C.POINTER(vp)]),
    "b200asr_ipc_close": (i32, [vp]),
    "b200asr_conv_out_lengths": (i32, [vp, vp, i32, vp, vp, i32, vp]),
    "b200asr_normalize_pad_workspace_bytes": (sz, [i32, i64]),
    "b200asr_normalize_pad": (i32, [vp, vp, vp, vp, vp, vp, sz, i32, i32, vp, vp, i32, vp]),
    "b200asr_words_from_ids": (i32, [vp, vp, i32, i32, i32, vp, vp, i32, vp]),
    "b200asr_edit_distance": (i32, [vp, vp, vp, vp, vp, i32, i32, i32, vp]),
    "b200asr_adamw_step": (i32, [vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, i32, vp, f32, vp]),
    "b200asr_adamw_step_dev": (i32, [vp, vp, vp, vp, vp, i64, vp, f32, f32, f32, f32, vp, f32, vp]),
    "b200asr_adamw_step_chunks": (i32, [vp, vp, vp, vp, vp, vp, i32, vp, f32, f32, f32, f32, vp, f32, vp]),
    "b200asr_sumsq": (i32, [vp, i64, vp, vp]),
    "b200asr_sumsq_workspace_bytes": (sz, []),
    "b200asr_sumsq_ws": (i32, [vp, i64, vp, vp, sz, vp]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = header and library out of sync
    _fn.restype = _res
    _fn.argtypes = _args


class B200AsrError(RuntimeError):
    pass


def last_error() -> str:
    return (lib.b200asr_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str = "") -> None:
    if rc == 0:
        return
    msg = f"{what}: {last_error()} (status {rc})"
    if rc == -1:
        raise ValueError(msg)
    if rc == -2:
        raise NotImplementedError(msg)
    raise B200AsrError(msg)


def launch_count() -> int:
    return int(lib.b200asr_launch_count())
