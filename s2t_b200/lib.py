"""ctypes binding of libs2t_b200.so — the only way the Python host code reaches the CUDA kernels.

There is no CPU or library fallback: if the shared library is missing, or a call is made without a
B200-class device, the call raises.  Signatures mirror ``include/s2t_b200.h`` one to one.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "libs2t_b200.so"

_vp, _i, _ll, _f, _ull, _u = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong, C.c_uint

ABI_VERSION = 5
ACT_NONE, ACT_RELU, ACT_SWISH = 0, 1, 2
GATE_NONE, GATE_POSITIVE, GATE_SWISH_GRAD, GATE_MUL = 0, 1, 2, 3


class GemmArgs(C.Structure):
    """struct s2t_gemm_args (include/s2t_b200.h)."""

    _fields_ = [
        ("M", _i), ("N", _i), ("K", _i), ("nb1", _i), ("nb2", _i),
        ("a", _vp), ("a_mn_major", _i), ("lda", _ll), ("a_bs1", _ll), ("a_bs2", _ll),
        ("b", _vp), ("b_mn_major", _i), ("ldb", _ll), ("b_bs1", _ll), ("b_bs2", _ll),
        ("out", _vp), ("out_fp32", _i), ("ldc", _ll), ("c_bs1", _ll), ("c_bs2", _ll),
        ("alpha", _f), ("bias", _vp), ("act", _i),
        ("gate", _vp), ("gate_mode", _i), ("ldg", _ll), ("g_bs1", _ll), ("g_bs2", _ll),
        ("dropout_p", _f), ("dropout_seed", _ull), ("dropout_stream", _u),
        ("out_scale", _f),
        ("residual", _vp), ("ldr", _ll), ("r_bs1", _ll), ("r_bs2", _ll),
        ("tile_n", _i),
        ("dropout_seed_offset", _vp),
        ("split_k", _i),
        ("max_ctas", _i),
        ("colsum", _vp), ("colsum_accumulate", _i), ("late_trigger", _i),
    ]


class LnReduceJob(C.Structure):
    """struct s2t_ln_reduce_job (include/s2t_b200.h)."""

    _fields_ = [("partials", _vp), ("rows", _i), ("C", _i), ("dgamma", _vp), ("dbeta", _vp), ("accumulate", _i)]


# name -> (restype, argtypes); every symbol declared in include/s2t_b200.h must appear here
SIGNATURES = {
    "s2t_last_error": (C.c_char_p, []),
    "s2t_abi_version": (_i, []),
    "s2t_device_arch": (_i, []),
    "s2t_set_pdl": (_i, [_i]),
    "s2t_gemm_trace_read": (_i, [_vp]),
    "s2t_gemm_bf16": (_i, [C.POINTER(GemmArgs), _vp]),
    "s2t_fbank_num_frames": (_i, [_ll]),
    "s2t_fbank_fwd": (_i, [_vp, _ll, _vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _ll, _ll, _vp, _vp]),
    "s2t_cmvn_utt": (_i, [_vp, _ll, _ll, _vp, _i, _ll, _ll, _vp, _i, _i, _i, _i, _i, _vp, _vp]),
    "s2t_cmvn_global": (_i, [_vp, _ll, _ll, _vp, _i, _ll, _ll, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "s2t_layernorm_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "s2t_layernorm_bwd_partials": (_i, [_i, _i]),
    "s2t_layernorm_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _f, _f, _ull, _u, _vp, _vp]),
    "s2t_layernorm_bwd_reduce": (_i, [_vp, _i, _i, _vp, _vp, _i, _vp]),
    "s2t_layernorm_bwd_reduce_batched": (_i, [C.POINTER(LnReduceJob), _i, _vp]),
    "s2t_glu_fwd": (_i, [_vp, _vp, _ll, _ll, _i, _i, _vp, _i, _vp]),
    "s2t_glu_bwd": (_i, [_vp, _vp, _ll, _vp, _i, _i, _vp, _i, _vp]),
    "s2t_colsum": (_i, [_vp, _ll, _i, _i, _vp, _i, _vp, _ll, _vp]),
    "s2t_colsum_workspace_floats": (_ll, [_i, _i]),
    "s2t_col2im1d": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s2t_embed_fwd": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _i, _f, _f, _ull, _u, _vp, _vp]),
    "s2t_embed_bwd": (_i, [_vp, _vp, _i, _i, _i, _f, _f, _ull, _u, _vp, _vp]),
    "s2t_dropout": (_i, [_vp, _vp, _ll, _f, _f, _ull, _u, _vp, _vp]),
    "s2t_im2col1d": (_i, [_vp, _ll, _vp, _i, _i, _i, _i, _i, _vp]),
    "s2t_frame_rows": (_i, [_vp, _vp, _vp, _ll, _i, _i, _i, _i, _vp]),
    "s2t_attn_fused_fwd": (_i, [_vp, _ll, _i, _vp, _vp, _ll, _vp, _vp, _ll, _vp, _i, _i, _i, _i, _f, _f, _ull, _u, _vp, _vp]),
    "s2t_attn_fused_bwd": (_i, [_vp, _ll, _i, _vp, _vp, _ll, _vp, _vp, _ll, _vp, _i, _i, _i, _i, _f, _f, _ull, _u, _vp, _vp]),
    "s2t_im2col2d": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s2t_col2im2d": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s2t_colstats_parts": (_i, [_ll]),
    "s2t_colstats": (_i, [_vp, _vp, _ll, _i, _vp]),
    "s2t_ctc_viterbi_workspace_bytes": (_ll, [_i, _i, _i]),
    "s2t_ctc_viterbi": (_i, [_vp, _ll, _ll, _vp, _ll, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp, _ll, _vp]),
    "s2t_specaugment": (_i, [_vp, _i, _ll, _ll, _vp, _i, _i, _i, _vp, _i, _vp, _i, _f, _i, _vp, _vp]),
    "s2t_specaugment_static": (_i, [_vp, _i, _ll, _ll, _vp, _i, _i, _i, _vp, _i, _vp, _i, _f, _i, _vp]),
    "s2t_counter_add": (_i, [_vp, _ull, _vp]),
    "s2t_add_bf16": (_i, [_vp, _vp, _ll, _vp]),
    "s2t_f32_sum": (_i, [_vp, _i, _f, _vp, _vp]),
    "s2t_f32_weighted_sum": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "s2t_mask_rows": (_i, [_vp, _vp, _i, _i, _i, _ll, _vp]),
    "s2t_permute_conv_weight": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "s2t_cast_f32_to_bf16": (_i, [_vp, _vp, _ll, _vp]),
    "s2t_cast_bf16_to_f32": (_i, [_vp, _vp, _ll, _vp]),
    "s2t_attn_softmax_fwd": (_i, [_vp, _ll, _vp, _ll, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f, _f, _ull, _u, _vp, _vp]),
    "s2t_attn_softmax_bwd": (_i, [_vp, _vp, _vp, _vp, _ll, _i, _vp, _i, _i, _i, _i, _ll, _f, _f, _ull, _u, _vp, _vp]),
    "s2t_dwconv1d_tiles": (_ll, [_i, _i]),
    "s2t_dwconv1d_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "s2t_dwconv1d_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "s2t_batchnorm_fwd": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _f, _f, _i, _i, _vp]),
    "s2t_batchnorm_bwd_partials": (_i, [_ll]),
    "s2t_batchnorm_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _i, _vp]),
    "s2t_relpos_qbias_fwd": (_i, [_vp, _ll, _vp, _vp, _vp, _ll, _i, _vp]),
    "s2t_relpos_qbias_bwd": (_i, [_vp, _vp, _ll, _ll, _i, _vp]),
    "s2t_act_dropout": (_i, [_vp, _vp, _ll, _i, _f, _ull, _u, _vp, _vp]),
    "s2t_softmax_rows_fwd": (_i, [_vp, _ll, _vp, _vp, _ll, _i, _i, _vp]),
    "s2t_softmax_rows_bwd": (_i, [_vp, _vp, _ll, _vp, _ll, _i, _i, _i, _vp]),
    "s2t_rows_truth_mix": (_i, [_vp, _ll, _vp, _vp, _i, _i, _i, _vp]),
    "s2t_rows_zero_masked": (_i, [_vp, _ll, _vp, _i, _i, _vp]),
    "s2t_row_lse": (_i, [_vp, _i, _ll, _i, _i, _vp, _vp]),
    "s2t_ctc_workspace_bytes": (_ll, [_i, _i, _i]),
    "s2t_ctc_loss": (_i, [_vp, _i, _ll, _vp, _vp, _ll, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _f, _vp, _vp, _ll, _vp]),
    "s2t_ctc_loss_masked": (_i, [_vp, _i, _ll, _vp, _vp, _ll, _vp, _ll, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _f, _vp, _vp, _ll, _vp]),
    "s2t_grad_norm_partials": (_i, [_ll]),
    "s2t_grad_norm": (_i, [_vp, _ll, _f, _vp, _vp, _vp]),
    "s2t_fused_adam": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _f, C.c_double, C.c_double, _f, _f, _ll, _f, _vp, _f, _vp, _vp, _vp]),
    "s2t_collate_frames": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _ll, _vp, _vp]),
    "s2t_batch_by_size": (_ll, [_vp, _ll, _ll, _ll, _i, _vp]),
    "s2t_ctc_greedy": (_i, [_vp, _i, _ll, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
}


class S2TLibraryError(RuntimeError):
    pass


_lib = None


def load() -> C.CDLL:
    """Load the CUDA library (once).  Raises if it has not been built — never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    from . import build as _build
    if _build.is_stale():
        # a library built from other sources than the tree's (or none): rebuild in place, never run a stale ABI
        try:
            _build.build()
        except Exception as e:
            raise S2TLibraryError(
                f"{LIB_PATH} is missing or stale and could not be rebuilt ({e}); build it with "
                "`python -m s2t_b200.build` (or __graft_entry__.build()). s2t_b200 has no CPU fallback.") from e
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.s2t_abi_version() != ABI_VERSION:
        raise S2TLibraryError("libs2t_b200.so ABI version mismatch")
    _lib = lib
    return lib


def check(status: int) -> None:
    if status != 0:
        msg = load().s2t_last_error().decode(errors="replace")
        raise S2TLibraryError(f"s2t_b200 call failed ({status}): {msg}")
