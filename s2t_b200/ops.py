"""Thin torch-tensor wrappers over the C ABI (``include/s2t_b200.h``).

torch is used for device memory and streams only; every computation below is one of our own sm_100a
kernels called through ctypes.  Nothing here falls back to torch ops: a missing library or a non-CUDA
tensor raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence, Tuple

import torch

from . import lib as _lib
from .lib import ACT_NONE, ACT_RELU, ACT_SWISH, GATE_MUL, GATE_NONE, GATE_POSITIVE, GATE_SWISH_GRAD  # noqa: F401

BF16 = torch.bfloat16

#: number of kernel launches issued through this module (bench.py reports it as ``gpu_launches``)
launch_count = 0
_LAUNCHES = {
    "s2t_gemm_bf16": 1, "s2t_fbank_fwd": 1, "s2t_cmvn_utt": 2, "s2t_cmvn_global": 1,
    "s2t_layernorm_fwd": 1, "s2t_layernorm_bwd": 1, "s2t_layernorm_bwd_reduce": 1, "s2t_glu_fwd": 1, "s2t_glu_bwd": 1,
    "s2t_colsum": 1, "s2t_col2im1d": 1, "s2t_embed_fwd": 1, "s2t_embed_bwd": 1, "s2t_mask_rows": 1,
    "s2t_permute_conv_weight": 1, "s2t_cast_f32_to_bf16": 1, "s2t_cast_bf16_to_f32": 1,
    "s2t_attn_softmax_fwd": 1, "s2t_attn_fused_fwd": 1, "s2t_attn_fused_bwd": 1, "s2t_attn_softmax_bwd": 1, "s2t_row_lse": 1, "s2t_ctc_loss": 4, "s2t_ctc_loss_masked": 4,
    "s2t_ctc_greedy": 2, "s2t_dwconv1d_fwd": 1, "s2t_dwconv1d_bwd": 3, "s2t_batchnorm_fwd": 1,
    "s2t_batchnorm_bwd": 2, "s2t_im2col2d": 1, "s2t_col2im2d": 1, "s2t_colstats": 1, "s2t_relpos_qbias_fwd": 1, "s2t_relpos_qbias_bwd": 1, "s2t_act_dropout": 1,
}


def _call(name: str, *args) -> None:
    global launch_count
    fn = getattr(_lib.load(), name)
    _lib.check(fn(*args))
    launch_count += _LAUNCHES.get(name, 1)


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    if not t.is_cuda:
        raise _lib.S2TLibraryError("s2t_b200 kernels need CUDA tensors (no CPU fallback)")
    return t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _drop(d):
    """dropout spec (p, seed, stream_id[, seed_offset u64 device tensor]) -> 4 C arguments."""
    if d is None or d[0] <= 0.0:
        return 0.0, 0, 0, None
    off = d[3] if len(d) > 3 else None
    return float(d[0]), int(d[1]), int(d[2]), _p(off)


# --------------------------------------------------------------------------------------------------
# GEMM
# --------------------------------------------------------------------------------------------------
# SM shares for the duration of the engine's multi-stream backward pass (0 = the whole GPU): GEMMs of the dgrad chain are
# bounded to ``chain`` CTAs and weight-gradient GEMMs (side streams) to ``wgrad`` CTAs without split-K, chain + wgrad =
# the SM count, so that neither ever waits for an SM held by the other (every GEMM CTA needs a whole SM's shared memory).
_late_trigger_side = os.environ.get("S2T_LATE_TRIGGER", "0") == "1"   # measured: early release + lagged joins is faster (1.73 vs 2.03 ms at C2)
_default_max_ctas = 0
_wgrad_max_ctas = 0
_wgrad_split_k = 0


def set_sm_shares(chain: int = 0, wgrad: int = 0, wgrad_split_k: int = 0) -> Tuple[int, int, int]:
    """Returns the previous (chain, wgrad, wgrad_split_k)."""
    global _default_max_ctas, _wgrad_max_ctas, _wgrad_split_k
    old = (_default_max_ctas, _wgrad_max_ctas, _wgrad_split_k)
    _default_max_ctas, _wgrad_max_ctas, _wgrad_split_k = int(chain), int(wgrad), int(wgrad_split_k)
    return old


def gemm_raw(
    *, M: int, N: int, K: int,
    a: torch.Tensor, lda: int, a_mn: bool = False, a_bs: Tuple[int, int] = (0, 0), a_off: int = 0,
    b: torch.Tensor, ldb: int, b_mn: bool = False, b_bs: Tuple[int, int] = (0, 0), b_off: int = 0,
    out: torch.Tensor, ldc: int, c_bs: Tuple[int, int] = (0, 0), c_off: int = 0,
    batch: Tuple[int, int] = (1, 1), alpha: float = 1.0, bias: Optional[torch.Tensor] = None,
    act: int = ACT_NONE, gate: Optional[torch.Tensor] = None, gate_mode: int = GATE_NONE, ldg: int = 0,
    g_bs: Tuple[int, int] = (0, 0), dropout: Optional[Tuple[float, int, int]] = None,
    out_scale: float = 1.0, residual: Optional[torch.Tensor] = None, ldr: int = 0,
    r_bs: Tuple[int, int] = (0, 0), tile_n: int = 0, split_k: int = 0, max_ctas: int = 0,
    colsum: Optional[torch.Tensor] = None, colsum_accumulate: bool = False, late_trigger: Optional[bool] = None,
) -> None:
    """out[z] = residual + out_scale*dropout(gate(act(alpha * A[z] @ B[z]^T + bias))) — see s2t_b200.h.

    ``*_off`` are element offsets into the given tensors (views into fused buffers, heads, …).
    """
    for t in (a, b):
        if t.dtype != BF16:
            raise TypeError("GEMM operands must be bf16")
    if out.dtype not in (BF16, torch.float32):
        raise TypeError("GEMM output must be bf16 or fp32")
    g = _lib.GemmArgs()
    g.M, g.N, g.K = M, N, K
    g.nb1, g.nb2 = batch
    g.a, g.a_mn_major, g.lda = _p(a) + 2 * a_off, int(a_mn), lda
    g.a_bs1, g.a_bs2 = a_bs
    g.b, g.b_mn_major, g.ldb = _p(b) + 2 * b_off, int(b_mn), ldb
    g.b_bs1, g.b_bs2 = b_bs
    g.out_fp32 = int(out.dtype == torch.float32)
    g.out = _p(out) + out.element_size() * c_off
    g.ldc = ldc
    g.c_bs1, g.c_bs2 = c_bs
    g.alpha = alpha
    g.bias = _p(bias)
    g.act = act
    g.gate, g.gate_mode, g.ldg = _p(gate), gate_mode, ldg
    g.g_bs1, g.g_bs2 = g_bs
    g.dropout_p, g.dropout_seed, g.dropout_stream, g.dropout_seed_offset = _drop(dropout)
    g.out_scale = out_scale
    g.residual, g.ldr = _p(residual), ldr
    g.r_bs1, g.r_bs2 = r_bs
    g.tile_n = tile_n
    g.split_k = split_k
    g.max_ctas = max_ctas if max_ctas else _default_max_ctas
    g.colsum, g.colsum_accumulate = _p(colsum), int(colsum_accumulate)
    if late_trigger is None:     # default: late for GEMMs that run with the side-stream (weight-gradient) SM share
        late_trigger = _late_trigger_side and _wgrad_max_ctas > 0 and g.max_ctas == _wgrad_max_ctas
    g.late_trigger = int(bool(late_trigger))
    _call("s2t_gemm_bf16", C.byref(g), _stream())


def linear(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, *,
           out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = BF16, **epi) -> torch.Tensor:
    """y[R, N] = epilogue(x[R, K] @ w[N, K]^T) — the nn.Linear contraction."""
    R, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty((R, N), device=x.device, dtype=out_dtype)
    gemm_raw(M=R, N=N, K=K, a=x, lda=x.stride(0), b=w, ldb=w.stride(0), out=out, ldc=out.stride(0),
             bias=bias, **epi)
    return out


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, *, out: Optional[torch.Tensor] = None, **epi) -> torch.Tensor:
    """dx[R, K] = dy[R, N] @ w[N, K]  (B operand read MN-major: no transposed weight copy)."""
    R, N = dy.shape
    K = w.shape[1]
    if out is None:
        out = torch.empty((R, K), device=dy.device, dtype=BF16)
    gemm_raw(M=R, N=K, K=N, a=dy, lda=dy.stride(0), b=w, ldb=w.stride(0), b_mn=True, out=out,
             ldc=out.stride(0), **epi)
    return out


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, *, out: Optional[torch.Tensor] = None,
                 out_dtype: torch.dtype = BF16, **epi) -> torch.Tensor:
    """dw[N, K] = dy[R, N]^T @ x[R, K]  (both operands read MN-major).  ``colsum=`` (bf16 [N]) additionally receives the
    column sums of dy — the bias gradient of the same layer — from the dy tiles the kernel stages anyway."""
    R, N = dy.shape
    K = x.shape[1]
    if out is None:
        out = torch.empty((N, K), device=dy.device, dtype=out_dtype)
    epi.setdefault("split_k", _wgrad_split_k)
    epi.setdefault("max_ctas", _wgrad_max_ctas)
    gemm_raw(M=N, N=K, K=R, a=dy, lda=dy.stride(0), a_mn=True, b=x, ldb=x.stride(0), b_mn=True,
             out=out, ldc=out.stride(0), **epi)
    return out


# --------------------------------------------------------------------------------------------------
# fbank / CMVN
# --------------------------------------------------------------------------------------------------
def fbank_num_frames(n: int) -> int:
    return _lib.load().s2t_fbank_num_frames(n)


def fbank(wave: torch.Tensor, num_samples: torch.Tensor, mel_w: torch.Tensor, mel_range: torch.Tensor,
          window: torch.Tensor, max_frames: int, out: Optional[torch.Tensor] = None,
          out_dtype: torch.dtype = torch.float32, frames: Optional[torch.Tensor] = None
          ) -> Tuple[torch.Tensor, torch.Tensor]:
    """wave fp32 [B, N] (int16 scale) -> log-mel [B, max_frames, n_mels], frames int32 [B]
    (= min(1 + (n - 400) // 160, max_frames), 0 when n < 400)."""
    B = wave.shape[0]
    n_mels = mel_w.shape[0]
    if out is None:
        out = torch.empty((B, max_frames, n_mels), device=wave.device, dtype=out_dtype)
    if frames is None:
        frames = torch.empty((B,), device=wave.device, dtype=torch.int32)
    _call("s2t_fbank_fwd", _p(wave), wave.stride(0), _p(num_samples), B, _p(mel_w), _p(mel_range),
          _p(window), n_mels, _p(out), int(out.dtype == BF16), max_frames, out.stride(1), out.stride(0),
          _p(frames), _stream())
    return out, frames


def cmvn_utt(feats: torch.Tensor, num_frames: torch.Tensor, out: Optional[torch.Tensor] = None,
             norm_means: bool = True, norm_vars: bool = True, stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    B, T, F = feats.shape
    if out is None:
        out = torch.empty_like(feats)
    if stats is None and (norm_means or norm_vars):
        stats = torch.empty((B, 2, F), device=feats.device, dtype=torch.float32)
    _call("s2t_cmvn_utt", _p(feats), feats.stride(1), feats.stride(0), _p(out), int(out.dtype == BF16),
          out.stride(1), out.stride(0), _p(num_frames), B, T, F, int(norm_means), int(norm_vars), _p(stats),
          _stream())
    return out


def cmvn_global(feats: torch.Tensor, num_frames: torch.Tensor, mean: torch.Tensor, std: torch.Tensor,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    B, T, F = feats.shape
    if out is None:
        out = torch.empty_like(feats)
    _call("s2t_cmvn_global", _p(feats), feats.stride(1), feats.stride(0), _p(out), int(out.dtype == BF16),
          out.stride(1), out.stride(0), _p(num_frames), B, T, F, _p(mean), _p(std), _stream())
    return out


# --------------------------------------------------------------------------------------------------
# LayerNorm
# --------------------------------------------------------------------------------------------------
def layernorm_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5,
                  out: Optional[torch.Tensor] = None, stats: Optional[torch.Tensor] = None):
    rows, Cd = x.shape
    if out is None:
        out = torch.empty_like(x)
    if stats is None:
        stats = torch.empty((rows, 2), device=x.device, dtype=torch.float32)
    _call("s2t_layernorm_fwd", _p(x), _p(gamma), _p(beta), _p(out), _p(stats), rows, Cd, eps, _stream())
    return out, stats


def layernorm_bwd(dy: torch.Tensor, x: torch.Tensor, gamma: torch.Tensor, stats: torch.Tensor,
                  dgamma: Optional[torch.Tensor], dbeta: Optional[torch.Tensor], *,
                  dx_residual: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                  accumulate: bool = False, partials: Optional[torch.Tensor] = None,
                  out_drop: Optional[torch.Tensor] = None, drop=None, drop_scale: float = 1.0,
                  reduce: bool = True) -> torch.Tensor:
    """dx = LN'(dy) (+ dx_residual); ``out_drop`` (optional) = drop_scale * dropout(dx) for the next branch.  With
    ``reduce=False`` only the partial dgamma/dbeta sums are written (call :func:`layernorm_bwd_reduce` later, e.g. on a
    side stream)."""
    rows, Cd = x.shape
    if out is None:
        out = torch.empty_like(x)
    if partials is None:
        n = _lib.load().s2t_layernorm_bwd_partials(rows, Cd)
        partials = torch.empty((n, 2, Cd), device=x.device, dtype=torch.float32)
    p, seed, sid, off = _drop(drop)
    _call("s2t_layernorm_bwd", _p(dy), _p(x), _p(gamma), _p(stats), _p(dx_residual), _p(out), _p(partials), rows, Cd,
          _p(out_drop), p, drop_scale, seed, sid, off, _stream())
    if reduce:
        layernorm_bwd_reduce(partials, rows, Cd, dgamma, dbeta, accumulate)
    return out


def layernorm_bwd_reduce(partials: torch.Tensor, rows: int, Cd: int, dgamma: Optional[torch.Tensor],
                         dbeta: Optional[torch.Tensor], accumulate: bool = False) -> None:
    _call("s2t_layernorm_bwd_reduce", _p(partials), rows, Cd, _p(dgamma), _p(dbeta), int(accumulate), _stream())


def layernorm_bwd_reduce_batched(jobs) -> None:
    """jobs: sequence of (partials, rows, C, dgamma, dbeta[, accumulate]) — all dgamma / dbeta reductions in one launch per
    32 jobs."""
    global launch_count
    if not jobs:
        return
    arr = (_lib.LnReduceJob * len(jobs))()
    for a, j in zip(arr, jobs):
        a.partials, a.rows, a.C = _p(j[0]), int(j[1]), int(j[2])
        a.dgamma, a.dbeta = _p(j[3]), _p(j[4])
        a.accumulate = int(j[5]) if len(j) > 5 else 0
    _lib.check(_lib.load().s2t_layernorm_bwd_reduce_batched(arr, len(jobs), _stream()))
    launch_count += (len(jobs) + 31) // 32


# --------------------------------------------------------------------------------------------------
# elementwise
# --------------------------------------------------------------------------------------------------
def glu_fwd(x: torch.Tensor, out: Optional[torch.Tensor] = None, lengths: Optional[torch.Tensor] = None,
            T: int = 0, *, ldy: Optional[int] = None, y_bs: int = 0, y_off: int = 0) -> torch.Tensor:
    """x [rows, 2C] -> out = x[:, :C] * sigmoid(x[:, C:]).  With ``y_bs`` the row b*T+t lands at element
    ``y_off + b*y_bs + t*ldy`` of ``out`` (a per-utterance framed buffer)."""
    rows, C2 = x.shape
    if out is None:
        out = torch.empty((rows, C2 // 2), device=x.device, dtype=BF16)
    if ldy is None:
        ldy = out.stride(0)
    _call("s2t_glu_fwd", _p(x), _p(out) + 2 * y_off, ldy, y_bs, rows, C2 // 2, _p(lengths), T, _stream())
    return out


def glu_bwd(x: torch.Tensor, dy: torch.Tensor, out: Optional[torch.Tensor] = None,
            lengths: Optional[torch.Tensor] = None, T: int = 0) -> torch.Tensor:
    rows, C2 = x.shape
    if out is None:
        out = torch.empty_like(x)
    _call("s2t_glu_bwd", _p(x), _p(dy), dy.stride(0), _p(out), rows, C2 // 2, _p(lengths), T, _stream())
    return out


def colsum(x: torch.Tensor, out: torch.Tensor, accumulate: bool = False,
           workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
    rows, Cd = x.shape
    need = _lib.load().s2t_colsum_workspace_floats(rows, Cd)
    if workspace is None or workspace.numel() < need:
        workspace = torch.empty((need,), device=x.device, dtype=torch.float32)
    _call("s2t_colsum", _p(x), x.stride(0), rows, Cd, _p(out), int(accumulate), _p(workspace),
          workspace.numel(), _stream())
    return out


def col2im1d(dcol: torch.Tensor, batch: int, t_in: int, t_out: int, Cd: int, ksize: int, stride: int,
             pad: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if out is None:
        out = torch.empty((batch, t_in, Cd), device=dcol.device, dtype=BF16)
    _call("s2t_col2im1d", _p(dcol), _p(out), batch, t_in, t_out, Cd, ksize, stride, pad, _stream())
    return out


def embed_fwd(x: torch.Tensor, lengths: torch.Tensor, pos_table: Optional[torch.Tensor], scale: float,
              dropout: Tuple[float, int, int] = (0.0, 0, 0), out: Optional[torch.Tensor] = None) -> torch.Tensor:
    B, T, Cd = x.shape
    if out is None:
        out = torch.empty_like(x)
    n_pos = pos_table.shape[0] if pos_table is not None else 0
    _call("s2t_embed_fwd", _p(x), _p(lengths), _p(pos_table), n_pos, _p(out), B, T, Cd, scale,
          *_drop(dropout), _stream())
    return out


def embed_bwd(dy: torch.Tensor, scale: float, dropout: Tuple[float, int, int] = (0.0, 0, 0),
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
    B, T, Cd = dy.shape
    if out is None:
        out = torch.empty_like(dy)
    _call("s2t_embed_bwd", _p(dy), _p(out), B, T, Cd, scale, *_drop(dropout), _stream())
    return out


def mask_rows_(x: torch.Tensor, lengths: torch.Tensor) -> torch.Tensor:
    B, T, Cd = x.shape
    _call("s2t_mask_rows", _p(x), _p(lengths), B, T, Cd, x.stride(1), _stream())
    return x


def permute_conv_weight(w: torch.Tensor, inverse: bool = False, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[Cout, Cin, k] -> [Cout, k, Cin] (inverse=False) or back."""
    if not inverse:
        cout, cin, k = w.shape
        shape = (cout, k, cin)
    else:
        cout, k, cin = w.shape
        shape = (cout, cin, k)
    if out is None:
        out = torch.empty(shape, device=w.device, dtype=BF16)
    _call("s2t_permute_conv_weight", _p(w), _p(out), cout, cin, k, int(inverse), _stream())
    return out


def cast_f32_to_bf16(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=BF16)
    _call("s2t_cast_f32_to_bf16", _p(x), _p(out), x.numel(), _stream())
    return out


def cast_bf16_to_f32(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=torch.float32)
    _call("s2t_cast_bf16_to_f32", _p(x), _p(out), x.numel(), _stream())
    return out


def dropout(x: torch.Tensor, spec, out: Optional[torch.Tensor] = None, scale: float = 1.0) -> torch.Tensor:
    """y = scale * dropout(x): standalone regenerable dropout (same Philox numbering as the GEMM epilogue
    on a dense tensor); spec with p = 0 gives a plain scaled copy."""
    if out is None:
        out = torch.empty_like(x)
    p, seed, sid, off = _drop(spec)
    _call("s2t_dropout", _p(x), _p(out), x.numel(), p, scale, seed, sid, off, _stream())
    return out


def im2col1d(xp: torch.Tensor, x_bs: int, batch: int, t_out: int, Cd: int, ksize: int, stride: int,
             out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if out is None:
        out = torch.empty((batch * t_out, ksize * Cd), device=xp.device, dtype=BF16)
    _call("s2t_im2col1d", _p(xp), x_bs, _p(out), batch, t_out, Cd, ksize, stride, _stream())
    return out


def frame_rows(x: torch.Tensor, lengths: Optional[torch.Tensor], framed: torch.Tensor, pad: int) -> torch.Tensor:
    """x bf16 [B, T, C] -> framed [B, T + 2 pad, C] rows [pad, pad + T), frames >= lengths[b] zeroed."""
    B, T, Cd = x.shape
    _call("s2t_frame_rows", _p(x), _p(lengths), _p(framed), framed.stride(0), pad, B, T, Cd, _stream())
    return framed


def im2col2d(x: torch.Tensor, k: int, stride: int, Kpad: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x bf16 [B, T, D, Cin] -> patch matrix [B*T2*D2, Kpad] of Conv2d(k, stride, no padding); column (j*k+i)*Cin+c."""
    B, T, D, Cin = x.shape
    T2, D2 = (T - k) // stride + 1, (D - k) // stride + 1
    if out is None:
        out = torch.empty((B * T2 * D2, Kpad), device=x.device, dtype=BF16)
    _call("s2t_im2col2d", _p(x), _p(out), B, T, D, Cin, k, stride, Kpad, _stream())
    return out


def col2im2d(dcol: torch.Tensor, B: int, T: int, D: int, Cin: int, k: int, stride: int,
             out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if out is None:
        out = torch.empty((B, T, D, Cin), device=dcol.device, dtype=BF16)
    _call("s2t_col2im2d", _p(dcol), _p(out), B, T, D, Cin, k, stride, dcol.shape[1], _stream())
    return out


def colstats_parts(rows: int) -> int:
    return _lib.load().s2t_colstats_parts(rows)


def colstats(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x bf16 [rows, C] -> fp32 partials [parts, 2, C] (per-channel sum, sum of squares) for batchnorm_fwd."""
    rows, Cd = x.shape
    if out is None:
        out = torch.empty((colstats_parts(rows), 2, Cd), device=x.device, dtype=torch.float32)
    _call("s2t_colstats", _p(x), _p(out), rows, Cd, _stream())
    return out


def counter_add(counter: torch.Tensor, inc: int = 1) -> None:
    _call("s2t_counter_add", _p(counter), inc, _stream())


def add_(y: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    _call("s2t_add_bf16", _p(x), _p(y), y.numel(), _stream())
    return y


def f32_sum(x: torch.Tensor, out: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    _call("s2t_f32_sum", _p(x), x.numel(), scale, _p(out), _stream())
    return out


def weighted_sum(x: torch.Tensor, w: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """out[0] = sum_r w[r] * sum_c x[r, c]  (fp32)."""
    _call("s2t_f32_weighted_sum", _p(x), _p(w), x.shape[0], x.shape[1], _p(out), _stream())
    return out


# --------------------------------------------------------------------------------------------------
# attention softmax
# --------------------------------------------------------------------------------------------------
def attn_softmax_fwd(scores: torch.Tensor, klen: Optional[torch.Tensor], heads: int, Tk: int, *,
                     probs: Optional[torch.Tensor] = None, probs_drop: Optional[torch.Tensor] = None,
                     bias: Optional[torch.Tensor] = None, bias_mode: int = 0, scale: float = 1.0,
                     clamp: float = 0.0, dropout: Tuple[float, int, int] = (0.0, 0, 0),
                     bias_head_major: bool = False):
    """scores bf16 [nbh, Tq, ld]; returns (probs, probs_drop or None)."""
    nbh, Tq, ld = scores.shape
    if probs is None:
        probs = torch.empty_like(scores)
    ld_bias = bias.stride(1) if bias is not None else 0
    _call("s2t_attn_softmax_fwd", _p(scores), ld, _p(bias), ld_bias, bias_mode, int(bias_head_major), _p(probs),
          _p(probs_drop),
          _p(klen), nbh, heads, Tq, Tk, scale, clamp, *_drop(dropout), _stream())
    return probs, probs_drop


def attn_fused_fwd(qkv: torch.Tensor, C: int, klen: Optional[torch.Tensor], B: int, H: int, T: int, dk: int, *,
                   probs: Optional[torch.Tensor] = None, probs_drop: Optional[torch.Tensor] = None,
                   lse: Optional[torch.Tensor] = None, ctx: torch.Tensor, scale: float, dropout=None) -> None:
    """Fused self-attention forward (d_k == 64, T <= 256): qkv bf16 [B*T, >=3C] -> ctx [B*T, C] and either the
    probabilities (+ probs_drop) [B*H, T, ld] or the row log-sum-exp ``lse`` fp32 [B*H*T] (for ``attn_fused_bwd``)."""
    _call("s2t_attn_fused_fwd", _p(qkv), qkv.stride(0), C, _p(probs), _p(probs_drop),
          probs.stride(1) if probs is not None else 0, _p(lse), _p(ctx), ctx.stride(0), _p(klen), B, H, T, dk, scale,
          *_drop(dropout), _stream())


def attn_fused_bwd(qkv: torch.Tensor, C: int, ctx: torch.Tensor, dctx: torch.Tensor, lse: torch.Tensor,
                   klen: Optional[torch.Tensor], B: int, H: int, T: int, dk: int, *, out: torch.Tensor, scale: float,
                   dropout=None) -> torch.Tensor:
    """Fused self-attention backward: (qkv, ctx, dctx, lse) -> out = dq | dk | dv, bf16 [B*T, >=3C] (one launch)."""
    assert ctx.stride(0) == dctx.stride(0)
    _call("s2t_attn_fused_bwd", _p(qkv), qkv.stride(0), C, _p(ctx), _p(dctx), ctx.stride(0), _p(lse), _p(out),
          out.stride(0), _p(klen), B, H, T, dk, scale, *_drop(dropout), _stream())
    return out


def attn_softmax_bwd(probs: torch.Tensor, dprobs_drop: torch.Tensor, klen: Optional[torch.Tensor],
                     heads: int, Tk: int, *, out: Optional[torch.Tensor] = None,
                     dbias: Optional[torch.Tensor] = None, scale: float = 1.0,
                     dropout: Tuple[float, int, int] = (0.0, 0, 0), bias_head_major: bool = False) -> torch.Tensor:
    nbh, Tq, ld = probs.shape
    if out is None:
        out = torch.empty_like(probs)
    ld_bias = dbias.stride(1) if dbias is not None else 0
    _call("s2t_attn_softmax_bwd", _p(probs), _p(dprobs_drop), _p(out), _p(dbias), ld_bias, int(bias_head_major),
          _p(klen), nbh,
          heads, Tq, Tk, ld, scale, *_drop(dropout), _stream())
    return out


# --------------------------------------------------------------------------------------------------
# Conformer convolution module, rel-pos query biases
# --------------------------------------------------------------------------------------------------
def dwconv1d_tiles(B: int, T: int) -> int:
    return int(_lib.load().s2t_dwconv1d_tiles(B, T))


def dwconv1d_fwd(x: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None,
                 bn_partials: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x bf16 [B, T, C], w bf16 [C, K] -> depthwise conv along T (zero padded); bn_partials fp32 [tiles, 2, C]."""
    B, T, Cd = x.shape
    if out is None:
        out = torch.empty_like(x)
    _call("s2t_dwconv1d_fwd", _p(x), _p(w), _p(out), _p(bn_partials), B, T, Cd, w.shape[-1], _stream())
    return out


def dwconv1d_bwd(x: torch.Tensor, w: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor,
                 out: Optional[torch.Tensor] = None, workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
    B, T, Cd = x.shape
    K = w.shape[-1]
    if out is None:
        out = torch.empty_like(x)
    need = dwconv1d_tiles(B, T) * K * Cd
    if workspace is None or workspace.numel() < need:
        workspace = torch.empty((need,), device=x.device, dtype=torch.float32)
    _call("s2t_dwconv1d_bwd", _p(x), _p(w), _p(dy), _p(out), _p(dw), _p(workspace), B, T, Cd, K, _stream())
    return out


def batchnorm_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, *, partials: Optional[torch.Tensor],
                  running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor],
                  num_batches_tracked: Optional[torch.Tensor], save_stats: torch.Tensor,
                  out: Optional[torch.Tensor] = None, eps: float = 1e-5, momentum: float = 0.1,
                  act: int = ACT_NONE, training: bool = True) -> torch.Tensor:
    """x bf16 [rows, C] -> act(BatchNorm1d(x)); partials fp32 [n, 2, C] from dwconv1d_fwd (training)."""
    rows, Cd = x.shape
    if out is None:
        out = torch.empty_like(x)
    _call("s2t_batchnorm_fwd", _p(x), _p(partials), partials.shape[0] if partials is not None else 0, _p(gamma),
          _p(beta), _p(running_mean), _p(running_var), _p(num_batches_tracked), _p(save_stats), _p(out), rows, Cd,
          eps, momentum, act, int(training), _stream())
    return out


def batchnorm_bwd_partials(rows: int) -> int:
    return _lib.load().s2t_batchnorm_bwd_partials(rows)


def batchnorm_bwd(da: torch.Tensor, x: torch.Tensor, save_stats: torch.Tensor, gamma: torch.Tensor,
                  beta: torch.Tensor, dgamma: torch.Tensor, dbeta: torch.Tensor, *,
                  out: Optional[torch.Tensor] = None, partials: Optional[torch.Tensor] = None,
                  act: int = ACT_NONE, training: bool = True) -> torch.Tensor:
    rows, Cd = x.shape
    if out is None:
        out = torch.empty_like(x)
    need = batchnorm_bwd_partials(rows) * 2 * Cd
    if partials is None or partials.numel() < need:
        partials = torch.empty((need,), device=x.device, dtype=torch.float32)
    _call("s2t_batchnorm_bwd", _p(da), _p(x), _p(save_stats), _p(gamma), _p(beta), _p(partials), _p(out),
          _p(dgamma), _p(dbeta), rows, Cd, act, int(training), _stream())
    return out


def relpos_qbias_fwd(q: torch.Tensor, u: torch.Tensor, v: torch.Tensor, C: int,
                     out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q = first C columns of a [rows, ld] tensor -> [rows, 2C] = (q + u | q + v)."""
    rows = q.shape[0]
    if out is None:
        out = torch.empty((rows, 2 * C), device=q.device, dtype=BF16)
    _call("s2t_relpos_qbias_fwd", _p(q), q.stride(0), _p(u), _p(v), _p(out), rows, C, _stream())
    return out


def relpos_qbias_bwd(dquv: torch.Tensor, dq: torch.Tensor, C: int) -> torch.Tensor:
    _call("s2t_relpos_qbias_bwd", _p(dquv), _p(dq), dq.stride(0), dquv.shape[0], C, _stream())
    return dq


def act_dropout(x: torch.Tensor, act: int, spec, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if out is None:
        out = torch.empty_like(x)
    p, seed, sid, off = _drop(spec)
    _call("s2t_act_dropout", _p(x), _p(out), x.numel(), act, p, seed, sid, off, _stream())
    return out


# --------------------------------------------------------------------------------------------------
# CTC
# --------------------------------------------------------------------------------------------------
def row_lse(logits: torch.Tensor, V: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """logits [rows, ldv] (bf16 or fp32) -> fp32 [rows] log-sum-exp over the first V columns."""
    rows = logits.shape[0]
    if out is None:
        out = torch.empty((rows,), device=logits.device, dtype=torch.float32)
    _call("s2t_row_lse", _p(logits), int(logits.dtype == BF16), logits.stride(0), rows, V, _p(out), _stream())
    return out


def softmax_rows_fwd(logits: torch.Tensor, lse: torch.Tensor, V: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """probs = exp(logits - lse) over the first V columns (bf16 [rows, ld]); padding columns zeroed."""
    rows = logits.shape[0]
    if out is None:
        out = torch.empty_like(logits)
    _call("s2t_softmax_rows_fwd", _p(logits), logits.stride(0), _p(lse), _p(out), out.stride(0), rows, V, _stream())
    return out


def rows_truth_mix_(probs: torch.Tensor, truth_tok: torch.Tensor, mask: torch.Tensor, V: int, smooth: bool) -> None:
    """In place: rows with mask != 0 become the (smoothed) one-hot distribution of their aligned ground-truth token (int32 [rows])."""
    _call("s2t_rows_truth_mix", _p(probs), probs.stride(0), _p(truth_tok), _p(mask), probs.shape[0], V, int(smooth),
          _stream())


def rows_zero_masked_(x: torch.Tensor, mask: torch.Tensor, ncols: int) -> None:
    _call("s2t_rows_zero_masked", _p(x), x.stride(0), _p(mask), x.shape[0], ncols, _stream())


def softmax_rows_bwd(probs: torch.Tensor, dprobs: torch.Tensor, dlogits: torch.Tensor, V: int,
                     accumulate: bool = False) -> torch.Tensor:
    _call("s2t_softmax_rows_bwd", _p(probs), _p(dprobs), probs.stride(0), _p(dlogits), dlogits.stride(0),
          probs.shape[0], V, int(accumulate), _stream())
    return dlogits


def ctc_workspace_bytes(B: int, T: int, max_target_len: int) -> int:
    return _lib.load().s2t_ctc_workspace_bytes(B, T, max_target_len)


def ctc_loss(logits: torch.Tensor, targets: torch.Tensor, input_lengths: torch.Tensor,
             target_lengths: torch.Tensor, V: int, blank: int = 0, *, need_grad: bool = True,
             grad_scale: float = 1.0, lse: Optional[torch.Tensor] = None,
             dlogits: Optional[torch.Tensor] = None, workspace: Optional[torch.Tensor] = None,
             nll: Optional[torch.Tensor] = None, grad_scale_dev: Optional[torch.Tensor] = None,
             force_emit: Optional[torch.Tensor] = None):
    """logits [B, T, ldv] utterance-major; targets int32 [B, Smax].  Returns (nll fp32 [B], dlogits|None).
    ``grad_scale_dev``: optional fp32 [1] device scalar multiplied into ``grad_scale`` on the device (batch weight)."""
    B, T, ldv = logits.shape
    smax = targets.shape[1]
    flat = logits.view(B * T, ldv)
    if lse is None:
        lse = row_lse(flat, V)
    need = ctc_workspace_bytes(B, T, smax)
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty(((need + 3) // 4,), device=logits.device, dtype=torch.float32)
    if nll is None:
        nll = torch.empty((B,), device=logits.device, dtype=torch.float32)
    if need_grad and dlogits is None:
        dlogits = torch.empty_like(logits)
    if force_emit is not None:
        # masked ("imputer") loss: force_emit int32 [B, T] = forced extended-label state per frame, -1 = free
        if force_emit.dtype != torch.int32 or force_emit.shape[0] != B or force_emit.shape[1] < T:
            raise ValueError("ctc_loss: force_emit must be int32 [B, >= T]")
        _call("s2t_ctc_loss_masked", _p(logits), int(logits.dtype == BF16), ldv, _p(lse), _p(targets),
              targets.stride(0), _p(force_emit), force_emit.stride(0), _p(input_lengths), _p(target_lengths), B, T, V,
              smax, blank, _p(nll), _p(dlogits) if need_grad else None, grad_scale, _p(grad_scale_dev), _p(workspace),
              workspace.numel() * workspace.element_size(), _stream())
        return nll, (dlogits if need_grad else None)
    _call("s2t_ctc_loss", _p(logits), int(logits.dtype == BF16), ldv, _p(lse), _p(targets),
          targets.stride(0), _p(input_lengths), _p(target_lengths), B, T, V, smax, blank, _p(nll),
          _p(dlogits) if need_grad else None, grad_scale, _p(grad_scale_dev), _p(workspace),
          workspace.numel() * workspace.element_size(), _stream())
    return nll, (dlogits if need_grad else None)


def ctc_greedy(logits: torch.Tensor, input_lengths: torch.Tensor, V: int, blank: int = 0, with_scores: bool = False):
    """logits [B, T, ldv] -> (tokens int32 [B, T] padded with -1, lengths int32 [B], frame_ids int32 [B, T]) and, with
    ``with_scores``, the hypothesis scores fp32 [B] of CTCDecoder.generate (s2t_ctc.py:325-332)."""
    B, T, ldv = logits.shape
    tokens = torch.empty((B, T), device=logits.device, dtype=torch.int32)
    lens = torch.empty((B,), device=logits.device, dtype=torch.int32)
    frame_ids = torch.empty((B, T), device=logits.device, dtype=torch.int32)
    scores = torch.empty((B,), device=logits.device, dtype=torch.float32) if with_scores else None
    frame_lp = torch.empty((B, T), device=logits.device, dtype=torch.float32) if with_scores else None
    _call("s2t_ctc_greedy", _p(logits), int(logits.dtype == BF16), ldv, _p(input_lengths), B, T, V, blank,
          _p(tokens), _p(lens), _p(frame_ids), _p(scores), _p(frame_lp), _stream())
    if with_scores:
        return tokens, lens, frame_ids, scores
    return tokens, lens, frame_ids


# --------------------------------------------------------------------------------------------------
# SpecAugment (apply half; the draws are host-side, s2t_b200/specaugment.py)
# --------------------------------------------------------------------------------------------------
def specaugment_(feats: torch.Tensor, num_frames: torch.Tensor, freq_masks: Optional[torch.Tensor],
                 time_masks: Optional[torch.Tensor], mask_value: Optional[float] = 0.0, n_freqs: Optional[int] = None,
                 means: Optional[torch.Tensor] = None) -> torch.Tensor:
    """In place on feats [B, T, ld] (fp32 or bf16, last dim contiguous): freq_masks int32 [B, Nf, 2] = (f0, f),
    time_masks int32 [B, Nt, 2] = (t0, t); ``mask_value=None`` fills with the utterance mean (specaugment.py:88-89)."""
    B, T, ld = feats.shape
    assert feats.stride(2) == 1 and feats.dtype in (torch.float32, BF16)
    nfm = 0 if freq_masks is None else freq_masks.shape[1]
    ntm = 0 if time_masks is None else time_masks.shape[1]
    use_mean = mask_value is None
    if use_mean and means is None:
        means = torch.empty((B * 32,), device=feats.device, dtype=torch.float32)      # partial sums (scratch)
    _call("s2t_specaugment", _p(feats), int(feats.dtype == BF16), feats.stride(1), feats.stride(0), _p(num_frames), B, T,
          n_freqs if n_freqs is not None else ld, _p(freq_masks), nfm, _p(time_masks), ntm,
          0.0 if use_mean else float(mask_value), int(use_mean), _p(means), _stream())
    return feats


# --------------------------------------------------------------------------------------------------
# CTC best alignment (Viterbi)
# --------------------------------------------------------------------------------------------------
def ctc_viterbi(log_probs: torch.Tensor, targets: torch.Tensor, input_lengths: torch.Tensor,
                target_lengths: torch.Tensor, blank: int = 0, batch_first: bool = True,
                with_nll: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """torch_imputer.best_alignment on device, backtrack included: log_probs fp32 [B, T, V] (or [T, B, V] with
    ``batch_first=False``, the reference's layout), targets int32 [B, S]; returns int32 states [B, T] (-1 beyond the
    utterance) and the fp32 nll [B] of the max-recursion."""
    assert log_probs.dtype == torch.float32 and log_probs.stride(2) == 1
    if batch_first:
        B, T, V = log_probs.shape
        ts, bs = log_probs.stride(1), log_probs.stride(0)
    else:
        T, B, V = log_probs.shape
        ts, bs = log_probs.stride(0), log_probs.stride(1)
    smax = targets.shape[1]
    states = torch.empty((B, T), device=log_probs.device, dtype=torch.int32)
    nll = torch.empty((B,), device=log_probs.device, dtype=torch.float32) if with_nll else None
    need = _lib.load().s2t_ctc_viterbi_workspace_bytes(B, T, smax)
    ws = torch.empty((need,), device=log_probs.device, dtype=torch.uint8) if need else None
    _call("s2t_ctc_viterbi", _p(log_probs), ts, bs, _p(targets), targets.stride(0), _p(input_lengths),
          _p(target_lengths), B, T, V, smax, blank, _p(states), _p(nll), _p(ws), need, _stream())
    return states, nll
