"""S2TEncoderEngine — the B200 execution engine of the speech-encoder training hot path.

One engine = one encoder (+ CTC head) resident on one GPU.  It owns
  * the flat bf16 parameter / gradient buffers (``params.ParamStore``; reference parameter names),
  * one *plan* per (batch, frames, max-target-length) bucket: every activation, gradient scratch and
    workspace pre-allocated, so a training step allocates nothing, never syncs with the host and can be
    captured once into a CUDA graph and replayed (the reference path has a dozen host syncs per step,
    SURVEY.md Appendix A16),
  * the explicit forward and backward kernel sequences (no autograd): every launch is one of our own
    sm_100a kernels reached through the C ABI (``ops``).

Row layout: activations are [b*T + t][C] (utterance-major), bf16 storage, fp32 accumulation.

Two encoder families share the layer code:
  * ``encoder_type="s2t"`` — ``S2TTransformerEncoder`` (fairseq/models/speech_to_text/s2t_transformer.py:883,
    forward :1714-2154): Conv1dSubsampling stem (modules/speech_to_text/subsampling.py:106-159), one stage of
    Transformer / Conformer layers (modules/s2t_transformer_layer.py:229-322);
  * ``encoder_type="pds"`` — ``PDSS2TTransformerEncoder`` (models/speech_to_text/pdss2t_transformer.py:1042-1281):
    per stage a ``Downsampling`` Conv1d (+LayerNorm) stem (:53-144), positions, and ``PDSTransformerEncoderLayer``s
    (modules/pds_layer.py:263-359).
Attention is fairseq ``MultiheadAttention`` (modules/multihead_attention.py) or ESPnet
``RelPositionMultiHeadedAttention`` (modules/espnet_multihead_attention.py:269-356); the conv module is
modules/convolution.py:8-120; heads are ``CTC`` (modules/speech_to_text/ctc.py:17-66) with the ``ctc`` criterion
arithmetic (criterions/ctc.py:435-474,516-540).
"""
from __future__ import annotations

import contextlib
import math
import os
import re
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import frontend, ops
from .params import BF16, ParamStore

F32 = torch.float32


@dataclass
class EncoderConfig:
    """The subset of the reference's args Namespace the path reads (s2t_transformer.py:2256-2405,
    pdss2t_transformer.py:1361-1560)."""
    encoder_type: str = "s2t"              # s2t | pds
    input_dim: int = 80
    embed_dim: int = 256
    ffn_dim: int = 2048
    heads: int = 4
    layers: int = 12
    subsampling_layers: int = 2
    subsampling_filter: int = 1024
    subsampling_kernel: int = 5
    subsampling_stride: int = 2
    subsampling_type: str = "conv1d"       # conv1d (Conv1d + GLU stack) | conv2d (Conv2dSubsampling, subsampling.py:162-228)
    subsampling_norm: str = "none"         # conv2d: none | batch2d
    subsampling_activation: str = "glu"    # conv1d: glu; conv2d: swish | relu | none
    attn_type: str = "selfattn"            # selfattn | rel_pos
    encoder_embed_norm: bool = True
    no_scale_embedding: bool = True
    macaron_style: bool = False
    use_cnn_module: bool = False
    cnn_module_kernel: int = 15
    cnn_module_norm: str = "batch_norm"
    encoder_activation_fn: str = "relu"    # FFN activation
    activation_fn: str = "relu"            # conv-module activation of the S2T layer (quirk A3)
    layer_padding_mask: bool = False
    dropout: float = 0.1
    max_source_positions: int = 6000
    vocab: int = 10000                     # CTC head (source dictionary); 0 = none
    ctc_weight: float = 1.0
    blank: int = 0
    # --- dual heads / inter-CTC / prediction-aware encoding (s2t_transformer.py:951-1244; C5) -------------
    tgt_vocab: int = 0                     # XCTC head (target dictionary); 0 = none
    xctc_weight: float = 0.0
    inter_ctc_layers: Tuple[int, ...] = ()     # 1-based layer indices (inter-ctc-layers: 6,9,12,15)
    inter_ctc_weight: float = 0.0
    inter_xctc_layers: Tuple[int, ...] = ()
    inter_xctc_weight: float = 0.0
    share_inter_ctc: bool = True
    ctc_pae: str = "none"                  # none | inter_league
    xctc_pae: str = "none"
    pae_unnorm_input: bool = False
    pae_gt_smooth: bool = False            # --pae-[o]racle-smooth: smoothed one-hot of the aligned ground-truth token (adapter.py:256-257)
    # --- PDS (pdss2t_transformer.py:1361-1560; defaults = pdss2t_transformer_s_8) -------------------------
    pds_layers: Tuple[int, ...] = (3, 3, 3, 3)
    pds_ratios: Tuple[int, ...] = (2, 2, 1, 2)
    pds_dims: Tuple[int, ...] = (256, 256, 256, 256)
    pds_kernels: Tuple[int, ...] = (5, 5, 5, 5)
    pds_ffn_ratios: Tuple[int, ...] = (8, 8, 8, 8)
    pds_heads: Tuple[int, ...] = (4, 4, 4, 4)
    pds_position_embed: Tuple[int, ...] = (1, 1, 1, 1)
    pds_embed_norm: bool = True
    pds_dropout: float = 0.0               # base_architecture default (pdss2t_transformer.py:1489)

    @staticmethod
    def from_args(args, vocab: int, family: str = "s2t") -> "EncoderConfig":
        g = lambda k, d=None: getattr(args, k, d)
        ints = lambda s: tuple(int(v) for v in str(s).split("_"))
        csv = lambda s: tuple(int(v) for v in str(s).split(",")) if s else ()
        kw = dict(
            input_dim=g("input_feat_per_channel", 80) * g("input_channels", 1),
            embed_dim=args.encoder_embed_dim, ffn_dim=args.encoder_ffn_embed_dim,
            heads=args.encoder_attention_heads, layers=args.encoder_layers,
            subsampling_layers=g("subsampling_layers", 2), subsampling_filter=g("subsampling_filter", 1024),
            subsampling_kernel=g("subsampling_kernel", 5), subsampling_stride=g("subsampling_stride", 2),
            subsampling_type=g("subsampling_type", "conv1d"), subsampling_norm=g("subsampling_norm", "none"),
            subsampling_activation=g("subsampling_activation", "glu"),
            attn_type=g("encoder_attention_type", "selfattn"), encoder_embed_norm=g("encoder_embed_norm", False),
            no_scale_embedding=g("encoder_no_scale_embedding", False), macaron_style=g("macaron_style", False),
            use_cnn_module=g("use_cnn_module", False), cnn_module_kernel=g("cnn_module_kernel", 31),
            cnn_module_norm=g("cnn_module_norm", "batch_norm"),
            encoder_activation_fn=g("encoder_activation_fn", "relu"), activation_fn=g("activation_fn", "relu"),
            layer_padding_mask=g("layer_padding_mask", False), dropout=args.dropout,
            max_source_positions=g("max_source_positions", 6000),
            vocab=vocab if g("ctc_weight", 0) > 0 else 0, ctc_weight=g("ctc_weight", 1.0) or 1.0,
            tgt_vocab=g("_tgt_vocab", vocab) if g("xctc_weight", 0) > 0 else 0, xctc_weight=g("xctc_weight", 0.0),
            inter_ctc_layers=csv(g("inter_ctc_layers")), inter_ctc_weight=g("inter_ctc_weight", 0.0),
            inter_xctc_layers=csv(g("inter_xctc_layers")), inter_xctc_weight=g("inter_xctc_weight", 0.0),
            share_inter_ctc=g("share_inter_ctc", False), ctc_pae=g("ctc_pae", "none"),
            xctc_pae=g("xctc_pae", g("ctc_pae", "none")), pae_unnorm_input=g("pae_unnorm_input", False),
            pae_gt_smooth=g("pae_oracle_smooth", False))
        if family == "pds":
            n = int(args.pds_stages)
            kw.update(encoder_type="pds", pds_layers=ints(args.pds_layers)[:n], pds_ratios=ints(args.pds_ratios)[:n],
                      pds_dims=ints(args.pds_embed_dims)[:n], pds_kernels=ints(args.pds_kernel_sizes)[:n],
                      pds_ffn_ratios=ints(args.pds_ffn_ratios)[:n], pds_heads=ints(args.pds_attn_heads)[:n],
                      pds_position_embed=ints(g("pds_position_embed", "_".join(["1"] * n)))[:n],
                      pds_embed_norm=g("pds_embed_norm", False), pds_dropout=g("pds_dropout", args.dropout))
        return EncoderConfig(**kw)


def sinusoidal_table(num_embeddings: int, dim: int, padding_idx: int = 1) -> torch.Tensor:
    """fairseq/modules/sinusoidal_positional_embedding.py:34-58: [sin | cos] halves, frequency
    denominator dim/2 - 1, padding row zeroed (quirk A6)."""
    half = dim // 2
    f = torch.exp(torch.arange(half, dtype=torch.float) * -(math.log(10000) / (half - 1)))
    ang = torch.arange(num_embeddings, dtype=torch.float).unsqueeze(1) * f.unsqueeze(0)
    tab = torch.cat([torch.sin(ang), torch.cos(ang)], dim=1)
    if dim % 2 == 1:
        tab = torch.cat([tab, torch.zeros(num_embeddings, 1)], dim=1)
    tab[padding_idx, :] = 0
    return tab.contiguous()


def rel_position_table(T: int, dim: int) -> torch.Tensor:
    """RelPositionalEncoding (fairseq/modules/positional_encoding.py:104-166): the centre 2T-1 rows of the table,
    row k <-> relative position T-1-k, interleaved sin/cos, div = exp(2i * -ln(1e4)/dim).  fp32 [2T-1, dim]."""
    pos = torch.arange(0, T, dtype=torch.float32).unsqueeze(1)
    div = torch.exp(torch.arange(0, dim, 2, dtype=torch.float32) * -(math.log(10000.0) / dim))
    pe_pos, pe_neg = torch.zeros(T, dim), torch.zeros(T, dim)
    pe_pos[:, 0::2], pe_pos[:, 1::2] = torch.sin(pos * div), torch.cos(pos * div)
    pe_neg[:, 0::2], pe_neg[:, 1::2] = torch.sin(-pos * div), torch.cos(-pos * div)
    return torch.cat([torch.flip(pe_pos, [0]), pe_neg[1:]], dim=0).contiguous()


def _round8(n: int) -> int:
    return (n + 7) // 8 * 8


_ACT = {"relu": ops.ACT_RELU, "swish": ops.ACT_SWISH}
_NORM_W = re.compile(r"(norm\d*|embed_ln)\.weight$")
_NORM_B = re.compile(r"(norm\d*|embed_ln)\.bias$")


@dataclass
class _LayerSpec:
    """Static description of one encoder layer: parameter prefix, widths, owning stage, dropout-site base."""
    prefix: str
    stage: int
    index: int          # position inside its stage
    gidx: int           # global layer index
    C: int
    F: int
    H: int
    conv_act: int
    p_drop: float
    first: str = ""      # first parameter of the layer in the flat store

    @property
    def site(self) -> int:
        return 64 * (self.gidx + 1)


@dataclass
class _HeadSpec:
    """One CTC head (modules/speech_to_text/ctc.py:17-66) and its inter-CTC uses (s2t_transformer.py:1881-1946,2043-2103)."""
    name: str                      # parameter prefix: "ctc" (source vocabulary) | "xctc" (target vocabulary)
    V: int
    weight: float                  # weight of the top loss
    inter_layers: Tuple[int, ...]  # 1-based layer indices carrying an inter loss with this (shared) head
    inter_weight: float
    norm: str                      # "ctc_norm" | "xctc_norm" (+ layer index)
    pae: Optional[str]             # adapter parameter prefix ("pae" | "xctc_pae") or None
    tgt: int                       # target set: 0 = transcript (source text), 1 = target text
    site: int                      # dropout site base

    @property
    def ldv(self) -> int:
        return _round8(max(self.V, 8))


class _Bufs:
    """Attribute bag for saved activations / scratch."""


class _Stage(_Bufs):
    """Buffers shared by the layers of one stage (same T, C): lengths, attention scratch, gradient scratch pool."""

    def __init__(self, eng: "S2TEncoderEngine", B: int, T: int, C: int, F: int, H: int, n_layers: int, idx: int):
        c, dev = eng.cfg, eng.device
        z = lambda *s, dt=BF16: torch.zeros(s, device=dev, dtype=dt)
        e = lambda *s, dt=BF16: torch.empty(s, device=dev, dtype=dt)
        self.idx, self.B, self.T, self.C, self.F, self.H = idx, B, T, C, F, H
        self.dk = C // H
        R = self.R = B * T
        self.ld = _round8(T)
        self.lens = z(B, dt=torch.int32)
        self.x = [e(R, C) for _ in range(n_layers + 1)]     # x[j] = input of layer j, x[-1] = stage output
        self.scores = z(B * H, T, self.ld)
        self.g_p, self.g_s = z(B * H, T, self.ld), z(B * H, T, self.ld)
        self.pool = [e(R, C) for _ in range(18)]      # incl. one block's worth in quarantine (engine._end_block)
        self._pool_ids = {id(t) for t in self.pool}
        self.g_fs = [e(R, F), e(R, F)]       # FFN hidden gradient: two buffers, alternated per FFN block (engine._end_block)
        self._g_f_i = 0
        self.g_qkv = e(R, 3 * C)
        lib = ops._lib.load()
        cs = max(lib.s2t_colsum_workspace_floats(R, 3 * C), lib.s2t_colsum_workspace_floats(R, F))
        self.colsum_ws, self.colsum_ws2 = e(cs, dt=F32), e(cs, dt=F32)   # two: consecutive colsums overlap on a side stream
        # dgamma / dbeta partial sums: one buffer per LayerNorm instance (allocated on first use), reduced together by ONE
        # batched launch where a gradient bucket is shipped or at the end of backward (engine._flush_ln)
        self.ln_parts: Dict[str, torch.Tensor] = {}
        if c.attn_type == "rel_pos":
            n_rel = 2 * T - 1
            self.ldb = _round8(n_rel)
            self.pos_emb = rel_position_table(T, C).to(dev).to(BF16)          # (2T-1, C), .to(x.dtype) in the reference
            self.bd = z(H, B, T, self.ldb)                                      # head-major (see softmax.cu)
            self.g_bd = z(H, B, T, self.ldb)
            self.g_quv = e(R, 2 * C)
            self.g_pp = e(n_rel, C)
        if c.use_cnn_module:
            self.g_2c = e(R, 2 * C)
            self.bn_part = e(ops.dwconv1d_tiles(B, T), 2, C, dt=F32)
            self.bn_bwd_part = e(ops.batchnorm_bwd_partials(R), 2, C, dt=F32)
            self.dw_ws = e(ops.dwconv1d_tiles(B, T) * c.cnn_module_kernel * C, dt=F32)

    def take(self) -> torch.Tensor:
        return self.pool.pop()

    def ln_part(self, name: str, rows: int, C: int, dev) -> torch.Tensor:
        t = self.ln_parts.get(name)
        if t is None:
            n = ops._lib.load().s2t_layernorm_bwd_partials(rows, C)
            t = self.ln_parts[name] = torch.empty((n, 2, C), device=dev, dtype=F32)
        return t

    def give(self, *bufs: torch.Tensor) -> None:
        for t in bufs:
            if t is not None and id(t) in self._pool_ids and all(t is not f for f in self.pool):
                self.pool.append(t)


class Plan:
    """All device buffers of one (B, T, Smax) bucket."""

    def __init__(self, eng: "S2TEncoderEngine", B: int, T: int, smax: int, wave_len: int = 0):
        c, dev = eng.cfg, eng.device
        self.B, self.T, self.smax, self.wave_len = B, T, smax, wave_len
        z = lambda *s, dt=BF16: torch.zeros(s, device=dev, dtype=dt)
        e = lambda *s, dt=BF16: torch.empty(s, device=dev, dtype=dt)
        # --- static inputs --------------------------------------------------------------------
        self.feats_in = z(B, T, c.input_dim, dt=F32)
        self.wave = z(B, wave_len, dt=F32) if wave_len else None
        self.num_samples = z(B, dt=torch.int32) if wave_len else None
        self.fbank_raw = e(B, T, c.input_dim, dt=F32) if wave_len else None
        self.cmvn_stats = e(B, 2, c.input_dim, dt=F32) if wave_len else None
        self.sa_freq = self.sa_time = None      # SpecAugment masks of the current batch (set_inputs), int32 [B, N, 2]
        self.sa_means = z(B * 32, dt=F32)       # s2t_specaugment scratch (partial sums of the utterance mean)
        self.src_len = z(B, dt=torch.int32)
        # weight of this batch in the gradient (1 = real batch, 0 = the zero-weight dummy batch a data-parallel rank
        # runs when it has no batch left, trainer.py:1045-1058,687-691); a device scalar so that one captured graph serves both
        self.batch_weight = torch.ones(1, device=dev, dtype=F32)
        self.batch_w_host = 1.0                 # host copy of batch_weight: set_inputs refills the device scalar only on change
        self.targets = z(B, max(smax, 1), dt=torch.int32)
        self.tgt_len = z(B, dt=torch.int32)
        drop = c.dropout > 0
        self.stages: List[_Stage] = []
        lib = ops._lib.load()
        if c.encoder_type == "s2t" and c.subsampling_type == "conv2d":
            # Conv2dSubsampling stem: channels-last activations [B][T][D][C], explicit patch matrices
            k, st = c.subsampling_kernel, c.subsampling_stride
            self.x_in = z(B, T, c.input_dim)
            self.c2: List[_Bufs] = []
            Ti, Di = T, c.input_dim
            for i, (cin, cout, Kpad) in enumerate(eng.sub2d):
                L = _Bufs()
                L.Tin, L.Din, L.cin, L.cout, L.Kpad = Ti, Di, cin, cout, Kpad
                L.Tout, L.Dout = (Ti - k) // st + 1, (Di - k) // st + 1
                assert L.Tout > 0 and L.Dout > 0, "input shorter than the Conv2d subsampling receptive field"
                L.rows = B * L.Tout * L.Dout
                L.col = e(L.rows, Kpad)
                L.pre, L.y = e(L.rows, cout), e(L.rows, cout)
                L.g_y, L.g_pre = e(L.rows, cout), e(L.rows, cout)
                L.g_col = e(L.rows, Kpad) if i > 0 else None
                L.bn_part = e(ops.colstats_parts(L.rows), 2, cout, dt=F32)
                L.bn_save = e(2, cout, dt=F32)
                L.bn_bwd_part = e(ops.batchnorm_bwd_partials(L.rows) * 2 * cout, dt=F32)
                L.colsum_ws = e(lib.s2t_colsum_workspace_floats(L.rows, cout), dt=F32)
                self.c2.append(L)
                Ti, Di = L.Tout, L.Dout
            assert Di == eng.sub2d_dim
            self.sub_len = [z(B, dt=torch.int32) for _ in eng.sub2d]
            St = _Stage(eng, B, Ti, c.embed_dim, c.ffn_dim, c.heads, c.layers, 0)
            St.lens = self.sub_len[-1]
            self.stages.append(St)
            self.x0 = e(St.R, c.embed_dim)
            self.e_ln, self.e_st = e(St.R, c.embed_dim), e(St.R, 2, dt=F32)
            self.stem_colsum_ws = e(lib.s2t_colsum_workspace_floats(St.R, c.embed_dim), dt=F32)
        elif c.encoder_type == "s2t":
            k, st = c.subsampling_kernel, c.subsampling_stride
            pad = self.pad = (k - 1) // 2
            # Conv1dSubsampling stem: framed inputs (pad zero rows front/back), pre-GLU conv outputs
            self.t_sub = [T]
            for _ in range(c.subsampling_layers):
                self.t_sub.append((self.t_sub[-1] + 2 * pad - k) // st + 1)
            chans = [c.input_dim] + [f // 2 for f in eng.sub_filters]
            n = c.subsampling_layers
            self.framed = [z(B, self.t_sub[i] + 2 * pad, chans[i]) for i in range(n)]
            self.pre = [e(B * self.t_sub[i + 1], eng.sub_filters[i]) for i in range(n)]
            self.sub_len = [z(B, dt=torch.int32) for _ in range(n)]
            # frames the REFERENCE's intermediate conv outputs have for this batch (its padded length may be shorter than
            # the plan's frame bucket): rows beyond are zero there (the next Conv1d's zero padding), so the intermediate
            # GLUs cut at this count — set per batch by set_inputs
            self.batch_sub = [torch.full((B,), self.t_sub[i + 1], device=dev, dtype=torch.int32) for i in range(n - 1)]
            self.batch_T = T
            self.g_pre = [e(B * self.t_sub[i + 1], eng.sub_filters[i]) for i in range(n)]
            self.g_col = [e(B * self.t_sub[i + 1], k * chans[i]) for i in range(n)]
            self.xcol = [e(B * self.t_sub[i + 1], k * chans[i]) for i in range(n)]
            self.g_sub = [e(B, self.t_sub[i], chans[i]) for i in range(n)]
            T2 = self.t_sub[-1]
            St = _Stage(eng, B, T2, c.embed_dim, c.ffn_dim, c.heads, c.layers, 0)
            St.lens = self.sub_len[-1]
            self.stages.append(St)
            self.x0 = e(St.R, c.embed_dim)
            self.e_ln, self.e_st = e(St.R, c.embed_dim), e(St.R, 2, dt=F32)
            cs = max(lib.s2t_colsum_workspace_floats(p.shape[0], p.shape[1]) for p in self.pre)
            self.stem_colsum_ws = e(cs, dt=F32)
        else:
            # PDS: pad T to a multiple of prod(ratios) — a whole extra block when already a multiple (quirk A7)
            mult = 1
            for r in c.pds_ratios:
                mult *= max(1, r)
            self.T_pad = T + (mult - T % mult) if mult > 1 else T
            t_in, cin = self.T_pad, c.input_dim
            self.pds: List[_Bufs] = []
            for i, (r, d, k) in enumerate(zip(c.pds_ratios, c.pds_dims, c.pds_kernels)):
                pad = (k - 1) // 2
                t_out = (t_in + 2 * pad - k) // r + 1
                D = _Bufs()
                D.t_in, D.t_out, D.cin, D.cout, D.k, D.stride, D.pad = t_in, t_out, cin, d, k, r, pad
                D.framed = z(B, t_in + 2 * pad, cin)
                D.conv = e(B * t_out, d)
                D.ln, D.st = (e(B * t_out, d), e(B * t_out, 2, dt=F32)) if c.pds_embed_norm else (None, None)
                D.in_len = z(B, dt=torch.int32)
                D.xcol = e(B * t_out, k * cin)
                D.g_col = e(B * t_out, k * cin)
                D.colsum_ws = e(lib.s2t_colsum_workspace_floats(B * t_out, d), dt=F32)
                self.pds.append(D)
                self.stages.append(_Stage(eng, B, t_out, d, d * c.pds_ffn_ratios[i], c.pds_heads[i], c.pds_layers[i], i))
                t_in, cin = t_out, d
        last = self.stages[-1]
        T2 = self.T2 = last.T
        R = self.R = last.R
        C = last.C
        self.ld = last.ld
        # --- per-layer saved activations ---------------------------------------------------------
        self.layers: List[_Bufs] = []
        for spec in eng.layer_specs:
            St = self.stages[spec.stage]
            Rs, Cs, Fs, Hs = St.R, St.C, St.F, St.H
            L = _Bufs()
            pdrop = spec.p_drop > 0
            if c.macaron_style:
                L.m_ln, L.m_st = e(Rs, Cs), e(Rs, 2, dt=F32)
                L.m_f = e(Rs, Fs)
                L.m_h = e(Rs, Fs) if c.encoder_activation_fn != "relu" else None
                L.xa = e(Rs, Cs)
            L.ln1, L.st1 = e(Rs, Cs), e(Rs, 2, dt=F32)
            L.qkv = e(Rs, 3 * Cs)
            if eng.attn_is_fused(St):
                # fused attention (attn.cu): the backward pass recomputes the probabilities from the row log-sum-exp
                L.lse = e(B * Hs * St.T, dt=F32)
                L.probs = L.probs_d = None
            else:
                L.probs = z(B * Hs, St.T, St.ld)
                L.probs_d = z(B * Hs, St.T, St.ld) if pdrop else L.probs
            L.ctx = e(Rs, Cs)
            L.x1 = e(Rs, Cs)
            if c.attn_type == "rel_pos":
                L.quv = e(Rs, 2 * Cs)
                L.pp = e(2 * St.T - 1, Cs)
            if c.use_cnn_module:
                L.c_ln, L.c_st = e(Rs, Cs), e(Rs, 2, dt=F32)
                L.pw1 = e(Rs, 2 * Cs)
                L.glu = e(Rs, Cs)
                L.dw = e(Rs, Cs)
                L.bn = e(Rs, Cs)
                L.bn_save = e(2, Cs, dt=F32)
                L.xc = e(Rs, Cs)
                L.xd, L.f_st = e(Rs, Cs), e(Rs, 2, dt=F32)     # input / stats of final_norm
            L.ln2, L.st2 = e(Rs, Cs), e(Rs, 2, dt=F32)
            L.f = e(Rs, Fs)
            L.h = e(Rs, Fs) if c.encoder_activation_fn != "relu" else None
            self.layers.append(L)
        self.enc_out, self.enc_st = e(R, C), e(R, 2, dt=F32)
        # --- CTC heads: top logits per head, inter-CTC blocks, one nll row per loss term ------------------------
        self.tg = [self.targets, z(B, max(smax, 1), dt=torch.int32)]
        self.tl = [self.tgt_len, z(B, dt=torch.int32)]
        self.heads: List[_Bufs] = []
        self.inter: Dict[Tuple[int, int], _Bufs] = {}
        weights: List[float] = []
        for k, hs in enumerate(eng.heads):
            Hd = _Bufs()
            Hd.logits, Hd.dlogits = z(R, hs.ldv), z(R, hs.ldv)
            Hd.lse = e(R, dt=F32)
            Hd.drop = e(R, C) if drop else self.enc_out
            Hd.colsum_ws = e(lib.s2t_colsum_workspace_floats(R, hs.ldv), dt=F32)
            Hd.loss_idx = len(weights)
            weights.append(hs.weight)
            if hs.inter_layers:
                Hd.ddist = e(R, hs.ldv)                                  # scratch shared by the head's inter blocks
            self.heads.append(Hd)
            for l in hs.inter_layers:
                I = _Bufs()
                I.nx, I.st = e(R, C), e(R, 2, dt=F32)
                I.nxd = e(R, C) if drop else I.nx
                I.logits, I.dlogits = z(R, hs.ldv), z(R, hs.ldv)
                I.lse = e(R, dt=F32)
                if hs.pae:
                    I.dist = z(R, hs.ldv)
                I.loss_idx = len(weights)
                weights.append(hs.inter_weight / len(hs.inter_layers))
                self.inter[(l, k)] = I
        if eng.heads:
            self.nll_all = z(len(weights), B, dt=F32)
            self.loss_w = torch.tensor(weights, dtype=F32, device=dev)
            self.loss_w0 = self.loss_w.clone()       # loss_w = loss_w0 * batch weight (a dummy batch logs loss 0, trainer.py:687-691)
            self.loss_weights = weights
            self.loss = z(1, dt=F32)
            self.ctc_ws = e((ops.ctc_workspace_bytes(B, T2, max(smax, 1)) + 3) // 4, dt=F32)
            # single-head conveniences (C1-C4 callers): the first head's buffers
            h0 = self.heads[0]
            self.logits, self.dlogits, self.lse, self.ldv = h0.logits, h0.dlogits, h0.lse, eng.heads[0].ldv
            self.nll = self.nll_all[0]
        # x chain: y[j] = buffer layer j writes; the inter blocks after it map y[j] -> ... -> x[j+1] (next layer's input)
        St0 = self.stages[0]
        St0.y = list(St0.x[1:])
        for j in range(len(St0.y)):
            blocks = eng.inter_after.get(j + 1, [])
            pae_blocks = [k for k in blocks if eng.heads[k].pae]
            cur = St0.x[j + 1]
            if pae_blocks:
                cur = St0.y[j] = e(R, C)
            for k in blocks:
                I = self.inter[(j + 1, k)]
                I.xin = cur
                if eng.heads[k].pae:
                    I.xout = St0.x[j + 1] if k == pae_blocks[-1] else e(R, C)
                    cur = I.xout
                else:
                    I.xout = cur
        # ground-truth curriculum of the prediction-aware adapters (engine.set_ground_truth): per head (aligned token, mask,
        # forced state) per frame, int32 [B*T']; None = off
        self.gt: Dict[int, Tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = {}
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.graph_kind: Optional[str] = None
        self.nbytes = 0

    @property
    def out_len(self) -> torch.Tensor:
        """int32 [B] valid frames of the encoder output."""
        return self.stages[-1].lens


class S2TEncoderEngine:
    SEED = 0x5217B200

    def __init__(self, cfg: EncoderConfig, device: Optional[torch.device] = None):
        if not torch.cuda.is_available():
            raise ops._lib.S2TLibraryError("S2TEncoderEngine needs a CUDA device (no CPU fallback)")
        ops._lib.load()
        if cfg.encoder_type not in ("s2t", "pds"):
            raise NotImplementedError(f"engine: encoder_type {cfg.encoder_type!r}")
        if cfg.attn_type not in ("selfattn", "rel_pos"):
            raise NotImplementedError(f"engine: attention type {cfg.attn_type!r} is not built")
        if cfg.encoder_activation_fn not in _ACT or cfg.activation_fn not in _ACT:
            raise NotImplementedError("engine: activations are relu | swish")
        if cfg.use_cnn_module and cfg.cnn_module_norm != "batch_norm":
            raise NotImplementedError("engine: conv module norm is batch_norm")
        self.cfg = cfg
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        c = cfg
        S = self.store = ParamStore(self.device)
        self.conv_names: List[str] = []          # Conv1d weights stored [Cout][k][Cin] (reference: [Cout][Cin][k])
        self.ref_shapes: Dict[str, Tuple[int, ...]] = {}   # reference shape where it differs by a plain reshape
        self.conv2d_names: Dict[str, Tuple[int, int, int, int]] = {}   # Conv2d weights [Cout][Kpad]: (cout, cin, k, Kpad)
        self.lin_perm: Dict[str, Tuple[int, int, int]] = {}            # Conv2dSubsampling.linear: (C, channels, D')
        self.buffers: Dict[str, torch.Tensor] = {}         # BatchNorm running statistics (fp32 / int64)
        self.layer_specs: List[_LayerSpec] = []
        self.ffn_act = _ACT[c.encoder_activation_fn]
        self.heads: List[_HeadSpec] = []
        if c.vocab > 0:
            self.heads.append(_HeadSpec("ctc", c.vocab, c.ctc_weight, tuple(c.inter_ctc_layers), c.inter_ctc_weight,
                                        "ctc_norm", "pae" if c.ctc_pae != "none" else None, 0, 40))
        if c.tgt_vocab > 0:
            self.heads.append(_HeadSpec("xctc", c.tgt_vocab, c.xctc_weight, tuple(c.inter_xctc_layers),
                                        c.inter_xctc_weight, "xctc_norm", "xctc_pae" if c.xctc_pae != "none" else None,
                                        1, 50))
        for hs in self.heads:
            if hs.inter_layers and (c.encoder_type != "s2t" or not c.share_inter_ctc):
                raise NotImplementedError("engine: inter-CTC is built for the S2T encoder with share_inter_ctc")
            if hs.pae and (c.ctc_pae if hs.name == "ctc" else c.xctc_pae) != "inter_league":
                raise NotImplementedError("engine: the prediction-aware adapter is inter_league")
        # inter_after[l] = head indices with an inter block after (1-based) layer l, CTC before XCTC (forward order)
        self.inter_after: Dict[int, List[int]] = {}
        for k, hs in enumerate(self.heads):
            for l in hs.inter_layers:
                self.inter_after.setdefault(l, []).append(k)
        first_inter = min(self.inter_after) if self.inter_after else None
        if c.encoder_type == "s2t":
            C = c.embed_dim
            assert C % c.heads == 0 and (C // c.heads) % 8 == 0 and C % 8 == 0 and c.input_dim % 8 == 0
            # Conv1dSubsampling filters: [filter]*(L-1) + [d], last one doubled for the GLU (subsampling.py:231-247,118-121)
            filt = [c.subsampling_filter] * (c.subsampling_layers - 1) + [C]
            if c.subsampling_type == "conv2d":
                self._declare_conv2d_stem(filt)
            else:
                self.sub_filters = [f if i < c.subsampling_layers - 1 else 2 * f for i, f in enumerate(filt)]
                cin = c.input_dim
                for i, cout in enumerate(self.sub_filters):
                    n = f"subsample.layers.{i}.0.weight"
                    S.declare(n, (cout, c.subsampling_kernel, cin))
                    self.conv_names.append(n)
                    S.declare(f"subsample.layers.{i}.0.bias", (cout,))
                    cin = cout // 2
            if c.encoder_embed_norm:
                S.declare("embed_ln.weight", (C,)); S.declare("embed_ln.bias", (C,))
            for l in range(c.layers):
                # conv-module activation of S2TTransformerEncoderLayer = args.activation_fn (quirk A3)
                spec = _LayerSpec(f"layers.{l}.", 0, l, l, C, c.ffn_dim, c.heads, _ACT[c.activation_fn], c.dropout)
                self.layer_specs.append(spec)
                self._declare_layer(spec)
                # inter-CTC norms sit after their layer; the shared heads after the first inter layer, so the backward
                # pass completes them (and reports the range final) as early as possible
                for k in self.inter_after.get(l + 1, []):
                    n = f"{self.heads[k].norm}{l + 1}"
                    S.declare(n + ".weight", (C,)); S.declare(n + ".bias", (C,))
                if first_inter == l + 1:
                    self._declare_heads(C)
            self.out_dim = C
        else:
            cin, g = c.input_dim, 0
            for i, (d, k) in enumerate(zip(c.pds_dims, c.pds_kernels)):
                assert d % c.pds_heads[i] == 0 and (d // c.pds_heads[i]) % 8 == 0 and cin % 8 == 0
                n = f"downsampling{i + 1}.conv.0.weight"
                S.declare(n, (d, k, cin))
                self.conv_names.append(n)
                S.declare(f"downsampling{i + 1}.conv.0.bias", (d,))
                if c.pds_embed_norm:
                    S.declare(f"downsampling{i + 1}.norm.weight", (d,)); S.declare(f"downsampling{i + 1}.norm.bias", (d,))
                for j in range(c.pds_layers[i]):
                    # PDSTransformerEncoderLayer: conv-module activation = encoder_activation_fn (pds_layer.py:67,104)
                    spec = _LayerSpec(f"stage{i + 1}.{j}.", i, j, g, d, d * c.pds_ffn_ratios[i], c.pds_heads[i],
                                      _ACT[c.encoder_activation_fn], c.dropout)
                    self.layer_specs.append(spec)
                    self._declare_layer(spec)
                    g += 1
                cin = d
            self.out_dim = c.pds_dims[-1]
        C = self.out_dim
        S.declare("layer_norm.weight", (C,)); S.declare("layer_norm.bias", (C,))
        if first_inter is None:
            self._declare_heads(C)
        S.finalize()
        self.embed_scale = 1.0 if (c.no_scale_embedding or c.encoder_type == "pds") else math.sqrt(C)  # quirk A17
        self.pos_tables: Dict[int, torch.Tensor] = {}
        self.step_counter = torch.zeros(1, device=self.device, dtype=torch.int64)
        self.plans: Dict[Tuple[int, int, int, int], Plan] = {}      # insertion order = recency (plan())
        self.plan_cache_bytes = int(float(os.environ.get("S2T_PLAN_CACHE_GB", "48")) * 2 ** 30)
        self.plan_builds = self.plan_evictions = 0
        self.training = True
        self.grad_ready_hook = None  # callable(offset): store.grad[offset:] is final (dp.BucketedAllReduce.on_ready)
        # fork/join streams: leaf work of the backward pass (weight / bias gradients) and independent attention
        # branches run beside the critical dgrad chain; captured into the CUDA graph as parallel branches
        self.multi_stream = os.environ.get("S2T_MULTI_STREAM", "1") == "1"
        # attn.cu: one launch per layer and direction for selfattn with d_k 64 and T' <= 256 (S2T_FUSED_ATTN=0: the unfused
        # GEMM / softmax / GEMM form, which also serves every other shape)
        self.fused_attention = os.environ.get("S2T_FUSED_ATTN", "1") == "1"
        self._side = [torch.cuda.Stream(device=self.device) for _ in range(3)]
        # SM partition of the backward pass: (chain GEMM CTAs, weight-gradient GEMM CTAs, weight-gradient split-K policy);
        # measured on B200 (148 SMs): 100 + 48 without split-K, 2.14 -> 1.94 ms/step (ops.set_sm_shares)
        self.sm_shares = tuple(int(v) for v in os.environ.get("S2T_SM_SHARES", "100,48,-1").split(","))
        self.specaug = None      # specaugment.SpecAugmentConfig: masks given to set_inputs are applied after CMVN (training)
        self.fwd_share = 0       # SM bound of forward GEMMs (0 = whole GPU; MicroBatchEngine splits the GPU between groups)
        self._ln_stream = torch.cuda.Stream(device=self.device)   # LayerNorm dgamma/dbeta reductions (_ln_bwd)
        self._ln_pending = False
        self._ln_jobs: list = []
        self._forked: set = set()
        self.lag_join = os.environ.get("S2T_LAG_JOIN", "1") == "1"
        self._lag_events: list = []
        self._lag_stage, self._lag_bufs = None, ()
        self.dp = None               # dp.BucketedAllReduce when data parallel
        self.ctc_masked_loss = False  # criterion option: imputer loss on the inter terms under the curriculum (_loss)
        self.collective_free_warmup = False   # capture(): warm-up step without collectives (per-rank plans, see capture)
        self.init_parameters()

    def attn_is_fused(self, St: "_Stage") -> bool:
        """attn.cu serves selfattn with d_k 64 and T' <= 256.  One CTA (226 KB of shared memory: one per SM) per (utterance,
        head, 128-query tile): beyond two waves of CTAs (B * H * 2 > 296; measured at B = 40: 5.95 vs 5.63 ms/step) the
        batched GEMM form, which tiles over all (b, h) at once, is the faster one."""
        return (self.fused_attention and self.cfg.attn_type == "selfattn" and St.dk == 64 and St.T <= 256
                and St.B * St.H * 2 <= 296)

    def _pos_table(self, dim: int) -> torch.Tensor:
        if dim not in self.pos_tables:
            self.pos_tables[dim] = sinusoidal_table(self.cfg.max_source_positions + 2, dim).to(self.device)
        return self.pos_tables[dim]

    @property
    def pos_table(self) -> torch.Tensor:
        return self._pos_table(self.out_dim)

    def _declare_heads(self, C: int) -> None:
        """CTC / XCTC projections (bias padded to the logits' row stride) and the adapters' (V, C) embedding matrices."""
        S = self.store
        self.padded_bias: Dict[str, int] = {}
        for hs in self.heads:
            S.declare(hs.name + ".ctc_projection.weight", (hs.V, C))
            S.declare(hs.name + ".ctc_projection.bias", (hs.ldv,))
            self.padded_bias[hs.name + ".ctc_projection.bias"] = hs.V
            if hs.pae and hs.inter_layers:
                S.declare(hs.pae + ".embed_adapter.weight", (hs.V, C))

    def _declare_layer(self, sp: _LayerSpec) -> None:
        """Parameters of one layer in forward order (names: SURVEY.md §8b)."""
        c, S, p, C, F = self.cfg, self.store, sp.prefix, sp.C, sp.F
        ln = lambda n: (S.declare(p + n + ".weight", (C,)), S.declare(p + n + ".bias", (C,)))
        sp.first = p + ("macaron_norm" if c.macaron_style else "self_attn_layer_norm") + ".weight"
        if c.macaron_style:
            ln("macaron_norm")
            S.declare(p + "macaron_ffn.w_1.weight", (F, C)); S.declare(p + "macaron_ffn.w_1.bias", (F,))
            S.declare(p + "macaron_ffn.w_2.weight", (C, F)); S.declare(p + "macaron_ffn.w_2.bias", (C,))
        ln("self_attn_layer_norm")
        a = p + "self_attn."
        q, k, v, o = ("linear_q", "linear_k", "linear_v", "linear_out") if c.attn_type == "rel_pos" else \
                     ("q_proj", "k_proj", "v_proj", "out_proj")
        S.declare(a + q + ".weight", (C, C))
        S.declare(a + k + ".weight", (C, C), pack=True)
        S.declare(a + v + ".weight", (C, C), pack=True)
        S.declare(a + q + ".bias", (C,))
        S.declare(a + k + ".bias", (C,), pack=True)
        S.declare(a + v + ".bias", (C,), pack=True)
        S.declare(a + o + ".weight", (C, C)); S.declare(a + o + ".bias", (C,))
        if c.attn_type == "rel_pos":
            S.declare(a + "linear_pos.weight", (C, C))
            S.declare(a + "pos_bias_u", (C,)); S.declare(a + "pos_bias_v", (C,), pack=True)
            self.ref_shapes[a + "pos_bias_u"] = self.ref_shapes[a + "pos_bias_v"] = (sp.H, C // sp.H)
        if c.use_cnn_module:
            K = c.cnn_module_kernel
            ln("conv_norm")
            m = p + "conv_module."
            S.declare(m + "pointwise_conv1.weight", (2 * C, C)); self.ref_shapes[m + "pointwise_conv1.weight"] = (2 * C, C, 1)
            S.declare(m + "depthwise_conv.weight", (C, K)); self.ref_shapes[m + "depthwise_conv.weight"] = (C, 1, K)
            S.declare(m + "norm.weight", (C,)); S.declare(m + "norm.bias", (C,))
            S.declare(m + "pointwise_conv2.weight", (C, C)); self.ref_shapes[m + "pointwise_conv2.weight"] = (C, C, 1)
            self.buffers[m + "norm.running_mean"] = torch.zeros(C, device=self.device, dtype=F32)
            self.buffers[m + "norm.running_var"] = torch.ones(C, device=self.device, dtype=F32)
            self.buffers[m + "norm.num_batches_tracked"] = torch.zeros((), device=self.device, dtype=torch.int64)
        ln("ffn_norm")
        S.declare(p + "ffn.w_1.weight", (F, C)); S.declare(p + "ffn.w_1.bias", (F,))
        S.declare(p + "ffn.w_2.weight", (C, F)); S.declare(p + "ffn.w_2.bias", (C,))
        if c.use_cnn_module:
            ln("final_norm")

    def _declare_conv2d_stem(self, filt: List[int]) -> None:
        """Conv2dSubsampling parameters (subsampling.py:178-205): per layer Conv2d(k, stride) [+ BatchNorm2d]; then
        Linear(D' * C_last, C_last).  Conv weights are kept as GEMM operands [Cout][Kpad], column (j*k + i)*Cin + c for
        time tap j, feature tap i (reference layout [Cout][Cin][i][j]); the Linear's input index is d*C + c (channels-last
        activations) instead of the reference's c*D' + d."""
        c, S = self.cfg, self.store
        if c.subsampling_norm not in ("none", "batch2d") or c.subsampling_activation not in ("swish", "relu", "none"):
            raise NotImplementedError("engine: conv2d subsampling supports norm none|batch2d and activation swish|relu|none")
        if c.subsampling_norm == "none" and c.subsampling_activation != "none":
            raise NotImplementedError("engine: conv2d subsampling activation without batch2d norm")
        k = c.subsampling_kernel
        cin, D = 1, c.input_dim
        self.sub2d = []
        for i, cout in enumerate(filt):
            assert cout % 8 == 0
            K = k * k * cin
            Kpad = max(16, (K + 7) // 8 * 8)
            n = f"subsample.layers.{i}.0.weight"
            S.declare(n, (cout, Kpad))
            self.conv2d_names[n] = (cout, cin, k, Kpad)
            S.declare(f"subsample.layers.{i}.0.bias", (cout,))
            if c.subsampling_norm == "batch2d":
                b = f"subsample.layers.{i}.1.1."
                S.declare(b + "weight", (cout,)); S.declare(b + "bias", (cout,))
                self.buffers[b + "running_mean"] = torch.zeros(cout, device=self.device, dtype=F32)
                self.buffers[b + "running_var"] = torch.ones(cout, device=self.device, dtype=F32)
                self.buffers[b + "num_batches_tracked"] = torch.zeros((), device=self.device, dtype=torch.int64)
            self.sub2d.append((cin, cout, Kpad))
            cin = cout
            D = (D - 1) // 2 if (k, c.subsampling_stride) == (3, 2) else (D - k) // c.subsampling_stride + 1
        self.sub2d_dim = D
        C = filt[-1]
        S.declare("subsample.linear.weight", (C, D * C))
        self.lin_perm["subsample.linear.weight"] = (C, C, D)
        S.declare("subsample.linear.bias", (C,))

    def _to_internal(self, name: str, t: torch.Tensor) -> torch.Tensor:
        if name in self.conv2d_names:
            cout, cin, k, Kpad = self.conv2d_names[name]
            if tuple(t.shape) != (cout, cin, k, k):
                raise ValueError(f"{name}: shape {tuple(t.shape)} != {(cout, cin, k, k)}")
            full = torch.zeros((cout, Kpad), dtype=t.dtype)
            full[:, : k * k * cin] = t.permute(0, 3, 2, 1).reshape(cout, k * k * cin)     # [Cout][j][i][Cin]
            return full
        if name in self.lin_perm:
            C, ch, D = self.lin_perm[name]
            return t.reshape(C, ch, D).permute(0, 2, 1).reshape(C, D * ch)
        return t

    def _to_reference(self, name: str, t: torch.Tensor) -> torch.Tensor:
        if name in self.conv2d_names:
            cout, cin, k, Kpad = self.conv2d_names[name]
            return t[:, : k * k * cin].reshape(cout, k, k, cin).permute(0, 3, 2, 1).contiguous()
        if name in self.lin_perm:
            C, ch, D = self.lin_perm[name]
            return t.reshape(C, D, ch).permute(0, 2, 1).reshape(C, ch * D)
        return t

    # ------------------------------------------------------------------------------------------
    # parameters
    # ------------------------------------------------------------------------------------------
    def init_parameters(self, seed: int = 0) -> None:
        """Reference initialisers: Conv1d / Linear default (kaiming-uniform a=sqrt(5) == U(+-1/sqrt(fan_in))),
        xavier-uniform fairseq attention projections (multihead_attention.py:95-113, q/k/v gain 1/sqrt(2)) and
        pos_bias_u/v (espnet_multihead_attention.py:284-288), LayerNorm / BatchNorm 1/0, CTC projection
        N(0, C^-0.5) with zero bias (modules/speech_to_text/ctc.py:27-33)."""
        g = torch.Generator().manual_seed(seed)
        S = self.store
        for name in S.names():
            shape = S.shape(name)
            if _NORM_W.search(name) or (name.startswith("subsample.layers.") and name.endswith(".1.1.weight")):
                t = torch.ones(shape)
            elif _NORM_B.search(name) or (name.startswith("subsample.layers.") and name.endswith(".1.1.bias")):
                t = torch.zeros(shape)
            elif name in self.conv2d_names or name[:-4] + "weight" in self.conv2d_names:
                cout, cin, k, Kpad = self.conv2d_names[name if name in self.conv2d_names else name[:-4] + "weight"]
                bound = 1.0 / math.sqrt(cin * k * k)
                if name.endswith("bias"):
                    t = (torch.rand(shape, generator=g) * 2 - 1) * bound
                else:
                    t = self._to_internal(name, (torch.rand((cout, cin, k, k), generator=g) * 2 - 1) * bound)
            elif name.endswith(("ctc_projection.weight", "embed_adapter.weight")):
                t = torch.randn(shape, generator=g) * shape[1] ** -0.5
            elif name.endswith("ctc_projection.bias"):
                t = torch.zeros(shape)
            elif name.endswith("_proj.weight"):
                gain = 1.0 if "out_proj" in name else 2 ** -0.5
                bound = gain * math.sqrt(6.0 / (shape[0] + shape[1]))
                t = (torch.rand(shape, generator=g) * 2 - 1) * bound
            elif name.endswith("out_proj.bias"):
                t = torch.zeros(shape)
            elif name.endswith(("pos_bias_u", "pos_bias_v")):
                h, dk = self.ref_shapes[name]
                t = (torch.rand(shape, generator=g) * 2 - 1) * math.sqrt(6.0 / (h + dk))
            else:
                if name.endswith("bias"):
                    w = S.shape(name[:-4] + "weight")
                    fan_in = int(math.prod(w[1:]))
                else:
                    fan_in = int(math.prod(shape[1:]))
                bound = 1.0 / math.sqrt(fan_in)
                t = (torch.rand(shape, generator=g) * 2 - 1) * bound
            S.p(name).copy_(t.to(BF16))

    def _ref_to_store(self, name: str, t: torch.Tensor) -> torch.Tensor:
        """One reference-layout tensor (fp32, CPU) in the layout and shape of its slot in the flat buffers."""
        S = self.store
        t = self._to_internal(name, t)
        if name in self.conv_names:
            t = t.permute(0, 2, 1).contiguous()               # [Cout][Cin][k] -> [Cout][k][Cin]
        if name in self.ref_shapes:
            if tuple(t.shape) != self.ref_shapes[name]:
                raise ValueError(f"{name}: shape {tuple(t.shape)} != {self.ref_shapes[name]}")
            t = t.reshape(S.shape(name))
        if name in self.padded_bias:
            full = torch.zeros(S.shape(name))
            full[: t.numel()] = t
            t = full
        if tuple(t.shape) != S.shape(name):
            raise ValueError(f"{name}: shape {tuple(t.shape)} != {S.shape(name)}")
        return t

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        """Load a reference-layout state_dict (names of SURVEY.md §8b)."""
        S = self.store
        for name in S.names():
            if name not in sd:
                raise KeyError(f"state_dict is missing {name}")
            S.p(name).copy_(self._ref_to_store(name, sd[name].detach().to(F32).cpu()).to(BF16))
        for name, buf in self.buffers.items():
            if name in sd:
                buf.copy_(sd[name].to(buf.dtype))

    def export_flat(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        """Any buffer laid out like the flat parameter buffer (gradients, fp32 master weights, optimizer moments) cut into
        the reference's named tensors, in the reference's shapes and layouts (fp32 copies)."""
        S = self.store
        if flat.numel() != S.total:
            raise ValueError(f"buffer of {flat.numel()} elements does not match the parameter layout ({S.total})")
        out = {}
        for name in S.names():
            off = S.offset(name)
            t = flat[off:off + S.numel(name)].view(S.shape(name)).detach().float()
            if name in self.conv_names:
                t = t.permute(0, 2, 1).contiguous()
            if name in self.ref_shapes:
                t = t.reshape(self.ref_shapes[name])
            if name in self.padded_bias:
                t = t[: self.padded_bias[name]]
            out[name] = self._to_reference(name, t)
        return out

    def import_flat(self, tensors: Dict[str, torch.Tensor], flat: torch.Tensor) -> None:
        """Inverse of export_flat: reference-layout tensors written into their slots of ``flat`` (alignment gaps and
        padding untouched)."""
        S = self.store
        if flat.numel() != S.total:
            raise ValueError(f"buffer of {flat.numel()} elements does not match the parameter layout ({S.total})")
        for name in S.names():
            if name not in tensors:
                raise KeyError(f"missing {name}")
            off = S.offset(name)
            t = self._ref_to_store(name, tensors[name].detach().to(F32).cpu())
            flat[off:off + S.numel(name)].view(S.shape(name)).copy_(t.to(flat.dtype))

    def _export(self, grad: bool) -> Dict[str, torch.Tensor]:
        return self.export_flat(self.store.grad if grad else self.store.flat)

    def state_dict(self) -> Dict[str, torch.Tensor]:
        sd = self._export(False)
        for name, buf in self.buffers.items():
            sd[name] = buf.detach().clone()
        return sd

    def grads(self) -> Dict[str, torch.Tensor]:
        """Gradients in the reference layout (fp32 copies)."""
        return self._export(True)

    def num_params(self) -> int:
        return self.store.num_params() - sum(hs.ldv - hs.V for hs in self.heads)

    def train(self, mode: bool = True):
        self.training = mode
        return self

    def eval(self):
        return self.train(False)

    # ------------------------------------------------------------------------------------------
    # plans
    # ------------------------------------------------------------------------------------------
    def plan(self, B: int, T: int, smax: int = 0, wave_len: int = 0) -> Plan:
        """The plan of one (batch, frames, max target length, samples) bucket, from an LRU cache bounded in bytes
        (``plan_cache_bytes``, default 48 GiB or S2T_PLAN_CACHE_GB): under max-tokens batching every new (B, T) would
        otherwise pin 0.5-2.3 GB of activations plus a captured graph for ever.  Callers quantise (B, T) with
        ``batching.BucketSet`` so that the cache sees a small fixed set of keys."""
        key = (B, T, smax, wave_len)
        P = self.plans.pop(key, None)
        if P is None:
            before = torch.cuda.memory_allocated(self.device)
            P = Plan(self, B, T, smax, wave_len)
            P.nbytes = max(0, torch.cuda.memory_allocated(self.device) - before)
            self.plan_builds += 1
            # make room: drop least-recently-used plans (their graphs and buffers) until the new one fits the cap
            total = P.nbytes + sum(q.nbytes for q in self.plans.values())
            while self.plans and total > self.plan_cache_bytes:
                old_key = next(iter(self.plans))
                old = self.plans.pop(old_key)
                total -= old.nbytes
                old.graph = None
                self.plan_evictions += 1
                del old
        self.plans[key] = P          # most recently used last
        return P

    def plan_cache_stats(self) -> Dict[str, int]:
        return {"plans": len(self.plans), "bytes": sum(q.nbytes for q in self.plans.values()),
                "cap_bytes": self.plan_cache_bytes, "builds": self.plan_builds, "evictions": self.plan_evictions}

    def _drop(self, site: int, p: Optional[float] = None):
        """Dropout spec of call site ``site`` (unique stream id per site; mask regenerable in backward)."""
        p = (self.cfg.dropout if p is None else p) if self.training else 0.0
        if p <= 0.0:
            return None
        return (p, self.SEED, site, self.step_counter)

    # ------------------------------------------------------------------------------------------
    # forward: stems
    # ------------------------------------------------------------------------------------------
    def _features(self, P: Plan, dst: torch.Tensor, pad: int) -> None:
        """Features (or waveform -> fbank -> utterance CMVN) into a zero-framed bf16 conv input [B][T+2pad][D];
        frames >= src_len zero filled (collater padding, speech_to_text_dataset.py:267-285)."""
        self._fbank(P)
        self._cmvn(P, dst, pad)

    def _fbank(self, P: Plan) -> None:
        """Waveform input: log-mel features into P.fbank_raw; the kernel also writes the frame counts (P.src_len)."""
        if P.wave is not None:
            mel_w, rng, win = frontend._tables(str(self.device), self.cfg.input_dim)
            ops.fbank(P.wave, P.num_samples, mel_w, rng, win, P.T, out=P.fbank_raw, frames=P.src_len)

    def _cmvn(self, P: Plan, dst: torch.Tensor, pad: int) -> None:
        c = self.cfg
        src = P.fbank_raw if P.wave is not None else P.feats_in
        cm = 1 if P.wave is not None else 0
        ops._call("s2t_cmvn_utt", ops._p(src), src.stride(1), src.stride(0), ops._p(dst) + 2 * pad * c.input_dim, 1,
                  dst.stride(1), dst.stride(0), ops._p(P.src_len), P.B, P.T, c.input_dim, cm, cm,
                  ops._p(P.cmvn_stats), ops._stream())
        if self.specaug is not None and self.training and (P.sa_freq is not None or P.sa_time is not None):
            # SpecAugment on the CMVN output (the reference's _train transform order: utterance_cmvn, specaugment)
            ops.specaugment_(dst.view(P.B, -1, c.input_dim)[:, pad:pad + P.T], P.src_len, P.sa_freq, P.sa_time, self.specaug.mask_value,
                             c.input_dim, P.sa_means)

    def _embed_block_fwd(self, P: Plan) -> None:
        """embed_ln -> x * embed_scale (+ sinusoidal positions for abs attention) -> dropout (s2t_transformer.py:1768-1794)."""
        c, S = self.cfg, self.store
        St = P.stages[0]
        src = P.x0
        if c.encoder_embed_norm:
            ops.layernorm_fwd(P.x0, S.p("embed_ln.weight"), S.p("embed_ln.bias"), out=P.e_ln, stats=P.e_st)
            src = P.e_ln
        C = c.embed_dim
        table = self._pos_table(C) if c.attn_type == "selfattn" else None       # rel_pos: x is not modified (A5)
        ops.embed_fwd(src.view(P.B, St.T, C), St.lens, table, self.embed_scale, self._drop(1),
                      out=St.x[0].view(P.B, St.T, C))

    def _stem_conv2d_fwd(self, P: Plan) -> None:
        """Conv2dSubsampling (subsampling.py:207-228): per layer Conv2d(k, stride, no padding) [+ BatchNorm2d over
        (B, D', T') in training mode: batch statistics include the zero-padded frames] + activation; lengths (len-1)//2
        per layer (:218, quirk A9); (B, C, D', T') -> (T', B, C*D') -> Linear.  The reference's host-side assert
        T' == max(lengths) (:222, quirk A16) is not evaluated.  Then the encoder zeroes padded frames (:1765)."""
        c, S = self.cfg, self.store
        k, st = c.subsampling_kernel, c.subsampling_stride
        self._features(P, P.x_in, 0)
        ln = P.src_len
        for i in range(len(P.c2)):
            torch.div(ln - 1, 2, rounding_mode="floor", out=P.sub_len[i])
            ln = P.sub_len[i]
        x = P.x_in.view(P.B, P.T, c.input_dim, 1)
        act = _ACT.get(c.subsampling_activation, ops.ACT_NONE)
        for i, L in enumerate(P.c2):
            n = f"subsample.layers.{i}."
            ops.im2col2d(x, k, st, L.Kpad, out=L.col)
            ops.linear(L.col, S.p(n + "0.weight"), S.p(n + "0.bias"), out=L.pre)
            if c.subsampling_norm == "batch2d":
                b, bufs = n + "1.1.", self.buffers
                if self.training:
                    ops.colstats(L.pre, out=L.bn_part)
                ops.batchnorm_fwd(L.pre, S.p(b + "weight"), S.p(b + "bias"), partials=L.bn_part if self.training else None,
                                  running_mean=bufs[b + "running_mean"], running_var=bufs[b + "running_var"],
                                  num_batches_tracked=bufs[b + "num_batches_tracked"], save_stats=L.bn_save, out=L.y,
                                  act=act, training=self.training)
                y = L.y
            else:
                y = L.pre
            x = y.view(P.B, L.Tout, L.Dout, L.cout)
        St = P.stages[0]
        L = P.c2[-1]
        ops.linear(y.view(P.B * L.Tout, L.Dout * L.cout), S.p("subsample.linear.weight"), S.p("subsample.linear.bias"),
                   out=P.x0)
        ops.mask_rows_(P.x0.view(P.B, St.T, c.embed_dim), St.lens)
        self._embed_block_fwd(P)

    def _stem_conv2d_bwd(self, P: Plan, dy: torch.Tensor) -> None:
        """Backward of the embed block and Conv2dSubsampling: Linear, then per layer BatchNorm2d + activation,
        bias = column sums, weight = dY^T patches, input = col2im(dY W)."""
        c, S = self.cfg, self.store
        St = P.stages[0]
        C = c.embed_dim
        k, st = c.subsampling_kernel, c.subsampling_stride
        t = St.take()
        ops.embed_bwd(dy.view(P.B, St.T, C), self.embed_scale, self._drop(1), out=t.view(P.B, St.T, C))
        St.give(dy)
        if c.encoder_embed_norm:
            o = St.take()
            self._ln_bwd_plain(St, t, P.x0, "embed_ln", P.e_st, o)
            St.give(t)
            self._ready("embed_ln.weight")
            t = o
        ops.mask_rows_(t.view(P.B, St.T, C), St.lens)                   # padded frames were zeroed in forward
        L = P.c2[-1]
        y_last = (L.y if c.subsampling_norm == "batch2d" else L.pre).view(P.B * L.Tout, L.Dout * L.cout)
        with self._fork(0):
            ops.linear_wgrad(t, y_last, out=S.g("subsample.linear.weight"), colsum=S.g("subsample.linear.bias"))
        ops.linear_dgrad(t, S.p("subsample.linear.weight"), out=L.g_y.view(P.B * L.Tout, L.Dout * L.cout))
        self._ready("subsample.linear.weight")
        St.give(t)
        act = _ACT.get(c.subsampling_activation, ops.ACT_NONE)
        for i in reversed(range(len(P.c2))):
            L = P.c2[i]
            n = f"subsample.layers.{i}."
            if c.subsampling_norm == "batch2d":
                b = n + "1.1."
                ops.batchnorm_bwd(L.g_y, L.pre, L.bn_save, S.p(b + "weight"), S.p(b + "bias"), S.g(b + "weight"),
                                  S.g(b + "bias"), out=L.g_pre, partials=L.bn_bwd_part, act=act, training=self.training)
                g = L.g_pre
            else:
                g = L.g_y
            with self._fork(0):
                ops.linear_wgrad(g, L.col, out=S.g(n + "0.weight"), colsum=S.g(n + "0.bias"))
            if i > 0:
                ops.linear_dgrad(g, S.p(n + "0.weight"), out=L.g_col)
                prev = P.c2[i - 1]
                ops.col2im2d(L.g_col, P.B, L.Tin, L.Din, L.cin, k, st, out=prev.g_y.view(P.B, L.Tin, L.Din, L.cin))
            self._ready(n + "0.weight")

    def _stem_s2t_fwd(self, P: Plan) -> None:
        """Conv1dSubsampling (subsampling.py:145-159): [Conv1d(k, stride, pad) -> GLU] per layer as an
        implicit GEMM whose A rows overlap in the zero-framed channels-last input; lengths (len-1)//2+1 (:155);
        the last GLU also applies the encoder's padding mask (s2t_transformer.py:1765).  Then the embed block:
        embed_ln -> x * embed_scale (+ sinusoidal positions for abs attention) -> dropout (:1768-1794)."""
        c, S = self.cfg, self.store
        k, st, n = c.subsampling_kernel, c.subsampling_stride, c.subsampling_layers
        self._fbank(P)
        with self._fork(0):          # integer length bookkeeping beside the feature kernels (joined before the last GLU)
            ln = P.src_len
            for i in range(n):
                torch.div(ln + (2 * P.pad - k), st, rounding_mode="floor", out=P.sub_len[i])
                P.sub_len[i] += 1
                ln = P.sub_len[i]
        self._cmvn(P, P.framed[0], P.pad)
        for i in range(n):
            xin = P.framed[i]
            cin, cout, t_out = xin.shape[2], self.sub_filters[i], P.t_sub[i + 1]
            w = S.p(f"subsample.layers.{i}.0.weight").view(cout, k * cin)
            ops.gemm_raw(M=t_out, N=cout, K=k * cin, a=xin, lda=st * cin, a_bs=(xin.stride(0), 0),
                         b=w, ldb=k * cin, out=P.pre[i], ldc=cout, c_bs=(t_out * cout, 0), batch=(P.B, 1),
                         bias=S.p(f"subsample.layers.{i}.0.bias"))
            if i < n - 1:
                nxt = P.framed[i + 1]
                ops.glu_fwd(P.pre[i], out=nxt, lengths=P.batch_sub[i], T=t_out, ldy=cout // 2, y_bs=nxt.stride(0),
                            y_off=P.pad * (cout // 2))
            else:
                self._join()
                ops.glu_fwd(P.pre[i], out=P.x0, lengths=P.sub_len[i], T=t_out)
        self._embed_block_fwd(P)

    def _stem_pds_fwd(self, P: Plan, i: int) -> None:
        """PDS stage stem (pdss2t_transformer.py:1088-1121): Downsampling = zero the padded input frames, Conv1d(k,
        stride=ratio, pad) [+ LayerNorm], zero the padded output frames, lens = floor((len-1)/stride+1) (:53-144);
        then += positions (abs) and dropout (stage 0: dropout, later stages: pds_dropout)."""
        c, S = self.cfg, self.store
        D, St = P.pds[i], P.stages[i]
        if i == 0:
            self._features(P, D.framed, D.pad)
            D.in_len.copy_(P.src_len)
        else:
            prev = P.stages[i - 1]
            D.in_len.copy_(prev.lens)
            ops.frame_rows(prev.x[-1].view(P.B, prev.T, prev.C), prev.lens, D.framed, D.pad)
        torch.div(D.in_len - 1, D.stride, rounding_mode="floor", out=St.lens)
        St.lens += 1
        n = f"downsampling{i + 1}."
        w = S.p(n + "conv.0.weight").view(D.cout, D.k * D.cin)
        ops.gemm_raw(M=D.t_out, N=D.cout, K=D.k * D.cin, a=D.framed, lda=D.stride * D.cin,
                     a_bs=(D.framed.stride(0), 0), b=w, ldb=D.k * D.cin, out=D.conv, ldc=D.cout,
                     c_bs=(D.t_out * D.cout, 0), batch=(P.B, 1), bias=S.p(n + "conv.0.bias"))
        src = D.conv
        if c.pds_embed_norm:
            ops.layernorm_fwd(D.conv, S.p(n + "norm.weight"), S.p(n + "norm.bias"), out=D.ln, stats=D.st)
            src = D.ln
        table = self._pos_table(D.cout) if (c.attn_type == "selfattn" and c.pds_position_embed[i]) else None
        p = c.dropout if i == 0 else c.pds_dropout
        ops.embed_fwd(src.view(P.B, St.T, St.C), St.lens, table, 1.0, self._drop(2 + i, p),
                      out=St.x[0].view(P.B, St.T, St.C))

    # ------------------------------------------------------------------------------------------
    # forward: layer blocks
    # ------------------------------------------------------------------------------------------
    def _ffn_fwd(self, St: _Stage, x: torch.Tensor, out: torch.Tensor, ln: torch.Tensor, st: torch.Tensor,
                 h: Optional[torch.Tensor], f: torch.Tensor, norm: str, ffn: str, scale: float, site: int) -> None:
        """out = x + scale * dropout(w_2(dropout(act(w_1(LN(x))))))  (FeedForwardModule, s2t_transformer_layer.py:26-66;
        scale 0.5 in macaron style, :270-276,303-309)."""
        S = self.store
        ops.layernorm_fwd(x, S.p(norm + ".weight"), S.p(norm + ".bias"), out=ln, stats=st)
        if self.ffn_act == ops.ACT_RELU:
            ops.linear(ln, S.p(ffn + ".w_1.weight"), S.p(ffn + ".w_1.bias"), out=f, act=ops.ACT_RELU,
                       dropout=self._drop(site))
        else:   # keep the pre-activation for the backward pass
            ops.linear(ln, S.p(ffn + ".w_1.weight"), S.p(ffn + ".w_1.bias"), out=h)
            ops.act_dropout(h, self.ffn_act, self._drop(site), out=f)
        ops.linear(f, S.p(ffn + ".w_2.weight"), S.p(ffn + ".w_2.bias"), out=out, dropout=self._drop(site + 1),
                   out_scale=scale, residual=x, ldr=St.C)

    def _attn_fwd(self, P: Plan, St: _Stage, sp: _LayerSpec, L: _Bufs, x: torch.Tensor) -> None:
        """x1 = x + dropout(attention(LN(x))).
        selfattn: fairseq MultiheadAttention (multihead_attention.py:246-265 projections, q * d_k^-0.5; :383-395 key
        padding; fp32 softmax + dropout on the probabilities; :372-416 bmm + out_proj).
        rel_pos: RelPositionMultiHeadedAttention (espnet_multihead_attention.py:313-356): AC = (q+u) k^T,
        BD = rel_shift((q+v) p^T), softmax((AC+BD)/sqrt(d_k)) with -inf key mask and the +-1e4 clamp (:119-129)."""
        c, S = self.cfg, self.store
        B, T, C, H, dk, ld = P.B, St.T, St.C, St.H, St.dk, St.ld
        a = sp.prefix + "self_attn."
        rel = c.attn_type == "rel_pos"
        q, o = ("linear_q", "linear_out") if rel else ("q_proj", "out_proj")
        if rel:   # the position projection does not depend on x: beside LayerNorm + QKV
            with self._fork(2):
                ops.linear(St.pos_emb, S.p(a + "linear_pos.weight"), out=L.pp)
        ops.layernorm_fwd(x, S.p(sp.prefix + "self_attn_layer_norm.weight"), S.p(sp.prefix + "self_attn_layer_norm.bias"),
                          out=L.ln1, stats=L.st1)
        ops.linear(L.ln1, S.fused(a + q + ".weight", 3 * C, C), S.fused(a + q + ".bias", 3 * C, 0), out=L.qkv)
        bs_s = (T * ld, H * T * ld)
        dspec = self._drop(sp.site + 0, sp.p_drop)
        if L.probs is None:
            # one launch: S = q k^T (TMEM) -> masked fp32 softmax -> dropout -> P v; keeps the row log-sum-exp  (attn.cu)
            ops.attn_fused_fwd(L.qkv, C, St.lens, B, H, T, dk, lse=L.lse, ctx=L.ctx, scale=dk ** -0.5, dropout=dspec)
            ops.linear(L.ctx, S.p(a + o + ".weight"), S.p(a + o + ".bias"), out=L.x1,
                       dropout=self._drop(sp.site + 1, sp.p_drop), residual=x, ldr=C)
            return
        if not rel:
            ops.gemm_raw(M=T, N=T, K=dk, a=L.qkv, lda=3 * C, a_bs=(dk, T * 3 * C), b=L.qkv, ldb=3 * C, b_off=C,
                         b_bs=(dk, T * 3 * C), out=St.scores, ldc=ld, c_bs=bs_s, batch=(H, B), alpha=dk ** -0.5)
            ops.attn_softmax_fwd(St.scores, St.lens, H, T, probs=L.probs, probs_drop=L.probs_d if dspec else None,
                                 dropout=dspec)
        else:
            ops.relpos_qbias_fwd(L.qkv, S.p(a + "pos_bias_u"), S.p(a + "pos_bias_v"), C, out=L.quv)
            self._join()                                                  # L.pp
            with self._fork(2):                                           # BD beside AC
                ops.gemm_raw(M=T, N=2 * T - 1, K=dk, a=L.quv, a_off=C, lda=2 * C, a_bs=(dk, T * 2 * C), b=L.pp, ldb=C,
                             b_bs=(dk, 0), out=St.bd, ldc=St.ldb, c_bs=(B * T * St.ldb, T * St.ldb), batch=(H, B))
            ops.gemm_raw(M=T, N=T, K=dk, a=L.quv, lda=2 * C, a_bs=(dk, T * 2 * C), b=L.qkv, ldb=3 * C, b_off=C,
                         b_bs=(dk, T * 3 * C), out=St.scores, ldc=ld, c_bs=bs_s, batch=(H, B))
            self._join()
            ops.attn_softmax_fwd(St.scores, St.lens, H, T, probs=L.probs, probs_drop=L.probs_d if dspec else None,
                                 bias=St.bd.view(H * B, T, St.ldb), bias_mode=2, bias_head_major=True,
                                 scale=dk ** -0.5, clamp=1e4, dropout=dspec)
        pd = L.probs_d if dspec else L.probs
        ops.gemm_raw(M=T, N=dk, K=T, a=pd, lda=ld, a_bs=bs_s, b=L.qkv, ldb=3 * C, b_off=2 * C,
                     b_mn=True, b_bs=(dk, T * 3 * C), out=L.ctx, ldc=C, c_bs=(dk, T * C), batch=(H, B))
        ops.linear(L.ctx, S.p(a + o + ".weight"), S.p(a + o + ".bias"), out=L.x1,
                   dropout=self._drop(sp.site + 1, sp.p_drop), residual=x, ldr=C)

    def _conv_fwd(self, P: Plan, St: _Stage, sp: _LayerSpec, L: _Bufs, x: torch.Tensor, out: torch.Tensor) -> None:
        """out = x + dropout(ConvolutionModule(LN(x)))  (convolution.py:76-120; s2t_transformer_layer.py:290-300):
        zero padded frames, pointwise Conv1d(C,2C) + GLU, depthwise Conv1d(k), BatchNorm1d + activation,
        pointwise Conv1d(C,C).  (Padded output rows are left unmasked: they only ever feed padded rows.)"""
        S = self.store
        p, m = sp.prefix, sp.prefix + "conv_module."
        B, T, C = P.B, St.T, St.C
        ops.layernorm_fwd(x, S.p(p + "conv_norm.weight"), S.p(p + "conv_norm.bias"), out=L.c_ln, stats=L.c_st)
        ops.mask_rows_(L.c_ln.view(B, T, C), St.lens)
        ops.linear(L.c_ln, S.p(m + "pointwise_conv1.weight"), out=L.pw1)
        ops.glu_fwd(L.pw1, out=L.glu)
        ops.dwconv1d_fwd(L.glu.view(B, T, C), S.p(m + "depthwise_conv.weight"), out=L.dw.view(B, T, C),
                         bn_partials=St.bn_part if self.training else None)
        bufs = self.buffers
        ops.batchnorm_fwd(L.dw, S.p(m + "norm.weight"), S.p(m + "norm.bias"),
                          partials=St.bn_part if self.training else None,
                          running_mean=bufs[m + "norm.running_mean"], running_var=bufs[m + "norm.running_var"],
                          num_batches_tracked=bufs[m + "norm.num_batches_tracked"], save_stats=L.bn_save, out=L.bn,
                          act=sp.conv_act, training=self.training)
        ops.linear(L.bn, S.p(m + "pointwise_conv2.weight"), out=out, dropout=self._drop(sp.site + 6, sp.p_drop),
                   residual=x, ldr=C)

    def _layer_fwd(self, P: Plan, g: int) -> None:
        """S2TTransformerEncoderLayer.forward / PDSTransformerEncoderLayer.forward, pre-norm
        (s2t_transformer_layer.py:267-317, pds_layer.py:263-359)."""
        c = self.cfg
        sp, L = self.layer_specs[g], P.layers[g]
        St = P.stages[sp.stage]
        x, y = St.x[sp.index], (St.y[sp.index] if sp.stage == 0 else St.x[sp.index + 1])
        p = sp.prefix
        if c.layer_padding_mask and c.encoder_type == "s2t":
            ops.mask_rows_(x.view(P.B, St.T, St.C), St.lens)                  # s2t_transformer.py:1828-1836
        s = 0.5 if c.macaron_style else 1.0
        if c.macaron_style:
            self._ffn_fwd(St, x, L.xa, L.m_ln, L.m_st, L.m_h, L.m_f, p + "macaron_norm", p + "macaron_ffn", s, sp.site + 2)
            x = L.xa
        self._attn_fwd(P, St, sp, L, x)
        x = L.x1
        if c.use_cnn_module:
            self._conv_fwd(P, St, sp, L, x, L.xc)
            x = L.xc
            self._ffn_fwd(St, x, L.xd, L.ln2, L.st2, L.h, L.f, p + "ffn_norm", p + "ffn", s, sp.site + 4)
            S = self.store
            ops.layernorm_fwd(L.xd, S.p(p + "final_norm.weight"), S.p(p + "final_norm.bias"), out=y, stats=L.f_st)
        else:
            self._ffn_fwd(St, x, y, L.ln2, L.st2, L.h, L.f, p + "ffn_norm", p + "ffn", s, sp.site + 4)

    def _ctc_proj_fwd(self, hs: _HeadSpec, x: torch.Tensor, xd: torch.Tensor, logits: torch.Tensor, site: int) -> None:
        """CTC.forward (modules/speech_to_text/ctc.py:55-66, quirk A11): dropout -> Linear(C, V)."""
        S = self.store
        spec = self._drop(site)
        src = x
        if spec:
            ops.dropout(x, spec, out=xd)
            src = xd
        ops.linear(src, S.p(hs.name + ".ctc_projection.weight"), S.p(hs.name + ".ctc_projection.bias"), out=logits)

    def _head_fwd(self, P: Plan) -> None:
        """Final LayerNorm (s2t_transformer.py:2112) and the top CTC / XCTC heads (:2124-2133)."""
        S = self.store
        ops.layernorm_fwd(P.stages[-1].x[-1], S.p("layer_norm.weight"), S.p("layer_norm.bias"), out=P.enc_out,
                          stats=P.enc_st)
        for hs, Hd in zip(self.heads, P.heads):
            self._ctc_proj_fwd(hs, P.enc_out, Hd.drop, Hd.logits, hs.site)

    def _inter_fwd(self, P: Plan, l: int, k: int) -> None:
        """Inter-CTC block after layer l with head k (s2t_transformer.py:1881-1946 / 2043-2103): norm_x = LN_l(x);
        logit = CTC(norm_x); with a prediction-aware adapter (adapter.py:189-217,264-266,296-297, inter_league):
        x <- (x | norm_x) + softmax(logit) @ embed_adapter.weight."""
        c, S = self.cfg, self.store
        hs, I = self.heads[k], P.inter[(l, k)]
        n = f"{hs.norm}{l}"
        ops.layernorm_fwd(I.xin, S.p(n + ".weight"), S.p(n + ".bias"), out=I.nx, stats=I.st)
        self._ctc_proj_fwd(hs, I.nx, I.nxd, I.logits, hs.site + 1 + hs.inter_layers.index(l))
        ops.row_lse(I.logits, hs.V, out=I.lse)
        if hs.pae:
            ops.softmax_rows_fwd(I.logits, I.lse, hs.V, out=I.dist)
            if k in P.gt and self.training:
                # curriculum: masked frames feed the adapter the (smoothed) one-hot of their aligned ground-truth token
                # instead of the model's prediction (adapter.py:244-262)
                tok, msk, _ = P.gt[k]
                ops.rows_truth_mix_(I.dist, tok, msk, hs.V, self.cfg.pae_gt_smooth and hs.name == "xctc")   # sic: only the XCTC adapter gets the smoothing option (s2t_transformer.py:1066 vs 1233)
            E = S.p(hs.pae + ".embed_adapter.weight")
            C = E.shape[1]
            res = I.xin if c.pae_unnorm_input else I.nx
            ops.gemm_raw(M=P.R, N=C, K=hs.V, a=I.dist, lda=hs.ldv, b=E, ldb=C, b_mn=True, out=I.xout, ldc=C,
                         residual=res, ldr=C)

    def _forward(self, P: Plan) -> None:
        prev = ops.set_sm_shares(self.fwd_share, 0, 0)
        try:
            self._forward_body(P)
        finally:
            ops.set_sm_shares(*prev)

    def _forward_body(self, P: Plan) -> None:
        g = 0
        for i, St in enumerate(P.stages):
            if self.cfg.encoder_type == "s2t" and self.cfg.subsampling_type == "conv2d":
                self._stem_conv2d_fwd(P)
            elif self.cfg.encoder_type == "s2t":
                self._stem_s2t_fwd(P)
            else:
                self._stem_pds_fwd(P, i)
            n = self.cfg.layers if self.cfg.encoder_type == "s2t" else self.cfg.pds_layers[i]
            for _ in range(n):
                self._layer_fwd(P, g)
                g += 1
                if i == 0:
                    for k in self.inter_after.get(g, []):
                        self._inter_fwd(P, g, k)
        self._head_fwd(P)

    def _loss(self, P: Plan, need_grad: bool) -> None:
        """fp32 log-softmax + CTCLoss(blank, 'none', zero_infinity=True).sum() per loss term and its gradient w.r.t. the
        logits (criterions/ctc.py:435-474); loss = ctc_w ctc + xctc_w xctc + inter_ctc_w mean_l(inter) +
        inter_xctc_w mean_l(inter_x) (:1004-1015) — every term's weight is folded into its gradient."""
        c = self.cfg
        B, T = P.B, P.T2
        for k, (hs, Hd) in enumerate(zip(self.heads, P.heads)):
            terms = [(Hd.logits, Hd.dlogits, Hd.lse, Hd.loss_idx, True)]
            terms += [(P.inter[(l, k)].logits, P.inter[(l, k)].dlogits, P.inter[(l, k)].lse, P.inter[(l, k)].loss_idx, False)
                      for l in hs.inter_layers]
            # masked ("imputer") loss when the criterion asked for it and the curriculum supplies the forced states: the
            # encoder attaches force_emit to the inter logits AND to the top logits (s2t_transformer.py:1936, 2136-2140;
            # criterions/ctc.py:455-465, 586-627, 646-666)
            fe = P.gt[k][2].view(B, T) if (self.ctc_masked_loss and k in P.gt and self.training) else None
            for logits, dlogits, lse, idx, need_lse in terms:
                if need_lse:
                    ops.row_lse(logits, hs.V, out=lse)
                w = P.loss_weights[idx]
                ops.ctc_loss(logits.view(B, T, hs.ldv), P.tg[hs.tgt], P.out_len, P.tl[hs.tgt], hs.V, c.blank,
                             need_grad=need_grad, grad_scale=w, grad_scale_dev=P.batch_weight, lse=lse,
                             dlogits=dlogits.view(B, T, hs.ldv) if need_grad else None, workspace=P.ctc_ws,
                             nll=P.nll_all[idx], force_emit=fe)
        ops.weighted_sum(P.nll_all, P.loss_w, P.loss)

    # ------------------------------------------------------------------------------------------
    # backward
    # ------------------------------------------------------------------------------------------
    def _fork(self, i: int = 0):
        """Context: the enclosed launches run on side stream ``i`` after everything issued so far on the main stream.
        Buffers they read must not be overwritten by the main stream before the next ``_join``."""
        if not self.multi_stream:
            return contextlib.nullcontext()
        side = self._side[i]
        side.wait_stream(torch.cuda.current_stream())
        self._forked.add(i)
        return torch.cuda.stream(side)

    def _join(self, only: Optional[Tuple[int, ...]] = None) -> None:
        """The main stream waits for the side-stream work forked since the last join (``only``: just those side streams —
        a data dependency on one of them must not wait for the weight gradients on another)."""
        main = torch.cuda.current_stream()
        for i in sorted(self._forked):
            if only is None or i in only:
                main.wait_stream(self._side[i])
        self._forked = set() if only is None else {i for i in self._forked if i not in only}

    def _end_block(self, St: _Stage, *bufs: torch.Tensor) -> None:
        """End of a residual block's backward.  Its side-stream work (weight / bias gradients) only READS buffers of the
        block, so the dgrad chain need not wait for it: the join is LAGGED by one block — here the chain waits for the side
        work of the PREVIOUS block (normally long finished), and this block's pool buffers stay quarantined until the
        next call.  Stage scratch read by side work (g_f, g_qkv, g_2c) is not rewritten before two blocks later (g_f
        alternates between two buffers: macaron FFN blocks are adjacent).  S2T_LAG_JOIN=0: join at once."""
        # (large batches: the chain's kernels fill the GPU by themselves and overlapping more side work only adds
        # contention — measured at R = 10000 rows: 5.95 vs 5.66 ms/step — so the lag applies up to 4096 rows)
        if not (self.multi_stream and self.lag_join and St.R <= 4096):
            self._join()
            St.give(*bufs)
            return
        main = torch.cuda.current_stream()
        for ev in self._lag_events:
            main.wait_event(ev)
        if self._lag_stage is not None:
            self._lag_stage.give(*self._lag_bufs)
        self._lag_events = []
        for i in sorted(self._forked):
            ev = torch.cuda.Event()
            ev.record(self._side[i])
            self._lag_events.append(ev)
        self._forked.clear()
        self._lag_stage, self._lag_bufs = St, bufs

    def _join_all(self) -> None:
        """Immediate join of everything outstanding (lagged block joins included)."""
        self._join()
        main = torch.cuda.current_stream()
        for ev in self._lag_events:
            main.wait_event(ev)
        self._lag_events = []
        if self._lag_stage is not None:
            self._lag_stage.give(*self._lag_bufs)
        self._lag_stage, self._lag_bufs = None, ()

    def _ready(self, first_name: str) -> None:
        """Everything at or after ``first_name`` in the flat gradient buffer is final."""
        if self.grad_ready_hook is None:
            return                       # single GPU: side work and the LayerNorm reductions are collected at the end of backward
        off = self.store.offset(first_name)
        if self.dp is not None and self.grad_ready_hook == self.dp.on_ready and not self.dp.would_ship(off):
            return                       # no bucket completes here: nothing has to be final yet
        self._join_all()
        self._join_ln()
        self.grad_ready_hook(off)

    def _join_ln(self) -> None:
        self._flush_ln()
        if self._ln_pending:
            torch.cuda.current_stream().wait_stream(self._ln_stream)
            self._ln_pending = False

    def _ln_bwd(self, St: _Stage, dy: torch.Tensor, x: torch.Tensor, norm: str, stats: torch.Tensor, *,
                dx_residual: Optional[torch.Tensor] = None, nxt=None):
        """LayerNorm backward on the main stream, its dgamma/dbeta reduction on a side stream.  ``nxt`` = (dropout spec,
        scale) of the residual branch whose backward consumes the result next: the kernel then also emits
        scale * dropout(dx) (second return value), so that branch needs no separate dropout pass."""
        S = self.store
        part = St.ln_part(norm, x.shape[0], x.shape[1], self.device)
        dx, dd, spec, scale = St.take(), None, None, 1.0
        if nxt is not None and (nxt[0] is not None or nxt[1] != 1.0):
            dd, (spec, scale) = St.take(), nxt
        ops.layernorm_bwd(dy, x, S.p(norm + ".weight"), stats, None, None, dx_residual=dx_residual, out=dx,
                          partials=part, out_drop=dd, drop=spec, drop_scale=scale, reduce=False)
        self._ln_jobs.append((part, x.shape[0], x.shape[1], S.g(norm + ".weight"), S.g(norm + ".bias")))
        return dx, dd

    def _ln_bwd_plain(self, St: _Stage, dy: torch.Tensor, x: torch.Tensor, norm: str, stats: torch.Tensor,
                      out: torch.Tensor) -> None:
        """LayerNorm backward of a stem norm (embed_ln, downsampling norms) into ``out``; reduction deferred likewise."""
        S = self.store
        part = St.ln_part(norm, x.shape[0], x.shape[1], self.device)
        ops.layernorm_bwd(dy, x, S.p(norm + ".weight"), stats, None, None, out=out, partials=part, reduce=False)
        self._ln_jobs.append((part, x.shape[0], x.shape[1], S.g(norm + ".weight"), S.g(norm + ".bias")))

    def _flush_ln(self) -> None:
        """Run the collected dgamma / dbeta reductions as one batched launch, beside the dgrad chain."""
        if not self._ln_jobs:
            return
        jobs, self._ln_jobs = self._ln_jobs, []
        if self.multi_stream:
            ln = self._ln_stream
            ln.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(ln):
                ops.layernorm_bwd_reduce_batched(jobs)
            self._ln_pending = True
        else:
            ops.layernorm_bwd_reduce_batched(jobs)

    def _branch_grad(self, dy: torch.Tensor, spec, scale: float, out: torch.Tensor) -> torch.Tensor:
        """Gradient entering a residual branch y = x + scale * dropout(branch): scale * mask * dy / (1-p)."""
        if spec is None and scale == 1.0:
            return dy
        return ops.dropout(dy, spec, out=out, scale=scale)

    def _acc(self, name: str) -> bool:
        """True if the gradient of ``name`` was already written in this backward pass (shared heads accumulate)."""
        seen = name in self._touched
        self._touched.add(name)
        return seen

    def _ctc_proj_bwd(self, P: Plan, hs: _HeadSpec, dlogits: torch.Tensor, src: torch.Tensor, ws: torch.Tensor,
                      out: torch.Tensor, residual: Optional[torch.Tensor]) -> None:
        """Backward of Linear(C, V): bias / weight gradients (accumulated across the uses of a shared head) and
        out = dlogits @ W (+ residual)."""
        S = self.store
        wn, bn = hs.name + ".ctc_projection.weight", hs.name + ".ctc_projection.bias"
        C = out.shape[1]
        gw = S.g(wn)
        dl = dlogits[:, : hs.V] if hs.V != hs.ldv else dlogits
        with self._fork(0):      # weight gradient; the bias gradient (column sums of dlogits) rides on its A tiles
            acc_b = self._acc(bn)
            if self._acc(wn):
                ops.linear_wgrad(dl, src, out=gw, residual=gw, ldr=C, colsum=S.g(bn), colsum_accumulate=acc_b)
            else:
                ops.linear_wgrad(dl, src, out=gw, colsum=S.g(bn), colsum_accumulate=acc_b)
        ops.gemm_raw(M=P.R, N=C, K=hs.V, a=dlogits, lda=hs.ldv, b=S.p(wn), ldb=C, b_mn=True, out=out, ldc=C,
                     residual=residual, ldr=C if residual is not None else 0)

    def _head_bwd(self, P: Plan, extra: Optional[torch.Tensor] = None):
        """Backward of _head_fwd.  ``extra``: an external gradient w.r.t. the encoder output (P.enc_out layout) added to
        the heads' contribution — the AR decoder's gradient when the encoder is trained through ``backward_external``."""
        S = self.store
        St = P.stages[-1]
        self._touched = set()
        d = None
        for hs, Hd in zip(self.heads, P.heads):
            spec = self._drop(hs.site)
            t0 = St.take()
            self._ctc_proj_bwd(P, hs, Hd.dlogits, Hd.drop if spec else P.enc_out, Hd.colsum_ws, t0, None)
            if spec:
                t1 = St.take()
                ops.dropout(t0, spec, out=t1)
                St.give(t0)
                t0 = t1
            if d is None:
                d = t0
            else:
                ops.add_(d, t0)
                St.give(t0)
        if extra is not None:
            if d is None:
                d = St.take()
                d.copy_(extra)
            else:
                ops.add_(d, extra)
        dx, dd = self._ln_bwd(St, d, St.x[-1], "layer_norm", P.enc_st, nxt=self._next_branch(len(self.layer_specs)))
        self._ready("layer_norm.weight")
        St.give(d)
        return dx, dd

    def _inter_bwd(self, P: Plan, l: int, k: int, dy: torch.Tensor, nxt=None):
        """Backward of _inter_fwd: ``dy`` = gradient w.r.t. the block output (== input when the head has no adapter);
        adds the block's own CTC-loss gradient (already in I.dlogits) and returns the gradient w.r.t. the block input."""
        c, S = self.cfg, self.store
        hs, Hd, I = self.heads[k], P.heads[k], P.inter[(l, k)]
        St = P.stages[0]
        C = St.C
        if hs.pae:
            en = hs.pae + ".embed_adapter.weight"
            E, gE = S.p(en), S.g(en)
            # d dist = dy E^T ; dE (+)= dist^T dy ; dlogits += softmax'(d dist)
            ops.gemm_raw(M=P.R, N=hs.V, K=C, a=dy, lda=C, b=E, ldb=C, out=Hd.ddist, ldc=hs.ldv)
            dist = I.dist[:, : hs.V] if hs.V != hs.ldv else I.dist
            with self._fork(2):
                if self._acc(en):
                    ops.linear_wgrad(dist, dy, out=gE, residual=gE, ldr=C)
                else:
                    ops.linear_wgrad(dist, dy, out=gE)
            if k in P.gt and self.training:
                ops.rows_zero_masked_(Hd.ddist, P.gt[k][1], hs.V)       # replaced frames: no gradient to their logits
            ops.softmax_rows_bwd(I.dist, Hd.ddist, I.dlogits, hs.V, accumulate=True)
        spec = self._drop(hs.site + 1 + hs.inter_layers.index(l))
        t0 = St.take()
        # d norm_x = dropout'(dlogits W) (+ dy when the adapter adds to the normalised input)
        to_norm = dy if (hs.pae and not c.pae_unnorm_input) else None
        self._ctc_proj_bwd(P, hs, I.dlogits, I.nxd if spec else I.nx, Hd.colsum_ws, t0, None if spec else to_norm)
        if spec:
            t1 = St.take()
            ops.dropout(t0, spec, out=t1)
            St.give(t0)
            t0 = t1
            if to_norm is not None:
                ops.add_(t0, to_norm)
        # gradient that bypasses the norm: dy flows to x unless the adapter consumed norm_x instead of x
        skip = None if to_norm is not None else dy
        dx, dd = self._ln_bwd(St, t0, I.xin, f"{hs.norm}{l}", I.st, dx_residual=skip, nxt=nxt)
        self._join()
        St.give(t0, dy)
        return dx, dd

    def _ffn_bwd(self, St: _Stage, dy: torch.Tensor, dyd: Optional[torch.Tensor], x: torch.Tensor, ln: torch.Tensor,
                 st: torch.Tensor, h: Optional[torch.Tensor], f: torch.Tensor, norm: str, ffn: str, scale: float,
                 site: int, nxt=None):
        """Backward of _ffn_fwd; consumes ``dy`` (gradient w.r.t. the block output), returns the gradient w.r.t. x.
        Critical chain: dropout' -> dgrad w_2 -> dgrad w_1 -> LayerNorm'; weight / bias gradients on side streams."""
        S = self.store
        t0, t1 = St.take(), St.take()
        St._g_f_i ^= 1
        g_f = St.g_fs[St._g_f_i]
        d2 = dyd if dyd is not None else self._branch_grad(dy, self._drop(site + 1), scale, t0)
        with self._fork(0):
            ops.linear_wgrad(d2, f, out=S.g(ffn + ".w_2.weight"), colsum=S.g(ffn + ".w_2.bias"))
        if self.ffn_act == ops.ACT_RELU:
            # f = dropout(relu(h)) is positive exactly where h > 0 AND the element was kept, so the gate on f covers the
            # dropout mask too: no Philox regeneration in this epilogue, only the 1/(1-p) scale (folded into alpha)
            spec = self._drop(site)
            ops.linear_dgrad(d2, S.p(ffn + ".w_2.weight"), out=g_f, gate=f, gate_mode=ops.GATE_POSITIVE, ldg=St.F,
                             alpha=1.0 / (1.0 - spec[0]) if spec else 1.0)
        else:
            ops.linear_dgrad(d2, S.p(ffn + ".w_2.weight"), out=g_f, gate=h, gate_mode=ops.GATE_SWISH_GRAD, ldg=St.F,
                             dropout=self._drop(site))
        with self._fork(0):
            ops.linear_wgrad(g_f, ln, out=S.g(ffn + ".w_1.weight"), colsum=S.g(ffn + ".w_1.bias"))
        ops.linear_dgrad(g_f, S.p(ffn + ".w_1.weight"), out=t1)
        dx, dd = self._ln_bwd(St, t1, x, norm, st, dx_residual=dy, nxt=nxt)
        self._end_block(St, t0, t1, dy, dyd)
        return dx, dd

    def _conv_bwd(self, P: Plan, St: _Stage, sp: _LayerSpec, L: _Bufs, dy: torch.Tensor, dyd: Optional[torch.Tensor],
                  x: torch.Tensor, nxt=None):
        S = self.store
        p, m = sp.prefix, sp.prefix + "conv_module."
        B, T, C = P.B, St.T, St.C
        t0, t1, t2 = St.take(), St.take(), St.take()
        d = dyd if dyd is not None else self._branch_grad(dy, self._drop(sp.site + 6, sp.p_drop), 1.0, t0)
        with self._fork(0):
            ops.linear_wgrad(d, L.bn, out=S.g(m + "pointwise_conv2.weight"))
        ops.linear_dgrad(d, S.p(m + "pointwise_conv2.weight"), out=t1)                     # d act(bn)
        ops.batchnorm_bwd(t1, L.dw, L.bn_save, S.p(m + "norm.weight"), S.p(m + "norm.bias"), S.g(m + "norm.weight"),
                          S.g(m + "norm.bias"), out=t2, partials=St.bn_bwd_part, act=sp.conv_act,
                          training=self.training)                                            # d dw
        ops.dwconv1d_bwd(L.glu.view(B, T, C), S.p(m + "depthwise_conv.weight"), t2.view(B, T, C),
                         S.g(m + "depthwise_conv.weight"), out=t1.view(B, T, C), workspace=St.dw_ws)   # d glu
        ops.glu_bwd(L.pw1, t1, out=St.g_2c)
        with self._fork(0):
            ops.linear_wgrad(St.g_2c, L.c_ln, out=S.g(m + "pointwise_conv1.weight"))
        ops.linear_dgrad(St.g_2c, S.p(m + "pointwise_conv1.weight"), out=t2)
        ops.mask_rows_(t2.view(B, T, C), St.lens)
        dx, dd = self._ln_bwd(St, t2, x, p + "conv_norm", L.c_st, dx_residual=dy, nxt=nxt)
        self._end_block(St, t0, t1, t2, dy, dyd)
        return dx, dd

    def _attn_bwd(self, P: Plan, St: _Stage, sp: _LayerSpec, L: _Bufs, dy: torch.Tensor, dyd: Optional[torch.Tensor],
                  x: torch.Tensor, nxt=None):
        """Backward of _attn_fwd.  Critical chain: dropout' -> dgrad out -> dP -> softmax' -> dQ -> dgrad qkv -> LayerNorm';
        dV, dK, the position branch and all weight / bias gradients run on side streams."""
        c, S = self.cfg, self.store
        B, T, C, H, dk, ld = P.B, St.T, St.C, St.H, St.dk, St.ld
        a = sp.prefix + "self_attn."
        rel = c.attn_type == "rel_pos"
        q, o = ("linear_q", "linear_out") if rel else ("q_proj", "out_proj")
        t0, t1, dctx = St.take(), St.take(), St.take()
        do = dyd if dyd is not None else self._branch_grad(dy, self._drop(sp.site + 1, sp.p_drop), 1.0, t0)
        with self._fork(0):
            ops.linear_wgrad(do, L.ctx, out=S.g(a + o + ".weight"), colsum=S.g(a + o + ".bias"))
        ops.linear_dgrad(do, S.p(a + o + ".weight"), out=dctx)
        dspec = self._drop(sp.site + 0, sp.p_drop)
        g_qkv = St.g_qkv
        if L.probs is None:
            # one launch: P recomputed from the saved log-sum-exp, dS, dQ / dK / dV  (attn.cu)
            ops.attn_fused_bwd(L.qkv, C, L.ctx, dctx, L.lse, St.lens, B, H, T, dk, out=g_qkv, scale=dk ** -0.5,
                               dropout=dspec)
            return self._attn_bwd_tail(P, St, sp, L, dy, dyd, x, nxt, t0, t1, dctx)
        pd = L.probs_d if dspec else L.probs
        bs_x, bs_qkv, bs_s = (dk, T * C), (dk, T * 3 * C), (T * ld, H * T * ld)
        # dV = P_drop^T dctx (side) ; dP_drop = dctx V^T (main)
        with self._fork(2):
            ops.gemm_raw(M=T, N=dk, K=T, a=pd, lda=ld, a_mn=True, a_bs=bs_s, b=dctx, ldb=C, b_mn=True, b_bs=bs_x,
                         out=g_qkv, c_off=2 * C, ldc=3 * C, c_bs=bs_qkv, batch=(H, B))
        ops.gemm_raw(M=T, N=T, K=dk, a=dctx, lda=C, a_bs=bs_x, b=L.qkv, ldb=3 * C, b_off=2 * C, b_bs=bs_qkv,
                     out=St.g_p, ldc=ld, c_bs=bs_s, batch=(H, B))
        if not rel:
            ops.attn_softmax_bwd(L.probs, St.g_p, St.lens, H, T, out=St.g_s, dropout=dspec)
            # dK = s dS^T Q (side) ; dQ = s dS K (main)
            with self._fork(2):
                ops.gemm_raw(M=T, N=dk, K=T, a=St.g_s, lda=ld, a_mn=True, a_bs=bs_s, b=L.qkv, ldb=3 * C, b_off=0,
                             b_mn=True, b_bs=bs_qkv, out=g_qkv, c_off=C, ldc=3 * C, c_bs=bs_qkv, batch=(H, B),
                             alpha=dk ** -0.5)
            ops.gemm_raw(M=T, N=dk, K=T, a=St.g_s, lda=ld, a_bs=bs_s, b=L.qkv, ldb=3 * C, b_off=C, b_mn=True,
                         b_bs=bs_qkv, out=g_qkv, c_off=0, ldc=3 * C, c_bs=bs_qkv, batch=(H, B), alpha=dk ** -0.5)
        else:
            ldb, nrel = St.ldb, 2 * T - 1
            bs_quv, bs_bd = (dk, T * 2 * C), (B * T * ldb, T * ldb)
            ops.attn_softmax_bwd(L.probs, St.g_p, St.lens, H, T, out=St.g_s, dbias=St.g_bd.view(H * B, T, ldb),
                                 bias_head_major=True, scale=dk ** -0.5, dropout=dspec)
            # side 2: dK = dS^T (q+u) ; dp[h] = sum_b dBD[h,b]^T (q+v)[b,h] (one GEMM per head over K = B*T: head-major
            # dBD) and the linear_pos weight gradient
            with self._fork(2):
                ops.gemm_raw(M=T, N=dk, K=T, a=St.g_s, lda=ld, a_mn=True, a_bs=bs_s, b=L.quv, ldb=2 * C, b_off=0,
                             b_mn=True, b_bs=bs_quv, out=g_qkv, c_off=C, ldc=3 * C, c_bs=bs_qkv, batch=(H, B))
                ops.gemm_raw(M=nrel, N=dk, K=B * T, a=St.g_bd, lda=ldb, a_mn=True, a_bs=(B * T * ldb, 0), b=L.quv,
                             ldb=2 * C, b_off=C, b_mn=True, b_bs=(dk, 0), out=St.g_pp, ldc=C, c_bs=(dk, 0), batch=(H, 1))
                ops.linear_wgrad(St.g_pp, St.pos_emb, out=S.g(a + "linear_pos.weight"))
            # main: d(q+u) = dS K ; d(q+v) = dBD p ; dq = d(q+u) + d(q+v)
            ops.gemm_raw(M=T, N=dk, K=T, a=St.g_s, lda=ld, a_bs=bs_s, b=L.qkv, ldb=3 * C, b_off=C, b_mn=True,
                         b_bs=bs_qkv, out=St.g_quv, c_off=0, ldc=2 * C, c_bs=bs_quv, batch=(H, B))
            ops.gemm_raw(M=T, N=dk, K=nrel, a=St.g_bd, lda=ldb, a_bs=bs_bd, b=L.pp, ldb=C, b_mn=True, b_bs=(dk, 0),
                         out=St.g_quv, c_off=C, ldc=2 * C, c_bs=bs_quv, batch=(H, B))
            with self._fork(1):
                ops.colsum(St.g_quv, S.fused(a + "pos_bias_u", 2 * C, 0, grad=True), workspace=St.colsum_ws)
            ops.relpos_qbias_bwd(St.g_quv, g_qkv, C)
        return self._attn_bwd_tail(P, St, sp, L, dy, dyd, x, nxt, t0, t1, dctx)

    def _attn_bwd_tail(self, P, St, sp, L, dy, dyd, x, nxt, t0, t1, dctx):
        """Backward of the fused q/k/v projection and the attention LayerNorm (g_qkv holds dq | dk | dv)."""
        c, S = self.cfg, self.store
        C = St.C
        a = sp.prefix + "self_attn."
        q = "linear_q" if c.attn_type == "rel_pos" else "q_proj"
        g_qkv = St.g_qkv
        self._join(only=(1, 2))                           # g_qkv is complete (unfused path: dV / dK came from side stream 2)
        with self._fork(0):
            ops.linear_wgrad(g_qkv, L.ln1, out=S.fused(a + q + ".weight", 3 * C, C, grad=True),
                             colsum=S.fused(a + q + ".bias", 3 * C, 0, grad=True))
        ops.linear_dgrad(g_qkv, S.fused(a + q + ".weight", 3 * C, C), out=t1)
        dx, dd = self._ln_bwd(St, t1, x, sp.prefix + "self_attn_layer_norm", L.st1, dx_residual=dy, nxt=nxt)
        self._end_block(St, t0, t1, dctx, dy, dyd)
        return dx, dd

    def _branch_spec(self, g: int, block: str):
        """(dropout spec, scale) of the residual branch ``block`` of layer g: y = x + scale * dropout(branch(x))."""
        c, sp = self.cfg, self.layer_specs[g]
        s = 0.5 if c.macaron_style else 1.0
        return {"ffn": (self._drop(sp.site + 5), s), "conv": (self._drop(sp.site + 6, sp.p_drop), 1.0),
                "attn": (self._drop(sp.site + 1, sp.p_drop), 1.0), "macaron": (self._drop(sp.site + 3), s)}[block]

    def _next_branch(self, n_done: int):
        """Branch spec of the first block the backward pass meets below layer index ``n_done`` (= the LAST forward block
        of layer n_done - 1), or None when something else sits in between (inter-CTC block, final_norm, stage stem,
        layer padding mask)."""
        c = self.cfg
        g = n_done - 1
        if g < 0 or n_done in self.inter_after or c.use_cnn_module:
            return None
        if n_done < len(self.layer_specs) and self.layer_specs[n_done].stage != self.layer_specs[g].stage:
            return None
        return self._branch_spec(g, "ffn")

    def _layer_bwd(self, P: Plan, g: int, dy: torch.Tensor, dyd: Optional[torch.Tensor] = None):
        """Backward of _layer_fwd.  ``dy`` = dL/d(layer output) in a pool buffer of the stage (``dyd``: its pre-masked
        copy for the FFN branch, when the producer emitted one); returns (dL/d(input), pre-masked copy or None)."""
        c, S = self.cfg, self.store
        sp, L = self.layer_specs[g], P.layers[g]
        St = P.stages[sp.stage]
        x_in = St.x[sp.index]
        p = sp.prefix
        s = 0.5 if c.macaron_style else 1.0
        x_attn = L.xa if c.macaron_style else x_in
        mask_out = c.layer_padding_mask and c.encoder_type == "s2t"
        # what consumes this layer's input gradient: the previous layer's FFN branch, unless something sits in between
        below = None if (mask_out or sp.index == 0) else self._next_branch(g)
        after_attn = self._branch_spec(g, "macaron") if c.macaron_style else below
        if c.use_cnn_module:
            d, dd = self._ln_bwd(St, dy, L.xd, p + "final_norm", L.f_st, nxt=self._branch_spec(g, "ffn"))
            St.give(dy, dyd)
            d, dd = self._ffn_bwd(St, d, dd, L.xc, L.ln2, L.st2, L.h, L.f, p + "ffn_norm", p + "ffn", s, sp.site + 4,
                                  nxt=self._branch_spec(g, "conv"))
            d, dd = self._conv_bwd(P, St, sp, L, d, dd, L.x1, nxt=self._branch_spec(g, "attn"))
        else:
            d, dd = self._ffn_bwd(St, dy, dyd, L.x1, L.ln2, L.st2, L.h, L.f, p + "ffn_norm", p + "ffn", s, sp.site + 4,
                                  nxt=self._branch_spec(g, "attn"))
        d, dd = self._attn_bwd(P, St, sp, L, d, dd, x_attn, nxt=after_attn)
        if c.macaron_style:
            d, dd = self._ffn_bwd(St, d, dd, x_in, L.m_ln, L.m_st, L.m_h, L.m_f, p + "macaron_norm", p + "macaron_ffn", s,
                                  sp.site + 2, nxt=below)
        if mask_out:
            ops.mask_rows_(d.view(P.B, St.T, St.C), St.lens)
        self._ready(sp.first)
        return d, dd

    def _stem_s2t_bwd(self, P: Plan, dy: torch.Tensor) -> None:
        """Backward of the embed block and Conv1dSubsampling: GLU', bias = column sums, weight = dY^T im2col(X),
        input = col2im(dY W)."""
        c, S = self.cfg, self.store
        St = P.stages[0]
        C = c.embed_dim
        t = St.take()
        ops.embed_bwd(dy.view(P.B, St.T, C), self.embed_scale, self._drop(1), out=t.view(P.B, St.T, C))
        St.give(dy)
        if c.encoder_embed_norm:
            o = St.take()
            self._ln_bwd_plain(St, t, P.x0, "embed_ln", P.e_st, o)
            St.give(t)
            self._ready("embed_ln.weight")
            t = o
        k, st, n = c.subsampling_kernel, c.subsampling_stride, c.subsampling_layers
        dy = t
        for i in reversed(range(n)):
            xin = P.framed[i]
            cin, cout, t_out = xin.shape[2], self.sub_filters[i], P.t_sub[i + 1]
            if i == n - 1:
                ops.glu_bwd(P.pre[i], dy, out=P.g_pre[i], lengths=P.sub_len[i], T=t_out)
            else:
                ops.glu_bwd(P.pre[i], dy, out=P.g_pre[i], lengths=P.batch_sub[i], T=t_out)
            with self._fork(0):
                ops.im2col1d(xin, xin.stride(0), P.B, t_out, cin, k, st, out=P.xcol[i])
                ops.linear_wgrad(P.g_pre[i], P.xcol[i], out=S.g(f"subsample.layers.{i}.0.weight").view(cout, k * cin),
                                 colsum=S.g(f"subsample.layers.{i}.0.bias"))
            if i > 0:
                ops.linear_dgrad(P.g_pre[i], S.p(f"subsample.layers.{i}.0.weight").view(cout, k * cin), out=P.g_col[i])
                ops.col2im1d(P.g_col[i], P.B, P.t_sub[i], t_out, cin, k, st, P.pad, out=P.g_sub[i])
                dy = P.g_sub[i].view(P.B * P.t_sub[i], cin)
            self._ready(f"subsample.layers.{i}.0.weight")
        St.give(t)

    def _stem_pds_bwd(self, P: Plan, i: int, dy: torch.Tensor) -> Optional[torch.Tensor]:
        """Backward of _stem_pds_fwd; returns the gradient w.r.t. the previous stage's output (None for stage 0)."""
        c, S = self.cfg, self.store
        D, St = P.pds[i], P.stages[i]
        n = f"downsampling{i + 1}."
        t = St.take()
        p = c.dropout if i == 0 else c.pds_dropout
        ops.embed_bwd(dy.view(P.B, St.T, St.C), 1.0, self._drop(2 + i, p), out=t.view(P.B, St.T, St.C))
        ops.mask_rows_(t.view(P.B, St.T, St.C), St.lens)           # the stem zeroes padded output frames (:136-142)
        St.give(dy)
        if c.pds_embed_norm:
            o = St.take()
            self._ln_bwd_plain(St, t, D.conv, n + "norm", D.st, o)
            St.give(t)
            t = o
        w = S.p(n + "conv.0.weight").view(D.cout, D.k * D.cin)
        with self._fork(0):
            ops.im2col1d(D.framed, D.framed.stride(0), P.B, D.t_out, D.cin, D.k, D.stride, out=D.xcol)
            ops.linear_wgrad(t, D.xcol, out=S.g(n + "conv.0.weight").view(D.cout, D.k * D.cin),
                             colsum=S.g(n + "conv.0.bias"))
        out = None
        if i > 0:
            prev = P.stages[i - 1]
            ops.linear_dgrad(t, w, out=D.g_col)
            out = prev.take()
            g_in = out.view(P.B, D.t_in, D.cin)
            ops.col2im1d(D.g_col, P.B, D.t_in, D.t_out, D.cin, D.k, D.stride, D.pad, out=g_in)
            ops.mask_rows_(g_in, prev.lens)                         # the stem zeroes padded input frames (:107-116)
        self._ready(n + "conv.0.weight")
        St.give(t)
        return out

    def _backward(self, P: Plan, extra: Optional[torch.Tensor] = None) -> None:
        c = self.cfg
        self._ln_jobs = []
        self._lag_events, self._lag_stage, self._lag_bufs = [], None, ()
        prev = ops.set_sm_shares(*(self.sm_shares if self.multi_stream else (0, 0, 0)))
        try:
            self._backward_body(P, extra)
        finally:
            ops.set_sm_shares(*prev)

    def _backward_body(self, P: Plan, extra: Optional[torch.Tensor] = None) -> None:
        c = self.cfg
        d, dd = self._head_bwd(P, extra)
        g = len(self.layer_specs)
        for i in reversed(range(len(P.stages))):
            n = c.layers if c.encoder_type == "s2t" else c.pds_layers[i]
            for _ in range(n):
                if i == 0 and (g in self.inter_after):
                    blocks = self.inter_after[g]
                    for j, k in enumerate(reversed(blocks)):
                        last = j == len(blocks) - 1
                        nxt = None if (not last or c.use_cnn_module) else self._branch_spec(g - 1, "ffn")
                        P.stages[0].give(dd)
                        d, dd = self._inter_bwd(P, g, k, d, nxt=nxt)
                    self._ready(f"{self.heads[blocks[0]].norm}{g}.weight")
                g -= 1
                d, dd = self._layer_bwd(P, g, d, dd)
            P.stages[i].give(dd)
            dd = None
            self._flush_ln()             # the layers' LayerNorm reductions run beside the stem's backward
            if c.encoder_type == "s2t" and c.subsampling_type == "conv2d":
                self._stem_conv2d_bwd(P, d)
            elif c.encoder_type == "s2t":
                self._stem_s2t_bwd(P, d)
            else:
                d = self._stem_pds_bwd(P, i, d)
        self._join_all()
        self._join_ln()

    # ------------------------------------------------------------------------------------------
    # public API
    # ------------------------------------------------------------------------------------------
    @staticmethod
    def strip_targets(tokens: torch.Tensor, pad: int = 1, eos: int = 2) -> Tuple[torch.Tensor, torch.Tensor]:
        """CtcCriterion.get_targets_for_ctc_loss (criterions/ctc.py:516-540): drop pad and eos; returns the
        left-packed int32 [B, S] target matrix and int32 lengths (host-side integer plumbing)."""
        tokens = tokens.cpu()
        keep = (tokens != pad) & (tokens != eos)
        lens = keep.sum(-1).to(torch.int32)
        out = torch.zeros((tokens.shape[0], max(int(lens.max()), 1)), dtype=torch.int32)
        for b in range(tokens.shape[0]):
            out[b, : int(lens[b])] = tokens[b][keep[b]].to(torch.int32)
        return out, lens

    def set_inputs(self, P: Plan, *, feats: Optional[torch.Tensor] = None, src_lengths: Optional[torch.Tensor] = None,
                   wave: Optional[torch.Tensor] = None, num_samples: Optional[torch.Tensor] = None,
                   targets: Optional[torch.Tensor] = None, target_lengths: Optional[torch.Tensor] = None,
                   xtargets: Optional[torch.Tensor] = None, xtarget_lengths: Optional[torch.Tensor] = None,
                   sa_freq: Optional[torch.Tensor] = None, sa_time: Optional[torch.Tensor] = None,
                   weight: Optional[float] = None) -> None:
        """Copy one batch into the plan's static input buffers (async on the current stream; host tensors
        should be pinned).  feats fp32 [B, T, D] + src_lengths, or wave fp32 [B, N] (int16 scale) + num_samples.
        sa_freq / sa_time: SpecAugment masks of this batch (specaugment.draw_masks), used when ``engine.specaug`` is set;
        their shapes are fixed by the first call (static buffers of a captured step)."""
        nb = True
        if hasattr(P, "batch_sub"):     # Conv1dSubsampling: intermediate frame counts of THIS batch's padded length
            from .frontend import num_frames
            t = feats.shape[1] if feats is not None else (num_frames(wave.shape[1]) if wave is not None else P.T)
            if t != P.batch_T:
                P.batch_T = t
                c = self.cfg
                for v in P.batch_sub:
                    t = (t + 2 * P.pad - c.subsampling_kernel) // c.subsampling_stride + 1
                    v.fill_(t)
        # 0.0 = dummy batch (gradient contribution and logged loss exactly zero), 1.0 = real batch.  Set on EVERY call: a
        # plan that ran a dummy batch and is reused by a caller that does not pass ``weight`` is a real batch again.
        w = 1.0 if weight is None else float(weight)
        if w != P.batch_w_host:
            P.batch_weight.fill_(w)
            if hasattr(P, "loss_w0"):
                torch.mul(P.loss_w0, w, out=P.loss_w)
            P.batch_w_host = w
        for name, m in (("sa_freq", sa_freq), ("sa_time", sa_time)):
            if m is not None and m.shape[1] > 0:
                if getattr(P, name) is None:
                    setattr(P, name, torch.zeros(m.shape, dtype=torch.int32, device=self.device))
                getattr(P, name).copy_(m, non_blocking=nb)
        if wave is not None:
            if wave.shape[1] == P.wave.shape[1]:
                P.wave.copy_(wave, non_blocking=nb)
            else:       # shorter than the plan's sample bucket: samples >= num_samples are never read
                P.wave[:, : wave.shape[1]].copy_(wave, non_blocking=nb)
            P.num_samples.copy_(num_samples, non_blocking=nb)
            # frames per utterance, snip_edges=True (torchaudio kaldi.fbank): 1 + (n - 400) // 160 — written into
            # P.src_len by the fbank kernel at the start of the step (_fbank)
        else:
            if feats.shape[1] == P.T:
                P.feats_in.copy_(feats, non_blocking=nb)
            else:       # a batch shorter than the plan's frame bucket: frames >= src_len are never read (zero-filled by CMVN)
                P.feats_in[:, : feats.shape[1]].copy_(feats, non_blocking=nb)
            P.src_len.copy_(src_lengths, non_blocking=nb)
        if targets is not None:
            P.targets[:, : targets.shape[1]].copy_(targets, non_blocking=nb)
            P.tgt_len.copy_(target_lengths, non_blocking=nb)
        if xtargets is not None:        # target-text tokens for the XCTC losses (criterions/ctc.py:763-767)
            P.tg[1][:, : xtargets.shape[1]].copy_(xtargets, non_blocking=nb)
            P.tl[1].copy_(xtarget_lengths, non_blocking=nb)

    def set_ground_truth(self, P: Plan, head: int, tokens: Optional[torch.Tensor], mask: Optional[torch.Tensor],
                   force_emit: Optional[torch.Tensor]) -> None:
        """Ground-truth curriculum inputs of head ``head`` (0 = CTC, 1 = XCTC) for this batch: per output frame the
        aligned ground-truth token, the mixing mask and the forced lattice state (-1 = free), each [B, T'] or [B*T'] —
        what the reference passes as ``ctc_alignment_oracle`` / builds at s2t_transformer.py:1904-1946, 2061-2103.
        ``tokens=None`` switches the curriculum of that head off.  Static int32 buffers: a captured graph stays valid as
        long as the set of heads with a curriculum does not change (re-capture otherwise)."""
        if tokens is None:
            P.gt.pop(head, None)
            return
        if head not in P.gt:
            z = lambda: torch.zeros(P.R, dtype=torch.int32, device=self.device)
            P.gt[head] = (z(), z(), torch.full((P.R,), -1, dtype=torch.int32, device=self.device))
        tok, msk, fe = P.gt[head]
        tok.copy_(tokens.reshape(-1).to(torch.int32), non_blocking=True)
        msk.copy_(mask.reshape(-1).to(torch.int32), non_blocking=True)
        fe.copy_(force_emit.reshape(-1).to(torch.int32), non_blocking=True)

    def forward(self, P: Plan) -> None:
        """Encoder forward on the plan's inputs; results in P.enc_out [B*T', C], P.logits [B*T', ldv], P.out_len."""
        self._forward(P)

    def out_frames(self, T: int) -> int:
        """Frames of the encoder output the REFERENCE produces for a batch padded to ``T`` input frames (the stem's length
        arithmetic on the padded length: subsampling.py:155,218; pdss2t_transformer.py:53-144 after quirk A7's padding).
        A plan of a larger frame bucket holds more (masked) rows; the plugin slices its outputs to this count."""
        c = self.cfg
        if c.encoder_type == "s2t" and c.subsampling_type == "conv2d":
            for _ in range(c.subsampling_layers):
                T = (T - c.subsampling_kernel) // c.subsampling_stride + 1
            return T
        if c.encoder_type == "s2t":
            pad = (c.subsampling_kernel - 1) // 2
            for _ in range(c.subsampling_layers):
                T = (T + 2 * pad - c.subsampling_kernel) // c.subsampling_stride + 1
            return T
        mult = 1
        for r in c.pds_ratios:
            mult *= max(1, r)
        T = T + (mult - T % mult) if mult > 1 else T
        for r, k in zip(c.pds_ratios, c.pds_kernels):
            T = (T + 2 * ((k - 1) // 2) - k) // r + 1
        return T

    def forward_train(self, P: Plan) -> None:
        """Training-mode forward only (fresh dropout streams): the first half of a step whose loss is computed OUTSIDE the
        engine (an AR decoder + cross-entropy on encoder_out, s2t_transformer.py:868-880); ``backward_external`` is the
        second half."""
        ops.counter_add(self.step_counter, 1)
        self._forward(P)

    def backward_external(self, P: Plan, d_enc_out: Optional[torch.Tensor] = None,
                          d_logits: Optional[Dict[tuple, torch.Tensor]] = None) -> None:
        """Backward pass driven by external gradients: ``d_enc_out`` [B*T', C] w.r.t. P.enc_out and ``d_logits`` keyed
        ("top", head) / ("inter", layer, head) w.r.t. the (unpadded, [B*T', V]) logits the forward pass produced.
        Missing entries are zero.  Gradients land in store.grad exactly as in ``train_step`` (bucketed all-reduce
        included)."""
        d_logits = d_logits or {}
        if self.dp is not None:
            self.dp.reset()
        for k, (hs, Hd) in enumerate(zip(self.heads, P.heads)):
            bufs = [(("top", k), Hd.dlogits)] + [(("inter", l, k), P.inter[(l, k)].dlogits) for l in hs.inter_layers]
            for key, buf in bufs:
                buf.zero_()
                g = d_logits.get(key)
                if g is not None:
                    buf[:, : hs.V].copy_(g.reshape(P.R, hs.V))
        extra = None
        if d_enc_out is not None:
            if not hasattr(P, "ext_grad"):
                P.ext_grad = torch.empty_like(P.enc_out)
            P.ext_grad.copy_(d_enc_out.reshape(P.enc_out.shape))
            extra = P.ext_grad
        self._backward(P, extra)
        if self.dp is not None:
            self.dp.finish()

    def train_step(self, P: Plan) -> None:
        """One fwd+bwd pass: loss in P.loss, gradients in store.grad.  No host sync, no allocation."""
        ops.counter_add(self.step_counter, 1)
        if self.dp is not None:
            self.dp.reset()
        self._forward(P)
        self._loss(P, need_grad=True)
        self._backward(P)
        if self.dp is not None:
            self.dp.finish()

    def enable_data_parallel(self, n_buckets: int = 4, group=None, tail_fraction: Optional[float] = None) -> None:
        """Bucketed gradient all-reduce overlapped with backward (dp.py); cut points = the offsets at which the
        backward pass reports progress (start of every layer / block in the flat gradient buffer)."""
        from .dp import BucketedAllReduce
        S = self.store
        firsts = {sp.first for sp in self.layer_specs}
        firsts |= {f"{self.heads[ks[0]].norm}{l}.weight" for l, ks in self.inter_after.items()}
        cuts = [S.offset(n) for n in S.names()
                if n in firsts or n in ("layer_norm.weight", "ctc.ctc_projection.weight", "embed_ln.weight")
                or (n.startswith("subsample.") and n.endswith("weight")) or n.endswith("conv.0.weight")]
        if tail_fraction is None:
            tail_fraction = float(os.environ.get("S2T_DP_TAIL", "0.12"))
        self.dp = BucketedAllReduce(S.grad, cuts, n_buckets, group, tail_fraction)
        self.grad_ready_hook = self.dp.on_ready

    def capture(self, P: Plan, kind: str = "train") -> None:
        """Capture train_step (or forward-only 'infer') of this plan into a CUDA graph."""
        fn = self.train_step if kind == "train" else self.forward
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        # the warm-up launch must leave no trace: the batch in the plan is usually the live one and will be run again
        # by the first replay, so the step counter (dropout streams) and the BatchNorm running statistics are restored
        saved = {k: v.clone() for k, v in self.buffers.items()}
        counter = self.step_counter.clone()
        # data parallel with per-rank plans (length-bucketed batches): ranks meet NEW plans at different steps, so the
        # eager warm-up must not issue collectives (the step count of all-reduces has to match across ranks) — it runs
        # without the gradient exchange; the captured graph contains it
        dp, hook = self.dp, self.grad_ready_hook
        if self.collective_free_warmup:
            self.dp, self.grad_ready_hook = None, None
        try:
            with torch.cuda.stream(s):
                fn(P)                                # warm-up: function attributes, tensor-map cache
        finally:
            self.dp, self.grad_ready_hook = dp, hook
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        for k, v in saved.items():
            self.buffers[k].copy_(v)
        self.step_counter.copy_(counter)
        g = torch.cuda.CUDAGraph()
        # the dgrad chain is captured on a high-priority stream: when a chain GEMM and a side-stream weight-gradient
        # GEMM become ready together (both 1 CTA/SM), the chain's CTAs are placed first
        prio = -1 if os.environ.get("S2T_CHAIN_PRIORITY", "1") == "1" else 0
        cs = torch.cuda.Stream(device=self.device, priority=prio)
        with torch.cuda.graph(g, stream=cs):
            fn(P)
        P.graph, P.graph_kind = g, kind

    def run(self, P: Plan, kind: str = "train") -> None:
        if P.graph is not None and P.graph_kind == kind:
            P.graph.replay()
        elif kind == "train":
            self.train_step(P)
        else:
            self.forward(P)

    def greedy_decode(self, P: Plan, head: int = 0, with_scores: bool = False):
        """CTCDecoder greedy (s2t_ctc.py:312-347) on the logits of head ``head`` (0 = CTC, 1 = XCTC): int32 tokens
        [B, T'] (-1 padded), lengths, per-frame ids."""
        hs = self.heads[head]
        return ops.ctc_greedy(P.heads[head].logits.view(P.B, P.T2, hs.ldv), P.out_len, hs.V, self.cfg.blank,
                              with_scores=with_scores)
