"""CPU oracle for the encoder + CTC head (TEST INFRASTRUCTURE — see oracle/__init__.py).

A plain-PyTorch fp32 restatement of the reference modules on the hot path.  Every class cites the
reference file:line it follows and keeps the reference's parameter names, so a ``state_dict`` moves
between the reference (through oracle/ref_shim.py), this oracle and the product unchanged.  Autograd
provides the backward pass.  Pinned against the reference itself by oracle/make_golden.py (run in the
build container) and by the committed fixtures tests/golden/enc_*.npz.

Also the timed CPU baseline (``bench.py --impl reference`` / ``cpu_baseline``): the structure mirrors the
reference op for op (separate q/k/v Linear, bmm, fp32 softmax, Dropout modules, nn.CTCLoss) so its cost
on host cores is the reference's cost.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class OracleConfig:
    """Subset of the reference's args Namespace that the hot path reads (s2t_transformer.py:2256-2405)."""
    input_dim: int = 80
    embed_dim: int = 256
    ffn_dim: int = 2048
    heads: int = 4
    layers: int = 12
    subsampling_layers: int = 2
    subsampling_filter: int = 1024
    subsampling_kernel: int = 5
    subsampling_stride: int = 2
    subsampling_type: str = "conv1d"      # conv1d | conv2d
    subsampling_norm: str = "none"        # conv2d: none | batch2d
    subsampling_activation: str = "glu"   # conv2d: swish | relu | none
    attn_type: str = "selfattn"          # selfattn | rel_pos
    encoder_embed_norm: bool = True
    no_scale_embedding: bool = True
    macaron_style: bool = False
    use_cnn_module: bool = False
    cnn_module_kernel: int = 15
    cnn_module_norm: str = "batch_norm"
    encoder_activation_fn: str = "relu"  # FFN activation
    activation_fn: str = "relu"          # conv-module activation in S2TTransformerEncoderLayer (quirk A3)
    layer_padding_mask: bool = False
    dropout: float = 0.1
    max_source_positions: int = 6000
    vocab: int = 10000                   # source dictionary size (CTC head); 0 = no CTC head
    tgt_vocab: int = 0                   # XCTC head (target vocabulary); 0 = none
    inter_ctc_layers: Tuple[int, ...] = ()
    inter_xctc_layers: Tuple[int, ...] = ()
    share_inter_ctc: bool = True
    ctc_pae: str = "none"                # none | inter_league
    xctc_pae: str = "none"
    pae_unnorm_input: bool = False
    # ground-truth curriculum of the prediction-aware adapters (s2t_transformer.py:1050-1052, 1129-1148, 1232-1233)
    ctc_pae_gt_ratio: float = 0.0        # --ctc-pae-ground-truth-ratio
    xctc_pae_gt_ratio: float = 0.0       # --xctc-pae-ground-truth-ratio
    pae_adaptive_gt: bool = False        # --xctc-pae-ground-truth-ratio-adaptive (applies to both heads, sic)
    pae_gt_only_mistake: bool = False    # --xctc-pae-ground-truth-only-mistake (applies to both heads, sic)
    pae_oracle_smooth: bool = False      # --pae-oracle-smooth (reaches the XCTC adapter only, sic :1066 vs :1233)
    # PDS (pdss2t_transformer.py:1361-1560)
    pds_layers: Tuple[int, ...] = (3, 3, 3, 3)
    pds_ratios: Tuple[int, ...] = (2, 2, 1, 2)
    pds_dims: Tuple[int, ...] = (256, 256, 256, 256)
    pds_kernels: Tuple[int, ...] = (5, 5, 5, 5)
    pds_ffn_ratios: Tuple[int, ...] = (8, 8, 8, 8)
    pds_heads: Tuple[int, ...] = (4, 4, 4, 4)
    pds_position_embed: Tuple[int, ...] = (1, 1, 1, 1)
    pds_embed_norm: bool = True
    pds_dropout: float = 0.0


def lengths_to_padding_mask(lens: torch.Tensor, max_len: Optional[int] = None) -> torch.Tensor:
    """fairseq/data/data_utils.py:518-522 — (B, T) bool, True = padding."""
    max_len = int(lens.max()) if max_len is None else max_len
    return torch.arange(max_len, device=lens.device)[None, :] >= lens[:, None]


def act_fn(name: str):
    return {"relu": F.relu, "swish": lambda x: x * torch.sigmoid(x), "gelu": F.gelu}[name]


class Conv1dSubsampling(nn.Module):
    """fairseq/modules/speech_to_text/subsampling.py:106-159 (act = glu, norm = none)."""

    def __init__(self, num_layers, in_dim, filters, kernel_size, stride=2):
        super().__init__()
        self.layers = nn.ModuleList()
        for i in range(num_layers):
            cin = in_dim if i == 0 else filters[i - 1] // 2
            cout = filters[i] * 2 if i == num_layers - 1 else filters[i]
            # Sequential(conv, norm=Identity, GLU) in the reference -> parameter name layers.{i}.0.*
            self.layers.append(nn.Sequential(
                nn.Conv1d(cin, cout, kernel_size, stride=stride, padding=(kernel_size - 1) // 2),
                nn.Identity(), nn.GLU(dim=1)))

    def forward(self, x, x_len):
        x = x.permute(1, 2, 0)                                   # (T,B,D) -> (B,D,T)
        for layer in self.layers:
            x = layer(x)
            x_len = torch.div(x_len - 1, 2, rounding_mode="floor") + 1   # :155
        return x.permute(2, 0, 1), x_len


class Conv2dSubsampling(nn.Module):
    """fairseq/modules/speech_to_text/subsampling.py:162-228 (norm = batch2d | none, act = swish | relu | none)."""

    class _Swish(nn.Module):
        def forward(self, x):
            return x * torch.sigmoid(x)

    def __init__(self, num_layers, in_dim, filters, kernel_size, stride=2, norm="none", act="swish"):
        super().__init__()
        acts = {"swish": Conv2dSubsampling._Swish, "relu": nn.ReLU, "none": nn.Identity}
        self.layers = nn.ModuleList()
        for i in range(num_layers):
            cin = 1 if i == 0 else filters[i - 1]
            # Sequential(conv, norm = Sequential(Identity, BatchNorm2d, Identity) | Identity, act): names layers.{i}.0.*, .1.1.*
            nrm = nn.Sequential(nn.Identity(), nn.BatchNorm2d(filters[i]), nn.Identity()) if norm == "batch2d" else nn.Identity()
            self.layers.append(nn.Sequential(nn.Conv2d(cin, filters[i], kernel_size, stride=stride), nrm, acts[act]()))
        dim = in_dim
        for _ in range(num_layers):
            dim = (dim - 1) // 2                                  # :202-204
        self.linear = nn.Linear(dim * filters[-1], filters[-1])

    def forward(self, x, x_len):
        x = x.permute(1, 2, 0).unsqueeze(1)                      # (T,B,D) -> (B,1,D,T)
        for layer in self.layers:
            x = layer(x)
            x_len = torch.div(x_len - 1, 2, rounding_mode="floor")   # :218 (quirk A9)
        b, ch, d, t = x.size()
        assert t == int(x_len.max()), (t, int(x_len.max()))       # :222
        x = x.reshape(b, ch * d, t).permute(2, 0, 1)
        return self.linear(x), x_len


def sinusoidal_table(num_embeddings: int, dim: int, padding_idx: int = 1) -> torch.Tensor:
    """fairseq/modules/sinusoidal_positional_embedding.py:34-58 — [sin | cos] halves, pad row zeroed."""
    half = dim // 2
    emb = math.log(10000) / (half - 1)
    emb = torch.exp(torch.arange(half, dtype=torch.float) * -emb)
    emb = torch.arange(num_embeddings, dtype=torch.float).unsqueeze(1) * emb.unsqueeze(0)
    emb = torch.cat([torch.sin(emb), torch.cos(emb)], dim=1).view(num_embeddings, -1)
    if dim % 2 == 1:
        emb = torch.cat([emb, torch.zeros(num_embeddings, 1)], dim=1)
    emb[padding_idx, :] = 0
    return emb


def abs_positions(padding_mask: torch.Tensor, dim: int, padding_idx: int = 1) -> torch.Tensor:
    """PositionalEmbedding called on the bool padding mask (s2t_transformer.py:1784; quirk A6):
    positions = cumsum(non-pad) * non-pad + padding_idx  (fairseq/utils.py:286-298)."""
    nonpad = padding_mask.ne(padding_idx).int()               # True(=1) == padding_idx -> pad
    pos = (torch.cumsum(nonpad, dim=1) * nonpad).long() + padding_idx
    table = sinusoidal_table(int(pos.max()) + 1 + padding_idx, dim, padding_idx)
    return table.index_select(0, pos.view(-1)).view(*pos.shape, dim)   # (B, T, C)


def rel_positional_encoding(T: int, dim: int) -> torch.Tensor:
    """fairseq/modules/positional_encoding.py:104-166 — (2T-1, 1, C); row k <-> relative position T-1-k,
    interleaved sin/cos."""
    position = torch.arange(0, T, dtype=torch.float32).unsqueeze(1)
    div = torch.exp(torch.arange(0, dim, 2, dtype=torch.float32) * -(math.log(10000.0) / dim))
    pos = torch.zeros(T, dim)
    neg = torch.zeros(T, dim)
    pos[:, 0::2], pos[:, 1::2] = torch.sin(position * div), torch.cos(position * div)
    neg[:, 0::2], neg[:, 1::2] = torch.sin(-position * div), torch.cos(-position * div)
    pe = torch.cat([torch.flip(pos, [0]), neg[1:]], dim=0)     # (2T-1, C)
    return pe.unsqueeze(1)


class FeedForwardModule(nn.Module):
    """fairseq/modules/s2t_transformer_layer.py:26-66."""

    def __init__(self, d, hidden, p, activation):
        super().__init__()
        self.w_1, self.w_2 = nn.Linear(d, hidden), nn.Linear(hidden, d)
        self.dropout1, self.dropout2 = nn.Dropout(p), nn.Dropout(p)
        self.activation = act_fn(activation)

    def forward(self, x):
        return self.dropout2(self.w_2(self.dropout1(self.activation(self.w_1(x)))))


class MultiheadAttention(nn.Module):
    """fairseq/modules/multihead_attention.py:24-431, encoder self-attention path only
    (q scaled by d_k^-0.5 BEFORE QK^T :265, key-padding -inf :383-395, fp32 softmax, dropout on probs)."""

    def __init__(self, d, heads, p):
        super().__init__()
        self.h, self.dk = heads, d // heads
        self.q_proj, self.k_proj, self.v_proj, self.out_proj = (nn.Linear(d, d) for _ in range(4))
        self.dropout = nn.Dropout(p)

    def forward(self, x, padding_mask):                       # x (T,B,C); mask (B,T) True = pad
        T, B, C = x.shape
        q = self.q_proj(x) * self.dk ** -0.5
        k, v = self.k_proj(x), self.v_proj(x)
        q = q.contiguous().view(T, B * self.h, self.dk).transpose(0, 1)
        k = k.contiguous().view(T, B * self.h, self.dk).transpose(0, 1)
        v = v.contiguous().view(T, B * self.h, self.dk).transpose(0, 1)
        w = torch.bmm(q, k.transpose(1, 2)).view(B, self.h, T, T)
        w = w.masked_fill(padding_mask[:, None, None, :], float("-inf")).view(B * self.h, T, T)
        w = self.dropout(F.softmax(w, dim=-1, dtype=torch.float32).type_as(w))
        a = torch.bmm(w, v).transpose(0, 1).contiguous().view(T, B, C)
        return self.out_proj(a)


class RelPositionMultiHeadedAttention(nn.Module):
    """fairseq/modules/espnet_multihead_attention.py:23-175 + :269-356 (rel_shift :292-311, quirk A5)."""

    def __init__(self, d, heads, p):
        super().__init__()
        self.h, self.dk = heads, d // heads
        self.linear_q, self.linear_k, self.linear_v, self.linear_out = (nn.Linear(d, d) for _ in range(4))
        self.linear_pos = nn.Linear(d, d, bias=False)
        self.pos_bias_u = nn.Parameter(torch.empty(heads, self.dk))
        self.pos_bias_v = nn.Parameter(torch.empty(heads, self.dk))
        nn.init.xavier_uniform_(self.pos_bias_u)
        nn.init.xavier_uniform_(self.pos_bias_v)
        self.dropout = nn.Dropout(p)

    @staticmethod
    def rel_shift(x):
        b, h, t, n = x.shape
        xp = torch.cat([x.new_zeros(b, h, t, 1), x], dim=-1).view(b, h, n + 1, t)
        return xp[:, :, 1:].view_as(x)[:, :, :, : n // 2 + 1]

    def forward(self, x, pos_emb, padding_mask):              # x (T,B,C); pos_emb (2T-1,1,C)
        x = x.transpose(0, 1)
        B, T, C = x.shape
        q = self.linear_q(x).view(B, T, self.h, self.dk)
        k = self.linear_k(x).view(B, T, self.h, self.dk).transpose(1, 2)
        v = self.linear_v(x).view(B, T, self.h, self.dk).transpose(1, 2)
        p = self.linear_pos(pos_emb.transpose(0, 1)).view(1, -1, self.h, self.dk).transpose(1, 2)
        ac = torch.matmul((q + self.pos_bias_u).transpose(1, 2), k.transpose(-2, -1))
        bd = self.rel_shift(torch.matmul((q + self.pos_bias_v).transpose(1, 2), p.transpose(-2, -1)))
        s = (ac + bd) / math.sqrt(self.dk)
        s = s.masked_fill(padding_mask[:, None, None, :], float("-inf"))
        lim = 1e8 if s.dtype == torch.float32 else 1e4
        s = s.clamp(min=-lim, max=lim)
        a = self.dropout(F.softmax(s, dim=-1, dtype=torch.float32).type_as(s))
        o = torch.matmul(a, v).transpose(1, 2).contiguous().view(B, T, C)
        return self.linear_out(o).transpose(0, 1)


class ConvolutionModule(nn.Module):
    """fairseq/modules/convolution.py:8-120 (bias-free convs; BatchNorm1d over B*T incl. padded frames, A10)."""

    def __init__(self, d, kernel, p, activation, norm_type="batch_norm"):
        super().__init__()
        self.pointwise_conv1 = nn.Conv1d(d, 2 * d, 1, bias=False)
        self.depthwise_conv = nn.Conv1d(d, d, kernel, padding=(kernel - 1) // 2, groups=d, bias=False)
        self.norm_type = norm_type
        self.norm = nn.BatchNorm1d(d) if norm_type == "batch_norm" else nn.LayerNorm(d)
        self.activation = act_fn(activation)
        self.pointwise_conv2 = nn.Conv1d(d, d, 1, bias=False)
        self.dropout = nn.Dropout(p)

    def forward(self, x, padding_mask):                       # x (B,T,C)
        x = x.transpose(1, 2)
        zero = padding_mask.unsqueeze(1)
        x = x.masked_fill(zero, 0.0)
        x = F.glu(self.pointwise_conv1(x), dim=1)
        x = self.depthwise_conv(x)
        if self.norm_type == "layer_norm":
            x = self.norm(x.transpose(1, 2)).transpose(1, 2)
        else:
            x = self.norm(x)
        x = self.pointwise_conv2(self.activation(x))
        x = x.masked_fill(zero, 0.0)
        return self.dropout(x.transpose(1, 2))


class EncoderLayer(nn.Module):
    """fairseq/modules/s2t_transformer_layer.py:69-322, encoder_normalize_before=True (pre-norm)."""

    def __init__(self, cfg: OracleConfig):
        super().__init__()
        d, p = cfg.embed_dim, cfg.dropout
        self.attn_type = cfg.attn_type
        if cfg.attn_type == "rel_pos":
            self.self_attn = RelPositionMultiHeadedAttention(d, cfg.heads, p)
        else:
            self.self_attn = MultiheadAttention(d, cfg.heads, p)
        self.self_attn_layer_norm = nn.LayerNorm(d)
        self.dropout_module = nn.Dropout(p)
        self.ffn_scale = 0.5 if cfg.macaron_style else 1.0
        if cfg.macaron_style:
            self.macaron_ffn = FeedForwardModule(d, cfg.ffn_dim, p, cfg.encoder_activation_fn)
            self.macaron_norm = nn.LayerNorm(d)
        else:
            self.macaron_ffn = self.macaron_norm = None
        if cfg.use_cnn_module:
            self.conv_norm = nn.LayerNorm(d)
            self.conv_module = ConvolutionModule(d, cfg.cnn_module_kernel, p, cfg.activation_fn, cfg.cnn_module_norm)
            self.final_norm = nn.LayerNorm(d)
        else:
            self.conv_norm = self.conv_module = self.final_norm = None
        self.ffn = FeedForwardModule(d, cfg.ffn_dim, p, cfg.encoder_activation_fn)
        self.ffn_norm = nn.LayerNorm(d)

    def forward(self, x, padding_mask, pos_emb=None):
        if self.macaron_norm is not None:
            x = x + self.ffn_scale * self.macaron_ffn(self.macaron_norm(x))
        h = self.self_attn_layer_norm(x)
        h = self.self_attn(h, pos_emb, padding_mask) if self.attn_type == "rel_pos" else self.self_attn(h, padding_mask)
        x = x + self.dropout_module(h)
        if self.conv_module is not None:
            h = self.conv_module(self.conv_norm(x.transpose(0, 1)), padding_mask).transpose(0, 1)
            x = x + h
        x = x + self.ffn_scale * self.ffn(self.ffn_norm(x))
        if self.conv_module is not None:
            x = self.final_norm(x)
        return x


class CTC(nn.Module):
    """fairseq/modules/speech_to_text/ctc.py:17-66 — [LayerNorm] -> dropout -> Linear(C, V) (quirk A11)."""

    def __init__(self, d, vocab, p, need_layernorm=False):
        super().__init__()
        self.ctc_projection = nn.Linear(d, vocab)
        nn.init.normal_(self.ctc_projection.weight, mean=0, std=d ** -0.5)
        nn.init.constant_(self.ctc_projection.bias, 0.0)
        self.ctc_dropout_module = nn.Dropout(p)
        self.need_layernorm = need_layernorm
        if need_layernorm:
            self.LayerNorm = nn.LayerNorm(d)

    def forward(self, x):
        if self.need_layernorm:
            x = self.LayerNorm(x)
        return self.ctc_projection(self.ctc_dropout_module(x))


class Adapter(nn.Module):
    """fairseq/modules/speech_to_text/adapter.py:76-299, adapter_type = "inter_league": out = representation +
    softmax(logit) @ embed_adapter.weight; the softmax runs in the logits' dtype (:215-217, quirk A13)."""

    def __init__(self, d, vocab):
        super().__init__()
        self.embed_adapter = nn.Linear(d, vocab, bias=False)              # used transposed as a (V, C) matrix
        nn.init.normal_(self.embed_adapter.weight, mean=0, std=d ** -0.5)

    def forward(self, representation, logit, oracle=None, oracle_mask=None, gt_ratio: float = 0.0, smooth: bool = False):
        T, B, _ = representation.shape
        V = logit.size(-1)
        dist = F.softmax(logit, dim=-1)
        if gt_ratio > 0 and oracle is not None:                           # adapter.py:244-262 (oracle, mask: (B, T))
            oracle_dis = (oracle.unsqueeze(-1) == torch.arange(V).unsqueeze(0)).to(dist.dtype).transpose(0, 1)
            if smooth:
                oracle_dis = torch.where(oracle_dis == 1, 0.9 + 0.1 / V, 0.1 / V).to(dist.dtype)
            m = oracle_mask.transpose(0, 1).unsqueeze(-1).repeat(1, 1, V)
            dist = m * oracle_dis + ~m * dist
        soft = torch.mm(dist.contiguous().view(-1, V), self.embed_adapter.weight).view(T, B, -1)
        return representation + soft


class S2TTransformerEncoder(nn.Module):
    """fairseq/models/speech_to_text/s2t_transformer.py:883 (init :887-1244, forward :1714-2154), for the
    options of configs C1/C2/C4: conv1d subsampler, embed norm, abs or rel positions, pre-norm layers
    (optionally macaron + conv module), final LayerNorm, CTC / XCTC heads on the top layer."""

    def __init__(self, cfg: OracleConfig):
        super().__init__()
        self.cfg = cfg
        d = cfg.embed_dim
        self.dropout_module = nn.Dropout(cfg.dropout)
        self.embed_scale = 1.0 if cfg.no_scale_embedding else math.sqrt(d)
        filters = [cfg.subsampling_filter] * (cfg.subsampling_layers - 1) + [d]      # subsampling.py:231-247
        if cfg.subsampling_type == "conv2d":
            self.subsample = Conv2dSubsampling(cfg.subsampling_layers, cfg.input_dim, filters, cfg.subsampling_kernel,
                                               cfg.subsampling_stride, cfg.subsampling_norm, cfg.subsampling_activation)
        else:
            self.subsample = Conv1dSubsampling(cfg.subsampling_layers, cfg.input_dim, filters,
                                               cfg.subsampling_kernel, cfg.subsampling_stride)
        if cfg.encoder_embed_norm:
            self.embed_ln = nn.LayerNorm(d)
        self.layers = nn.ModuleList([EncoderLayer(cfg) for _ in range(cfg.layers)])
        self.layer_norm = nn.LayerNorm(d)
        if cfg.vocab > 0:
            self.ctc = CTC(d, cfg.vocab, cfg.dropout)
        if cfg.tgt_vocab > 0:
            self.xctc = CTC(d, cfg.tgt_vocab, cfg.dropout)
        # inter-CTC / inter-XCTC with shared heads (s2t_transformer.py:1001-1100, 1135-1244)
        assert cfg.share_inter_ctc or not cfg.inter_ctc_layers, "oracle: only share_inter_ctc is restated"
        for l in cfg.inter_ctc_layers:
            setattr(self, f"ctc_norm{l}", nn.LayerNorm(d))
        for l in cfg.inter_xctc_layers:
            setattr(self, f"xctc_norm{l}", nn.LayerNorm(d))
        if cfg.inter_ctc_layers and cfg.ctc_pae != "none":
            self.pae = Adapter(d, cfg.vocab)
        if cfg.inter_xctc_layers and cfg.xctc_pae != "none":
            self.xctc_pae = Adapter(d, cfg.tgt_vocab)

    def _draw(self, entry, ratio: float, masks, key):
        """(oracle tokens, mask, force_emit) of one head from a ``ctc_alignment_oracle`` entry (oracle, best_aligns_pad,
        mistake_flag, mistake_ratio), s2t_transformer.py:1908-1936 / 2065-2092; ``masks[key]`` injects the mask instead of
        drawing it (torch.rand) — the parity tests pass the same mask to the CUDA path."""
        cfg = self.cfg
        oracle, best_aligns_pad, mistake_flag, mistake_ratio = entry
        if masks is not None and key in masks:
            mask = masks[key].bool().clone()
        else:
            prob = ratio * mistake_ratio.unsqueeze(-1) if cfg.pae_adaptive_gt else ratio
            mask = (torch.rand(oracle.size()) < prob).bool()
            if cfg.pae_gt_only_mistake:
                mask.masked_fill_(~mistake_flag, False)
        return oracle, mask, best_aligns_pad.masked_fill(~mask, -1)

    def forward(self, src_tokens: torch.Tensor, src_lengths: torch.Tensor, ctc_alignment_oracle=None,
                oracle_masks=None) -> Dict[str, list]:
        cfg = self.cfg
        ctc_or = xctc_or = None                                            # drawn at the first inter layer, then reused
        x = src_tokens.transpose(0, 1)                                     # :1728
        x, lens = self.subsample(x, src_lengths)                           # :1748
        pad = lengths_to_padding_mask(lens)                                # :1751
        if pad.size(1) < x.size(0):                                        # :1752-1762 (quirk A9)
            miss = torch.zeros(pad.size(0), x.size(0) - pad.size(1), dtype=torch.bool)
            pad = torch.cat([miss, pad], dim=1)
        x = x * (1 - pad.transpose(0, 1).unsqueeze(-1).type_as(x))        # :1765
        if cfg.encoder_embed_norm:
            x = self.embed_ln(x)                                           # :1768
        x = self.embed_scale * x                                           # :1773
        positions = None
        if cfg.attn_type == "rel_pos":
            positions = rel_positional_encoding(x.size(0), cfg.embed_dim).to(x.dtype)   # :1777
        else:
            x = x + abs_positions(pad, cfg.embed_dim).transpose(0, 1).to(x.dtype)       # :1784-1786
        x = self.dropout_module(x)                                         # :1794
        inter_ctc, inter_xctc = [], []
        for idx, layer in enumerate(self.layers, start=1):                 # :1824
            if cfg.layer_padding_mask and not bool(pad.all()):
                x = x.transpose(0, 1).masked_fill(pad.unsqueeze(2), 0.0).transpose(0, 1)  # :1828-1836
            x = layer(x, pad, pos_emb=positions)                           # :1846
            if idx in cfg.inter_ctc_layers:                                # :1881-1946
                norm_x = getattr(self, f"ctc_norm{idx}")(x)
                logit = self.ctc(norm_x)
                inter_logit = [logit, pad]
                if cfg.ctc_pae_gt_ratio > 0 and ctc_alignment_oracle is not None and ctc_alignment_oracle.get("ctc") is not None:
                    if ctc_or is None:
                        ctc_or = self._draw(ctc_alignment_oracle["ctc"], cfg.ctc_pae_gt_ratio, oracle_masks, "ctc")
                    inter_logit = [logit, None, ctc_or[2]]
                if cfg.ctc_pae != "none":
                    o = ctc_or if ctc_or is not None else (None, None, None)
                    x = self.pae(x if cfg.pae_unnorm_input else norm_x, logit, o[0], o[1], cfg.ctc_pae_gt_ratio, False)
                inter_ctc.append(inter_logit)
            if idx in cfg.inter_xctc_layers:                               # :2043-2103
                norm_x = getattr(self, f"xctc_norm{idx}")(x)
                logit = self.xctc(norm_x)
                inter_logit = logit
                if cfg.xctc_pae_gt_ratio > 0 and ctc_alignment_oracle is not None and ctc_alignment_oracle.get("xctc") is not None:
                    if xctc_or is None:
                        xctc_or = self._draw(ctc_alignment_oracle["xctc"], cfg.xctc_pae_gt_ratio, oracle_masks, "xctc")
                    inter_logit = [logit, None, xctc_or[2]]
                if cfg.xctc_pae != "none":
                    o = xctc_or if xctc_or is not None else (None, None, None)
                    x = self.xctc_pae(x if cfg.pae_unnorm_input else norm_x, logit, o[0], o[1], cfg.xctc_pae_gt_ratio,
                                      cfg.pae_oracle_smooth)
                inter_xctc.append(inter_logit)
        x = self.layer_norm(x)                                             # :2112
        out = {"encoder_out": [x], "ctc_logit": [], "xctc_logit": [], "inter_ctc_logits": inter_ctc,
               "inter_xctc_logits": inter_xctc, "encoder_padding_mask": [pad], "mixup": None,
               "encoder_embedding": [], "encoder_states": [], "src_tokens": [], "src_lengths": []}
        if cfg.vocab > 0:
            out["ctc_logit"] = [self.ctc(x)]                               # :2126
            if ctc_or is not None:
                out["ctc_logit"] = [[out["ctc_logit"][0], None, ctc_or[2]]]          # :2136-2137
        if cfg.tgt_vocab > 0:
            out["xctc_logit"] = [self.xctc(x)]                             # :2130
            if xctc_or is not None:
                out["xctc_logit"] = [[out["xctc_logit"][0], None, xctc_or[2]]]       # :2139-2140
        return out


class Downsampling(nn.Module):
    """fairseq/models/speech_to_text/pdss2t_transformer.py:53-144, reduced_way = "conv"."""

    def __init__(self, embed_norm, cin, cout, kernel, stride):
        super().__init__()
        self.stride = stride
        self.conv = nn.Sequential(nn.Conv1d(cin, cout, kernel, stride=stride, padding=(kernel - 1) // 2))
        self.embed_norm = embed_norm
        if embed_norm:
            self.norm = nn.LayerNorm(cout)

    def forward(self, x, lengths):                              # x (T,B,C)
        T = x.size(0)
        assert T % self.stride == 0
        x = x.masked_fill(lengths_to_padding_mask(lengths, T).transpose(0, 1).unsqueeze(2), 0.0)       # :107-116
        lengths = ((lengths.float() - 1) / self.stride + 1).floor().long()                             # :121
        x = self.conv(x.permute(1, 2, 0)).permute(2, 0, 1)
        if self.embed_norm:
            x = self.norm(x)
        x = x.masked_fill(lengths_to_padding_mask(lengths, x.size(0)).transpose(0, 1).unsqueeze(2), 0.0)  # :136-142 (A8)
        return x, lengths


class PDSS2TTransformerEncoder(nn.Module):
    """fairseq/models/speech_to_text/pdss2t_transformer.py:226 (forward :1042-1281) for the options of config C3:
    conv down-sampling per stage, absolute or relative positions per stage, pre-norm layers, no fusion, final
    LayerNorm, CTC head on top.  embed_scale is computed but never applied (quirk A17)."""

    def __init__(self, cfg: OracleConfig):
        super().__init__()
        self.cfg = cfg
        self.n = len(cfg.pds_dims)
        cin = cfg.input_dim
        for i in range(self.n):
            d = cfg.pds_dims[i]
            setattr(self, f"downsampling{i + 1}", Downsampling(cfg.pds_embed_norm, cin, d, cfg.pds_kernels[i], cfg.pds_ratios[i]))
            from dataclasses import replace
            lcfg = replace(cfg, embed_dim=d, ffn_dim=d * cfg.pds_ffn_ratios[i], heads=cfg.pds_heads[i],
                           activation_fn=cfg.encoder_activation_fn)           # pds_layer.py:67,104 (quirk A3)
            setattr(self, f"stage{i + 1}", nn.ModuleList([EncoderLayer(lcfg) for _ in range(cfg.pds_layers[i])]))
            cin = d
        self.layer_norm = nn.LayerNorm(cfg.pds_dims[-1])
        if cfg.vocab > 0:
            self.ctc = CTC(cfg.pds_dims[-1], cfg.vocab, cfg.dropout)

    def forward(self, src_tokens, src_lengths):
        cfg = self.cfg
        x = src_tokens.transpose(0, 1)
        lens = src_lengths
        mult = 1
        for r in cfg.pds_ratios:
            mult *= max(1, r)
        pad_to = mult - x.size(0) % mult                                      # :1050-1055 (quirk A7)
        if mult > 1 and pad_to > 0:
            x = torch.cat([x, x.new_zeros((pad_to, x.size(1), x.size(2)))], dim=0)
        for i in range(self.n):
            x, lens = getattr(self, f"downsampling{i + 1}")(x, lens)
            pad = lengths_to_padding_mask(lens, x.size(0))
            positions = None
            if cfg.pds_position_embed[i]:
                if cfg.attn_type == "rel_pos":
                    positions = rel_positional_encoding(x.size(0), cfg.pds_dims[i]).to(x.dtype)
                else:
                    x = x + abs_positions(pad, cfg.pds_dims[i]).transpose(0, 1).to(x.dtype)
            x = F.dropout(x, cfg.dropout if i == 0 else cfg.pds_dropout, self.training)          # :1118-1121
            for layer in getattr(self, f"stage{i + 1}"):
                x = layer(x, pad, pos_emb=positions)
        x = self.layer_norm(x)
        out = {"encoder_out": [x], "ctc_logit": [], "xctc_logit": [], "inter_ctc_logits": [],
               "inter_xctc_logits": [], "encoder_padding_mask": [pad], "mixup": None,
               "encoder_embedding": [], "encoder_states": [], "src_tokens": [], "src_lengths": []}
        if cfg.vocab > 0:
            out["ctc_logit"] = [self.ctc(x)]
        return out


def ctc_criterion_loss(logit_tbv: torch.Tensor, padding_mask: torch.Tensor, tokens: torch.Tensor,
                       pad: int = 1, eos: int = 2, blank: int = 0) -> torch.Tensor:
    """CtcCriterion arithmetic (fairseq/criterions/ctc.py): input_lengths :556, targets :516-540,
    fp32 log_softmax (s2t_ctc.py:98-112) and nn.CTCLoss(blank, 'none', zero_infinity=True).sum() :435-474."""
    input_lengths = (~padding_mask).long().sum(-1)
    mask = (tokens != pad) & (tokens != eos)
    targets_flat = tokens.masked_select(mask)
    target_lengths = mask.sum(-1)
    lprobs = F.log_softmax(logit_tbv, dim=-1, dtype=torch.float32)
    with torch.backends.cudnn.flags(enabled=False):
        loss = F.ctc_loss(lprobs, targets_flat, input_lengths, target_lengths, blank=blank,
                          reduction="none", zero_infinity=True)
    return loss.sum()


class _ImputerLoss(torch.autograd.Function):
    """torch_imputer.imputer_loss (fairseq/torch_imputer/imputer.py:21-84 over imputer.cu) through the float64 numpy
    restatement oracle/ctc.py imputer_loss: sum over utterances of the masked CTC nll, gradient w.r.t. the LOGITS."""

    @staticmethod
    def forward(ctx, logit_tbv, targets_list, force_emit, input_lengths, blank):
        from oracle import ctc as OC
        x = logit_tbv.detach().transpose(0, 1).double().numpy()                 # (B, T, V)
        nll, grad = OC.imputer_loss(x, targets_list, force_emit.numpy(), input_lengths.tolist(), blank)
        ctx.grad = torch.from_numpy(grad).transpose(0, 1).to(logit_tbv.dtype)
        return torch.tensor(float(nll.sum()), dtype=torch.float32)

    @staticmethod
    def backward(ctx, g):
        return ctx.grad * g, None, None, None, None


def imputer_criterion_loss(logit_tbv: torch.Tensor, padding_mask: torch.Tensor, tokens: torch.Tensor, force_emit: torch.Tensor,
                           pad: int = 1, eos: int = 2, blank: int = 0) -> torch.Tensor:
    """CtcCriterion.get_ctc_loss with ``ctc_masked_loss`` and a force_emit matrix (criterions/ctc.py:455-465): fp32
    log_softmax + imputer_loss(..., reduction 'none', zero_infinity=True).sum()."""
    input_lengths = (~padding_mask).long().sum(-1)
    mask = (tokens != pad) & (tokens != eos)
    targets_list = [tokens[b][mask[b]].numpy() for b in range(tokens.shape[0])]
    return _ImputerLoss.apply(logit_tbv.float(), targets_list, force_emit, input_lengths, blank)


def ground_truth_alignment(out: Dict[str, list], src_tokens: torch.Tensor, tgt_tokens: Optional[torch.Tensor],
                           pad: int = 1, eos: int = 2, blank: int = 0) -> Dict[str, tuple]:
    """CtcCriterion.get_ground_truth_alignment / get_ctc_align (criterions/ctc.py:283-433) on an encoder output dict:
    per head (oracle tokens (B, T'), best_aligns_pad (B, T'), mistake_flag (B, T'), mistake_ratio (B)) from the best CTC
    alignment (torch_imputer.best_alignment -> oracle/viterbi.py) of the TOP logits."""
    from oracle import viterbi as OV
    res = {}
    pmask = out["encoder_padding_mask"][0]
    input_lengths = (~pmask).long().sum(-1)
    for key, logits, tokens in (("ctc", out["ctc_logit"], src_tokens), ("xctc", out["xctc_logit"], tgt_tokens)):
        if not logits or tokens is None:
            continue
        top = logits[0][0] if isinstance(logits[0], list) else logits[0]
        lp = F.log_softmax(top.detach(), dim=-1, dtype=torch.float32)              # (T, B, V)
        tl = ((tokens != pad) & (tokens != eos)).sum(-1)
        paths, _ = OV.best_alignment(lp.transpose(0, 1).numpy(), tokens.numpy(), input_lengths.tolist(), tl.tolist(), blank)
        T = lp.size(0)
        bap = torch.tensor([a + [0] * (T - len(a)) for a in paths], dtype=tokens.dtype)
        pos = torch.div(bap, 2, rounding_mode="floor").clip(max=tokens.shape[1] - 1)
        oracle = tokens.gather(-1, pos)
        oracle.masked_fill_(bap % 2 == 0, blank)
        mistake = lp.argmax(dim=-1).transpose(0, 1) != oracle
        res[key] = (oracle, bap, mistake, mistake.sum(-1) / input_lengths)
    return res


def ctc_criterion_total(out: Dict[str, list], src_tokens: torch.Tensor, tgt_tokens: Optional[torch.Tensor], *,
                        ctc_weight: float, xctc_weight: float = 0.0, inter_ctc_weight: float = 0.0,
                        inter_xctc_weight: float = 0.0, masked_loss: bool = False) -> torch.Tensor:
    """CtcCriterion.compute_ctc_loss weighting (fairseq/criterions/ctc.py:542-682, 759-833, 1004-1015):
    ctc_w * ctc + xctc_w * xctc + inter_ctc_w * mean_l(inter) + inter_xctc_w * mean_l(inter_x); source-side losses use
    the transcript tokens, target-side (XCTC) losses the target tokens."""
    pad = out["encoder_padding_mask"][0]
    loss = src_tokens.new_zeros((), dtype=torch.float32)

    def term(lg, tokens):
        """one inter term: bare logit, [logit, pad_mask] or, under the curriculum, [logit, None, force_emit]
        (criterions/ctc.py:586-627); the masked loss only with ``masked_loss`` (ctc_masked_loss)."""
        if not isinstance(lg, list):
            return ctc_criterion_loss(lg, pad, tokens)
        pm = lg[1] if lg[1] is not None else pad
        if len(lg) >= 3 and masked_loss:
            return imputer_criterion_loss(lg[0], pm, tokens, lg[2])
        return ctc_criterion_loss(lg[0], pm, tokens)

    if inter_ctc_weight > 0 and out["inter_ctc_logits"]:
        it = [term(lg, src_tokens) for lg in out["inter_ctc_logits"]]
        loss = loss + inter_ctc_weight * sum(it) / len(it)
    if ctc_weight > 0 and out["ctc_logit"]:
        loss = loss + ctc_weight * term(out["ctc_logit"][0], src_tokens)
    if tgt_tokens is not None:
        if inter_xctc_weight > 0 and out["inter_xctc_logits"]:
            it = [term(lg, tgt_tokens) for lg in out["inter_xctc_logits"]]
            loss = loss + inter_xctc_weight * sum(it) / len(it)
        if xctc_weight > 0 and out["xctc_logit"]:
            loss = loss + xctc_weight * term(out["xctc_logit"][0], tgt_tokens)
    return loss


def config_from_args(args) -> OracleConfig:
    """Map a reference-style args Namespace (after the arch function ran) onto OracleConfig."""
    g = lambda k, d=None: getattr(args, k, d)
    split = lambda s: tuple(int(v) for v in str(s).split(",")) if s else ()
    return OracleConfig(
        input_dim=g("input_feat_per_channel", 80) * g("input_channels", 1), embed_dim=args.encoder_embed_dim,
        ffn_dim=args.encoder_ffn_embed_dim, heads=args.encoder_attention_heads, layers=g("encoder_layers", 12),
        subsampling_layers=g("subsampling_layers", 2), subsampling_filter=g("subsampling_filter", 1024),
        subsampling_kernel=g("subsampling_kernel", 5), subsampling_stride=g("subsampling_stride", 2),
        subsampling_type=g("subsampling_type", "conv1d"), subsampling_norm=g("subsampling_norm", "none"),
        subsampling_activation=g("subsampling_activation", "glu"),
        attn_type=g("encoder_attention_type", "selfattn"), encoder_embed_norm=g("encoder_embed_norm", False),
        no_scale_embedding=g("encoder_no_scale_embedding", False), macaron_style=g("macaron_style", False),
        use_cnn_module=g("use_cnn_module", False), cnn_module_kernel=g("cnn_module_kernel", 31),
        cnn_module_norm=g("cnn_module_norm", "batch_norm"), encoder_activation_fn=g("encoder_activation_fn", "relu"),
        activation_fn=g("activation_fn", "relu"), layer_padding_mask=g("layer_padding_mask", False),
        dropout=args.dropout, max_source_positions=g("max_source_positions", 6000),
        vocab=g("_src_vocab", 0) if g("ctc_weight", 0) > 0 else 0,
        tgt_vocab=g("_tgt_vocab", 0) if g("xctc_weight", 0) > 0 else 0,
        inter_ctc_layers=split(g("inter_ctc_layers")), inter_xctc_layers=split(g("inter_xctc_layers")),
        share_inter_ctc=g("share_inter_ctc", False), ctc_pae=g("ctc_pae", "none"),
        xctc_pae=g("xctc_pae", g("ctc_pae", "none")),
        pae_unnorm_input=g("pae_unnorm_input", False), ctc_pae_gt_ratio=g("ctc_pae_ground_truth_ratio", 0.0),
        xctc_pae_gt_ratio=g("xctc_pae_ground_truth_ratio", 0.0),
        pae_adaptive_gt=g("xctc_pae_ground_truth_ratio_adaptive", False),
        pae_gt_only_mistake=g("xctc_pae_ground_truth_only_mistake", False),
        pae_oracle_smooth=g("pae_oracle_smooth", False), **_pds_from_args(args))


def _pds_from_args(args) -> dict:
    if getattr(args, "pds_stages", None) is None:
        return {}
    n = int(args.pds_stages)
    ints = lambda k, d=None: tuple(int(v) for v in str(getattr(args, k, d)).split("_"))[:n]
    return dict(pds_layers=ints("pds_layers"), pds_ratios=ints("pds_ratios"), pds_dims=ints("pds_embed_dims"),
                pds_kernels=ints("pds_kernel_sizes"), pds_ffn_ratios=ints("pds_ffn_ratios"),
                pds_heads=ints("pds_attn_heads"), pds_position_embed=ints("pds_position_embed", "_".join(["1"] * n)),
                pds_embed_norm=getattr(args, "pds_embed_norm", False),
                pds_dropout=getattr(args, "pds_dropout", args.dropout))
