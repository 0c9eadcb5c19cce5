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

Reference behaviour followed (file:line in the method docstrings): ``S2TTransformerEncoder.forward``
(fairseq/models/speech_to_text/s2t_transformer.py:1714-2154), ``Conv1dSubsampling``
(modules/speech_to_text/subsampling.py:106-159), ``S2TTransformerEncoderLayer``
(modules/s2t_transformer_layer.py:229-322), ``MultiheadAttention`` (modules/multihead_attention.py),
``CTC`` (modules/speech_to_text/ctc.py:17-66) and the ``ctc`` criterion arithmetic
(criterions/ctc.py:435-474,516-540).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import torch

from . import frontend, ops
from .params import BF16, ParamStore


@dataclass
class EncoderConfig:
    """The subset of the reference's args Namespace the path reads (s2t_transformer.py:2256-2405)."""
    input_dim: int = 80
    embed_dim: int = 256
    ffn_dim: int = 2048
    heads: int = 4
    layers: int = 12
    subsampling_layers: int = 2
    subsampling_filter: int = 1024
    subsampling_kernel: int = 5
    subsampling_stride: int = 2
    attn_type: str = "selfattn"            # selfattn | rel_pos
    encoder_embed_norm: bool = True
    no_scale_embedding: bool = True
    macaron_style: bool = False
    use_cnn_module: bool = False
    cnn_module_kernel: int = 15
    cnn_module_norm: str = "batch_norm"
    encoder_activation_fn: str = "relu"    # FFN activation
    activation_fn: str = "relu"            # conv-module activation (quirk A3)
    layer_padding_mask: bool = False
    dropout: float = 0.1
    max_source_positions: int = 6000
    vocab: int = 10000                     # CTC head (source dictionary); 0 = none
    ctc_weight: float = 1.0
    blank: int = 0

    @staticmethod
    def from_args(args, vocab: int) -> "EncoderConfig":
        g = lambda k, d=None: getattr(args, k, d)
        return EncoderConfig(
            input_dim=g("input_feat_per_channel", 80) * g("input_channels", 1),
            embed_dim=args.encoder_embed_dim, ffn_dim=args.encoder_ffn_embed_dim,
            heads=args.encoder_attention_heads, layers=args.encoder_layers,
            subsampling_layers=g("subsampling_layers", 2), subsampling_filter=g("subsampling_filter", 1024),
            subsampling_kernel=g("subsampling_kernel", 5), subsampling_stride=g("subsampling_stride", 2),
            attn_type=g("encoder_attention_type", "selfattn"), encoder_embed_norm=g("encoder_embed_norm", False),
            no_scale_embedding=g("encoder_no_scale_embedding", False), macaron_style=g("macaron_style", False),
            use_cnn_module=g("use_cnn_module", False), cnn_module_kernel=g("cnn_module_kernel", 31),
            cnn_module_norm=g("cnn_module_norm", "batch_norm"),
            encoder_activation_fn=g("encoder_activation_fn", "relu"), activation_fn=g("activation_fn", "relu"),
            layer_padding_mask=g("layer_padding_mask", False), dropout=args.dropout,
            max_source_positions=g("max_source_positions", 6000),
            vocab=vocab if g("ctc_weight", 0) > 0 else 0, ctc_weight=g("ctc_weight", 1.0) or 1.0)


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


def _round8(n: int) -> int:
    return (n + 7) // 8 * 8


class _LayerBufs:
    """Saved activations of one encoder layer."""
    __slots__ = ("ln1", "st1", "qkv", "probs", "probs_d", "ctx", "x1", "ln2", "st2", "f")


class Plan:
    """All device buffers of one (B, T, Smax) bucket."""

    def __init__(self, eng: "S2TEncoderEngine", B: int, T: int, smax: int, wave_len: int = 0):
        c, dev = eng.cfg, eng.device
        self.B, self.T, self.smax, self.wave_len = B, T, smax, wave_len
        z = lambda *s, dt=BF16: torch.zeros(s, device=dev, dtype=dt)
        e = lambda *s, dt=BF16: torch.empty(s, device=dev, dtype=dt)
        k, st = c.subsampling_kernel, c.subsampling_stride
        pad = (k - 1) // 2
        self.pad = pad
        # --- static inputs --------------------------------------------------------------------
        self.feats_in = z(B, T, c.input_dim, dt=torch.float32)
        self.wave = z(B, wave_len, dt=torch.float32) if wave_len else None
        self.num_samples = z(B, dt=torch.int32) if wave_len else None
        self.fbank_raw = e(B, T, c.input_dim, dt=torch.float32) if wave_len else None
        self.src_len = z(B, dt=torch.int32)
        self.targets = z(B, max(smax, 1), dt=torch.int32)
        self.tgt_len = z(B, dt=torch.int32)
        # --- subsampler: framed inputs (pad zero rows front/back), pre-GLU conv outputs --------------
        self.t_sub = [T]
        for _ in range(c.subsampling_layers):
            self.t_sub.append((self.t_sub[-1] + 2 * pad - k) // st + 1)
        chans = [c.input_dim] + [f // 2 for f in eng.sub_filters]
        self.framed = [z(B, self.t_sub[i] + 2 * pad, chans[i]) for i in range(c.subsampling_layers)]
        self.pre = [e(B * self.t_sub[i + 1], eng.sub_filters[i]) for i in range(c.subsampling_layers)]
        self.sub_len = [z(B, dt=torch.int32) for _ in range(c.subsampling_layers)]
        T2 = self.T2 = self.t_sub[-1]
        R = self.R = B * T2
        C, F, H = c.embed_dim, c.ffn_dim, c.heads
        self.ld = _round8(T2)
        self.x0 = e(R, C)
        self.e_ln, self.e_st = e(R, C), e(R, 2, dt=torch.float32)
        self.x = [e(R, C) for _ in range(c.layers + 1)]
        self.layers: List[_LayerBufs] = []
        drop = c.dropout > 0
        for _ in range(c.layers):
            L = _LayerBufs()
            L.ln1, L.st1 = e(R, C), e(R, 2, dt=torch.float32)
            L.qkv = e(R, 3 * C)
            L.probs = z(B * H, T2, self.ld)
            L.probs_d = z(B * H, T2, self.ld) if drop else L.probs
            L.ctx = e(R, C)
            L.x1 = e(R, C)
            L.ln2, L.st2 = e(R, C), e(R, 2, dt=torch.float32)
            L.f = e(R, F)
            self.layers.append(L)
        self.scores = z(B * H, T2, self.ld)
        self.enc_out, self.enc_st = e(R, C), e(R, 2, dt=torch.float32)
        self.enc_drop = e(R, C) if drop else self.enc_out
        V = c.vocab
        self.ldv = _round8(max(V, 8))
        if V > 0:
            self.logits = z(R, self.ldv)
            self.dlogits = z(R, self.ldv)
            self.lse = e(R, dt=torch.float32)
            self.nll = z(B, dt=torch.float32)
            self.loss = z(1, dt=torch.float32)
            self.ctc_ws = e((ops.ctc_workspace_bytes(B, T2, max(smax, 1)) + 3) // 4, dt=torch.float32)
        # --- backward scratch -----------------------------------------------------------------
        self.g_a, self.g_b, self.g_c, self.g_d = e(R, C), e(R, C), e(R, C), e(R, C)
        self.g_f = e(R, F)
        self.g_qkv = e(R, 3 * C)
        self.g_p = z(B * H, T2, self.ld)
        self.g_s = z(B * H, T2, self.ld)
        self.g_pre = [e(B * self.t_sub[i + 1], eng.sub_filters[i]) for i in range(c.subsampling_layers)]
        self.g_col = [e(B * self.t_sub[i + 1], k * chans[i]) for i in range(c.subsampling_layers)]
        self.xcol = [e(B * self.t_sub[i + 1], k * chans[i]) for i in range(c.subsampling_layers)]
        self.g_sub = [e(B, self.t_sub[i], chans[i]) for i in range(c.subsampling_layers)]
        lib = ops._lib.load()
        cs = 0
        for rows, cols in [(R, 3 * C), (R, F), (R, self.ldv)] + [(p.shape[0], p.shape[1]) for p in self.pre]:
            cs = max(cs, lib.s2t_colsum_workspace_floats(rows, cols))
        self.colsum_ws = e(cs, dt=torch.float32)
        self.ln_part = e(lib.s2t_layernorm_bwd_partials(R, C), 2, C, dt=torch.float32)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.graph_kind: Optional[str] = None


class S2TEncoderEngine:
    SEED = 0x5217B200

    def __init__(self, cfg: EncoderConfig, device: Optional[torch.device] = None):
        if not torch.cuda.is_available():
            raise ops._lib.S2TLibraryError("S2TEncoderEngine needs a CUDA device (no CPU fallback)")
        ops._lib.load()
        if cfg.attn_type != "selfattn" or cfg.macaron_style or cfg.use_cnn_module:
            raise NotImplementedError("engine: rel_pos / macaron / conv-module blocks are not built yet")
        if cfg.encoder_activation_fn != "relu":
            raise NotImplementedError("engine: only the relu FFN is built")
        self.cfg = cfg
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        c = cfg
        C = c.embed_dim
        assert C % c.heads == 0 and (C // c.heads) % 8 == 0 and C % 8 == 0 and c.input_dim % 8 == 0
        # Conv1dSubsampling filters: [filter]*(L-1) + [d], last one doubled for the GLU (subsampling.py:231-247,118-121)
        filt = [c.subsampling_filter] * (c.subsampling_layers - 1) + [C]
        self.sub_filters = [f if i < c.subsampling_layers - 1 else 2 * f for i, f in enumerate(filt)]
        S = self.store = ParamStore(self.device)
        self.conv_names: List[str] = []
        cin = c.input_dim
        for i, cout in enumerate(self.sub_filters):
            n = f"subsample.layers.{i}.0.weight"
            S.declare(n, (cout, c.subsampling_kernel, cin))  # kernel layout [Cout][k][Cin]; reference is [Cout][Cin][k]
            self.conv_names.append(n)
            S.declare(f"subsample.layers.{i}.0.bias", (cout,))
            cin = cout // 2
        if c.encoder_embed_norm:
            S.declare("embed_ln.weight", (C,)); S.declare("embed_ln.bias", (C,))
        for l in range(c.layers):
            p = f"layers.{l}."
            S.declare(p + "self_attn_layer_norm.weight", (C,)); S.declare(p + "self_attn_layer_norm.bias", (C,))
            S.declare(p + "self_attn.q_proj.weight", (C, C))
            S.declare(p + "self_attn.k_proj.weight", (C, C), pack=True)
            S.declare(p + "self_attn.v_proj.weight", (C, C), pack=True)
            S.declare(p + "self_attn.q_proj.bias", (C,))
            S.declare(p + "self_attn.k_proj.bias", (C,), pack=True)
            S.declare(p + "self_attn.v_proj.bias", (C,), pack=True)
            S.declare(p + "self_attn.out_proj.weight", (C, C)); S.declare(p + "self_attn.out_proj.bias", (C,))
            S.declare(p + "ffn_norm.weight", (C,)); S.declare(p + "ffn_norm.bias", (C,))
            S.declare(p + "ffn.w_1.weight", (c.ffn_dim, C)); S.declare(p + "ffn.w_1.bias", (c.ffn_dim,))
            S.declare(p + "ffn.w_2.weight", (C, c.ffn_dim)); S.declare(p + "ffn.w_2.bias", (C,))
        S.declare("layer_norm.weight", (C,)); S.declare("layer_norm.bias", (C,))
        if c.vocab > 0:
            S.declare("ctc.ctc_projection.weight", (c.vocab, C))
            S.declare("ctc.ctc_projection.bias", (_round8(c.vocab),))  # padded to the logits' row stride
        S.finalize()
        self.embed_scale = 1.0 if c.no_scale_embedding else math.sqrt(C)
        self.pos_table = sinusoidal_table(c.max_source_positions + 2, C).to(self.device)
        self.step_counter = torch.zeros(1, device=self.device, dtype=torch.int64)
        self.plans: Dict[Tuple[int, int, int, int], Plan] = {}
        self.training = True
        self.grad_ready_hook = None  # callable(offset): store.grad[offset:] is final (dp.BucketedAllReduce.on_ready)
        self.dp = None               # dp.BucketedAllReduce when data parallel
        self.init_parameters()

    # ------------------------------------------------------------------------------------------
    # parameters
    # ------------------------------------------------------------------------------------------
    def init_parameters(self, seed: int = 0) -> None:
        """Reference initialisers: Conv1d / Linear default (kaiming-uniform a=sqrt(5) == U(+-1/sqrt(fan_in))),
        xavier-uniform attention projections (multihead_attention.py:95-113, q/k/v gain 1/sqrt(2)), LayerNorm 1/0,
        CTC projection N(0, C^-0.5) with zero bias (modules/speech_to_text/ctc.py:27-33)."""
        g = torch.Generator().manual_seed(seed)
        S = self.store
        for name in S.names():
            shape = S.shape(name)
            if name.endswith("norm.weight") or name == "embed_ln.weight":
                t = torch.ones(shape)
            elif name.endswith("norm.bias") or name == "embed_ln.bias":
                t = torch.zeros(shape)
            elif name == "ctc.ctc_projection.weight":
                t = torch.randn(shape, generator=g) * self.cfg.embed_dim ** -0.5
            elif name == "ctc.ctc_projection.bias":
                t = torch.zeros(shape)
            elif name.endswith("_proj.weight"):
                gain = 1.0 if "out_proj" in name else 2 ** -0.5
                bound = gain * math.sqrt(6.0 / (shape[0] + shape[1]))
                t = (torch.rand(shape, generator=g) * 2 - 1) * bound
            elif name.endswith("out_proj.bias"):
                t = torch.zeros(shape)
            else:
                if name.endswith("bias"):
                    w = S.shape(name[:-4] + "weight")
                    fan_in = int(math.prod(w[1:]))
                else:
                    fan_in = int(math.prod(shape[1:]))
                bound = 1.0 / math.sqrt(fan_in)
                t = (torch.rand(shape, generator=g) * 2 - 1) * bound
            S.p(name).copy_(t.to(BF16))

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        """Load a reference-layout state_dict (names of SURVEY.md §8b)."""
        S = self.store
        for name in S.names():
            if name not in sd:
                raise KeyError(f"state_dict is missing {name}")
            t = sd[name].detach().to(torch.float32)
            if name in self.conv_names:
                t = t.permute(0, 2, 1).contiguous()               # [Cout][Cin][k] -> [Cout][k][Cin]
            if name == "ctc.ctc_projection.bias":
                full = torch.zeros(S.shape(name))
                full[: t.numel()] = t
                t = full
            if tuple(t.shape) != S.shape(name):
                raise ValueError(f"{name}: shape {tuple(t.shape)} != {S.shape(name)}")
            S.p(name).copy_(t.to(BF16))

    def _export(self, grad: bool) -> Dict[str, torch.Tensor]:
        S = self.store
        out = {}
        for name in S.names():
            t = (S.g(name) if grad else S.p(name)).detach().float()
            if name in self.conv_names:
                t = t.permute(0, 2, 1).contiguous()
            if name == "ctc.ctc_projection.bias":
                t = t[: self.cfg.vocab]
            out[name] = t
        return out

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return self._export(False)

    def grads(self) -> Dict[str, torch.Tensor]:
        """Gradients in the reference layout (fp32 copies)."""
        return self._export(True)

    def num_params(self) -> int:
        return self.store.num_params() - (_round8(self.cfg.vocab) - self.cfg.vocab if self.cfg.vocab else 0)

    def train(self, mode: bool = True):
        self.training = mode
        return self

    def eval(self):
        return self.train(False)

    # ------------------------------------------------------------------------------------------
    # plans
    # ------------------------------------------------------------------------------------------
    def plan(self, B: int, T: int, smax: int = 0, wave_len: int = 0) -> Plan:
        key = (B, T, smax, wave_len)
        if key not in self.plans:
            self.plans[key] = Plan(self, B, T, smax, wave_len)
        return self.plans[key]

    def _drop(self, site: int):
        """Dropout spec of call site ``site`` (unique stream id per site; mask regenerable in backward)."""
        p = self.cfg.dropout if self.training else 0.0
        if p <= 0.0:
            return None
        return (p, self.SEED, site, self.step_counter)

    # ------------------------------------------------------------------------------------------
    # forward
    # ------------------------------------------------------------------------------------------
    def _stage_inputs(self, P: Plan) -> None:
        """Features (or waveform -> fbank -> utterance CMVN) into the zero-framed bf16 conv input, and the
        length arithmetic of the subsampler (subsampling.py:155: len = (len - 1) // 2 + 1 per layer)."""
        c = self.cfg
        fr = P.framed[0]
        dst_off = P.pad * c.input_dim
        if P.wave is not None:
            mel_w, rng, win = frontend._tables(str(self.device), c.input_dim)
            ops.fbank(P.wave, P.num_samples, mel_w, rng, win, P.T, out=P.fbank_raw)
            src = P.fbank_raw
        else:
            src = P.feats_in
        cm = 1 if P.wave is not None else 0
        # utterance CMVN (wave path) or a plain cast (feature path): fp32 -> bf16 into the framed buffer,
        # frames >= src_len zero filled (collater padding, speech_to_text_dataset.py:267-285)
        ops._call("s2t_cmvn_utt", ops._p(src), src.stride(1), src.stride(0), ops._p(fr) + 2 * dst_off, 1,
                  fr.stride(1), fr.stride(0), ops._p(P.src_len), P.B, P.T, c.input_dim, cm, cm, ops._stream())
        ln = P.src_len
        k, st = c.subsampling_kernel, c.subsampling_stride
        for i in range(c.subsampling_layers):
            torch.div(ln + (2 * P.pad - k), st, rounding_mode="floor", out=P.sub_len[i])
            P.sub_len[i] += 1
            ln = P.sub_len[i]

    def _subsample_fwd(self, P: Plan) -> None:
        """Conv1dSubsampling (subsampling.py:145-159): [Conv1d(k, stride, pad) -> GLU] per layer as an
        implicit GEMM whose A rows overlap in the zero-framed channels-last input; the last GLU also applies
        the encoder's padding mask (s2t_transformer.py:1765)."""
        c, S = self.cfg, self.store
        k, st, n = c.subsampling_kernel, c.subsampling_stride, c.subsampling_layers
        for i in range(n):
            xin = P.framed[i]
            cin, cout, t_out = xin.shape[2], self.sub_filters[i], P.t_sub[i + 1]
            w = S.p(f"subsample.layers.{i}.0.weight").view(cout, k * cin)
            ops.gemm_raw(M=t_out, N=cout, K=k * cin, a=xin, lda=st * cin, a_bs=(xin.stride(0), 0),
                         b=w, ldb=k * cin, out=P.pre[i], ldc=cout, c_bs=(t_out * cout, 0), batch=(P.B, 1),
                         bias=S.p(f"subsample.layers.{i}.0.bias"))
            if i < n - 1:
                nxt = P.framed[i + 1]
                ops.glu_fwd(P.pre[i], out=nxt, T=t_out, ldy=cout // 2, y_bs=nxt.stride(0), y_off=P.pad * (cout // 2))
            else:
                ops.glu_fwd(P.pre[i], out=P.x0, lengths=P.sub_len[i], T=t_out)

    def _embed_fwd(self, P: Plan) -> None:
        """embed_ln -> x * embed_scale + sinusoidal positions -> dropout (s2t_transformer.py:1768-1794)."""
        c, S = self.cfg, self.store
        src = P.x0
        if c.encoder_embed_norm:
            ops.layernorm_fwd(P.x0, S.p("embed_ln.weight"), S.p("embed_ln.bias"), out=P.e_ln, stats=P.e_st)
            src = P.e_ln
        C = c.embed_dim
        ops.embed_fwd(src.view(P.B, P.T2, C), P.sub_len[-1], self.pos_table, self.embed_scale,
                      self._drop(1), out=P.x[0].view(P.B, P.T2, C))

    def _layer_fwd(self, P: Plan, l: int) -> None:
        """S2TTransformerEncoderLayer.forward, pre-norm (s2t_transformer_layer.py:267-317) with fairseq
        MultiheadAttention (multihead_attention.py:246-265 projections, q * d_k^-0.5; :383-395 key padding;
        fp32 softmax + dropout on the probabilities; :372-416 bmm + out_proj)."""
        c, S = self.cfg, self.store
        L, x = P.layers[l], P.x[l]
        B, T, C, H, F = P.B, P.T2, c.embed_dim, c.heads, c.ffn_dim
        dk, ld = C // H, P.ld
        p, site = f"layers.{l}.", 16 * (l + 1)
        ops.layernorm_fwd(x, S.p(p + "self_attn_layer_norm.weight"), S.p(p + "self_attn_layer_norm.bias"),
                          out=L.ln1, stats=L.st1)
        ops.linear(L.ln1, S.fused(p + "self_attn.q_proj.weight", 3 * C, C), S.fused(p + "self_attn.q_proj.bias", 3 * C, 0),
                   out=L.qkv)
        # scores[b,h] = d_k^-0.5 * q k^T  (bf16, like the reference's bf16 bmm), softmax in fp32
        ops.gemm_raw(M=T, N=T, K=dk, a=L.qkv, lda=3 * C, a_bs=(dk, T * 3 * C), b=L.qkv, ldb=3 * C, b_off=C,
                     b_bs=(dk, T * 3 * C), out=P.scores, ldc=ld, c_bs=(T * ld, H * T * ld), batch=(H, B),
                     alpha=dk ** -0.5)
        dspec = self._drop(site + 0)
        ops.attn_softmax_fwd(P.scores, P.sub_len[-1], H, T, probs=L.probs,
                             probs_drop=L.probs_d if dspec else None, dropout=dspec)
        pd = L.probs_d if dspec else L.probs
        ops.gemm_raw(M=T, N=dk, K=T, a=pd, lda=ld, a_bs=(T * ld, H * T * ld), b=L.qkv, ldb=3 * C, b_off=2 * C,
                     b_mn=True, b_bs=(dk, T * 3 * C), out=L.ctx, ldc=C, c_bs=(dk, T * C), batch=(H, B))
        ops.linear(L.ctx, S.p(p + "self_attn.out_proj.weight"), S.p(p + "self_attn.out_proj.bias"), out=L.x1,
                   dropout=self._drop(site + 1), residual=x, ldr=C)
        ops.layernorm_fwd(L.x1, S.p(p + "ffn_norm.weight"), S.p(p + "ffn_norm.bias"), out=L.ln2, stats=L.st2)
        ops.linear(L.ln2, S.p(p + "ffn.w_1.weight"), S.p(p + "ffn.w_1.bias"), out=L.f, act=ops.ACT_RELU,
                   dropout=self._drop(site + 2))
        ops.linear(L.f, S.p(p + "ffn.w_2.weight"), S.p(p + "ffn.w_2.bias"), out=P.x[l + 1],
                   dropout=self._drop(site + 3), residual=L.x1, ldr=C)

    def _head_fwd(self, P: Plan) -> None:
        """Final LayerNorm (s2t_transformer.py:2112) and the CTC head: dropout -> Linear(C, V)
        (modules/speech_to_text/ctc.py:55-66, quirk A11)."""
        c, S = self.cfg, self.store
        ops.layernorm_fwd(P.x[c.layers], S.p("layer_norm.weight"), S.p("layer_norm.bias"), out=P.enc_out,
                          stats=P.enc_st)
        if c.vocab <= 0:
            return
        spec = self._drop(2)
        src = P.enc_out
        if spec:
            ops.dropout(P.enc_out, spec, out=P.enc_drop)
            src = P.enc_drop
        ops.linear(src, S.p("ctc.ctc_projection.weight"), S.p("ctc.ctc_projection.bias"), out=P.logits)

    def _forward(self, P: Plan) -> None:
        self._stage_inputs(P)
        self._subsample_fwd(P)
        self._embed_fwd(P)
        for l in range(self.cfg.layers):
            self._layer_fwd(P, l)
        self._head_fwd(P)

    def _loss(self, P: Plan, need_grad: bool) -> None:
        """fp32 log-softmax + CTCLoss(blank, 'none', zero_infinity=True).sum() and its gradient w.r.t. the
        logits (criterions/ctc.py:435-474); loss = ctc_weight * sum_b nll_b (:1004)."""
        c = self.cfg
        B, T, V = P.B, P.T2, c.vocab
        ops.row_lse(P.logits, V, out=P.lse)
        ops.ctc_loss(P.logits.view(B, T, P.ldv), P.targets, P.sub_len[-1], P.tgt_len, V, c.blank,
                     need_grad=need_grad, grad_scale=c.ctc_weight, lse=P.lse,
                     dlogits=P.dlogits.view(B, T, P.ldv) if need_grad else None, workspace=P.ctc_ws, nll=P.nll)
        ops.f32_sum(P.nll, P.loss, c.ctc_weight)

    # ------------------------------------------------------------------------------------------
    # backward
    # ------------------------------------------------------------------------------------------
    def _ready(self, first_name: str) -> None:
        """Everything at or after ``first_name`` in the flat gradient buffer is final."""
        if self.grad_ready_hook is not None:
            self.grad_ready_hook(self.store.offset(first_name))

    def _branch_grad(self, dy: torch.Tensor, spec, scale: float, out: torch.Tensor) -> torch.Tensor:
        """Gradient entering a residual branch y = x + scale * dropout(branch): scale * mask * dy / (1-p)."""
        if spec is None and scale == 1.0:
            return dy
        return ops.dropout(dy, spec, out=out, scale=scale)

    def _head_bwd(self, P: Plan) -> torch.Tensor:
        c, S = self.cfg, self.store
        src = P.enc_drop if self._drop(2) else P.enc_out
        ops.colsum(P.dlogits, S.g("ctc.ctc_projection.bias"), workspace=P.colsum_ws)
        ops.linear_wgrad(P.dlogits[:, : c.vocab] if c.vocab != P.ldv else P.dlogits, src,
                         out=S.g("ctc.ctc_projection.weight"))
        ops.gemm_raw(M=P.R, N=c.embed_dim, K=c.vocab, a=P.dlogits, lda=P.ldv,
                     b=S.p("ctc.ctc_projection.weight"), ldb=c.embed_dim, b_mn=True, out=P.g_a, ldc=c.embed_dim)
        d = self._branch_grad(P.g_a, self._drop(2), 1.0, P.g_b)
        self._ready("ctc.ctc_projection.weight")
        dx = P.g_c
        ops.layernorm_bwd(d, P.x[c.layers], S.p("layer_norm.weight"), P.enc_st, S.g("layer_norm.weight"),
                          S.g("layer_norm.bias"), out=dx, partials=P.ln_part)
        self._ready("layer_norm.weight")
        return dx

    def _layer_bwd(self, P: Plan, l: int, dy: torch.Tensor) -> torch.Tensor:
        """Backward of _layer_fwd.  ``dy`` = dL/dx[l+1]; returns dL/dx[l] (in one of the g_* scratch buffers)."""
        c, S = self.cfg, self.store
        L, x = P.layers[l], P.x[l]
        B, T, C, H, F = P.B, P.T2, c.embed_dim, c.heads, c.ffn_dim
        dk, ld = C // H, P.ld
        p, site = f"layers.{l}.", 16 * (l + 1)
        free = [b for b in (P.g_a, P.g_b, P.g_c, P.g_d) if b is not dy]
        t0, t1, t2 = free[0], free[1], free[2]
        # ---- FFN: x2 = x1 + drop3(w_2 drop2(relu(w_1 LN(x1)))) ------------------------------------------
        d2 = self._branch_grad(dy, self._drop(site + 3), 1.0, t0)
        ops.colsum(d2, S.g(p + "ffn.w_2.bias"), workspace=P.colsum_ws)
        ops.linear_wgrad(d2, L.f, out=S.g(p + "ffn.w_2.weight"))
        ops.linear_dgrad(d2, S.p(p + "ffn.w_2.weight"), out=P.g_f, gate=L.f, gate_mode=ops.GATE_POSITIVE, ldg=F,
                         dropout=self._drop(site + 2))
        ops.colsum(P.g_f, S.g(p + "ffn.w_1.bias"), workspace=P.colsum_ws)
        ops.linear_wgrad(P.g_f, L.ln2, out=S.g(p + "ffn.w_1.weight"))
        ops.linear_dgrad(P.g_f, S.p(p + "ffn.w_1.weight"), out=t0)
        dx1 = t1
        ops.layernorm_bwd(t0, L.x1, S.p(p + "ffn_norm.weight"), L.st2, S.g(p + "ffn_norm.weight"),
                          S.g(p + "ffn_norm.bias"), dx_residual=dy, out=dx1, partials=P.ln_part)
        # ---- attention: x1 = x + drop1(out_proj(P_drop V)) ----------------------------------------------
        do = self._branch_grad(dx1, self._drop(site + 1), 1.0, t0)
        ops.colsum(do, S.g(p + "self_attn.out_proj.bias"), workspace=P.colsum_ws)
        ops.linear_wgrad(do, L.ctx, out=S.g(p + "self_attn.out_proj.weight"))
        dctx = t2
        ops.linear_dgrad(do, S.p(p + "self_attn.out_proj.weight"), out=dctx)
        dspec = self._drop(site + 0)
        pd = L.probs_d if dspec else L.probs
        bs_x, bs_qkv, bs_s = (dk, T * C), (dk, T * 3 * C), (T * ld, H * T * ld)
        # dP_drop = dctx V^T
        ops.gemm_raw(M=T, N=T, K=dk, a=dctx, lda=C, a_bs=bs_x, b=L.qkv, ldb=3 * C, b_off=2 * C, b_bs=bs_qkv,
                     out=P.g_p, ldc=ld, c_bs=bs_s, batch=(H, B))
        # dV = P_drop^T dctx
        ops.gemm_raw(M=T, N=dk, K=T, a=pd, lda=ld, a_mn=True, a_bs=bs_s, b=dctx, ldb=C, b_mn=True, b_bs=bs_x,
                     out=P.g_qkv, c_off=2 * C, ldc=3 * C, c_bs=bs_qkv, batch=(H, B))
        ops.attn_softmax_bwd(L.probs, P.g_p, P.sub_len[-1], H, T, out=P.g_s, dropout=dspec)
        # dQ = s dS K ; dK = s dS^T Q
        ops.gemm_raw(M=T, N=dk, K=T, a=P.g_s, lda=ld, a_bs=bs_s, b=L.qkv, ldb=3 * C, b_off=C, b_mn=True,
                     b_bs=bs_qkv, out=P.g_qkv, c_off=0, ldc=3 * C, c_bs=bs_qkv, batch=(H, B), alpha=dk ** -0.5)
        ops.gemm_raw(M=T, N=dk, K=T, a=P.g_s, lda=ld, a_mn=True, a_bs=bs_s, b=L.qkv, ldb=3 * C, b_off=0,
                     b_mn=True, b_bs=bs_qkv, out=P.g_qkv, c_off=C, ldc=3 * C, c_bs=bs_qkv, batch=(H, B),
                     alpha=dk ** -0.5)
        ops.colsum(P.g_qkv, S.fused(p + "self_attn.q_proj.bias", 3 * C, 0, grad=True), workspace=P.colsum_ws)
        ops.linear_wgrad(P.g_qkv, L.ln1, out=S.fused(p + "self_attn.q_proj.weight", 3 * C, C, grad=True))
        ops.linear_dgrad(P.g_qkv, S.fused(p + "self_attn.q_proj.weight", 3 * C, C), out=t0)
        dx = t2
        ops.layernorm_bwd(t0, x, S.p(p + "self_attn_layer_norm.weight"), L.st1,
                          S.g(p + "self_attn_layer_norm.weight"), S.g(p + "self_attn_layer_norm.bias"),
                          dx_residual=dx1, out=dx, partials=P.ln_part)
        self._ready(p + "self_attn_layer_norm.weight")
        return dx

    def _embed_bwd(self, P: Plan, dy: torch.Tensor) -> torch.Tensor:
        c, S = self.cfg, self.store
        C = c.embed_dim
        t = P.g_a if dy is not P.g_a else P.g_b
        ops.embed_bwd(dy.view(P.B, P.T2, C), self.embed_scale, self._drop(1), out=t.view(P.B, P.T2, C))
        if not c.encoder_embed_norm:
            return t
        o = P.g_c if dy is not P.g_c else P.g_d
        ops.layernorm_bwd(t, P.x0, S.p("embed_ln.weight"), P.e_st, S.g("embed_ln.weight"), S.g("embed_ln.bias"),
                          out=o, partials=P.ln_part)
        self._ready("embed_ln.weight")
        return o

    def _subsample_bwd(self, P: Plan, dy: torch.Tensor) -> None:
        """Backward of Conv1dSubsampling: GLU', bias = column sums, weight = dY^T im2col(X), input = col2im(dY W)."""
        c, S = self.cfg, self.store
        k, st, n = c.subsampling_kernel, c.subsampling_stride, c.subsampling_layers
        for i in reversed(range(n)):
            xin = P.framed[i]
            cin, cout, t_out = xin.shape[2], self.sub_filters[i], P.t_sub[i + 1]
            if i == n - 1:
                ops.glu_bwd(P.pre[i], dy, out=P.g_pre[i], lengths=P.sub_len[i], T=t_out)
            else:
                ops.glu_bwd(P.pre[i], dy, out=P.g_pre[i])
            ops.colsum(P.g_pre[i], S.g(f"subsample.layers.{i}.0.bias"), workspace=P.colsum_ws)
            ops.im2col1d(xin, xin.stride(0), P.B, t_out, cin, k, st, out=P.xcol[i])
            ops.linear_wgrad(P.g_pre[i], P.xcol[i], out=S.g(f"subsample.layers.{i}.0.weight").view(cout, k * cin))
            if i > 0:
                ops.linear_dgrad(P.g_pre[i], S.p(f"subsample.layers.{i}.0.weight").view(cout, k * cin), out=P.g_col[i])
                ops.col2im1d(P.g_col[i], P.B, P.t_sub[i], t_out, cin, k, st, P.pad, out=P.g_sub[i])
                dy = P.g_sub[i].view(P.B * P.t_sub[i], cin)
            self._ready(f"subsample.layers.{i}.0.weight")

    def _backward(self, P: Plan) -> None:
        d = self._head_bwd(P)
        for l in reversed(range(self.cfg.layers)):
            d = self._layer_bwd(P, l, d)
        d = self._embed_bwd(P, d)
        self._subsample_bwd(P, d)

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
                   targets: Optional[torch.Tensor] = None, target_lengths: Optional[torch.Tensor] = None) -> None:
        """Copy one batch into the plan's static input buffers (async on the current stream; host tensors
        should be pinned).  feats fp32 [B, T, D] + src_lengths, or wave fp32 [B, N] (int16 scale) + num_samples."""
        nb = True
        if wave is not None:
            P.wave.copy_(wave, non_blocking=nb)
            P.num_samples.copy_(num_samples, non_blocking=nb)
            # frames per utterance, snip_edges=True (torchaudio kaldi.fbank): 1 + (n - 400) // 160
            torch.div(P.num_samples - frontend.FRAME_LEN, frontend.FRAME_SHIFT, rounding_mode="floor", out=P.src_len)
            P.src_len += 1
            P.src_len.clamp_(min=0)
        else:
            P.feats_in.copy_(feats, non_blocking=nb)
            P.src_len.copy_(src_lengths, non_blocking=nb)
        if targets is not None:
            P.targets[:, : targets.shape[1]].copy_(targets, non_blocking=nb)
            P.tgt_len.copy_(target_lengths, non_blocking=nb)

    def forward(self, P: Plan) -> None:
        """Encoder forward on the plan's inputs; results in P.enc_out [B*T', C], P.logits [B*T', ldv], P.sub_len[-1]."""
        self._forward(P)

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

    def enable_data_parallel(self, n_buckets: int = 4, group=None) -> None:
        """Bucketed gradient all-reduce overlapped with backward (dp.py); cut points = the offsets at which the
        backward pass reports progress (start of every layer / block in the flat gradient buffer)."""
        from .dp import BucketedAllReduce
        S = self.store
        cuts = [S.offset(n) for n in S.names()
                if n.endswith("self_attn_layer_norm.weight") or n in ("layer_norm.weight", "ctc.ctc_projection.weight",
                                                                      "embed_ln.weight")
                or (n.startswith("subsample.") and n.endswith("weight"))]
        self.dp = BucketedAllReduce(S.grad, cuts, n_buckets, group)
        self.grad_ready_hook = self.dp.on_ready

    def capture(self, P: Plan, kind: str = "train") -> None:
        """Capture train_step (or forward-only 'infer') of this plan into a CUDA graph."""
        fn = self.train_step if kind == "train" else self.forward
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn(P)                                # warm-up: function attributes, tensor-map cache
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn(P)
        P.graph, P.graph_kind = g, kind

    def run(self, P: Plan, kind: str = "train") -> None:
        if P.graph is not None and P.graph_kind == kind:
            P.graph.replay()
        elif kind == "train":
            self.train_step(P)
        else:
            self.forward(P)

    def greedy_decode(self, P: Plan):
        """CTCDecoder greedy (s2t_ctc.py:312-347) on the plan's logits: int32 tokens [B, T'] (-1 padded), lengths."""
        c = self.cfg
        return ops.ctc_greedy(P.logits.view(P.B, P.T2, P.ldv), P.sub_len[-1], c.vocab, c.blank)
