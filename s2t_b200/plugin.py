"""The reference's plugin surface, backed by the B200 engine.

A fairseq ``--user-dir`` package imports this module and calls :func:`register`; the arch names
``s2t_transformer*`` / ``pdss2t_transformer*`` / ``s2t_ctc*`` then build encoders that keep the reference contracts:

* ``S2TTransformerEncoderB200`` / ``PDSS2TTransformerEncoderB200`` — ``FairseqEncoder`` with
  ``__init__(args, task=None, embed_tokens=None)`` and ``forward(src_tokens (B,T,80), src_lengths (B), **kw) -> dict``
  with exactly the keys of ``s2t_transformer.py:2142-2154`` / ``pdss2t_transformer.py:1267-1281``, plus
  ``reorder_encoder_out``, ``set_num_updates``, ``set_ctc_infer``, ``ctc_valid``, ``set_flag``, ``dump``,
  ``max_positions``, ``pooling_ratio`` and the ``pae_ground_truth_ratio`` attribute the criterion reads
  (``criterions/ctc.py:262``).
* ``S2TCTCModelB200`` — ``forward(src_tokens, src_lengths, prev_output_tokens=None)`` returns the encoder dict,
  ``get_normalized_probs`` the fp32 (log-)softmax (``s2t_ctc.py:98-126``).
* ``CtcCriterionB200`` — ``forward(model, sample, reduce=True) -> (loss, sample_size, logging_output)``
  (``criterions/ctc.py:258-281``), un-normalised summed loss (quirk A12), same target plumbing
  (``get_targets_for_ctc_loss``, ``:516-540``) and weighting (``:1004-1015``).
* ``CTCDecoderB200.generate(models, sample) -> List[List[dict]]`` (``s2t_ctc.py:236-349``, greedy).

Training integration: the whole model is ONE flat bf16 ``nn.Parameter`` (``encoder.flat``, the engine's parameter
store; ``state_dict()`` / ``load_state_dict()`` speak the reference's per-tensor names).  The criterion runs the engine's
fused forward + CTC loss + backward and returns the loss through an autograd node whose backward hands the finished
flat gradient to ``flat.grad`` (scaled by the incoming gradient, i.e. fairseq's loss scale) — so the reference trainer's
``optimizer.backward(loss)`` / ``multiply_grads`` / ``clip_grad_norm`` / ``step`` sequence runs unchanged.

No CPU fallback: constructing an encoder without a CUDA device raises.
"""
from __future__ import annotations

import math
from argparse import Namespace
from typing import Dict, List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .batching import BucketSet
from .engine import EncoderConfig, Plan, S2TEncoderEngine

try:  # the reference package, when this runs inside a fairseq process
    from fairseq.models import FairseqEncoder as _EncoderBase  # type: ignore
    _HAVE_FAIRSEQ = True
except Exception:  # pragma: no cover - fairseq is not importable in the build container
    _HAVE_FAIRSEQ = False

    class _EncoderBase(nn.Module):  # the two things FairseqEncoder adds that the path touches
        def __init__(self, dictionary=None):
            super().__init__()
            self.dictionary = dictionary

        def set_num_updates(self, num_updates):
            self.num_updates = num_updates


S2T_ARCHS = ("s2t_transformer", "s2t_transformer_s", "s2t_transformer_s_relative", "s2t_transformer_xs",
             "s2t_transformer_sp", "s2t_transformer_m", "s2t_transformer_mp", "s2t_transformer_l", "s2t_transformer_lp")
PDS_ARCHS = ("pdss2t_transformer", "pdss2t_transformer_s", "pdss2t_transformer_s_8", "pdss2t_transformer_s_16",
             "pdss2t_transformer_s_32", "pdss2t_transformer_sd", "pdss2t_transformer_sd_8", "pdss2t_transformer_sd_16",
             "pdss2t_transformer_sd_32", "pdss2t_transformer_m", "pdss2t_transformer_m_8", "pdss2t_transformer_m_16",
             "pdss2t_transformer_m_32", "pdss2t_transformer_l", "pdss2t_transformer_l_8", "pdss2t_transformer_l_16",
             "pdss2t_transformer_l_32")
CTC_ARCHS = ("s2t_ctc", "s2t_ctc_s", "s2t_ctc_s_relative", "s2t_ctc_xs", "s2t_ctc_sp", "s2t_ctc_m", "s2t_ctc_mp",
             "s2t_ctc_l", "s2t_ctc_lp")


def _bucket(n: int, mult: int = 32) -> int:
    """Target-length bucket of a plan (so a handful of CUDA-graph-able plans serve all batches)."""
    return max(mult, (int(n) + mult - 1) // mult * mult)


class _EngineStep(torch.autograd.Function):
    """loss = fused(forward + CTC losses + backward) of the engine; backward publishes the flat gradient."""

    @staticmethod
    def forward(ctx, flat: torch.Tensor, encoder: "S2TTransformerEncoderB200", plan: Plan):
        eng = encoder.engine
        eng.run(plan, "train")
        ctx.encoder = encoder
        return plan.loss[0].clone()

    @staticmethod
    def backward(ctx, grad_out):
        g = ctx.encoder.engine.store.grad
        # fairseq scales the loss (fp16 loss scale / update_freq): d loss_scaled / d flat = grad_out * d loss / d flat
        return g * grad_out.to(g.dtype), None, None


class _EncoderForward(torch.autograd.Function):
    """Differentiable encoder forward for losses computed OUTSIDE the engine — the AR decoder + (label-smoothed CE [+
    CTC]) criterion of the ``s2t_transformer*`` / ``pdss2t_transformer*`` models (s2t_transformer.py:868-880).
    Outputs: encoder_out [B*T', C] and every CTC logit set [B*T', V] (top heads, then inter blocks) as fresh tensors;
    backward copies their gradients into the engine (``backward_external``) and returns the flat parameter gradient."""

    @staticmethod
    def forward(ctx, flat: torch.Tensor, encoder: "S2TTransformerEncoderB200", plan: Plan):
        eng = encoder.engine
        eng.forward_train(plan)
        keys, outs = [], [plan.enc_out.clone()]
        for k, hs in enumerate(eng.heads):
            keys.append(("top", k))
            outs.append(plan.heads[k].logits[:, : hs.V].clone())
        for k, hs in enumerate(eng.heads):
            for l in hs.inter_layers:
                keys.append(("inter", l, k))
                outs.append(plan.inter[(l, k)].logits[:, : hs.V].clone())
        ctx.encoder, ctx.plan, ctx.keys = encoder, plan, keys
        return tuple(outs)

    @staticmethod
    def backward(ctx, d_enc, *d_logits):
        eng = ctx.encoder.engine
        eng.backward_external(ctx.plan, d_enc, {k: g for k, g in zip(ctx.keys, d_logits) if g is not None})
        return eng.store.grad.clone(), None, None


class S2TTransformerEncoderB200(_EncoderBase):
    """Drop-in for ``S2TTransformerEncoder`` (fairseq/models/speech_to_text/s2t_transformer.py:883)."""

    FAMILY = "s2t"

    def __init__(self, args: Namespace, task=None, embed_tokens=None):
        super().__init__(None)
        src_dict = getattr(task, "source_dictionary", None) if task is not None else None
        tgt_dict = getattr(task, "target_dictionary", None) if task is not None else None
        vocab = len(src_dict) if src_dict is not None else int(getattr(args, "_src_vocab", 0))
        if tgt_dict is not None and not hasattr(args, "_tgt_vocab"):
            args._tgt_vocab = len(tgt_dict)
        self.args = args
        self.padding_idx = 1
        self.cfg = EncoderConfig.from_args(args, vocab, family=self.FAMILY)
        self.engine = S2TEncoderEngine(self.cfg)          # raises without a CUDA device (no CPU fallback)
        # the whole model is one flat bf16 parameter; .grad is the engine's flat gradient buffer
        self.flat = nn.Parameter(self.engine.store.flat, requires_grad=True)
        # ground-truth curriculum of the prediction-aware adapters (s2t_transformer.py:947-949, 1050-1052, 1129-1148):
        # the criterion reads the sum (criterions/ctc.py:262) and, when non-zero, aligns the batch first (_curriculum)
        g = lambda k, d: getattr(args, k, d)
        self.gt_ratio = {"ctc": float(g("ctc_pae_ground_truth_ratio", 0) or 0), "xctc": float(g("xctc_pae_ground_truth_ratio", 0) or 0)}
        self.pae_adaptive_gt = bool(g("xctc_pae_ground_truth_ratio_adaptive", False))
        self.pae_gt_only_mistake = bool(g("xctc_pae_ground_truth_only_mistake", False))
        self.pae_ground_truth_ratio = self.gt_ratio["ctc"] + self.gt_ratio["xctc"]
        self.num_updates = 0
        self.ctc_infer_cfg = None
        self._last_plan: Optional[Plan] = None
        self.use_graph = bool(getattr(args, "b200_cuda_graph", True))
        # frame buckets of the plan cache (batching.BucketSet).  BatchNorm statistics in training mode run over the padded
        # frames of the batch (convolution.py:60, quirk A4), so models with BatchNorm keep the batch's exact frame count.
        bn = self.cfg.use_cnn_module or (self.cfg.subsampling_type == "conv2d" and self.cfg.subsampling_norm == "batch2d")
        self.buckets: Optional[BucketSet] = None if bn else BucketSet(max_frames=self.cfg.max_source_positions)

    # ---- reference state_dict names ---------------------------------------------------------------------
    def state_dict(self, destination=None, prefix="", keep_vars=False):
        out = destination if destination is not None else {}
        for k, v in self.engine.state_dict().items():
            out[prefix + k] = v
        return out

    def load_state_dict(self, state_dict, strict=True, **kw):
        self.engine.load_state_dict(state_dict)
        return nn.modules.module._IncompatibleKeys([], [])

    def upgrade_state_dict_named(self, state_dict, name):
        return state_dict

    # ---- the FairseqEncoder contract -----------------------------------------------------------------------
    def _plan_for(self, src_tokens: torch.Tensor, smax: int) -> Plan:
        B, T, D = src_tokens.shape
        if D != self.cfg.input_dim:
            raise ValueError(f"src_tokens has {D} features per frame, the encoder was built for {self.cfg.input_dim}")
        Tq = self.buckets.frames(T) if self.buckets is not None else T
        P = self.engine.plan(B, Tq, _bucket(smax))
        P.batch_frames = T          # frames of THIS batch: outputs are sliced to what the reference would return for it
        return P

    def output_dict(self, P: Plan, diff: Optional[tuple] = None) -> Dict[str, list]:
        """The encoder_out dict (s2t_transformer.py:2142-2154) as stride views of the plan's buffers (or, with ``diff``,
        of the autograd-tracked tensors ``_EncoderForward`` returned), cut to the output frames of the batch."""
        eng = self.engine
        B, T2, C = P.B, P.T2, eng.out_dim
        Tv = min(T2, eng.out_frames(getattr(P, "batch_frames", P.T)))
        tbc = lambda t, n: t.view(B, T2, -1)[:, :Tv, :n].transpose(0, 1)
        pad = torch.arange(Tv, device=P.out_len.device)[None, :] >= P.out_len[:, None]
        enc_src = diff[0] if diff is not None else P.enc_out
        out = {"encoder_out": [tbc(enc_src, C)], "ctc_logit": [], "inter_ctc_logits": [], "xctc_logit": [],
               "inter_xctc_logits": [], "encoder_padding_mask": [pad], "mixup": None, "encoder_embedding": [],
               "encoder_states": [], "src_tokens": [], "src_lengths": []}
        it = iter(diff[1:]) if diff is not None else None
        tops = [next(it) if it is not None else P.heads[k].logits for k in range(len(eng.heads))]
        for k, hs in enumerate(eng.heads):
            top = tbc(tops[k], hs.V)
            inter = [tbc(next(it) if it is not None else P.inter[(l, k)].logits, hs.V) for l in hs.inter_layers]
            if hs.name == "ctc":
                out["ctc_logit"] = [top]
                out["inter_ctc_logits"] = [[lg, pad] for lg in inter]       # [logit, pad_mask] (quirk A13)
            else:
                out["xctc_logit"] = [top]
                out["inter_xctc_logits"] = inter                            # bare logits (quirk A13)
        if self.FAMILY == "pds":
            out["inter_xctc_logits"] = out["inter_ctc_logits"]              # sic, pdss2t_transformer.py:1274 (A8)
        return out

    def forward(self, src_tokens, src_lengths, **kwargs):
        """src_tokens (B, T, 80*channels) float, src_lengths (B) -> encoder_out dict.
        In training mode with autograd enabled the returned ``encoder_out`` / ``*ctc_logit*`` tensors carry an autograd
        graph back to the flat parameter (``_EncoderForward``): a decoder + criterion stacked on top trains the encoder
        exactly as decoder(encoder(.)) does in the reference (s2t_transformer.py:868-880).  Otherwise (eval / no_grad) the
        tensors are plain views.  The pure-CTC models train through ``CtcCriterionB200``'s fused step instead."""
        if not src_tokens.is_cuda:
            raise ops._lib.S2TLibraryError("S2TTransformerEncoderB200.forward needs CUDA tensors (no CPU fallback)")
        P = self._plan_for(src_tokens, 0)
        self.engine.train(self.training)
        self.engine.set_inputs(P, feats=src_tokens.float(), src_lengths=src_lengths.to(torch.int32))
        self._last_plan = P
        if self.training and torch.is_grad_enabled():
            return self.output_dict(P, diff=_EncoderForward.apply(self.flat, self, P))
        self.engine.forward(P)
        return self.output_dict(P)

    def reorder_encoder_out(self, encoder_out, new_order):
        """s2t_transformer.py:2156-2208: index_select dim 1 for T x B x *, dim 0 for masks."""
        sel1 = lambda xs: [x.index_select(1, new_order) for x in xs]
        sel0 = lambda xs: [x.index_select(0, new_order) for x in xs]
        return {
            "encoder_out": sel1(encoder_out["encoder_out"]),
            "ctc_logit": sel1(encoder_out["ctc_logit"]),
            "xctc_logit": sel1(encoder_out.get("xctc_logit", [])),
            "inter_ctc_logits": encoder_out.get("inter_ctc_logits", []),
            "inter_xctc_logits": encoder_out.get("inter_xctc_logits", []),
            "encoder_padding_mask": sel0(encoder_out["encoder_padding_mask"]),
            "mixup": encoder_out.get("mixup"),
            "encoder_embedding": sel0(encoder_out["encoder_embedding"]),
            "encoder_states": sel1(encoder_out["encoder_states"]),
            "src_tokens": [], "src_lengths": [],
        }

    def set_num_updates(self, num_updates):
        self.num_updates = num_updates

    def set_ctc_infer(self, ctc_infer, post_process, src_dict=None, tgt_dict=None, path=None):
        self.ctc_infer_cfg = dict(ctc_infer=ctc_infer, post_process=post_process, src_dict=src_dict, tgt_dict=tgt_dict,
                                  path=path)

    def ctc_valid(self, lprobs, targets, input_lengths, dictionary, lang="source"):
        """CTC.valid (modules/speech_to_text/ctc.py:96-136): greedy-decode lprobs (T, B, V) and count token errors."""
        from .decode import ctc_valid_counts
        return ctc_valid_counts(lprobs, targets, input_lengths, dictionary)

    def set_flag(self, **kwargs):
        for k, v in kwargs.items():
            setattr(self, k, v)

    def dump(self, fstream, info=""):
        print(f"{info} s2t_b200 encoder: {self.engine.num_params()} parameters, plans {sorted(self.engine.plans)}",
              file=fstream)

    def max_positions(self):
        return self.cfg.max_source_positions

    @staticmethod
    def pooling_ratio():
        return 4


class PDSS2TTransformerEncoderB200(S2TTransformerEncoderB200):
    """Drop-in for ``PDSS2TTransformerEncoder`` (fairseq/models/speech_to_text/pdss2t_transformer.py:226)."""

    FAMILY = "pds"

    @staticmethod
    def pooling_ratio():
        return 8


class S2TCTCModelB200(nn.Module):
    """Drop-in for ``S2TCTCModel`` (fairseq/models/speech_to_text/s2t_ctc.py:36-126): encoder-only model whose
    ``--encoder-type`` picks the S2T or PDS encoder (``:50-82``)."""

    def __init__(self, encoder: S2TTransformerEncoderB200):
        super().__init__()
        self.encoder = encoder

    @classmethod
    def build_encoder(cls, args, task=None, embed_tokens=None):
        kind = getattr(args, "encoder_type", "transformer")
        if kind == "pds":
            return PDSS2TTransformerEncoderB200(args, task, embed_tokens)
        if kind in ("transformer", "s2t", None):
            return S2TTransformerEncoderB200(args, task, embed_tokens)
        raise NotImplementedError(f"s2t_b200: --encoder-type {kind!r} (sate: SURVEY.md §8f rank 3)")

    @classmethod
    def build_model(cls, args, task):
        return cls(cls.build_encoder(args, task))

    def forward(self, src_tokens, src_lengths, prev_output_tokens=None, **kwargs):
        return self.encoder(src_tokens, src_lengths, **kwargs)

    def get_normalized_probs(self, net_output, log_probs, sample=None):
        """s2t_ctc.py:98-112: fp32 (log-)softmax of ctc_logit (a dict) or of a bare [logit] list."""
        logits = net_output["ctc_logit"][0] if isinstance(net_output, dict) else net_output[0]
        if isinstance(logits, list):
            logits = logits[0]
        fn = torch.log_softmax if log_probs else torch.softmax
        return fn(logits, dim=-1, dtype=torch.float32)

    def max_positions(self):
        return self.encoder.max_positions()

    def set_num_updates(self, num_updates):
        self.encoder.set_num_updates(num_updates)


class CtcCriterionB200(nn.Module):
    """Drop-in for ``CtcCriterion`` (fairseq/criterions/ctc.py:156-281) on a B200 encoder.

    ``forward(model, sample, reduce=True) -> (loss, sample_size, logging_output)``: loss is the un-normalised weighted
    sum over utterances (quirk A12; the trainer divides by sample_size, trainer.py:729-734).  In training mode the
    engine's fused step runs forward + every CTC term + backward in one pass (a CUDA graph per batch shape)."""

    def __init__(self, cfg=None, task=None, ctc_weight: float = 1.0, save_dir=None):
        super().__init__()
        g = lambda k, d: getattr(cfg, k, d) if cfg is not None else d
        self.task = task
        tgt = getattr(task, "target_dictionary", None) if task is not None else None
        self.blank_idx = 0
        self.pad_idx = tgt.pad() if tgt is not None else 1
        self.eos_idx = tgt.eos() if tgt is not None else 2
        self.sentence_avg = bool(g("sentence_avg", False))
        self.ctc_masked_loss = bool(g("ctc_masked_loss", False))          # criterions/ctc.py:92-95, 214
        self.post_process = g("post_process", "sentencepiece")
        self.ctc_weight = ctc_weight

    @staticmethod
    def logging_outputs_can_be_summed() -> bool:
        return True

    # criterions/ctc.py:516-540 — pad / eos stripped, left-packed int32 [B, S] + lengths
    def get_targets_for_ctc_loss(self, tokens: torch.Tensor):
        return S2TEncoderEngine.strip_targets(tokens, self.pad_idx, self.eos_idx)

    def _curriculum(self, enc, eng, P, Tv: Optional[int] = None) -> None:
        """``get_ground_truth_alignment`` (criterions/ctc.py:262-265, 283-433) and the mask draw of the encoder
        (s2t_transformer.py:1904-1936, 2061-2092), all on the device: a first forward pass without gradient, the best CTC
        alignment of its TOP logits against the targets (``s2t_ctc_viterbi`` = torch_imputer.best_alignment + backtrack),
        per frame the aligned ground-truth token, the mistake flags (argmax != aligned token) and the random mixing
        mask; the forced states are the alignment where the mask is set, -1 elsewhere.

        ``Tv``: output frames of THIS batch (``eng.out_frames(batch frames)``).  A plan of a larger frame bucket has more
        (masked) rows; they are not part of the reference's tensors, so mistake counts, the adaptive ratio and the shape
        of the random draw are taken over the first Tv frames only and the rest is filled with token 0 / mask 0 / free."""
        eng.forward_train(P)
        B, T2 = P.B, P.T2
        Tv = T2 if Tv is None else min(int(Tv), T2)
        for k, hs in enumerate(eng.heads):
            ratio = enc.gt_ratio.get(hs.name, 0.0)
            if ratio <= 0 or not hs.pae or not hs.inter_layers:
                eng.set_ground_truth(P, k, None, None, None)
                continue
            lp = torch.log_softmax(P.heads[k].logits.view(B, T2, hs.ldv)[:, :Tv, : hs.V], dim=-1, dtype=torch.float32)
            states, _ = ops.ctc_viterbi(lp.contiguous(), P.tg[hs.tgt], P.out_len, P.tl[hs.tgt], self.blank_idx,
                                        with_nll=False)
            bap = states.clamp(min=0).long()                              # best_aligns_pad: 0 beyond the utterance
            tgt = P.tg[hs.tgt].long()
            pos = torch.div(bap, 2, rounding_mode="floor").clamp(max=tgt.shape[1] - 1)
            tok = tgt.gather(1, pos).masked_fill(bap % 2 == 0, self.blank_idx)
            mistake = lp.argmax(dim=-1) != tok
            prob = ratio * (mistake.sum(-1) / P.out_len).unsqueeze(-1) if enc.pae_adaptive_gt else ratio
            mask = torch.rand(tok.shape, device=tok.device) < prob
            if enc.pae_gt_only_mistake:
                mask = mask & mistake
            fe = bap.masked_fill(~mask, -1)
            if Tv < T2:
                padc = (0, T2 - Tv)
                tok, mask, fe = F.pad(tok, padc, value=0), F.pad(mask, padc, value=False), F.pad(fe, padc, value=-1)
            eng.set_ground_truth(P, k, tok, mask, fe)

    def forward(self, model, sample, reduce=True):
        enc: S2TTransformerEncoderB200 = getattr(model, "encoder", model)
        eng = enc.engine
        net_input = sample["net_input"]
        src_tokens, src_lengths = net_input["src_tokens"], net_input["src_lengths"]
        # source-side targets: transcript when present, else target (criterions/ctc.py:543-546)
        src_text = sample["transcript"]["tokens"] if "transcript" in sample else sample["target"]
        tg, tl = self.get_targets_for_ctc_loss(src_text)
        xt = xl = None
        if len(eng.heads) > 1:      # XCTC terms use the target text (get_ctc_target_text, :763-767)
            tgt_text = sample["ctc_target"]["tokens"] if "ctc_target" in sample else sample["target"]
            xt, xl = self.get_targets_for_ctc_loss(tgt_text)
        P = enc._plan_for(src_tokens, max(tg.shape[1], xt.shape[1] if xt is not None else 0))
        eng.train(model.training)
        eng.set_inputs(P, feats=src_tokens.float(), src_lengths=src_lengths.to(torch.int32), targets=tg,
                       target_lengths=tl, xtargets=xt, xtarget_lengths=xl)
        if model.training and torch.is_grad_enabled():
            if getattr(enc, "pae_ground_truth_ratio", 0) != 0:
                eng.ctc_masked_loss = self.ctc_masked_loss
                self._curriculum(enc, eng, P, eng.out_frames(src_tokens.shape[1]))
            if enc.use_graph and P.graph is None:
                eng.capture(P, "train")
            loss = _EngineStep.apply(enc.flat, enc, P)
        else:
            eng.forward(P)
            eng._loss(P, need_grad=False)
            loss = P.loss[0].clone()
        enc._last_plan = P
        ntokens = sample.get("ntokens", int(tl.sum()))
        nsent = src_tokens.size(0)
        sample_size = nsent if self.sentence_avg else ntokens
        logging_output = {"loss": loss.detach(), "ntokens": ntokens, "nsentences": nsent, "sample_size": sample_size,
                          "nfeatures": P.out_len.sum()}
        # per-term losses for the reference's logging keys (device scalars: no host sync on the training path)
        for k, hs in enumerate(eng.heads):
            logging_output[f"{hs.name}_loss"] = P.nll_all[P.heads[k].loss_idx].sum()
            if hs.inter_layers:
                idx = [P.inter[(l, k)].loss_idx for l in hs.inter_layers]
                logging_output[f"inter_{hs.name}_loss"] = P.nll_all[idx].sum() / len(idx)
        if not model.training:
            out = enc.output_dict(P)
            lp = torch.log_softmax(out["ctc_logit"][0], -1, dtype=torch.float32)
            c_err, c_len, w_errs, w_len, wv = enc.ctc_valid(lp, src_text, P.out_len, _DictShim(self.pad_idx, self.eos_idx))
            logging_output.update(c_errors=c_err, c_total=c_len, w_errors=w_errs, w_total=w_len, wv_errors=wv)
        return loss, sample_size, logging_output

    @staticmethod
    def reduce_metrics(logging_outputs) -> None:
        """criterions/ctc.py:1104-1333: sums over workers, loss / sample_size / ln 2.  Logs through fairseq.metrics when
        the reference package is importable; a no-op otherwise."""
        try:
            from fairseq import metrics  # type: ignore
        except Exception:
            return
        s = lambda k: sum(float(log.get(k, 0)) for log in logging_outputs)
        sample_size, ntokens = s("sample_size"), s("ntokens")
        metrics.log_scalar("loss", s("loss") / max(sample_size, 1) / math.log(2), sample_size, round=3)
        metrics.log_scalar("ntokens", ntokens)
        metrics.log_scalar("nsentences", s("nsentences"))
        for key in ("ctc_loss", "xctc_loss", "inter_ctc_loss", "inter_xctc_loss"):
            if any(key in log for log in logging_outputs):
                metrics.log_scalar(key, s(key) / max(sample_size, 1) / math.log(2), sample_size, round=3)
        if s("c_total") > 0:
            metrics.log_scalar("cer", 100.0 * s("c_errors") / s("c_total"), round=3)
        if s("w_total") > 0:
            metrics.log_scalar("wer", 100.0 * s("w_errors") / s("w_total"), round=3)


class _DictShim:
    def __init__(self, pad, eos):
        self._pad, self._eos = pad, eos

    def pad(self):
        return self._pad

    def eos(self):
        return self._eos


def register() -> Dict[str, List[str]]:
    """Assign the B200 classes into the reference's registries (fairseq/models/__init__.py:30-35,101-198;
    criterions/__init__.py:18-25) — what a ``--user-dir`` module does at import.  Encoders are swapped by overriding
    ``build_encoder`` on the registered model classes (s2t_transformer.py:776-796, pdss2t_transformer.py:287-301,
    s2t_ctc.py:50-82), so the AR-decoder models keep their own decoder.  Returns what was patched."""
    done: Dict[str, List[str]] = {"models": [], "criterions": []}
    if not _HAVE_FAIRSEQ:
        return done
    from fairseq import criterions, models  # type: ignore
    for arch, model_cls in list(models.ARCH_MODEL_REGISTRY.items()):
        if arch in S2T_ARCHS:
            model_cls.build_encoder = classmethod(lambda cls, args, task=None, embed_tokens=None:
                                                  S2TTransformerEncoderB200(args, task, embed_tokens))
        elif arch in PDS_ARCHS:
            model_cls.build_encoder = classmethod(lambda cls, args, task=None, embed_tokens=None:
                                                  PDSS2TTransformerEncoderB200(args, task, embed_tokens))
        elif arch in CTC_ARCHS:
            model_cls.build_encoder = classmethod(lambda cls, args, task=None, embed_tokens=None:
                                                  S2TCTCModelB200.build_encoder(args, task, embed_tokens))
        else:
            continue
        done["models"].append(arch)
    if "ctc" in getattr(criterions, "CRITERION_REGISTRY", {}):
        criterions.CRITERION_REGISTRY["ctc"] = CtcCriterionB200
        done["criterions"].append("ctc")
    return done
