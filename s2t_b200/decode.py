"""Greedy CTC decoding behind the reference's generator / validation interfaces.

* ``CTCDecoderB200`` — ``task.build_generator`` replacement for ``CTCDecoder`` (fairseq/models/speech_to_text/
  s2t_ctc.py:131-349, ``infer == "greedy"``): ``generate(models, sample) -> List[List[dict(tokens, score, attention,
  alignment, positional_scores)]]``.  argmax / pad->blank / collapse / score all run in the ``s2t_ctc_greedy`` kernels;
  only the final per-utterance slicing into Python lists happens on the host (one device->host copy per batch, where the
  reference loops over utterances with ``unique_consecutive``).
* ``ctc_valid_counts`` — ``CTC.valid`` (modules/speech_to_text/ctc.py:96-136): unit / word error counts of the greedy
  path against the targets (edit distance is host-side integer work, as in the reference).
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch

from . import ops


def edit_distance(a: Sequence, b: Sequence) -> int:
    """Levenshtein distance (the reference calls the ``editdistance`` package, ctc.py:119,128)."""
    if len(a) < len(b):
        a, b = b, a
    prev = list(range(len(b) + 1))
    for i, x in enumerate(a, 1):
        cur = [i]
        for j, y in enumerate(b, 1):
            cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y)))
        prev = cur
    return prev[-1]


def post_process(sentence: str, symbol: str) -> str:
    """fairseq/data/data_utils.py:361-367 for the symbols the S2T recipes use."""
    if symbol == "sentencepiece":
        return sentence.replace(" ", "").replace("▁", " ").strip()
    if symbol in ("wordpiece", "letter", "_EOW", "none", None):
        return sentence
    if symbol is not None and symbol != "none":
        return (sentence + " ").replace(symbol, "").rstrip()
    return sentence


def greedy_batch(logits_btv: torch.Tensor, lengths: torch.Tensor, V: int, blank: int = 0):
    """logits [B, T, ldv] (bf16 / fp32, CUDA) -> (list of int64 token tensors on the CPU, scores fp32 [B] on the CPU)."""
    tokens, lens, _, scores = ops.ctc_greedy(logits_btv, lengths.to(torch.int32), V, blank, with_scores=True)
    tokens, lens, scores = tokens.cpu(), lens.cpu(), scores.cpu()
    return [tokens[b, : int(lens[b])].long() for b in range(tokens.shape[0])], scores


class CTCDecoderB200:
    """Greedy ``CTCDecoder`` (s2t_ctc.py:131-349).  ``models`` = [a model whose encoder is an
    ``S2TTransformerEncoderB200``]; ``blank_idx`` as in ``tasks/speech_to_text.py:507-513``."""

    def __init__(self, models, args=None, dictionary=None, blank_idx: int = 0):
        self.model = models[0]
        self.dictionary = dictionary
        self.blank = blank_idx
        self.infer = getattr(args, "ctc_infer", "greedy") if args is not None else "greedy"
        if self.infer != "greedy":
            raise NotImplementedError("CTCDecoderB200 implements the greedy inference of the reference (ctc_infer=greedy)")

    @torch.no_grad()
    def generate(self, models, sample: Dict[str, Dict[str, torch.Tensor]], **kwargs) -> List[List[dict]]:
        net_input = sample["net_input"]
        enc = getattr(self.model, "encoder", self.model)
        enc.eval()
        enc(net_input["src_tokens"], net_input["src_lengths"])
        P, eng = enc._last_plan, enc.engine
        # the target-side head decodes when it exists (s2t_ctc.py:265-271)
        head = len(eng.heads) - 1 if len(eng.heads) > 1 else 0
        hs = eng.heads[head]
        hyps, scores = greedy_batch(P.heads[head].logits.view(P.B, P.T2, hs.ldv), P.out_len, hs.V, self.blank)
        finalized = []
        for b, hyp in enumerate(hyps):
            n = len(hyp)
            score = scores[b : b + 1]
            finalized.append([{
                "tokens": hyp, "score": score, "attention": None, "alignment": None,
                "positional_scores": torch.Tensor([float(score) / n] * n) if n else torch.Tensor([]),
            }])
        return finalized


def ctc_valid_counts(lprobs_tbv: torch.Tensor, targets: torch.Tensor, input_lengths: torch.Tensor, dictionary,
                     blank: int = 0, post: str = "sentencepiece"):
    """CTC.valid (ctc.py:96-136): (c_err, c_len, w_errs, w_len, wv_errs) of the greedy path."""
    x = lprobs_tbv.transpose(0, 1).contiguous()
    B, T, V = x.shape
    if V % 8:                                        # the kernel wants 16-byte aligned bf16 rows
        x = torch.nn.functional.pad(x, (0, 8 - V % 8))
    if x.dtype not in (torch.bfloat16, torch.float32):
        x = x.float()
    hyps, _ = greedy_batch(x, input_lengths, V, blank)
    pad, eos = dictionary.pad(), dictionary.eos()
    c_err = c_len = w_errs = w_len = 0
    for hyp, t in zip(hyps, targets.cpu()):
        targ = t[(t != pad) & (t != eos)].tolist()
        pred = hyp.tolist()
        c_err += edit_distance(pred, targ)
        c_len += len(targ)
        if hasattr(dictionary, "string"):
            tw = post_process(dictionary.string(targ), post).split()
            pw = post_process(dictionary.string(pred), post).split()
        else:
            tw, pw = [str(v) for v in targ], [str(v) for v in pred]
        w_errs += edit_distance(pw, tw)
        w_len += len(tw)
    return c_err, c_len, w_errs, w_len, w_errs
