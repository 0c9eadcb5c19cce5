"""CPU oracle for the CTC criterion arithmetic and greedy CTC decoding (TEST INFRASTRUCTURE).

Restates
  * ``torch.nn.CTCLoss(blank, reduction="none", zero_infinity=True)`` — the third-party ATen alpha/beta
    recursion the reference calls at ``fairseq/criterions/ctc.py:243-245`` and ``:467-472`` (cuDNN
    disabled ``:451``) — in float64 numpy, including the gradient w.r.t. the LOGITS through
    ``utils.log_softmax(dtype=float32)`` (``fairseq/utils.py:477-481``);
  * ``CtcCriterion.get_targets_for_ctc_loss`` (``fairseq/criterions/ctc.py:516-540``): strip pad and eos;
  * ``CtcCriterion.get_ctc_loss`` (``:435-474``): sum over utterances of loss * coef;
  * greedy ``CTCDecoder.generate`` (``fairseq/models/speech_to_text/s2t_ctc.py:312-347``):
    argmax -> padded frames to blank -> unique_consecutive -> drop blank.
Pinned against torch.nn.CTCLoss (torch 2.11 CPU) in tests/test_oracle_cpu.py.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def log_softmax(logits: np.ndarray) -> np.ndarray:
    x = np.asarray(logits, dtype=np.float64)
    m = x.max(axis=-1, keepdims=True)
    return x - m - np.log(np.exp(x - m).sum(axis=-1, keepdims=True))


def _logaddexp3(a, b, c):
    m = max(a, b, c)
    if m == -np.inf:
        return -np.inf
    return m + np.log(np.exp(a - m) + np.exp(b - m) + np.exp(c - m))


def ctc_alpha_beta(lp: np.ndarray, target: np.ndarray, blank: int = 0):
    """lp: [T, V] log-probs of one utterance (float64); target [S].  Returns (nll, alpha, beta, ext)."""
    T = lp.shape[0]
    S = len(target)
    ext = np.full(2 * S + 1, blank, dtype=np.int64)
    ext[1::2] = target
    Se = 2 * S + 1
    alpha = np.full((T, Se), -np.inf)
    beta = np.full((T, Se), -np.inf)
    if T == 0:
        return (0.0 if S == 0 else np.inf), alpha, beta, ext
    alpha[0, 0] = lp[0, blank]
    if Se > 1:
        alpha[0, 1] = lp[0, ext[1]]
    for t in range(1, T):
        for s in range(Se):
            a = alpha[t - 1, s]
            b = alpha[t - 1, s - 1] if s >= 1 else -np.inf
            c = alpha[t - 1, s - 2] if (s >= 2 and ext[s] != blank and ext[s] != ext[s - 2]) else -np.inf
            alpha[t, s] = _logaddexp3(a, b, c) + lp[t, ext[s]]
    beta[T - 1, Se - 1] = lp[T - 1, blank]
    if Se > 1:
        beta[T - 1, Se - 2] = lp[T - 1, ext[Se - 2]]
    for t in range(T - 2, -1, -1):
        for s in range(Se):
            a = beta[t + 1, s]
            b = beta[t + 1, s + 1] if s + 1 < Se else -np.inf
            c = beta[t + 1, s + 2] if (s + 2 < Se and ext[s] != blank and ext[s] != ext[s + 2]) else -np.inf
            beta[t, s] = _logaddexp3(a, b, c) + lp[t, ext[s]]
    tail = alpha[T - 1, Se - 2] if Se > 1 else -np.inf
    ll = np.logaddexp(alpha[T - 1, Se - 1], tail)
    return -ll, alpha, beta, ext


def ctc_loss(logits: np.ndarray, targets: List[np.ndarray], input_lengths, blank: int = 0,
             zero_infinity: bool = True) -> Tuple[np.ndarray, np.ndarray]:
    """logits [B, T, V] (utterance-major).  Returns (nll [B], d sum(nll) / d logits [B, T, V]) in float64."""
    logits = np.asarray(logits, dtype=np.float64)
    B, T, V = logits.shape
    nll = np.zeros(B)
    grad = np.zeros_like(logits)
    lp_all = log_softmax(logits)
    for b in range(B):
        Tb = int(input_lengths[b])
        lp = lp_all[b, :Tb]
        loss, alpha, beta, ext = ctc_alpha_beta(lp, np.asarray(targets[b], dtype=np.int64), blank)
        if not np.isfinite(loss):
            nll[b] = 0.0 if zero_infinity else loss
            continue
        nll[b] = loss
        occ = np.zeros((Tb, V))
        g = alpha + beta + loss  # log gamma_t(s) + lp_t(ext_s)
        for s in range(len(ext)):
            col = g[:, s] - lp[:, ext[s]]
            occ[:, ext[s]] += np.where(np.isfinite(col), np.exp(col), 0.0)
        grad[b, :Tb] = np.exp(lp) - occ
    return nll, grad


def targets_for_ctc_loss(tokens: np.ndarray, pad: int = 1, eos: int = 2) -> List[np.ndarray]:
    """criterions/ctc.py:516-540 — drop pad and eos from each row of a collated [B, S] token matrix."""
    return [row[(row != pad) & (row != eos)] for row in np.asarray(tokens)]


def criterion_ctc(logits: np.ndarray, tokens: np.ndarray, input_lengths, blank: int = 0, pad: int = 1,
                  eos: int = 2, coef: float = 1.0):
    """criterions/ctc.py:435-474: (loss * coef).sum() over the batch, zero_infinity always on (A12)."""
    tgts = targets_for_ctc_loss(tokens, pad, eos)
    nll, grad = ctc_loss(logits, tgts, input_lengths, blank, zero_infinity=True)
    return float((nll * coef).sum()), grad * coef, nll


def greedy_decode(logits: np.ndarray, input_lengths, blank: int = 0) -> List[np.ndarray]:
    """s2t_ctc.py:312-347 — per-utterance token ids (int64) after collapse; lowest index wins ties."""
    logits = np.asarray(logits)
    B, T, _ = logits.shape
    ids = logits.argmax(axis=-1)  # numpy argmax returns the first maximal index
    out = []
    for b in range(B):
        row = ids[b].copy()
        row[int(input_lengths[b]):] = blank           # pad -> blank BEFORE collapsing (A14)
        keep = np.ones(T, dtype=bool)
        keep[1:] = row[1:] != row[:-1]                # unique_consecutive
        row = row[keep]
        out.append(row[row != blank].astype(np.int64))
    return out


def greedy_scores(logits: np.ndarray, blank: int = 0) -> np.ndarray:
    """Hypothesis score of CTCDecoder.generate (s2t_ctc.py:322-332, quirk A14): topk(1) of the fp32 log-softmax over
    ALL frames (the padding mask is applied to the indices only), blank positions filled with the VALUE ``blank``
    (0), score = -sum_t.  float64 here."""
    lp = log_softmax(np.asarray(logits, dtype=np.float64))
    top = lp.max(axis=-1)
    idx = lp.argmax(axis=-1)
    real = np.where(idx == blank, float(blank), top)
    return -real.sum(axis=-1)


# ---------------------------------------------------------------------------------------------------------------------
# Masked ("imputer") CTC loss — fairseq/torch_imputer/imputer.cu:57-225 (log_alpha), :364-520 (log_beta), :535-640
# (gradient), called from fairseq/criterions/ctc.py:455-465 when ``ctc_masked_loss`` is set.
# force_emit[t] = index of the EXTENDED-label state (0 .. 2S) that must be occupied at frame t, or -1 (free): every other
# state of that frame is removed from the lattice (alpha = beta = -inf).  With all entries -1 it is the plain CTC loss.
# Parity status: the reference path exists only as a CUDA extension that cannot run in the CPU-only build container and
# the reference holds no test vectors for it ("parity unpinned"); the restatement is checked against brute-force
# enumeration of all alignments and against the plain CTC oracle (itself pinned to torch.nn.CTCLoss).
# ---------------------------------------------------------------------------------------------------------------------
def imputer_alpha_beta(lp: np.ndarray, target: np.ndarray, force_emit: np.ndarray, blank: int = 0):
    """lp [T, V] log-probs (float64), target [S], force_emit [T] int.  Returns (nll, alpha, beta, ext)."""
    T = lp.shape[0]
    S = len(target)
    ext = np.full(2 * S + 1, blank, dtype=np.int64)
    ext[1::2] = target
    Se = 2 * S + 1
    alpha = np.full((T, Se), -np.inf)
    beta = np.full((T, Se), -np.inf)
    if T == 0:
        return (0.0 if S == 0 else np.inf), alpha, beta, ext
    allowed = lambda t, s: force_emit[t] < 0 or int(force_emit[t]) == s          # imputer.cu:117-119, 146-151
    if allowed(0, 0):
        alpha[0, 0] = lp[0, blank]
    if Se > 1 and allowed(0, 1):
        alpha[0, 1] = lp[0, ext[1]]
    for t in range(1, T):
        for s in range(Se):
            if not allowed(t, s):
                continue
            a = alpha[t - 1, s]
            b = alpha[t - 1, s - 1] if s >= 1 else -np.inf
            c = alpha[t - 1, s - 2] if (s >= 2 and ext[s] != blank and ext[s] != ext[s - 2]) else -np.inf
            alpha[t, s] = _logaddexp3(a, b, c) + lp[t, ext[s]]
    if allowed(T - 1, Se - 1):
        beta[T - 1, Se - 1] = lp[T - 1, blank]
    if Se > 1 and allowed(T - 1, Se - 2):
        beta[T - 1, Se - 2] = lp[T - 1, ext[Se - 2]]
    for t in range(T - 2, -1, -1):
        for s in range(Se):
            if not allowed(t, s):
                continue
            a = beta[t + 1, s]
            b = beta[t + 1, s + 1] if s + 1 < Se else -np.inf
            c = beta[t + 1, s + 2] if (s + 2 < Se and ext[s] != blank and ext[s] != ext[s + 2]) else -np.inf
            beta[t, s] = _logaddexp3(a, b, c) + lp[t, ext[s]]
    tail = alpha[T - 1, Se - 2] if Se > 1 else -np.inf
    ll = np.logaddexp(alpha[T - 1, Se - 1], tail)
    return -ll, alpha, beta, ext


def imputer_loss(logits: np.ndarray, targets: List[np.ndarray], force_emits, input_lengths, blank: int = 0,
                 zero_infinity: bool = True) -> Tuple[np.ndarray, np.ndarray]:
    """logits [B, T, V]; force_emits [B, T] (state indices or -1).  Returns (nll [B], d sum(nll) / d logits) in float64."""
    logits = np.asarray(logits, dtype=np.float64)
    B, T, V = logits.shape
    nll = np.zeros(B)
    grad = np.zeros_like(logits)
    lp_all = log_softmax(logits)
    for b in range(B):
        Tb = int(input_lengths[b])
        lp = lp_all[b, :Tb]
        loss, alpha, beta, ext = imputer_alpha_beta(lp, np.asarray(targets[b], dtype=np.int64),
                                                    np.asarray(force_emits[b][:Tb]), blank)
        if not np.isfinite(loss):
            nll[b] = 0.0 if zero_infinity else loss
            continue
        nll[b] = loss
        occ = np.zeros((Tb, V))
        g = alpha + beta + loss
        for s in range(len(ext)):
            col = g[:, s] - lp[:, ext[s]]
            occ[:, ext[s]] += np.where(np.isfinite(col), np.exp(col), 0.0)
        grad[b, :Tb] = np.exp(lp) - occ
    return nll, grad


def imputer_nll_bruteforce(lp: np.ndarray, target, force_emit, blank: int = 0) -> float:
    """-log of the summed probability of every frame-level state path through the extended target (start in state 0 or
    1, stay / advance by one / skip a blank between different labels, end in the last two states) that occupies the
    forced states.  Exponential in T: tiny cases only."""
    import itertools
    T = lp.shape[0]
    S = len(target)
    ext = [blank] * (2 * S + 1)
    ext[1::2] = list(target)
    Se = len(ext)
    total = 0.0
    for path in itertools.product(range(Se), repeat=T):
        if path[0] > 1 or path[-1] < Se - 2:
            continue
        ok = True
        for t in range(T):
            if force_emit[t] >= 0 and path[t] != force_emit[t]:
                ok = False
                break
            if t > 0:
                d = path[t] - path[t - 1]
                if d < 0 or d > 2 or (d == 2 and (ext[path[t]] == blank or ext[path[t]] == ext[path[t - 1]])):
                    ok = False
                    break
        if ok:
            total += float(np.exp(sum(lp[t, ext[path[t]]] for t in range(T))))
    return float(-np.log(total)) if total > 0 else float("inf")
