"""CPU restatement of the reference's CTC best-alignment (Viterbi) path (TEST INFRASTRUCTURE).

PARITY UNPINNED: the reference implements this path only as a CUDA extension
(fairseq/torch_imputer/best_alignment.cu, bound through imputer.cpp) that cannot be built or run in the CPU-only build
container, and its tests hold no vectors for it.  The restatement below follows the kernel statement by statement and
is additionally checked against a brute-force enumeration of all alignments on tiny cases (tests/test_viterbi_cpu.py).

  * recursion: ``ctc_alignment_log_alpha_gpu_kernel`` (best_alignment.cu:62-201) — alpha_0 = (lp[0][blank], lp[0][l_1]
    or -inf when the target is empty, -inf ...); for t >= 1 the best predecessor among s, s-1 (if s > 0), s-2 (if
    l_s != l_{s-2}) with STRICT '>' replacing in that order (so ties keep the smaller jump), alpha = best + lp[t][l_s]
    in the input dtype (fp32), paths[t][s] = argmax state;
  * nll = -logsumexp(alpha[T-1][2S], alpha[T-1][2S-1]) (best_alignment.cu:186-199);
  * backtrack: ``get_alignment_path`` (imputer.py:245-259) — start at argmax over the last two states (first maximum
    on ties, i.e. state 2S-1), state 0 for an empty target, then follow paths[t][state] for t = T-1 .. 1.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def best_alignment(log_probs: np.ndarray, targets: np.ndarray, input_lengths, target_lengths, blank: int = 0
                   ) -> Tuple[List[List[int]], np.ndarray]:
    """log_probs fp32 [B, T, V] (batch-major), targets int [B, S_max].  Returns the per-utterance state sequences
    (length input_length, states 0..2S) and the fp32 nll of the kernel."""
    B = log_probs.shape[0]
    neg = np.float32(-np.inf)
    out: List[List[int]] = []
    nll = np.zeros((B,), dtype=np.float32)
    for b in range(B):
        T, S = int(input_lengths[b]), int(target_lengths[b])
        Se = 2 * S + 1
        lab = [blank if s % 2 == 0 else int(targets[b, (s - 1) // 2]) for s in range(Se)]
        lp = log_probs[b].astype(np.float32, copy=False)
        alpha = np.full((T, Se), neg, dtype=np.float32)
        path = np.zeros((T, Se), dtype=np.int64)
        alpha[0, 0] = lp[0, blank]
        if S > 0:
            alpha[0, 1] = lp[0, lab[1]]
        for t in range(1, T):
            for s in range(Se):
                lamax, mp = alpha[t - 1, s], s
                if s > 0 and alpha[t - 1, s - 1] > lamax:
                    lamax, mp = alpha[t - 1, s - 1], s - 1
                if s > 1 and lab[s - 2] != lab[s] and alpha[t - 1, s - 2] > lamax:
                    lamax, mp = alpha[t - 1, s - 2], s - 2
                alpha[t, s] = np.float32(lamax + lp[t, lab[s]])
                path[t, s] = mp
        l1 = alpha[T - 1, 2 * S]
        l2 = alpha[T - 1, 2 * S - 1] if S > 0 else neg
        m = max(l1, l2)
        m = np.float32(0.0) if m == neg else m
        with np.errstate(all="ignore"):
            nll[b] = -(np.log(np.exp(l1 - m, dtype=np.float32) + np.exp(l2 - m, dtype=np.float32), dtype=np.float32) + m)
        cur = 0 if Se == 1 else int(np.argmax(alpha[T - 1, Se - 2:])) + (Se - 2)
        states = [cur]
        for t in range(T - 1, 0, -1):
            cur = int(path[t, cur])
            states.append(cur)
        out.append(states[::-1])
    return out, nll


def states_to_tokens(states, target_row, blank: int = 0) -> List[int]:
    """Frame-level token sequence of an alignment: even states emit blank, odd state s emits target[(s-1)//2]."""
    return [blank if s % 2 == 0 else int(target_row[(s - 1) // 2]) for s in states]
