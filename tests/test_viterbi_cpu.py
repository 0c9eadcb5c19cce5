"""oracle/viterbi.py (restatement of torch_imputer/best_alignment.cu + imputer.get_alignment_path; parity UNPINNED, see
its header) against a brute-force search over all monotone CTC state paths on tiny cases."""
import itertools

import numpy as np
import pytest


def brute_force(lp, target, blank):
    T = lp.shape[0]
    S = len(target)
    Se = 2 * S + 1
    lab = [blank if s % 2 == 0 else target[(s - 1) // 2] for s in range(Se)]
    best = -np.inf
    for path in itertools.product(range(Se), repeat=T):
        if path[0] not in (0, 1) or (S == 0 and path[0] != 0):
            continue
        if path[-1] not in (Se - 1, Se - 2):
            continue
        ok = True
        for a, b in zip(path, path[1:]):
            d = b - a
            if d < 0 or d > 2 or (d == 2 and (lab[b] == blank or lab[b] == lab[a])):
                ok = False
                break
        if ok:
            best = max(best, float(sum(np.float64(lp[t, lab[s]]) for t, s in enumerate(path))))
    return best


@pytest.mark.parametrize("seed", range(6))
def test_oracle_path_is_a_maximum_probability_alignment(seed):
    from oracle.viterbi import best_alignment
    rng = np.random.RandomState(seed)
    T, V = 6, 5
    S = [0, 1, 2, 3, 2, 3][seed]
    target = rng.randint(1, V, size=(1, max(S, 1)))
    if seed == 4:
        target[0, :2] = 3                      # repeated label: the s-2 jump is forbidden
    logits = rng.randn(1, T, V).astype(np.float32)
    lp = logits - np.log(np.exp(logits).sum(-1, keepdims=True))
    states, nll = best_alignment(lp.astype(np.float32), target, [T], [S], blank=0)
    st = states[0]
    assert len(st) == T and st[0] in (0, 1) and st[-1] in (2 * S, max(2 * S - 1, 0))
    lab = [0 if s % 2 == 0 else int(target[0, (s - 1) // 2]) for s in range(2 * S + 1)]
    score = float(sum(np.float64(lp[0, t, lab[s]]) for t, s in enumerate(st)))
    assert abs(score - brute_force(lp[0], [int(v) for v in target[0, :S]], 0)) < 1e-4
    assert np.isfinite(nll[0])
