"""CPU tests of the masked ("imputer") CTC loss oracle (oracle/ctc.py, restating fairseq/torch_imputer/imputer.cu): parity
unpinned against the reference (CUDA-only extension, no test vectors), so the restatement is checked against brute-force
enumeration of all alignments, against the plain CTC oracle (pinned to torch.nn.CTCLoss) when nothing is forced, and its
gradient against finite differences."""
import numpy as np
import pytest


def _case(seed, T, S, V, n_forced):
    from oracle import ctc as O
    rng = np.random.default_rng(seed)
    logits = rng.normal(0, 1.0, size=(T, V))
    target = rng.integers(1, V, size=S)
    if S >= 2 and seed % 2 == 0:
        target[1] = target[0]                                 # repeated label: needs the blank in between
    # force states along a VALID alignment so that the loss stays finite
    lp = O.log_softmax(logits)
    from oracle import viterbi as OV
    path = None
    if T >= S:                      # a valid alignment to force states along (oracle/viterbi.py: best path)
        paths, _ = OV.best_alignment(lp[None].astype(np.float32), target[None] if S else np.zeros((1, 1), dtype=np.int64),
                                     [T], [S])
        path = paths[0]
    force = np.full(T, -1, dtype=np.int64)
    if path is not None and n_forced > 0:
        idx = rng.choice(T, size=min(n_forced, T), replace=False)
        force[idx] = np.asarray(path)[idx]
    return logits, target, force


@pytest.mark.parametrize("seed,T,S,V,n_forced", [(0, 5, 2, 4, 2), (1, 6, 2, 5, 3), (2, 4, 1, 3, 1), (3, 6, 3, 4, 0),
                                                  (4, 5, 2, 4, 5), (5, 3, 0, 3, 1)])
def test_imputer_oracle_matches_bruteforce(seed, T, S, V, n_forced):
    from oracle import ctc as O
    logits, target, force = _case(seed, T, S, V, n_forced)
    lp = O.log_softmax(logits)
    nll, alpha, beta, ext = O.imputer_alpha_beta(lp, target, force)
    brute = O.imputer_nll_bruteforce(lp, target, force)
    if np.isfinite(brute):
        assert abs(nll - brute) <= 1e-9 * max(1.0, abs(brute)), (nll, brute)
    else:
        assert not np.isfinite(nll)
    # alpha and beta describe the same lattice: sum_s alpha_t(s) beta_t(s) / p_t(s) is the path mass at every frame
    if np.isfinite(nll):
        for t in range(T):
            g = alpha[t] + beta[t] - lp[t, ext]
            tot = np.log(np.exp(g[np.isfinite(g)]).sum())
            assert abs(tot + nll) <= 1e-9, (t, tot, nll)


def test_imputer_oracle_unforced_equals_ctc_and_infeasible_force_is_zeroed():
    from oracle import ctc as O
    rng = np.random.default_rng(7)
    B, T, V = 3, 12, 9
    logits = rng.normal(size=(B, T, V))
    targets = [rng.integers(1, V, size=n) for n in (4, 0, 6)]
    ilens = [12, 7, 11]
    free = np.full((B, T), -1)
    nll0, g0 = O.ctc_loss(logits, targets, ilens)
    nll1, g1 = O.imputer_loss(logits, targets, free, ilens)
    assert np.allclose(nll0, nll1, rtol=1e-12) and np.allclose(g0, g1, atol=1e-12)
    # forcing a state that no alignment can occupy at that frame: infeasible -> zero_infinity -> loss 0, gradient 0
    bad = free.copy()
    bad[0, 0] = 7                                   # state 7 at frame 0 (only states 0 and 1 can start)
    nll2, g2 = O.imputer_loss(logits, targets, bad, ilens)
    assert nll2[0] == 0.0 and np.all(g2[0] == 0.0) and np.allclose(nll2[1:], nll0[1:])


def test_imputer_oracle_gradient_matches_finite_differences():
    from oracle import ctc as O
    logits, target, force = _case(1, 6, 2, 5, 3)
    f = lambda x: O.imputer_loss(x[None], [target], force[None], [6])[0][0]
    _, g = O.imputer_loss(logits[None], [target], force[None], [6])
    eps = 1e-6
    num = np.zeros_like(logits)
    for i in range(logits.shape[0]):
        for j in range(logits.shape[1]):
            d = np.zeros_like(logits)
            d[i, j] = eps
            num[i, j] = (f(logits + d) - f(logits - d)) / (2 * eps)
    assert np.abs(num - g[0]).max() <= 1e-6
    # forced frames only receive gradient through the forced state's label and the softmax normaliser
    t = int(np.nonzero(force >= 0)[0][0])
    assert abs(g[0][t].sum()) <= 1e-9
