"""Device CTC best alignment (csrc/viterbi.cu through the C ABI) against the CPU restatement of
torch_imputer/best_alignment.cu + imputer.get_alignment_path (oracle/viterbi.py) on identical fp32 log-probabilities:
state sequences bit-exact (one fp32 add per state and frame, same comparison order); nll to 1e-5 relative (libm)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, B, T, V, S, ragged=True, repeats=False):
    g = torch.Generator().manual_seed(seed)
    lp = torch.log_softmax(torch.randn((B, T, V), generator=g) * 2.0, dim=-1)
    tg = torch.randint(1, V, (B, max(S, 1)), generator=g).to(torch.int32)
    if repeats:
        tg[:, 1::2] = tg[:, 0::2][:, : tg[:, 1::2].shape[1]]          # doubled labels: the s-2 jump is forbidden
    il = torch.full((B,), T, dtype=torch.int32)
    tl = torch.full((B,), S, dtype=torch.int32)
    if ragged and B > 1:
        il = torch.randint(max(2 * S + 1, T // 2), T + 1, (B,), generator=g).to(torch.int32)
        tl = torch.randint(0, S + 1, (B,), generator=g).to(torch.int32)
        il[0], tl[0] = T, S
    return lp, tg, il, tl


@pytest.mark.parametrize("B,T,V,S,repeats", [(4, 60, 30, 9, False), (3, 250, 500, 40, False), (5, 40, 12, 7, True),
                                             (2, 17, 9, 0, False), (2, 1, 9, 1, False), (8, 250, 10000, 60, False)])
def test_states_match_oracle(cuda_lib, B, T, V, S, repeats):
    from oracle.viterbi import best_alignment
    from s2t_b200 import ops
    lp, tg, il, tl = _case(B * 7 + T, B, T, V, S, ragged=T > 1, repeats=repeats)
    want, want_nll = best_alignment(lp.numpy(), tg.numpy(), il.tolist(), tl.tolist(), blank=0)
    states, nll = ops.ctc_viterbi(lp.cuda(), tg.cuda(), il.cuda(), tl.cuda(), blank=0)
    states, nll = states.cpu().numpy(), nll.cpu().numpy()
    for b in range(B):
        n = int(il[b])
        assert states[b, :n].tolist() == want[b], f"utterance {b}"
        assert (states[b, n:] == -1).all()
    assert np.allclose(nll, want_nll, rtol=1e-5, atol=1e-5)


def test_time_major_layout_and_long_targets_use_global_backpointers(cuda_lib):
    """The reference's (T, N, C) layout through strides; T * (2S+1) beyond the shared-memory budget takes the
    global-workspace path."""
    from oracle.viterbi import best_alignment
    from s2t_b200 import lib, ops
    B, T, V, S = 2, 1200, 40, 150
    assert lib.load().s2t_ctc_viterbi_workspace_bytes(B, T, S) > 0
    lp, tg, il, tl = _case(3, B, T, V, S, ragged=False)
    want, _ = best_alignment(lp.numpy(), tg.numpy(), il.tolist(), tl.tolist(), blank=0)
    tm = lp.transpose(0, 1).contiguous().cuda()                      # [T, B, V]
    states, _ = ops.ctc_viterbi(tm, tg.cuda(), il.cuda(), tl.cuda(), blank=0, batch_first=False)
    for b in range(B):
        assert states[b].cpu().tolist() == want[b]


def test_alignment_is_monotone_and_collapses_to_the_target(cuda_lib):
    """Size-independent property at the bench size (8 x 250 frames, V = 10000): states start in {0, 1}, end in
    {2S-1, 2S}, move by 0/1/2, and the emitted labels collapse (CTC rule) to the target."""
    from s2t_b200 import ops
    B, T, V, S = 8, 250, 10000, 60
    lp, tg, il, tl = _case(11, B, T, V, S, ragged=False)
    states, _ = ops.ctc_viterbi(lp.cuda(), tg.cuda(), il.cuda(), tl.cuda(), blank=0)
    st = states.cpu().numpy()
    for b in range(B):
        s = st[b]
        assert s[0] in (0, 1) and s[-1] in (2 * S - 1, 2 * S)
        d = np.diff(s)
        assert ((d >= 0) & (d <= 2)).all()
        toks = [0 if v % 2 == 0 else int(tg[b, (v - 1) // 2]) for v in s]
        out, prev = [], -1
        for v in toks:
            if v != 0 and v != prev:
                out.append(v)
            prev = v
        # repeated target labels need a blank between them, which the state walk guarantees
        assert out == tg[b, :S].tolist()
