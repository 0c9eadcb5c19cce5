"""CPU test of the HOST arithmetic of the criterion's ground-truth curriculum (plugin.CtcCriterionB200._curriculum): the
device kernel it calls (s2t_ctc_viterbi) is replaced by the Viterbi restatement and the engine by a recorder, so that the
token / mistake / mask / forced-state arithmetic is compared with oracle.encoder.ground_truth_alignment — the restatement
of criterions/ctc.py:283-433 that the pinned curriculum fixture was produced with."""
from types import SimpleNamespace

import numpy as np
import torch


def test_criterion_curriculum_host_arithmetic_matches_the_restatement(monkeypatch):
    from oracle import encoder as OE
    from oracle import viterbi as OV
    from s2t_b200 import plugin
    B, T2, V, ldv = 3, 17, 30, 32
    g = torch.Generator().manual_seed(0)
    lens = torch.tensor([17, 12, 9], dtype=torch.int32)
    tokens = {"ctc": torch.tensor([[5, 6, 6, 9, 2, 1, 1], [7, 8, 2, 1, 1, 1, 1], [4, 4, 4, 11, 12, 13, 2]]),
              "xctc": torch.tensor([[15, 16, 2, 1, 1], [17, 18, 19, 20, 2], [21, 2, 1, 1, 1]])}
    logits = {k: torch.randn((B * T2, ldv), generator=g) * 2 for k in tokens}
    strip = lambda t: (lambda keep: (torch.stack([torch.cat([r[m], torch.zeros(int((~m).sum()), dtype=r.dtype)]) for r, m in zip(t, keep)]).to(torch.int32),
                                     keep.sum(-1).to(torch.int32)))((t != 1) & (t != 2))
    tg, tl = zip(*(strip(tokens[k]) for k in ("ctc", "xctc")))

    def fake_viterbi(lp, targets, input_lengths, target_lengths, blank=0, batch_first=True, with_nll=True):
        paths, _ = OV.best_alignment(lp.numpy(), targets.numpy().astype(np.int64), input_lengths.tolist(), target_lengths.tolist(), blank)
        out = torch.full((lp.shape[0], lp.shape[1]), -1, dtype=torch.int32)
        for b, p in enumerate(paths):
            out[b, : len(p)] = torch.tensor(p, dtype=torch.int32)
        return out, None

    monkeypatch.setattr(plugin.ops, "ctc_viterbi", fake_viterbi)
    rec = {}
    heads = [SimpleNamespace(name="ctc", V=V, ldv=ldv, tgt=0, pae="pae", inter_layers=(2,)),
             SimpleNamespace(name="xctc", V=V, ldv=ldv, tgt=1, pae="xctc_pae", inter_layers=(2,))]
    eng = SimpleNamespace(heads=heads, forward_train=lambda P: None,
                          set_ground_truth=lambda P, k, tok, msk, fe: rec.__setitem__(k, (tok, msk, fe)))
    P = SimpleNamespace(B=B, T2=T2, out_len=lens, tg=list(tg), tl=list(tl),
                        heads=[SimpleNamespace(logits=logits["ctc"]), SimpleNamespace(logits=logits["xctc"])])
    enc = SimpleNamespace(gt_ratio={"ctc": 1.0, "xctc": 1.0}, pae_adaptive_gt=False, pae_gt_only_mistake=True)
    crit = plugin.CtcCriterionB200(None, None)
    crit._curriculum(enc, eng, P)
    # the restatement on the same logits, in the reference's (T, B, V) layout with its raw token matrices
    pad = torch.arange(T2)[None, :] >= lens[:, None]
    out = {"encoder_padding_mask": [pad],
           "ctc_logit": [logits["ctc"].view(B, T2, ldv)[..., :V].transpose(0, 1)],
           "xctc_logit": [logits["xctc"].view(B, T2, ldv)[..., :V].transpose(0, 1)]}
    ref = OE.ground_truth_alignment(out, tokens["ctc"], tokens["xctc"])
    for k, h in enumerate(("ctc", "xctc")):
        tok, msk, fe = rec[k]
        oracle_tok, bap, mistake, _ = ref[h]
        assert torch.equal(tok, oracle_tok.long()), h
        assert torch.equal(msk, mistake), h                               # ratio 1 + only-mistake: mask == mistake flags
        assert torch.equal(fe, bap.long().masked_fill(~mistake, -1)), h
    # ratio 0 for a head switches its curriculum off
    enc.gt_ratio = {"ctc": 0.0, "xctc": 1.0}
    rec.clear()
    crit._curriculum(enc, eng, P)
    assert rec[0] == (None, None, None) and rec[1][0] is not None


def test_criterion_curriculum_ignores_the_rows_of_a_larger_frame_bucket(monkeypatch):
    """With frame buckets the plan holds more output rows than the batch has (their logits are the head's bias: argmax is
    not blank).  The reference never sees those rows: mistake counts, the adaptive ratio
    (--xctc-pae-ground-truth-ratio-adaptive, criterions/ctc.py:283-330) and the shape of the random draw must be those of
    the batch's own T' frames.  Same seed, same logits: the padded plan gives the unpadded plan's tokens / masks / forced
    states on the first T' columns and 0 / 0 / -1 beyond."""
    from oracle import viterbi as OV
    from s2t_b200 import plugin
    B, Tv, extra, V, ldv = 3, 17, 6, 30, 32
    g = torch.Generator().manual_seed(1)
    lens = torch.tensor([17, 12, 9], dtype=torch.int32)
    tg = torch.tensor([[5, 6, 6, 9], [7, 8, 0, 0], [4, 4, 4, 11]], dtype=torch.int32)
    tl = torch.tensor([4, 2, 4], dtype=torch.int32)
    small = torch.randn((B, Tv, ldv), generator=g) * 2
    big = torch.zeros((B, Tv + extra, ldv))
    big[:, :Tv] = small
    big[:, Tv:, 7] = 9.0                                                   # padded rows: a confident non-blank argmax

    def fake_viterbi(lp, targets, input_lengths, target_lengths, blank=0, batch_first=True, with_nll=True):
        paths, _ = OV.best_alignment(lp.numpy(), targets.numpy().astype(np.int64), input_lengths.tolist(), target_lengths.tolist(), blank)
        out = torch.full((lp.shape[0], lp.shape[1]), -1, dtype=torch.int32)
        for b, p in enumerate(paths):
            out[b, : len(p)] = torch.tensor(p, dtype=torch.int32)
        return out, None

    monkeypatch.setattr(plugin.ops, "ctc_viterbi", fake_viterbi)
    heads = [SimpleNamespace(name="ctc", V=V, ldv=ldv, tgt=0, pae="pae", inter_layers=(2,))]
    enc = SimpleNamespace(gt_ratio={"ctc": 0.9}, pae_adaptive_gt=True, pae_gt_only_mistake=False)
    crit = plugin.CtcCriterionB200(None, None)
    got = {}
    for tag, logits, T2, tv in (("small", small, Tv, None), ("big", big, Tv + extra, Tv)):
        rec = {}
        eng = SimpleNamespace(heads=heads, forward_train=lambda P: None,
                              set_ground_truth=lambda P, k, tok, msk, fe: rec.__setitem__(k, (tok, msk, fe)))
        P = SimpleNamespace(B=B, T2=T2, out_len=lens, tg=[tg], tl=[tl], heads=[SimpleNamespace(logits=logits.reshape(B * T2, ldv))])
        torch.manual_seed(5)
        crit._curriculum(enc, eng, P, tv)
        got[tag] = rec[0]
    for a, b, fill in zip(got["small"], got["big"], (0, False, -1)):
        assert b.shape == (B, Tv + extra)
        assert torch.equal(b[:, :Tv], a)
        assert bool((b[:, Tv:] == fill).all())
    assert bool(got["small"][1].any())
