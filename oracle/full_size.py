"""TEST INFRASTRUCTURE — the full-size parity cases (BASELINE.json configs[2..4] at their stated size).

``case(name)`` builds, from fixed seeds, everything a full-size parity run needs: encoder keywords
(s2t_b200/configs.py), seeded weights, 8 ragged utterances of up to 998 frames and the token matrices.  Used by
  * oracle/make_golden_full.py (this container): runs the REFERENCE's own encoder classes on it, asserts the restatement
    equal, stores compact summaries in tests/golden/enc_full_<name>.npz;
  * tests/test_oracle_cpu.py: the restatement reproduces those summaries;
  * tests/test_engine_gpu.py::test_engine_matches_oracle_full_size: the live oracle run on the GPU box is checked against
    the same summaries before the CUDA path is compared with it.
"""
from __future__ import annotations

from dataclasses import fields
from typing import Dict, Optional

import torch

B, T, S, V = 8, 998, 60, 10000
NAMES = ("c3", "c4", "c5")


def case(name: str) -> Dict:
    from oracle import encoder as OE
    from oracle.make_golden import seeded_state_dict
    from s2t_b200.configs import encoder_kwargs
    kw = encoder_kwargs(name, dropout=0.0)
    okw = {k: v for k, v in kw.items() if k in {f.name for f in fields(OE.OracleConfig)}}
    pds = kw.get("encoder_type") == "pds"
    dual = kw.get("tgt_vocab", 0) > 0
    torch.manual_seed(0)
    cfg = OE.OracleConfig(**okw)
    orc = (OE.PDSS2TTransformerEncoder if pds else OE.S2TTransformerEncoder)(cfg)
    shapes = {k: tuple(v.shape) for k, v in orc.state_dict().items()}
    sd = seeded_state_dict(shapes, 11)
    orc.load_state_dict(sd)
    orc.train()
    g = torch.Generator().manual_seed(12)
    feats = torch.randn((B, T, 80), generator=g)
    lens = torch.tensor([998] + sorted([int(v) for v in torch.randint(500, 998, (B - 1,), generator=g)], reverse=True))
    for b in range(B):
        feats[b, lens[b]:] = 0

    def toks(seed, smax):
        gg = torch.Generator().manual_seed(seed)
        t = torch.full((B, smax + 1), 1, dtype=torch.long)
        for b in range(B):
            n = smax - 3 * b
            t[b, :n] = torch.randint(4, V, (n,), generator=gg)
            t[b, n] = 2
        return t

    tokens = toks(13, S)
    tgt_tokens: Optional[torch.Tensor] = toks(14, S - 6) if dual else None
    return dict(name=name, kw=kw, cfg=cfg, pds=pds, dual=dual, oracle=orc, sd=sd, feats=feats, lens=lens, tokens=tokens,
                tgt_tokens=tgt_tokens)


def loss_of(c: Dict, out) -> torch.Tensor:
    from oracle import encoder as OE
    kw = c["kw"]
    if c["dual"]:
        return OE.ctc_criterion_total(out, c["tokens"], c["tgt_tokens"], ctc_weight=kw["ctc_weight"],
                                      xctc_weight=kw["xctc_weight"], inter_ctc_weight=kw["inter_ctc_weight"],
                                      inter_xctc_weight=kw["inter_xctc_weight"])
    return OE.ctc_criterion_loss(out["ctc_logit"][0], out["encoder_padding_mask"][0], c["tokens"])


def summary(out, loss: torch.Tensor, grads: Dict[str, torch.Tensor]) -> Dict:
    """Compact, size-independent description of one run: loss, output lengths, per-utterance mean |encoder_out| over the
    valid frames, a fixed slice of encoder_out, and the gradient norm of every parameter."""
    import numpy as np
    pad = out["encoder_padding_mask"][0]
    enc = out["encoder_out"][0].detach()                                   # T' x B x C
    valid = (~pad).T.unsqueeze(-1).to(enc.dtype)
    per_utt = (enc.abs() * valid).sum((0, 2)) / valid.sum((0, 2)).clamp(min=1) / enc.shape[-1]
    names = sorted(grads)
    return dict(loss=np.float64(loss.item()), out_len=(~pad).sum(-1).numpy(), enc_abs_mean=per_utt.double().numpy(),
                enc_slice=enc[:4, :, :16].double().numpy(), grad_names=np.array(names),
                grad_norms=np.array([float(grads[k].double().norm()) for k in names]))
