"""Regenerate tests/golden/enc_full_{c3,c4,c5}.npz from the REFERENCE itself at the BASELINE sizes (TEST INFRASTRUCTURE).

    python -m oracle.make_golden_full          # needs /root/reference (or S2T_REFERENCE)

For every full-size case (oracle/full_size.py: pdss2t_transformer_s_8 3_3_3_3; Conformer 12 layers; BiL-CTC 18 layers
d 512 with dual heads, inter-CTC at 6,9,12,15 and inter_league adapters; 8 ragged utterances of up to 998 frames,
V = 10000) the reference's unmodified encoder class (oracle/ref_shim.py) is run forward + backward on the seeded inputs
and weights, the restatement oracle/encoder.py is asserted equal on the spot (outputs |d| < 2e-4, loss rel 1e-5, every
gradient rel-L2 < 1e-3), and a compact summary of the REFERENCE's results is stored (full gradients of a 140 M parameter
model would be 560 MB).
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
GOLDEN = ROOT / "tests" / "golden"


def run(name: str):
    from oracle import full_size as FS
    from oracle.make_golden import build_reference_encoder
    c = FS.case(name)
    ref = build_reference_encoder(c["cfg"], c["pds"])
    missing = ref.load_state_dict(c["sd"], strict=False)
    assert not missing.unexpected_keys and all(k.endswith("_float_tensor") for k in missing.missing_keys), missing
    ref.train()
    out = ref(c["feats"], c["lens"])
    loss = FS.loss_of(c, out)
    loss.backward()
    rg = {k: p.grad for k, p in ref.named_parameters() if p.grad is not None}
    orc = c["oracle"]
    o_out = orc(c["feats"], c["lens"])
    o_loss = FS.loss_of(c, o_out)
    o_loss.backward()
    og = {k: p.grad for k, p in orc.named_parameters() if p.grad is not None}
    assert set(og) == set(rg), sorted(set(og) ^ set(rg))[:5]
    dx = (o_out["encoder_out"][0] - out["encoder_out"][0]).abs().max().item()
    assert dx < 2e-4, dx
    assert abs(o_loss.item() - loss.item()) <= 1e-5 * abs(loss.item()), (o_loss.item(), loss.item())
    worst = max(((og[k] - rg[k]).norm().item() / (rg[k].norm().item() + 1e-12), k) for k in rg)
    assert worst[0] < 1e-3, worst
    print(f"[full {name}] oracle == reference: |d enc_out| {dx:.2e}, loss {loss.item():.4f}, worst gradient rel-L2 "
          f"{worst[0]:.2e} ({worst[1]}), {len(rg)} parameter tensors")
    fix = FS.summary(out, loss, rg)
    fix["torch_version"] = torch.__version__
    np.savez_compressed(GOLDEN / f"enc_full_{name}.npz", **fix)


def main():
    from oracle import full_size as FS
    for name in FS.NAMES:
        run(name)


if __name__ == "__main__":
    main()
