"""CPU tests of the batch collation: tests/golden/collate.npz holds the outputs of the reference's own _collate_frames
and collate_tokens (extracted from its files by source and executed, oracle/make_golden_collate.py).  The restatement,
the host-side packing (s2t_b200/collate.py: what is shipped to the device) and the token collation must agree with it."""
from pathlib import Path

import numpy as np
import torch

GOLDEN = Path(__file__).resolve().parent / "golden" / "collate.npz"
CASES = ("feat80", "feat83", "wave", "single")


def _frames(fx, name):
    F = int(fx[f"{name}::F"])
    lens = fx[f"{name}::lengths"].tolist()
    packed = torch.from_numpy(fx[f"{name}::packed"])
    out, o = [], 0
    for n in lens:
        out.append(packed[o:o + n * F].view(n, F) if F > 1 else packed[o:o + n])
        o += n * F
    return out, F


def test_collate_restatement_and_packing_match_reference_golden():
    from oracle import collate as OC
    from s2t_b200 import collate as PC
    fx = np.load(GOLDEN)
    for name in CASES:
        frames, F = _frames(fx, name)
        padded = OC.collate_frames(frames, F == 1)
        n_sorted, order = OC.collater_order(frames)
        assert np.array_equal(order.numpy(), fx[f"{name}::order"])
        assert np.array_equal(padded.index_select(0, order).numpy(), fx[f"{name}::collated"])
        packed, offsets, lengths, order2 = PC.pack_frames(frames, pin=False)
        assert np.array_equal(packed.numpy(), fx[f"{name}::packed"]) and torch.equal(order2, order)
        assert lengths.tolist() == [f.shape[0] for f in frames]
        # what the device kernel must produce from the packed buffer (pure index arithmetic, checked here on the host)
        T = int(lengths.max())
        ref = torch.zeros((len(frames), T * F))
        for b, src in enumerate(order.tolist()):
            n = int(lengths[src]) * F
            ref[b, :n] = packed[int(offsets[src]): int(offsets[src]) + n]
        assert np.array_equal(ref.view(fx[f"{name}::collated"].shape).numpy(), fx[f"{name}::collated"])


def test_collate_tokens_matches_reference_golden():
    from s2t_b200.collate import collate_tokens
    fx = np.load(GOLDEN)
    flat, lens = torch.from_numpy(fx["tokens::flat"]), fx["tokens::lengths"].tolist()
    toks, o = [], 0
    for n in lens:
        toks.append(flat[o:o + n])
        o += n
    for lp, mv in ((False, False), (False, True), (True, False)):
        assert np.array_equal(collate_tokens(toks, 1, 2, lp, mv).numpy(), fx[f"tokens::{int(lp)}{int(mv)}"])
