"""Regenerate tests/golden/collate.npz: the reference's own _collate_frames / collate_tokens (extracted by source,
oracle/collate.py) on seeded ragged batches; the restatements are asserted equal on the spot (TEST INFRASTRUCTURE).

    python -m oracle.make_golden_collate        # needs /root/reference
"""
from pathlib import Path

import numpy as np
import torch

GOLDEN = Path(__file__).resolve().parent.parent / "tests" / "golden" / "collate.npz"


def batch(seed: int, B: int, F: int, lo: int, hi: int):
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(lo, hi, (B,), generator=g).tolist()
    if B > 2:
        lens[1] = lens[0]                                     # a tie in the sort key
    return [torch.randn((n, F), generator=g) if F > 1 else torch.randn((n,), generator=g) for n in lens]


CASES = {"feat80": (0, 7, 80, 5, 60), "feat83": (1, 4, 83, 1, 30), "wave": (2, 5, 1, 400, 3000), "single": (3, 1, 80, 9, 10)}


def main():
    from oracle import collate as OC
    ref = OC.reference_functions()
    out = {}
    for name, (seed, B, F, lo, hi) in CASES.items():
        frames = batch(seed, B, F, lo, hi)
        audio = F == 1
        padded = ref["_collate_frames"](frames, audio)
        assert torch.equal(padded, OC.collate_frames(frames, audio))
        n_sorted, order = OC.collater_order(frames)
        out[f"{name}::lengths"] = np.array([f.shape[0] for f in frames])
        out[f"{name}::packed"] = torch.cat([f.reshape(-1) for f in frames]).numpy()
        out[f"{name}::order"] = order.numpy()
        out[f"{name}::collated"] = padded.index_select(0, order).numpy()
        out[f"{name}::F"] = np.int64(F)
        print(f"[{name}] B={B} F={F} max_len={padded.shape[1]}: restatement == reference")
    g = torch.Generator().manual_seed(9)
    toks = [torch.cat([torch.randint(4, 100, (n,), generator=g), torch.tensor([2])]) for n in (5, 1, 9, 3)]
    from s2t_b200.collate import collate_tokens
    for lp, mv in ((False, False), (False, True), (True, False)):
        a = ref["collate_tokens"](toks, 1, 2, left_pad=lp, move_eos_to_beginning=mv)
        assert torch.equal(a, collate_tokens(toks, 1, 2, lp, mv))
        out[f"tokens::{int(lp)}{int(mv)}"] = a.numpy()
    out["tokens::flat"] = torch.cat(toks).numpy()
    out["tokens::lengths"] = np.array([t.numel() for t in toks])
    np.savez_compressed(GOLDEN, **out)
    print("wrote", GOLDEN)


if __name__ == "__main__":
    main()
