"""CPU test of the update arithmetic: tests/golden/fused_adam.npz holds three updates of the reference's own Adam class
and clip_grad_norm_ (extracted from its files by source, oracle/make_golden_optim.py); the restatement oracle/optim.py —
the checker of the fused CUDA update — must reproduce them."""
from pathlib import Path

import numpy as np
import torch

GOLDEN = Path(__file__).resolve().parent / "golden" / "fused_adam.npz"


def test_update_restatement_matches_reference_golden():
    from oracle import optim as OO
    fx = np.load(GOLDEN)
    for name in ("recipe", "clipped_wd", "no_clip"):
        lr, b1, b2, eps, wd, mult, max_norm = (float(v) for v in fx[f"{name}::hyper"])
        master = torch.from_numpy(fx[f"{name}::p0"]).clone()
        ea, es = torch.zeros_like(master), torch.zeros_like(master)
        for step in (1, 2, 3):
            g = torch.from_numpy(fx[f"{name}::grads"][step - 1]).to(torch.bfloat16)
            tn, bf = OO.update(master, ea, es, g, step, lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd, grad_mult=mult,
                               max_norm=max_norm)
            assert abs(float(tn) - fx[f"{name}::norms"][step - 1]) <= 1e-5 * fx[f"{name}::norms"][step - 1]
            assert np.allclose(master.numpy(), fx[f"{name}::master{step}"], rtol=1e-6, atol=1e-9)
            assert torch.equal(bf, master.to(torch.bfloat16))
        assert np.allclose(ea.numpy(), fx[f"{name}::exp_avg"], rtol=1e-6, atol=1e-12)
        assert np.allclose(es.numpy(), fx[f"{name}::exp_avg_sq"], rtol=1e-6, atol=1e-12)
