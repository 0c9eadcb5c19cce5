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


def test_reference_optimizer_state_format_round_trip():
    """pack_reference_state / unpack_reference_state: the flat moments, cut into named tensors, become a torch
    ``Optimizer.state_dict()`` (the format fairseq's Adam, a torch.optim.Optimizer subclass, saves as
    last_optimizer_state).  A torch Optimizer over parameters of the same shapes must accept it (load_state_dict checks
    group sizes and maps state by position), the moments must land on the right parameters when the reference's
    parameter order differs from the engine's, and unpacking returns what was packed."""
    from s2t_b200.optim import pack_reference_state, unpack_reference_state
    g = torch.Generator().manual_seed(0)
    shapes = {"a.weight": (4, 3), "a.bias": (4,), "b.weight": (2, 4, 5)}
    ea = {n: torch.randn(s, generator=g) for n, s in shapes.items()}
    es = {n: torch.rand(s, generator=g) for n, s in shapes.items()}
    ref_order = ["b.weight", "a.weight", "a.bias"]                       # the reference model's own parameter order
    sd = pack_reference_state(ea, es, 7, ref_order, lr=2e-3, betas=(0.9, 0.98), eps=1e-8, weight_decay=1e-4)
    params = [torch.nn.Parameter(torch.zeros(shapes[n])) for n in ref_order]
    opt = torch.optim.Adam(params, lr=1.0)
    opt.load_state_dict(sd)
    assert opt.param_groups[0]["lr"] == 2e-3 and opt.param_groups[0]["betas"] == (0.9, 0.98)
    for p, n in zip(params, ref_order):
        assert torch.equal(opt.state[p]["exp_avg"], ea[n]) and torch.equal(opt.state[p]["exp_avg_sq"], es[n])
        assert int(opt.state[p]["step"]) == 7
    ea2, es2, step = unpack_reference_state(opt.state_dict(), ref_order)
    assert step == 7 and all(torch.equal(ea2[n], ea[n]) and torch.equal(es2[n], es[n]) for n in shapes)
    import pytest
    with pytest.raises(ValueError):
        unpack_reference_state(sd, ref_order[:2])
    sd["state"][1]["step"] = 8
    with pytest.raises(ValueError):
        unpack_reference_state(sd, ref_order)
