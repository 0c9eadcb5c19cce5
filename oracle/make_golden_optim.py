"""Regenerate tests/golden/fused_adam.npz: three updates of the reference's own Adam class + clip_grad_norm_ (extracted by
source, oracle/optim.py) on seeded bf16 parameters / gradients; the restatement is asserted equal (TEST INFRASTRUCTURE).

    python -m oracle.make_golden_optim          # needs /root/reference
"""
from pathlib import Path

import numpy as np
import torch

GOLDEN = Path(__file__).resolve().parent.parent / "tests" / "golden" / "fused_adam.npz"
CASES = {
    # name: (n, lr, betas, eps, weight_decay, grad_mult, max_norm)
    "recipe": (4099, 2e-3, (0.9, 0.98), 1e-8, 0.0, 1.0 / 480.0, 10.0),          # base.yaml: adam_betas (0.9, 0.98), clip-norm 10
    "clipped_wd": (1000, 1e-3, (0.9, 0.999), 1e-6, 0.01, 0.5, 0.05),
    "no_clip": (257, 5e-4, (0.8, 0.9), 1e-8, 0.0, 1.0, 0.0),
}


def main():
    from oracle import optim as OO
    ref = OO.reference_objects()
    out = {}
    for name, (n, lr, betas, eps, wd, mult, max_norm) in CASES.items():
        g = torch.Generator().manual_seed(len(name))
        p0 = (torch.randn(n, generator=g) * 0.1).to(torch.bfloat16)
        grads = [(torch.randn(n, generator=g) * (30.0 if name == "recipe" else 1.0)).to(torch.bfloat16) for _ in range(3)]
        # reference: fp32 master parameter, its .grad filled from the bf16 gradient (fp16_optimizer: copy + unscale)
        master = torch.nn.Parameter(p0.float().clone())
        opt = ref["Adam"]([master], lr=lr, betas=betas, eps=eps, weight_decay=wd)
        m2, ea, es = p0.float().clone(), torch.zeros(n), torch.zeros(n)
        norms = []
        for step, gr in enumerate(grads, start=1):
            master.grad = gr.float() * mult
            tn = ref["clip_grad_norm_"]([master], max_norm)
            opt.step()
            tn2, bf = OO.update(m2, ea, es, gr, step, lr=lr, betas=betas, eps=eps, weight_decay=wd, grad_mult=mult,
                                max_norm=max_norm)
            assert torch.allclose(tn, tn2, rtol=1e-6) and torch.allclose(master.data, m2, rtol=1e-6, atol=1e-9), (name, step)
            norms.append(float(tn))
            out[f"{name}::master{step}"] = master.data.clone().numpy()
        st = opt.state[master]
        assert torch.allclose(st["exp_avg"], ea, rtol=1e-6, atol=1e-12) and torch.allclose(st["exp_avg_sq"], es, rtol=1e-6, atol=1e-12)
        out[f"{name}::p0"] = p0.float().numpy()
        out[f"{name}::grads"] = torch.stack(grads).float().numpy()
        out[f"{name}::norms"] = np.array(norms)
        out[f"{name}::exp_avg"] = st["exp_avg"].numpy()
        out[f"{name}::exp_avg_sq"] = st["exp_avg_sq"].numpy()
        out[f"{name}::hyper"] = np.array([lr, betas[0], betas[1], eps, wd, mult, max_norm])
        print(f"[{name}] n={n}: restatement == reference over 3 updates, gnorm {norms}")
    np.savez_compressed(GOLDEN, **out)
    print("wrote", GOLDEN)


if __name__ == "__main__":
    main()
