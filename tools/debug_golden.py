"""Print per-parameter gradient rel-L2 of the engine against a golden fixture (debug aid)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from oracle.make_golden import ENC_CASES, seeded_state_dict
from oracle import encoder as OE
from s2t_b200.engine import EncoderConfig, S2TEncoderEngine
sys.path.insert(0, "tests")

case = sys.argv[1] if len(sys.argv) > 1 else "c3_small"
fx = np.load(f"tests/golden/enc_{case}.npz")
spec = ENC_CASES[case]
ints = lambda k: tuple(int(v) for v in spec[k].split("_"))
cfg = dict(encoder_type="pds", vocab=spec["vocab"], dropout=0.0, pds_layers=ints("pds_layers"),
           pds_ratios=ints("pds_ratios"), pds_dims=ints("pds_embed_dims"), pds_kernels=ints("pds_kernel_sizes"),
           pds_ffn_ratios=ints("pds_ffn_ratios"), pds_heads=ints("pds_attn_heads"), pds_embed_norm=True)
shapes = {n: eval(s) for n, s in zip(fx["param_names"], fx["param_shapes"])}
sd = seeded_state_dict(shapes, int(fx["seed"]))
eng = S2TEncoderEngine(EncoderConfig(**cfg)); eng.load_state_dict(sd)
feats, lens, tokens = (torch.from_numpy(fx[k]) for k in ("feats", "lens", "tokens"))
tg, tl = eng.strip_targets(tokens)
P = eng.plan(feats.shape[0], feats.shape[1], tg.shape[1])
eng.set_inputs(P, feats=feats, src_lengths=lens.to(torch.int32), targets=tg, target_lengths=tl)
eng.train_step(P); torch.cuda.synchronize()
okw = dict(cfg); okw.pop("encoder_type")
orc = OE.PDSS2TTransformerEncoder(OE.OracleConfig(**okw)); orc.load_state_dict(sd); orc.train()
out = orc(feats, lens)
loss = OE.ctc_criterion_loss(out["ctc_logit"][0], out["encoder_padding_mask"][0], tokens); loss.backward()
print("loss", P.loss.item(), loss.item())
g = eng.grads()
for k, p in orc.named_parameters():
    r = float((g[k].cpu().double() - p.grad.double()).norm() / (p.grad.double().norm() + 1e-12))
    print(f"{k:50s} {r:.4f}  |g|={p.grad.norm().item():.3e}")
