"""Diagnostic: full-size C5, gradients w.r.t. every CTC logit set (engine dlogits vs oracle autograd)."""
import sys
import torch
sys.path.insert(0, ".")
from dataclasses import fields
from oracle import encoder as OE
from oracle.make_golden import seeded_state_dict
from s2t_b200.configs import encoder_kwargs
from s2t_b200.engine import EncoderConfig, S2TEncoderEngine

kw = encoder_kwargs("c5", dropout=0.0)
okw = {k: v for k, v in kw.items() if k in {f.name for f in fields(OE.OracleConfig)}}
torch.manual_seed(0)
orc = OE.S2TTransformerEncoder(OE.OracleConfig(**okw))
shapes = {k: tuple(v.shape) for k, v in orc.state_dict().items()}
sd = seeded_state_dict(shapes, 11)
orc.load_state_dict(sd); orc.train()
B, T, S, V = 8, 998, 60, 10000
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
        t[b, :n] = torch.randint(4, V, (n,), generator=gg); t[b, n] = 2
    return t
tokens, tgt = toks(13, S), toks(14, S - 6)
out = orc(feats, lens)
logit_sets = {("top", 0): out["ctc_logit"][0], ("top", 1): out["xctc_logit"][0]}
for i, l in enumerate(kw["inter_ctc_layers"]):
    logit_sets[("inter", l, 0)] = out["inter_ctc_logits"][i][0]
    logit_sets[("inter", l, 1)] = out["inter_xctc_logits"][i]
for t in logit_sets.values():
    t.retain_grad()
loss = OE.ctc_criterion_total(out, tokens, tgt, ctc_weight=kw["ctc_weight"], xctc_weight=kw["xctc_weight"],
                              inter_ctc_weight=kw["inter_ctc_weight"], inter_xctc_weight=kw["inter_xctc_weight"])
loss.backward()
eng = S2TEncoderEngine(EncoderConfig(**kw)); eng.load_state_dict(sd)
tg, tl = eng.strip_targets(tokens); xt, xl = eng.strip_targets(tgt)
P = eng.plan(B, T, max(tg.shape[1], xt.shape[1]))
eng.set_inputs(P, feats=feats, src_lengths=lens.to(torch.int32), targets=tg, target_lengths=tl, xtargets=xt, xtarget_lengths=xl)
# CTC-loss part of every dlogits buffer alone (forward + loss only) against autograd of the single weighted term
eng._forward(P); eng._loss(P, need_grad=True); torch.cuda.synchronize()
import torch.nn.functional as F
wts = {("top", 0): kw["ctc_weight"], ("top", 1): kw["xctc_weight"]}
for l in kw["inter_ctc_layers"]:
    wts[("inter", l, 0)] = kw["inter_ctc_weight"] / 4
    wts[("inter", l, 1)] = kw["inter_xctc_weight"] / 4
pad0 = out["encoder_padding_mask"][0]
for key, ref in logit_sets.items():
    buf = P.heads[key[1]].dlogits if key[0] == "top" else P.inter[(key[1], key[2])].dlogits
    lg = (P.heads[key[1]].logits if key[0] == "top" else P.inter[(key[1], key[2])].logits)
    mine = buf.float().view(B, P.T2, -1)[..., :V].transpose(0, 1).cpu()
    x = lg.float().view(B, P.T2, -1)[..., :V].transpose(0, 1).cpu().clone().requires_grad_(True)   # the ENGINE's logits as a leaf
    term = wts[key] * OE.ctc_criterion_loss(x, pad0, tgt if key[-1] == 1 else tokens)
    gref, = torch.autograd.grad(term, x)
    v = ~pad0.T
    print("ctc-only", key, "rel", round(float((mine[v] - gref[v]).norm() / gref[v].norm()), 4), "|g|", float(gref[v].norm()))
    ctc_part = globals().setdefault("ctc_part", {})
    xo = ref.detach().clone().requires_grad_(True)          # the ORACLE's logits as a leaf: its own CTC-only gradient
    go, = torch.autograd.grad(wts[key] * OE.ctc_criterion_loss(xo, pad0, tgt if key[-1] == 1 else tokens), xo)
    ctc_part[key] = (mine.clone(), go)
eng.step_counter.zero_()
eng.train_step(P); torch.cuda.synchronize()
pad = out["encoder_padding_mask"][0]
valid = ~pad.T
rel = lambda a, b: float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
for key, ref in logit_sets.items():
    if key[0] == "top":
        buf, lg = P.heads[key[1]].dlogits, P.heads[key[1]].logits
    else:
        buf, lg = P.inter[(key[1], key[2])].dlogits, P.inter[(key[1], key[2])].logits
    mine = buf.float().view(B, P.T2, -1)[..., :V].transpose(0, 1).cpu()
    mlg = lg.float().view(B, P.T2, -1)[..., :V].transpose(0, 1).cpu()
    print(key, "logits rel", round(rel(mlg[valid], ref.detach()[valid]), 4), "dlogits rel", round(rel(mine[valid], ref.grad[valid]), 4),
          "|ref grad|", float(ref.grad[valid].norm()))
    if key[0] == "inter":
        a_mine = (mine - ctc_part[key][0])[valid]
        a_ref = (ref.grad - ctc_part[key][1])[valid]
        print("   adapter part: |mine|", float(a_mine.norm()), "|ref|", float(a_ref.norm()), "rel", round(rel(a_mine, a_ref), 4),
              "cos", float((a_mine.double() * a_ref.double()).sum() / (a_mine.double().norm() * a_ref.double().norm() + 1e-30)))
grads = eng.grads()
for n in ("ctc_norm6", "xctc_norm6", "ctc_norm9", "xctc_norm9", "ctc_norm12", "xctc_norm12", "ctc_norm15", "xctc_norm15"):
    p = dict(orc.named_parameters())[n + ".weight"]
    print(n, "weight grad rel", round(rel(grads[n + ".weight"].cpu(), p.grad), 4), "|ref|", float(p.grad.norm()))
