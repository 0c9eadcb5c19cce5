"""Experiment: step time with the side-stream work (weight/bias gradients, LayerNorm dgamma/dbeta reductions) suppressed,
i.e. the length of the forward + dgrad chain alone — the lower bound the multi-stream backward can approach.
The gradients of such a step are WRONG; this is a scheduling measurement only.  usage: python tools/chain_only.py"""
import sys
import torch
sys.path.insert(0, ".")
import bench
from s2t_b200 import frontend, ops
from s2t_b200.engine import EncoderConfig, S2TEncoderEngine


def timed(eng, P, n=50):
    for _ in range(5):
        eng.run(P)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        eng.run(P)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1000.0


def build(skip):
    eng = S2TEncoderEngine(EncoderConfig(vocab=bench.VOCAB, dropout=0.1))
    n = 160000
    P = eng.plan(8, frontend.num_frames(n), bench.TARGET_LEN, wave_len=n)
    w, ns, tg, tl = bench.synth_batch(8, 10.0, 0)
    eng.set_inputs(P, wave=w, num_samples=ns, targets=tg, target_lengths=tl)
    sides = set()
    if skip:
        names = ["gemm_raw", "colsum", "layernorm_bwd_reduce"]
        orig = {k: getattr(ops, k) for k in names}
        for k in names:
            def wrap(*a, _f=orig[k], **kw):
                if torch.cuda.current_stream() in sides:
                    return kw.get("out")
                return _f(*a, **kw)
            setattr(ops, k, wrap)
    eng.train_step(P)
    torch.cuda.synchronize()
    sides.update(eng._side)
    sides.add(eng._ln_stream)
    eng.capture(P)
    t = timed(eng, P)
    if skip:
        for k, f in orig.items():
            setattr(ops, k, f)
    return t


torch.cuda.set_device(0)
full = build(False)
chain = build(True)
print(f"full step {full:.1f} us   chain only {chain:.1f} us   contention + joins {full - chain:.1f} us")
