"""Isolated device times of the round-2 kernels at the C2 shapes (B = 8, T' = 250, C = 256, H = 4, F = 2048): each kernel
is captured 20x back to back into a CUDA graph and replayed between CUDA events (per-launch time incl. the launch gap
inside a graph).  usage: python tools/kernel_bench.py"""
import sys
import torch
sys.path.insert(0, ".")
from s2t_b200 import ops
from s2t_b200.params import BF16

torch.cuda.set_device(0)
B, T, C, H, dk, F, R = 8, 250, 256, 4, 64, 2048, 2000
g = torch.Generator().manual_seed(0)
rnd = lambda *s: (torch.randn(s, generator=g) * 0.5).to(BF16).cuda()


def timed(name, fn, n=20, reps=10):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(); torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            for _ in range(n):
                fn()
        gr.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            gr.replay()
        e1.record(s)
    torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / (n * reps)
    print(f"{name:48s} {us:8.2f} us")
    return us


qkv, ctx, dctx = rnd(R, 3 * C), rnd(R, C), rnd(R, C)
lens = torch.full((B,), T, dtype=torch.int32).cuda()
lse = torch.zeros(B * H * T, device="cuda")
out = torch.zeros(R, 3 * C, dtype=BF16, device="cuda")
cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
for p in (0.0, 0.1):
    spec = (p, 1, 7, cnt) if p > 0 else None
    timed(f"attn_fused_fwd p={p}", lambda: ops.attn_fused_fwd(qkv, C, lens, B, H, T, dk, lse=lse, ctx=ctx, scale=0.125, dropout=spec))
    timed(f"attn_fused_bwd p={p}", lambda: ops.attn_fused_bwd(qkv, C, ctx, dctx, lse, lens, B, H, T, dk, out=out, scale=0.125, dropout=spec))
dy256, x256, dyF, xF, dy768 = rnd(R, C), rnd(R, C), rnd(R, F), rnd(R, F), rnd(R, 3 * C)
for (nm, dy, x) in (("wgrad out_proj [256x256]", dy256, x256), ("wgrad w_2 [256x2048]", dy256, xF),
                    ("wgrad w_1 [2048x256]", dyF, x256), ("wgrad qkv [768x256]", dy768, x256)):
    N = dy.shape[1]
    dw = torch.zeros(N, x.shape[1], dtype=BF16, device="cuda")
    db = torch.zeros(N, dtype=BF16, device="cuda")
    for mc in (0, 48):
        a = timed(f"{nm} max_ctas={mc}", lambda: ops.linear_wgrad(dy, x, out=dw, max_ctas=mc, split_k=-1))
        b = timed(f"{nm} max_ctas={mc} + colsum", lambda: ops.linear_wgrad(dy, x, out=dw, colsum=db, max_ctas=mc, split_k=-1))
    ws = torch.zeros(ops._lib.load().s2t_colsum_workspace_floats(R, N), device="cuda")
    timed(f"standalone colsum [{N}]", lambda: ops.colsum(dy, db, workspace=ws))
