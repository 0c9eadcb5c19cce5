"""Phase timeline (us, relative to kernel entry) of CTA (0,0,0) of the fused attention backward kernel at the C2 shape:
S2T_GEMM_TRACE=1 python tools/attn_trace.py"""
import ctypes
import os
import sys
import torch
os.environ["S2T_GEMM_TRACE"] = "1"
sys.path.insert(0, ".")
from s2t_b200 import lib, ops
L = lib.load()
NAMES = ["entry", "setup done (barriers, TMEM)", "pdl_wait done", "operands landed", "S, dP ready (softmax warps)",
         "Pd, dS tiles written (MMA thread)", "dQ/dK/dV MMAs retired", "before cluster sync 1", "after cluster sync 1",
         "partials pushed", "after cluster sync 2", "exit"]
B, T, C, H, dk, R = 8, 250, 256, 4, 64, 2000
BF16 = torch.bfloat16
qkv, ctx, dctx = (torch.randn(R, n, device="cuda").mul(0.5).to(BF16) for n in (3 * C, C, C))
lens = torch.full((B,), T, dtype=torch.int32, device="cuda")
lse = torch.zeros(B * H * T, device="cuda")
out = torch.zeros(R, 3 * C, dtype=BF16, device="cuda")
ops.attn_fused_fwd(qkv, C, lens, B, H, T, dk, lse=lse, ctx=ctx, scale=0.125)
for p in (0.0, 0.1):
    spec = (p, 1, 7, None) if p > 0 else None
    best = None
    for _ in range(6):
        torch.cuda.synchronize()
        ops.attn_fused_bwd(qkv, C, ctx, dctx, lse, lens, B, H, T, dk, out=out, scale=0.125, dropout=spec)
        torch.cuda.synchronize()
        buf = (ctypes.c_ulonglong * 16)()
        assert L.s2t_gemm_trace_read(ctypes.cast(buf, ctypes.c_void_p)) == 0
        t = [int(v) for v in buf]
        if best is None or (t[11] - t[0]) < (best[11] - best[0]):
            best = t
    print(f"== attn_fused_bwd p={p}: CTA(0,0,0) entry->exit {(best[11] - best[0]) / 1e3:.2f} us")
    for i, n in enumerate(NAMES):
        if best[i]:
            print(f"   {n:40s} {(best[i] - best[0]) / 1e3:8.2f} us")
