"""Sweep (tile_n, split_k) for few-tile K-heavy GEMMs to calibrate the dispatch cost model."""
import os
import sys
import torch
sys.path.insert(0, ".")
sys.path.insert(0, "tools")
from s2t_b200 import ops
from gemm_bench import timeit
BF16 = torch.bfloat16
MC = int(os.environ.get("MC", "0"))     # SM share (max_ctas) applied to every launch
for name, M, N, K, kind in [("wgrad out", 256, 256, 2000, "wgrad"), ("wgrad qkv", 768, 256, 2000, "wgrad"),
                            ("wgrad ffn1", 2048, 256, 2000, "wgrad"), ("wgrad ffn2", 256, 2048, 2000, "wgrad"),
                            ("ffn2", 2000, 256, 2048, "lin"), ("dgrad ffn1", 2000, 256, 2048, "dgrad"),
                            ("dgrad qkv", 2000, 256, 768, "dgrad"),
                            ("out", 2000, 256, 256, "lin"), ("dgrad ctc", 2000, 256, 10000, "dgrad")]:
    if kind == "lin":
        a = torch.randn(M, K, device="cuda").to(BF16); b = torch.randn(N, K, device="cuda").to(BF16)
        out = torch.empty(M, N, device="cuda", dtype=BF16)
        mk = lambda tn, sk: (lambda: ops.linear(a, b, out=out, tile_n=tn, split_k=sk, max_ctas=MC))
    elif kind == "dgrad":
        a = torch.randn(M, K, device="cuda").to(BF16); b = torch.randn(K, N, device="cuda").to(BF16)
        out = torch.empty(M, N, device="cuda", dtype=BF16)
        mk = lambda tn, sk: (lambda: ops.linear_dgrad(a, b, out=out, tile_n=tn, split_k=sk, max_ctas=MC))
    else:
        a = torch.randn(K, M, device="cuda").to(BF16); b = torch.randn(K, N, device="cuda").to(BF16)
        out = torch.empty(M, N, device="cuda", dtype=BF16)
        mk = lambda tn, sk: (lambda: ops.linear_wgrad(a, b, out=out, tile_n=tn, split_k=sk, max_ctas=MC))
    print(f"{name} {M}x{N}x{K}: auto {timeit(mk(0, 0)):.1f}us")
    for tn in (64, 128, 256):
        print("   bn%3d:" % tn, "  ".join(f"s{sk}:{timeit(mk(tn, sk)):5.1f}" for sk in (-1, 2, 3, 4, 6, 8)))
