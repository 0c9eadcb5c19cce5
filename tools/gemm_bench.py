"""Micro-benchmark of s2t_gemm_bf16 at the shapes of the C1/C2 training step (CUDA events, warm, back to back).
Prints us/launch and TFLOP/s for each tile width, next to torch.matmul (cuBLAS) as the achievable bar."""
import sys
import torch
sys.path.insert(0, ".")
from s2t_b200 import ops

BF16 = torch.bfloat16
dev = "cuda"


def timeit(fn, n=50, reps=5):
    """us per launch with the host out of the way: n launches captured in a CUDA graph, replayed reps times."""
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (n * reps)


def main():
    shapes = [  # (name, M, N, K, kind)
        ("qkv 2000x768x256", 2000, 768, 256, "lin"), ("out 2000x256x256", 2000, 256, 256, "lin"),
        ("ffn1 2000x2048x256", 2000, 2048, 256, "lin"), ("ffn2 2000x256x2048", 2000, 256, 2048, "lin"),
        ("ctc 2000x10000x256", 2000, 10000, 256, "lin"), ("ffn1 10000x2048x256", 10000, 2048, 256, "lin"),
        ("dgrad ffn2 2000x2048x256", 2000, 2048, 256, "dgrad"), ("wgrad ffn1 2048x256x2000", 2048, 256, 2000, "wgrad"),
        ("wgrad ctc 10000x256x2000", 10000, 256, 2000, "wgrad"), ("dgrad ctc 2000x256x10000", 2000, 256, 10000, "dgrad"),
        ("wgrad ffn2 256x2048x2000", 256, 2048, 2000, "wgrad"), ("wgrad out 256x256x2000", 256, 256, 2000, "wgrad"),
        ("wgrad qkv 768x256x2000", 768, 256, 2000, "wgrad"), ("dgrad qkv 2000x256x768", 2000, 256, 768, "dgrad"),
    ]
    for name, M, N, K, kind in shapes:
        if kind == "lin":
            x = torch.randn(M, K, device=dev).to(BF16); w = torch.randn(N, K, device=dev).to(BF16)
            out = torch.empty(M, N, device=dev, dtype=BF16)
            mk = lambda tn: (lambda: ops.linear(x, w, out=out, tile_n=tn))
            ref = lambda: torch.matmul(x, w.t(), out=out)
        elif kind == "dgrad":
            x = torch.randn(M, K, device=dev).to(BF16); w = torch.randn(K, N, device=dev).to(BF16)
            out = torch.empty(M, N, device=dev, dtype=BF16)
            mk = lambda tn: (lambda: ops.linear_dgrad(x, w, out=out, tile_n=tn))
            ref = lambda: torch.matmul(x, w, out=out)
        else:
            dy = torch.randn(K, M, device=dev).to(BF16); x = torch.randn(K, N, device=dev).to(BF16)
            out = torch.empty(M, N, device=dev, dtype=BF16)
            mk = lambda tn: (lambda: ops.linear_wgrad(dy, x, out=out, tile_n=tn))
            ref = lambda: torch.matmul(dy.t(), x, out=out)
        fl = 2.0 * M * N * K
        row = [f"{name:28s}"]
        for tn in (0,):
            us = timeit(mk(tn))
            row.append(f"auto: {us:7.1f}us {fl / us / 1e6:7.1f}TF")
        us = timeit(ref)
        row.append(f"cublas: {us:7.1f}us {fl / us / 1e6:7.1f}TF")
        print("  ".join(row), flush=True)
    # attention-shaped batched GEMMs
    B, T, H, D = 8, 250, 4, 64
    C = H * D
    qkv = torch.randn(B * T, 3 * C, device=dev).to(BF16)
    sc = torch.zeros(B * H, T, 256, device=dev, dtype=BF16)
    f = lambda: ops.gemm_raw(M=T, N=T, K=D, a=qkv, lda=3 * C, a_bs=(D, T * 3 * C), b=qkv, ldb=3 * C, b_off=C,
                             b_bs=(D, T * 3 * C), out=sc, ldc=256, c_bs=(T * 256, H * T * 256), batch=(H, B), alpha=0.125)
    print(f"qk^T 32x(250x250x64): {timeit(f):7.1f}us")
    ctx = torch.zeros(B * T, C, device=dev, dtype=BF16)
    f = lambda: ops.gemm_raw(M=T, N=D, K=T, a=sc, lda=256, a_bs=(T * 256, H * T * 256), b=qkv, ldb=3 * C, b_off=2 * C,
                             b_mn=True, b_bs=(D, T * 3 * C), out=ctx, ldc=C, c_bs=(D, T * C), batch=(H, B))
    print(f"pv 32x(250x64x250): {timeit(f):7.1f}us")
    x = torch.randn(2000, 256, device=dev).to(BF16)
    g = torch.ones(256, device=dev, dtype=BF16)
    y = torch.empty_like(x)
    print(f"layernorm_fwd 2000x256: {timeit(lambda: ops.layernorm_fwd(x, g, g, out=y)):7.1f}us")
    ws = torch.empty(64 * 2048, device=dev, dtype=torch.float32)
    o = torch.empty(256, device=dev, dtype=BF16)
    print(f"colsum 2000x256: {timeit(lambda: ops.colsum(x, o, workspace=ws)):7.1f}us")


if __name__ == "__main__":
    main()
