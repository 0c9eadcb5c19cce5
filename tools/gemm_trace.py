"""Phase timeline (ns, relative to kernel entry) of CTA 0 of the persistent GEMM kernel: S2T_GEMM_TRACE=1 python tools/gemm_trace.py"""
import ctypes
import os
import sys
import torch
os.environ["S2T_GEMM_TRACE"] = "1"
sys.path.insert(0, ".")
from s2t_b200 import lib, ops
L = lib.load()
NAMES = ["entry", "setup done", "pdl_wait done", "producer: tile0 loads issued", "mma: first stage landed",
         "mma: tile0 committed", "epi: tile0 acc ready", "epi: tile0 stored", "epi: tile1 acc ready", "epi: tile1 stored",
         "epi: stores drained", "exit"]


def run(tag, M, N, K, **epi):
    x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = torch.randn(N, K, device="cuda").to(torch.bfloat16)
    b = torch.randn(N, device="cuda").to(torch.bfloat16)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    flush = torch.empty(64 << 20, device="cuda", dtype=torch.float32)
    best = None
    for _ in range(6):
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.linear(x, w, b, out=out, **epi)
        e1.record()
        torch.cuda.synchronize()
        buf = (ctypes.c_ulonglong * 16)()
        assert L.s2t_gemm_trace_read(ctypes.cast(buf, ctypes.c_void_p)) == 0
        t = [int(v) for v in buf]
        if best is None or (t[11] - t[0]) < (best[11] - best[0]):
            best = t
            ev = e0.elapsed_time(e1) * 1e3
    print(f"== {tag}: M={M} N={N} K={K}  event {ev:.1f} us, CTA0 entry->exit {(best[11] - best[0]) / 1e3:.2f} us")
    for i, n in enumerate(NAMES):
        if best[i]:
            print(f"   {n:32s} {(best[i] - best[0]) / 1e3:8.2f} us")


def run_raw(tag, persistent=False, **kw):
    flush = torch.empty(64 << 20, device="cuda", dtype=torch.float32)
    best = None
    for _ in range(6):
        flush.zero_()
        torch.cuda.synchronize()
        ops.gemm_raw(**kw)
        torch.cuda.synchronize()
        buf = (ctypes.c_ulonglong * 16)()
        assert L.s2t_gemm_trace_read(ctypes.cast(buf, ctypes.c_void_p)) == 0
        t = [int(v) for v in buf]
        if best is None or (t[11] - t[0]) < (best[11] - best[0]):
            best = t
    print(f"== {tag}: CTA0 entry->exit {(best[11] - best[0]) / 1e3:.2f} us")
    names = dict(enumerate(NAMES))
    if not persistent:
        names.update({5: "mma: all k-blocks committed", 6: "all partials visible (cluster sync)", 7: "slice: first group summed",
                      9: "slice: epilogue stored"})
    for i in range(12):
        if best[i]:
            print(f"   {names[i]:40s} {(best[i] - best[0]) / 1e3:8.2f} us")


bf = lambda *s: torch.randn(*s, device="cuda").to(torch.bfloat16)
R = 2000
dy, x256, f2048, w2 = bf(R, 256), bf(R, 256), bf(R, 2048), bf(256, 2048)
# FFN2 forward (split-K): out[R,256] = f[R,2048] @ w2[256,2048]^T + residual
o = torch.empty(R, 256, device="cuda", dtype=torch.bfloat16)
run_raw("ffn2 fwd (K=2048, residual)", M=R, N=256, K=2048, a=f2048, lda=2048, b=w2, ldb=2048, out=o, ldc=256,
        bias=bf(256), residual=x256, ldr=256)
# FFN1 dgrad: dx[R,256] = dh[R,2048] @ w1[2048,256]
w1 = bf(2048, 256)
run_raw("ffn1 dgrad (K=2048, B mn-major)", M=R, N=256, K=2048, a=f2048, lda=2048, b=w1, ldb=256, b_mn=True, out=o, ldc=256)
# wgrad: dw[2048,256] = dh[R,2048]^T @ x[R,256]
dw = torch.empty(2048, 256, device="cuda", dtype=torch.bfloat16)
run_raw("ffn1 wgrad (K=2000, both mn-major)", M=2048, N=256, K=R, a=f2048, lda=2048, a_mn=True, b=x256, ldb=256, b_mn=True,
        out=dw, ldc=256)
dw2 = torch.empty(256, 256, device="cuda", dtype=torch.bfloat16)
run_raw("out-proj wgrad 256x256 (K=2000)", M=256, N=256, K=R, a=dy, lda=256, a_mn=True, b=x256, ldb=256, b_mn=True,
        out=dw2, ldc=256)
run("ffn1 plain", 2000, 2048, 256)
run("ffn1 relu+dropout", 2000, 2048, 256, act=ops.ACT_RELU, dropout=(0.1, 1234, 7, None))
run("qkv", 2000, 768, 256)
run("out-proj", 2000, 256, 256)
run("ctc head", 2000, 10000, 256)

go = torch.empty(R, 2048, device="cuda", dtype=torch.bfloat16)
for cap in (0, 100):
    run_raw(f"ffn2 dgrad + relu gate (N=2048, K=256, B mn-major, max_ctas={cap})", True, M=R, N=2048, K=256, a=dy, lda=256,
            b=w2, ldb=2048, b_mn=True, out=go, ldc=2048, gate=f2048, gate_mode=ops.GATE_POSITIVE, ldg=2048, alpha=1.111,
            max_ctas=cap)
run_raw("ffn2 dgrad without gate", True, M=R, N=2048, K=256, a=dy, lda=256, b=w2, ldb=2048, b_mn=True, out=go, ldc=2048)
wq = bf(768, 256)
gq = bf(R, 768)
run_raw("qkv dgrad (K=768, no split)", True, M=R, N=256, K=768, a=gq, lda=768, b=wq, ldb=256, b_mn=True, out=o, ldc=256,
        split_k=-1)
