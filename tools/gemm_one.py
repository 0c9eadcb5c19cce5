"""One GEMM shape, a few launches (for ncu --set full)."""
import sys
import torch
sys.path.insert(0, ".")
from s2t_b200 import ops
M, N, K, tn = (int(v) for v in sys.argv[1:5])
x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
w = torch.randn(N, K, device="cuda").to(torch.bfloat16)
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(5):
    ops.linear(x, w, out=out, tile_n=tn)
torch.cuda.synchronize()
print("ok")
