import torch, sys
sys.path.insert(0, '.')
from s2t_b200 import ops
BF16 = torch.bfloat16
def run(B, T, Cin, Cout, k=5):
    g = torch.Generator().manual_seed(0)
    T_out = (T + 4 - k) // 2 + 1
    x = torch.randn((B, T, Cin), generator=g).to(BF16)
    w = (torch.randn((Cout, Cin, k), generator=g) * 0.1).to(BF16)
    Tp = T + 5
    xp = torch.zeros((B, Tp, Cin), dtype=BF16); xp[:, 2:2 + T] = x
    xpd = xp.cuda(); wp = ops.permute_conv_weight(w.cuda())
    out = torch.zeros((B * T_out, Cout), device="cuda", dtype=BF16)
    ops.gemm_raw(M=T_out, N=Cout, K=k * Cin, a=xpd, lda=2 * Cin, a_bs=(Tp * Cin, 0), b=wp.view(Cout, k * Cin),
                 ldb=k * Cin, out=out, ldc=Cout, c_bs=(T_out * Cout, 0), batch=(B, 1))
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv1d(x.float().transpose(1, 2), w.float(), None, stride=2, padding=2).transpose(1, 2).reshape(B * T_out, Cout)
    err = (out.float().cpu() - ref).abs().max(dim=1).values
    bad = (err > 0.1).nonzero().flatten().tolist()
    print(f"B={B} T={T} Cin={Cin} Cout={Cout} T_out={T_out} bad rows: {bad[:20]}{'...' if len(bad)>20 else ''} n={len(bad)}")
    # explicit im2col as control
    col = torch.stack([xp[:, 2 * t: 2 * t + k].reshape(B, -1) for t in range(T_out)], 1).reshape(B * T_out, k * Cin).contiguous()
    out2 = ops.linear(col.cuda(), wp.view(Cout, k * Cin))
    err2 = (out2.float().cpu() - ref).abs().max().item()
    print("   im2col control max err", err2)
for args in [(1, 101, 80, 256), (2, 101, 80, 256), (1, 101, 64, 64), (1, 300, 80, 256), (2, 100, 512, 512), (1, 40, 80, 64)]:
    run(*args)
