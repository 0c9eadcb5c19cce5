"""Device time of s2t_ctc_loss at the C1 shape (B=8, T'=250, V=10000, S=60), per kernel (torch profiler)."""
import sys
import torch
sys.path.insert(0, ".")
from s2t_b200 import ops
B, T, V, S = 8, 250, 10000, 60
g = torch.Generator().manual_seed(0)
logits = torch.randn((B, T, V), generator=g).to(torch.bfloat16).cuda()
tg = torch.randint(4, V, (B, S), generator=g).to(torch.int32).cuda()
il = torch.full((B,), T, dtype=torch.int32).cuda()
tl = torch.full((B,), S, dtype=torch.int32).cuda()
for _ in range(3):
    nll, dl = ops.ctc_loss(logits, tg, il, tl, V)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        nll, dl = ops.ctc_loss(logits, tg, il, tl, V)
    torch.cuda.synchronize()
import collections
agg = collections.defaultdict(list)
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        agg[e.name.split("(")[0][:60]].append(e.time_range.end - e.time_range.start)
for k, v in agg.items():
    print(f"{k:62s} n={len(v)} min {min(v):8.1f} us")
print("nll", nll[:3].tolist())
