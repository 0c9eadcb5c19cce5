"""CTC best alignment at the bench workload's size (8 utterances x 250 encoder frames, V = 10000, 60 target tokens):
device kernel incl. backtrack (CUDA events around a CUDA graph of 20 launches) against the CPU restatement.
Algorithmic bytes = gathered emissions B*T*(2S+1)*4 read + B*T*4 states written.  usage: python tools/viterbi_bench.py"""
import sys
import time
import torch
sys.path.insert(0, ".")
from oracle.viterbi import best_alignment
from s2t_b200 import ops

B, T, V, S = 8, 250, 10000, 60
g = torch.Generator().manual_seed(0)
lp = torch.log_softmax(torch.randn((B, T, V), generator=g), dim=-1)
tg = torch.randint(1, V, (B, S), generator=g).to(torch.int32)
il = torch.full((B,), T, dtype=torch.int32)
tl = torch.full((B,), S, dtype=torch.int32)
d = [t.cuda() for t in (lp, tg, il, tl)]
ops.ctc_viterbi(*d)
torch.cuda.synchronize()
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr):
    for _ in range(20):
        ops.ctc_viterbi(*d)
gr.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    gr.replay()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / 200
byts = B * T * (2 * S + 1) * 4 + B * T * 4
print(f"device: {us:.1f} us per batch ({T} dependent frames), {byts / 1e6:.2f} MB algorithmic -> {byts / us / 1e3:.2f} GB/s "
      f"(latency-bound: one barrier + one shared-memory round trip per frame)")
t0 = time.perf_counter()
best_alignment(lp.numpy(), tg.numpy(), il.tolist(), tl.tolist())
print(f"cpu restatement (pure-Python loops over T x (2S+1), 1 thread): {(time.perf_counter() - t0) * 1e3:.0f} ms per batch")
