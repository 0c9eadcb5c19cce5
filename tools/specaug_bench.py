"""SpecAugment at the bench workload's size (8 x 998 frames x 80 bins): device apply (C ABI, CUDA events around a
CUDA graph of 50 applications) against the CPU oracle (numpy, one utterance at a time as the reference's dataset does).
Algorithmic bytes: mean fill reads the valid block once (B*T*F*es) and both variants write the masked elements.
usage: python tools/specaug_bench.py"""
import sys
import time
import numpy as np
import torch
sys.path.insert(0, ".")
from oracle.specaugment import specaugment
from s2t_b200.specaugment import SpecAugmentConfig, apply_, draw_masks

B, T, F = 8, 998, 80
lens = [T] * B
g = torch.Generator().manual_seed(0)
x = torch.randn((B, T, F), generator=g)
for mean_fill in (False, True):
    cfg = SpecAugmentConfig(0, 2, 27, 2, 100, 1.0, None if mean_fill else 0.0)     # policy "ld"
    np.random.seed(1)
    fm, tm = draw_masks(cfg, lens, F)
    masked = 0
    for b in range(B):
        m = np.zeros((T, F), dtype=bool)
        for f0, f in fm[b].tolist():
            m[:, f0:f0 + f] = True
        for t0, t in tm[b].tolist():
            m[t0:t0 + t] = True
        masked += int(m.sum())
    for dt in (torch.float32, torch.bfloat16):
        dev = x.to(dt).cuda()
        nf, fmd, tmd = torch.tensor(lens, dtype=torch.int32).cuda(), fm.cuda(), tm.cuda()
        means = torch.empty(B * 32, device="cuda")
        apply_(cfg, dev, nf, fmd, tmd, means=means)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for _ in range(50):
                apply_(cfg, dev, nf, fmd, tmd, means=means)
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 500
        es = dev.element_size()
        byts = masked * es + (B * T * F * es if mean_fill else 0)
        print(f"device {'mean' if mean_fill else 'const'} fill {str(dt)[6:]:9s}: {us:6.2f} us per batch, "
              f"{byts / 1e6:.2f} MB algorithmic -> {byts / us / 1e3:.1f} GB/s")
    xs = [x[b].numpy() for b in range(B)]
    np.random.seed(1)
    t0 = time.perf_counter()
    for _ in range(20):
        for b in range(B):
            specaugment(xs[b], freq_mask_n=2, freq_mask_f=27, time_mask_n=2, time_mask_t=100, time_mask_p=1.0,
                        mask_value=cfg.mask_value)
    cpu_us = (time.perf_counter() - t0) / 20 * 1e6
    print(f"cpu oracle {'mean' if mean_fill else 'const'} fill (numpy, 1 thread): {cpu_us:8.1f} us per batch of {B}")
