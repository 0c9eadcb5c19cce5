"""Device SpecAugment (csrc/specaugment.cu through the C ABI) against the CPU oracle on the same masks.
Constant fill: bit-exact.  Mean fill (mask_value=None): the fill value is an fp32 mean over up to 80k elements whose
summation order differs from numpy's pairwise sum: |value_gpu - value_numpy| <= 1e-5 (features are O(1)); every
unmasked element is bit-identical to the input."""
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

G = np.load(Path(__file__).parent / "golden" / "specaugment.npz")


def _batch(names):
    xs = [G[f"{n}_x"] for n in names]
    T = max(x.shape[0] for x in xs)
    F = xs[0].shape[1]
    feats = np.zeros((len(xs), T, F), dtype=np.float32)
    for b, x in enumerate(xs):
        feats[b, : x.shape[0]] = x
    return feats, [x.shape[0] for x in xs]


@pytest.mark.parametrize("mean_fill", [False, True])
def test_device_apply_matches_oracle(cuda_lib, mean_fill):
    from oracle.specaugment import specaugment
    from s2t_b200.specaugment import SpecAugmentConfig, apply_, draw_masks
    names = ["lb", "ld", "sm", "short"]
    feats, lens = _batch(names)
    cfg = SpecAugmentConfig(0, 2, 27, 2, 100, 1.0, None if mean_fill else -0.75)
    np.random.seed(77)
    fm, tm = draw_masks(cfg, lens, feats.shape[2])
    np.random.seed(77)
    want = np.zeros_like(feats)
    for b, n in enumerate(lens):      # the oracle draws in the same per-utterance order
        want[b, :n] = specaugment(feats[b, :n], freq_mask_n=2, freq_mask_f=27, time_mask_n=2, time_mask_t=100,
                                  time_mask_p=1.0, mask_value=cfg.mask_value)
    dev = torch.from_numpy(feats).cuda()
    apply_(cfg, dev, torch.tensor(lens, dtype=torch.int32).cuda(), fm.cuda(), tm.cuda())
    got = dev.cpu().numpy()
    changed = want != feats
    assert changed.any()
    assert np.array_equal(got[~changed], feats[~changed])            # nothing outside the masks is touched
    if mean_fill:
        assert np.abs(got - want).max() <= 1e-5
    else:
        assert np.array_equal(got, want)
    assert np.array_equal(got[1, lens[1]:], feats[1, lens[1]:])       # padding frames stay zero
    dev2 = torch.from_numpy(got).cuda()                               # idempotent with a constant fill
    if not mean_fill:
        apply_(cfg, dev2, torch.tensor(lens, dtype=torch.int32).cuda(), fm.cuda(), tm.cuda())
        assert np.array_equal(dev2.cpu().numpy(), got)


def test_device_apply_bf16_strided_full_size(cuda_lib):
    """C2-sized batch (8 x 998 frames) inside a wider bf16 buffer (ld > n_freqs): masked elements equal the bf16 fill,
    everything else — including the columns beyond n_freqs — is untouched."""
    from s2t_b200.specaugment import SpecAugmentConfig, apply_, draw_masks
    B, T, F, ld = 8, 998, 80, 96
    g = torch.Generator().manual_seed(9)
    buf = torch.randn((B, T, ld), generator=g).to(torch.bfloat16).cuda()
    ref = buf.clone()
    lens = [998, 998, 900, 850, 700, 612, 600, 3]
    cfg = SpecAugmentConfig(0, 2, 27, 2, 100, 1.0, 0.5)
    np.random.seed(5)
    fm, tm = draw_masks(cfg, lens, F)
    apply_(cfg, buf, torch.tensor(lens, dtype=torch.int32).cuda(), fm.cuda(), tm.cuda(), n_freqs=F)
    got, ref = buf.float().cpu(), ref.float().cpu()
    mask = torch.zeros((B, T, ld), dtype=torch.bool)
    for b, n in enumerate(lens):
        for f0, f in fm[b].tolist():
            mask[b, :n, f0:f0 + f] = True
        for t0, t in tm[b].tolist():
            mask[b, t0:min(t0 + t, n), :F] = True
    assert mask.any()
    assert torch.equal(got[mask], torch.full_like(got[mask], 0.5))
    assert torch.equal(got[~mask], ref[~mask])


def test_engine_applies_masks_after_cmvn(cuda_lib):
    """Engine with ``specaug`` set + masks in set_inputs == engine without it fed the oracle-augmented features
    (constant fill 0.0 survives the bf16 cast exactly): same loss and gradients, bit for bit; eval mode ignores masks."""
    from oracle.specaugment import specaugment
    from s2t_b200.engine import EncoderConfig, S2TEncoderEngine
    from s2t_b200.specaugment import SpecAugmentConfig, draw_masks
    kw = dict(layers=2, ffn_dim=512, subsampling_filter=256, vocab=90, dropout=0.0)
    a, b = S2TEncoderEngine(EncoderConfig(**kw)), S2TEncoderEngine(EncoderConfig(**kw))
    b.load_state_dict(a.state_dict())
    g = torch.Generator().manual_seed(4)
    B, T, S = 4, 300, 6
    x = torch.randn((B, T, 80), generator=g)
    lens = [300, 280, 200, 151]
    tg = torch.randint(4, 90, (B, S), generator=g).to(torch.int32)
    tl = torch.tensor([6, 5, 6, 3], dtype=torch.int32)
    cfg = SpecAugmentConfig(0, 2, 27, 2, 100, 0.2, 0.0)
    np.random.seed(21)
    fm, tm = draw_masks(cfg, lens, 80)
    np.random.seed(21)
    xa = x.clone()
    for i, n in enumerate(lens):
        xa[i, :n] = torch.from_numpy(specaugment(x[i, :n].numpy(), freq_mask_n=2, freq_mask_f=27, time_mask_n=2,
                                                 time_mask_t=100, time_mask_p=0.2, mask_value=0.0))
    assert not torch.equal(xa, x)
    sl = torch.tensor(lens, dtype=torch.int32)
    a.specaug = cfg
    Pa, Pb = a.plan(B, T, S), b.plan(B, T, S)
    a.set_inputs(Pa, feats=x, src_lengths=sl, targets=tg, target_lengths=tl, sa_freq=fm, sa_time=tm)
    b.set_inputs(Pb, feats=xa, src_lengths=sl, targets=tg, target_lengths=tl)
    a.train_step(Pa)
    b.train_step(Pb)
    torch.cuda.synchronize()
    assert float(Pa.loss) == float(Pb.loss)
    assert torch.equal(a.store.grad, b.store.grad)
    a.capture(Pa)                       # the masks are read from static buffers: a replay gives the same step
    a.run(Pa)
    torch.cuda.synchronize()
    assert float(Pa.loss) == float(Pb.loss)
    b.set_inputs(Pb, feats=x, src_lengths=sl, targets=tg, target_lengths=tl)
    b.eval(); a.eval()
    a.forward(Pa); b.forward(Pb)
    torch.cuda.synchronize()
    assert torch.equal(Pa.logits, Pb.logits)            # inference: no augmentation
