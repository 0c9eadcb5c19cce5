"""The double-buffered input feed (s2t_b200/feeder.py) must hand every batch to the captured step unchanged: the loss
sequence of a pipelined loop equals the one of the plain copy -> step -> read loop, bit for bit (dropout off)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_pipelined_feed_matches_sequential_loop(cuda_lib):
    from s2t_b200.engine import EncoderConfig, S2TEncoderEngine
    from s2t_b200.feeder import InputFeeder
    eng = S2TEncoderEngine(EncoderConfig(layers=2, ffn_dim=512, subsampling_filter=256, vocab=120, dropout=0.0))
    g = torch.Generator().manual_seed(5)
    B, n, S = 4, 24000, 7
    batches = []
    for i in range(5):
        wave = (torch.randn((B, n), generator=g) * 3000.0).pin_memory()
        num = torch.tensor([24000, 23000 - 160 * i, 20000, 16000 + 80 * i], dtype=torch.int32).pin_memory()
        tg = torch.randint(4, 120, (B, S), generator=g).to(torch.int32).pin_memory()
        tl = torch.tensor([7, 5, 6, 3], dtype=torch.int32).pin_memory()
        batches.append(dict(wave=wave, num_samples=num, targets=tg, target_lengths=tl))
    P = eng.plan(B, 1 + (n - 400) // 160, S, wave_len=n)
    eng.set_inputs(P, **batches[0])
    eng.train_step(P)
    torch.cuda.synchronize()
    eng.capture(P)
    want = []
    for b in batches:
        eng.set_inputs(P, **b)
        eng.run(P)
        want.append(float(P.loss.item()))
    feeder = InputFeeder(eng, P)
    slots = [torch.zeros(1).pin_memory() for _ in range(2)]
    evs = [torch.cuda.Event(), torch.cuda.Event()]
    got = []
    feeder.stage(**batches[0])
    for i in range(len(batches)):
        feeder.commit()
        if i + 1 < len(batches):
            feeder.stage(**batches[i + 1])
        eng.run(P)
        slots[i & 1].copy_(P.loss, non_blocking=True)
        evs[i & 1].record()
        if i > 0:
            evs[(i - 1) & 1].synchronize()
            got.append(float(slots[(i - 1) & 1][0]))
    evs[(len(batches) - 1) & 1].synchronize()
    got.append(float(slots[(len(batches) - 1) & 1][0]))
    assert got == want
    assert len(set(want)) == len(want)          # the batches really differ
    with pytest.raises(RuntimeError):
        feeder.commit()                          # nothing staged
