"""Concurrent micro-batches (s2t_b200/experimental/microbatch.py) must give the loss and gradients of the one-batch engine:
gradient accumulation over utterance groups is exact up to bf16 summation order (trainer.py:716-734 semantics)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-12))


@pytest.mark.parametrize("n_micro", [2, 4])
@pytest.mark.parametrize("graph", [False, True])
def test_microbatch_matches_single_batch(cuda_lib, n_micro, graph):
    from s2t_b200.engine import EncoderConfig, S2TEncoderEngine
    from s2t_b200.experimental.microbatch import MicroBatchEngine
    kw = dict(layers=3, ffn_dim=512, subsampling_filter=256, vocab=120, dropout=0.0)
    ref = S2TEncoderEngine(EncoderConfig(**kw))
    mb = MicroBatchEngine(EncoderConfig(**kw), n_micro=n_micro)
    mb.load_state_dict(ref.state_dict())
    g = torch.Generator().manual_seed(3)
    B, n, S = 8, 24000, 9
    wave = torch.randn((B, n), generator=g) * 3000.0
    num = torch.tensor([24000, 23000, 22000, 20000, 19000, 18000, 16000, 12000], dtype=torch.int32)
    tg = torch.randint(4, 120, (B, S), generator=g).to(torch.int32)
    tl = torch.tensor([9, 8, 7, 9, 3, 5, 6, 4], dtype=torch.int32)
    T = 1 + (n - 400) // 160
    P = ref.plan(B, T, S, wave_len=n)
    ref.set_inputs(P, wave=wave, num_samples=num, targets=tg, target_lengths=tl)
    ref.train_step(P)
    MP = mb.plan(B, T, S, wave_len=n)
    mb.set_inputs(MP, wave=wave, num_samples=num, targets=tg, target_lengths=tl)
    if graph:
        mb.capture(MP)
        mb.run(MP)
        mb.run(MP)          # replays must not accumulate into the gradient buffer
    else:
        mb.train_step(MP)
    torch.cuda.synchronize()
    assert abs(float(MP.loss) - float(P.loss)) <= 2e-3 * abs(float(P.loss)), (float(MP.loss), float(P.loss))
    gr, gm = ref.grads(), mb.grads()
    worst = max((rel_l2(gm[k], gr[k]), k) for k in gr if float(gr[k].float().norm()) > 0)
    assert worst[0] <= 2e-2, worst
    assert torch.equal(MP.out_len.cpu(), P.out_len.cpu())


def test_microbatch_rejects_batchnorm_models(cuda_lib):
    from s2t_b200.engine import EncoderConfig
    from s2t_b200.experimental.microbatch import MicroBatchEngine
    with pytest.raises(ValueError):
        MicroBatchEngine(EncoderConfig(layers=1, vocab=50, use_cnn_module=True, macaron_style=True, attn_type="rel_pos"),
                         n_micro=2)
