"""The reference's plugin surface on the B200 engine: FairseqEncoder dict contract, the ctc criterion
triple with autograd hand-off of the flat gradient, greedy CTCDecoder.  Checked against the reference golden fixtures
and the CPU oracle."""
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"

REF_KEYS = {"encoder_out", "ctc_logit", "inter_ctc_logits", "xctc_logit", "inter_xctc_logits", "encoder_padding_mask",
            "mixup", "encoder_embedding", "encoder_states", "src_tokens", "src_lengths"}      # s2t_transformer.py:2142-2154


def _args(case):
    from oracle.make_golden import ENC_CASES, reference_args
    spec = dict(ENC_CASES[case])
    args = reference_args(**spec)
    for k, v in dict(encoder_embed_dim=256, encoder_ffn_embed_dim=2048, encoder_layers=12, encoder_attention_heads=4,
                     subsampling_layers=2, subsampling_filter=1024, subsampling_kernel=5, subsampling_stride=2).items():
        if not hasattr(args, k):
            setattr(args, k, v)
    return args, spec


def _model(case):
    from oracle.make_golden import seeded_state_dict
    from s2t_b200 import plugin
    fx = np.load(GOLDEN / f"enc_{case}.npz")
    args, spec = _args(case)
    if spec.get("family") == "pds":
        args.encoder_type = "pds"
    args.b200_cuda_graph = False
    model = plugin.S2TCTCModelB200.build_model(args, None)
    shapes = {n: eval(s) for n, s in zip(fx["param_names"], fx["param_shapes"])}
    model.encoder.load_state_dict(seeded_state_dict(shapes, int(fx["seed"])))
    return model, fx


def _sample(fx):
    s = {"net_input": {"src_tokens": torch.from_numpy(fx["feats"]).cuda(), "src_lengths": torch.from_numpy(fx["lens"]).cuda(),
                       "prev_output_tokens": None},
         "transcript": {"tokens": torch.from_numpy(fx["tokens"])}, "target": torch.from_numpy(fx["tokens"])}
    if "tgt_tokens" in fx.files:
        s["target"] = torch.from_numpy(fx["tgt_tokens"])
    return s


@pytest.mark.parametrize("case", ["c1_small", "c3_small", "c5_small"])
def test_encoder_forward_contract(cuda_lib, case):
    model, fx = _model(case)
    model.eval()
    s = _sample(fx)
    out = model(**s["net_input"])
    assert set(out) == REF_KEYS
    ref_enc, ref_logit = torch.from_numpy(fx["encoder_out"]), torch.from_numpy(fx["ctc_logit"])
    enc, logit, pad = out["encoder_out"][0], out["ctc_logit"][0], out["encoder_padding_mask"][0]
    assert tuple(enc.shape) == tuple(ref_enc.shape) and tuple(logit.shape) == tuple(ref_logit.shape)   # T' x B x C / V
    assert pad.dtype == torch.bool and torch.equal(pad.cpu(), torch.from_numpy(fx["padding_mask"]))
    valid = ~torch.from_numpy(fx["padding_mask"]).T
    rel = lambda a, b: float((a.float().cpu()[valid] - b[valid]).norm() / b[valid].norm())
    tol = 3e-2 if case == "c5_small" else 2e-2   # eval-mode BatchNorm (running stats 0/1) differs from the train-mode fixture
    if case != "c5_small":
        assert rel(enc, ref_enc) <= tol and rel(logit, ref_logit) <= 3e-2
    else:
        assert len(out["inter_ctc_logits"]) == 2 and isinstance(out["inter_ctc_logits"][0], list)
        assert len(out["inter_xctc_logits"]) == 2 and torch.is_tensor(out["inter_xctc_logits"][0])
        assert tuple(out["xctc_logit"][0].shape) == tuple(ref_logit.shape)
    lp = model.get_normalized_probs(out, log_probs=True)
    assert lp.dtype == torch.float32 and torch.allclose(lp.exp().sum(-1), torch.ones_like(lp[..., 0]), atol=1e-3)
    # reorder_encoder_out (s2t_transformer.py:2156-2208)
    order = torch.tensor([2, 0], device="cuda")
    ro = model.encoder.reorder_encoder_out(out, order)
    assert torch.equal(ro["encoder_out"][0], enc.index_select(1, order))
    assert torch.equal(ro["encoder_padding_mask"][0], pad.index_select(0, order))


@pytest.mark.parametrize("case", ["c1_small", "c5_small"])
def test_criterion_forward_backward(cuda_lib, case):
    from s2t_b200 import plugin
    model, fx = _model(case)
    crit = plugin.CtcCriterionB200(None, None)
    model.train()
    loss, sample_size, log = crit(model, _sample(fx))
    # the fixture's c1 loss is the bare CTC loss; the criterion applies ctc_weight (criterions/ctc.py:1004), which the
    # c5 fixture's total already contains
    ref = float(fx["loss"]) * (_args(case)[0].ctc_weight if case == "c1_small" else 1.0)
    assert loss.dim() == 0 and loss.requires_grad
    assert abs(loss.item() - ref) <= 1e-2 * abs(ref), (loss.item(), ref)
    assert sample_size == log["sample_size"] == log["ntokens"] and log["nsentences"] == fx["feats"].shape[0]
    assert {"loss", "ntokens", "nsentences", "sample_size", "ctc_loss"} <= set(log)
    (loss * 4.0).backward()                               # fairseq's loss scaling reaches the flat gradient
    enc = model.encoder
    g = enc.flat.grad
    assert g is not None and g.shape == enc.flat.shape
    assert torch.allclose(g.float(), enc.engine.store.grad.float() * 4.0, rtol=1e-2, atol=1e-6)
    name = "ctc.ctc_projection.weight"
    mine = enc.engine.grads()[name].cpu()
    want = torch.from_numpy(fx["grad::" + name]) * (_args(case)[0].ctc_weight if case == "c1_small" else 1.0)
    assert float((mine - want).norm() / want.norm()) <= 5e-2
    # eval: no graph, no gradient, validation error counts present
    model.eval()
    with torch.no_grad():
        l2, _, log2 = crit(model, _sample(fx))
    assert not l2.requires_grad and np.isfinite(l2.item())
    assert log2["c_total"] == int(((fx["tokens"] != 1) & (fx["tokens"] != 2)).sum()) and log2["c_errors"] >= 0


def test_criterion_cuda_graph_path(cuda_lib):
    """With b200_cuda_graph the criterion captures the fused step once per batch shape and replays it."""
    from s2t_b200 import plugin
    model, fx = _model("c1_small")
    model.encoder.use_graph = True
    crit = plugin.CtcCriterionB200(None, None)
    model.train()
    l1, _, _ = crit(model, _sample(fx))
    l1.backward()
    g1 = model.encoder.flat.grad.clone()
    model.encoder.flat.grad = None
    l2, _, _ = crit(model, _sample(fx))
    l2.backward()
    assert model.encoder._last_plan.graph is not None
    assert torch.equal(l1, l2) and torch.equal(g1, model.encoder.flat.grad)      # dropout 0: replays are bit-identical


def test_ctc_decoder_generate_bit_exact(cuda_lib):
    from oracle import ctc as OC
    from s2t_b200 import decode
    model, fx = _model("c5_small")
    gen = decode.CTCDecoderB200([model], None, None, blank_idx=0)
    s = _sample(fx)
    hyps = gen.generate([model], s)
    P, eng = model.encoder._last_plan, model.encoder.engine
    hs = eng.heads[-1]                                                     # target-side head decodes (s2t_ctc.py:265-271)
    lg = P.heads[-1].logits.float().view(P.B, P.T2, hs.ldv)[..., : hs.V].cpu().numpy()
    want = OC.greedy_decode(lg, P.out_len.cpu().numpy(), blank=0)
    scores = OC.greedy_scores(lg, blank=0)
    assert len(hyps) == P.B
    for b, h in enumerate(hyps):
        assert set(h[0]) == {"tokens", "score", "attention", "alignment", "positional_scores"}
        assert h[0]["tokens"].dtype == torch.int64 and h[0]["tokens"].tolist() == [int(v) for v in want[b]]
        assert abs(float(h[0]["score"]) - scores[b]) <= 1e-3 * abs(scores[b]) + 1e-3
        assert len(h[0]["positional_scores"]) == len(want[b])


@pytest.mark.parametrize("case", ["c1_small", "c5_small"])
def test_encoder_trains_through_external_loss(cuda_lib, case):
    """decoder(encoder(.)) training (s2t_transformer.py:868-880): in training mode ``forward`` returns encoder_out / logits
    with an autograd graph; a loss computed OUTSIDE the engine (here a fixed linear 'decoder' read-out of encoder_out
    plus a torch CTC loss on ctc_logit) must give the encoder the gradients the oracle gets for the same loss."""
    from oracle import encoder as OE
    from oracle.make_golden import seeded_state_dict
    model, fx = _model(case)
    args, spec = _args(case)
    shapes = {n: eval(s) for n, s in zip(fx["param_names"], fx["param_shapes"])}
    sd = seeded_state_dict(shapes, int(fx["seed"]))
    orc = OE.S2TTransformerEncoder(OE.config_from_args(args))
    orc.load_state_dict(sd)
    orc.train()
    feats, lens, tokens = (torch.from_numpy(fx[k]) for k in ("feats", "lens", "tokens"))
    g = torch.Generator().manual_seed(3)
    C = args.encoder_embed_dim
    readout = torch.randn((C,), generator=g) / C ** 0.5

    def toy_loss(out, dev):
        enc, pad = out["encoder_out"][0], out["encoder_padding_mask"][0]
        keep = (~pad).T.unsqueeze(-1).to(enc.dtype)                      # T' x B x 1
        dec = ((enc.float() * keep) @ readout.to(dev)).pow(2).sum()     # stands for the decoder's cross-entropy
        ctc = OE.ctc_criterion_loss(out["ctc_logit"][0].float(), pad, tokens.to(dev))
        extra = 0.0
        if out["inter_ctc_logits"]:                                     # an inter-CTC term through the autograd path too
            extra = 0.1 * OE.ctc_criterion_loss(out["inter_ctc_logits"][0][0].float(), pad, tokens.to(dev))
        return dec + 0.3 * ctc + extra

    ref_loss = toy_loss(orc(feats, lens), "cpu")
    ref_loss.backward()
    model.train()
    out = model.encoder(feats.cuda(), lens.cuda())
    assert out["encoder_out"][0].requires_grad and out["ctc_logit"][0].requires_grad
    assert tuple(out["encoder_out"][0].shape) == tuple(fx["encoder_out"].shape)
    loss = toy_loss(out, "cuda")
    assert abs(loss.item() - ref_loss.item()) <= 2e-2 * abs(ref_loss.item()), (loss.item(), ref_loss.item())
    loss.backward()
    enc = model.encoder
    assert enc.flat.grad is not None
    grads = enc.engine.grads()
    worst = max((float((grads[k].cpu() - p.grad).norm() / (p.grad.norm() + 1e-12)), k)
                for k, p in orc.named_parameters()
                if p.grad is not None and not k.endswith(("k_proj.bias", "linear_k.bias")) and float(p.grad.norm()) > 0)
    # 5e-2; the 4-layer Conformer's depthwise-conv / position weights carry more bf16 noise (the reference's own bf16 mode
    # deviates by 0.08-0.12 on them for the CTC loss, see test_engine_gpu golden_c5_small): 0.2 there
    assert worst[0] <= (5e-2 if case == "c1_small" else 0.2), worst
    # eval / no_grad: plain tensors, no graph
    model.eval()
    with torch.no_grad():
        out2 = model.encoder(feats.cuda(), lens.cuda())
    assert not out2["encoder_out"][0].requires_grad


def test_plan_cache_is_bounded_under_max_tokens_batching(cuda_lib):
    """50 batches of random (B, T) as max-tokens batching produces them: the frame buckets keep the number of plans
    small and the LRU byte cap holds (the cache used to grow by one plan + one graph per distinct (B, T))."""
    from s2t_b200.batching import BucketSet
    from s2t_b200.engine import EncoderConfig, S2TEncoderEngine
    eng = S2TEncoderEngine(EncoderConfig(layers=2, ffn_dim=512, subsampling_filter=256, vocab=200, dropout=0.0))
    cap = 1 << 30
    eng.plan_cache_bytes = cap
    bs = BucketSet()
    rng = np.random.RandomState(0)
    keys = set()
    for i in range(50):
        T = int(rng.randint(120, 1500))
        B = max(8, (8000 // T) // 8 * 8)
        Bq, Tq = bs.key(B, T)
        keys.add((Bq, Tq))
        P = eng.plan(Bq, Tq, 32)
        feats = torch.randn((B, T, 80))
        lens = torch.full((B,), T, dtype=torch.int32)
        tg = torch.randint(4, 200, (B, 8)).to(torch.int32)
        tl = torch.full((B,), 8, dtype=torch.int32)
        eng.set_inputs(P, feats=feats, src_lengths=lens, targets=tg, target_lengths=tl)
        eng.train_step(P)
        st = eng.plan_cache_stats()
        assert st["bytes"] <= cap, st
    torch.cuda.synchronize()
    st = eng.plan_cache_stats()
    assert st["builds"] <= len(keys) + st["evictions"] and st["evictions"] > 0, st      # 1 GiB cap forces evictions here
    assert np.isfinite(P.loss.item())
    # an evicted bucket is rebuilt on demand and gives the same result as before
    k0 = sorted(keys)[0]
    assert eng.plan(k0[0], k0[1], 32) is eng.plan(k0[0], k0[1], 32)


def test_multichannel_waveform_is_merged_to_mono(cuda_lib):
    """audio_utils.py:67-72: the (channels, N) array get_waveform returns is merged to mono before the fbank."""
    from s2t_b200 import frontend
    g = np.random.RandomState(0)
    stereo = (g.randn(2, 16000) * 3000).astype(np.float32)
    a = frontend._get_torchaudio_fbank(stereo, 16000, 80)
    b = frontend._get_torchaudio_fbank(stereo.mean(axis=0), 16000, 80)
    c = frontend._get_torchaudio_fbank(stereo[:1], 16000, 80)
    assert a.shape == b.shape == (98, 80) and np.array_equal(a, b)
    assert c.shape == (98, 80) and np.array_equal(c, frontend._get_torchaudio_fbank(stereo[0], 16000, 80))
    with pytest.raises(ValueError):
        frontend._get_torchaudio_fbank(np.zeros((2, 2, 100), dtype=np.float32), 16000, 80)


def test_criterion_curriculum_end_to_end(cuda_lib):
    """CtcCriterionB200 with the curriculum options of reproduction_bil_ctc_synchronous.yaml:62-65 switched on
    (xctc/ctc-pae-ground-truth-ratio, only-mistake, oracle-smooth, ctc_masked_loss): the alignment the criterion computes
    on the device equals the Viterbi restatement on the same logits, the drawn masks respect the mistake flags, the loss
    and the flat gradient are finite, and with ratio 0 the criterion gives the curriculum-free loss."""
    from pathlib import Path
    _GOLDEN = Path(__file__).resolve().parent / "golden"
    from argparse import Namespace
    from oracle import viterbi as OV
    from oracle.make_golden import CURR_CASE, reference_args, seeded_state_dict
    from s2t_b200 import plugin
    fx = np.load(_GOLDEN / "enc_c5_curr_small.npz")
    args = reference_args(**dict(CURR_CASE))
    for k, v in dict(encoder_embed_dim=256, encoder_ffn_embed_dim=2048, encoder_layers=12, encoder_attention_heads=4,
                     subsampling_layers=2, subsampling_filter=1024, subsampling_kernel=5, subsampling_stride=2).items():
        if not hasattr(args, k):
            setattr(args, k, v)
    args.b200_cuda_graph = False
    model = plugin.S2TCTCModelB200.build_model(args, None)
    shapes = {n: eval(s) for n, s in zip(fx["param_names"], fx["param_shapes"])}
    model.encoder.load_state_dict(seeded_state_dict(shapes, int(fx["seed"])))
    enc = model.encoder
    assert enc.pae_ground_truth_ratio > 0 and enc.pae_gt_only_mistake
    crit = plugin.CtcCriterionB200(Namespace(ctc_masked_loss=True), None)
    sample = {"net_input": {"src_tokens": torch.from_numpy(fx["feats"]).cuda(), "src_lengths": torch.from_numpy(fx["lens"]).cuda(),
                            "prev_output_tokens": None},
              "transcript": {"tokens": torch.from_numpy(fx["tokens"])}, "target": torch.from_numpy(fx["tgt_tokens"])}
    model.train()
    torch.manual_seed(0)
    loss, _, log = crit(model, sample)
    loss.backward()
    eng, P = enc.engine, enc._last_plan
    assert np.isfinite(loss.item()) and torch.isfinite(enc.flat.grad.float()).all()
    assert set(P.gt) == {0, 1}
    drawn = {k: tuple(t.view(P.B, P.T2).cpu().clone() for t in P.gt[k]) for k in (0, 1)}
    # the alignment was taken on the FIRST pass (no mixing yet): reproduce that pass (dropout 0, batch statistics)
    for k in (0, 1):
        eng.set_ground_truth(P, k, None, None, None)
    eng.forward(P)
    torch.cuda.synchronize()
    for k, (tokens_key, h) in enumerate((("tokens", "ctc"), ("tgt_tokens", "xctc"))):
        tok, msk, fe = drawn[k]
        assert int(msk.sum()) > 0 and bool(((fe >= 0) == (msk != 0)).all())
        # the forced states lie on the best alignment of the first pass's top logits
        hs = eng.heads[k]
        lp = torch.log_softmax(P.heads[k].logits.float().view(P.B, P.T2, hs.ldv)[..., : hs.V], -1).cpu().numpy()
        tg, tl = eng.strip_targets(torch.from_numpy(fx[tokens_key]))
        paths, _ = OV.best_alignment(lp, tg.numpy().astype(np.int64), P.out_len.cpu().tolist(), tl.tolist())
        for b in range(P.B):
            n = int(P.out_len[b])
            forced = fe[b, :n] >= 0
            assert np.array_equal(fe[b, :n][forced].numpy(), np.asarray(paths[b])[forced.numpy()]), (h, b)
    # ratio 0: the plain criterion
    enc.pae_ground_truth_ratio = 0
    plain, _, _ = crit(model, sample)
    assert abs(plain.item() - float(np.load(_GOLDEN / "enc_c5_small.npz")["loss"])) <= 1e-2 * abs(plain.item())
