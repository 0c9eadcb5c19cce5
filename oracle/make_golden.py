"""Generate tests/golden/*.npz from the REFERENCE ITSELF (run only in the build container).

    python -m oracle.make_golden            # needs /root/reference

For each case the unmodified reference modules (loaded through oracle/ref_shim.py) are run on seeded
synthetic inputs with seeded weights; the restatement in oracle/ is checked against them on the spot
(the script fails if they disagree) and the reference's outputs are stored.  Weights are NOT stored:
both sides regenerate them with ``seeded_state_dict`` (torch CPU generator, platform independent) and
the fixture keeps a checksum so a mismatch is detected.

Fixtures:
  fbank_cmvn.npz  : waveform -> torchaudio fbank -> reference UtteranceCMVN / GlobalCMVN
  enc_<case>.npz  : features -> reference S2TTransformerEncoder -> ctc criterion loss, selected grads
"""
from __future__ import annotations

import hashlib
import os
import sys
from argparse import Namespace
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
GOLDEN = ROOT / "tests" / "golden"

# name -> overrides of the reference args (on top of s2t_transformer_s + base.yaml-style flags)
ENC_CASES = {
    # C1 in miniature: s2t_transformer_s structure (conv1d x2 + GLU, embed norm, abs pos, pre-norm, relu FFN)
    "c1_small": dict(encoder_layers=2, encoder_embed_dim=256, encoder_ffn_embed_dim=512,
                     encoder_attention_heads=4, subsampling_filter=256, vocab=200, B=3, T=83, S=7),
    # C4 in miniature: conformer.yaml (macaron, conv module k15 BatchNorm, rel_pos, swish FFN, layer mask)
    "c4_small": dict(encoder_layers=2, encoder_embed_dim=256, encoder_ffn_embed_dim=512,
                     encoder_attention_heads=4, subsampling_filter=256, vocab=200, B=3, T=83, S=7,
                     macaron_style=True, use_cnn_module=True, cnn_module_kernel=15,
                     encoder_attention_type="rel_pos", encoder_activation_fn="swish",
                     layer_padding_mask=True),
    # C5 in miniature: BiL-CTC synchronous (reproduction_bil_ctc_synchronous.yaml): conformer + CTC (source vocab) +
    # XCTC (target vocab) + shared inter-CTC / inter-XCTC at layers 2,3 + inter_league adapters on un-normalised input
    "c5_small": dict(encoder_layers=4, encoder_embed_dim=256, encoder_ffn_embed_dim=512, encoder_attention_heads=4,
                     subsampling_filter=256, vocab=200, B=3, T=83, S=7, macaron_style=True, use_cnn_module=True,
                     cnn_module_kernel=15, encoder_attention_type="rel_pos", encoder_activation_fn="swish",
                     ctc_weight=0.2, xctc_weight=0.1, inter_ctc_weight=0.1, inter_ctc_layers="2,3",
                     share_inter_ctc=True, inter_xctc_weight=0.05, inter_xctc_layers="2,3",
                     ctc_pae="inter_league", xctc_pae="inter_league", pae_unnorm_input=True),
    # Conv2dSubsampling stem (egs/librispeech/asr/conf/ConformerCTCSmall.yaml:16-22: conv2d, kernel 3, stride 2,
    # batch2d norm, swish) in front of the plain C1 layers
    "conv2d_small": dict(encoder_layers=2, encoder_embed_dim=256, encoder_ffn_embed_dim=512, encoder_attention_heads=4,
                         subsampling_type="conv2d", subsampling_layers=2, subsampling_filter=64, subsampling_kernel=3,
                         subsampling_stride=2, subsampling_norm="batch2d", subsampling_activation="swish",
                         vocab=200, B=3, T=83, S=5),
    # C3 in miniature: pdss2t_transformer_s_8 structure (4 stages, ratios 2_2_1_2, conv down-sampling + LayerNorm,
    # abs positions per stage, relu FFN), one layer per stage
    "c3_small": dict(family="pds", pds_stages=4, pds_layers="1_1_1_1", pds_ratios="2_2_1_2", pds_fusion=False,
                     pds_embed_dims="256_256_256_256", pds_kernel_sizes="5_5_5_5", pds_ffn_ratios="2_2_2_2",
                     pds_attn_heads="4_4_4_4", pds_position_embed="1_1_1_1", pds_embed_norm=True,
                     pds_ds_method="conv", pds_attn_ds_ratios="1_1_1_1", vocab=200, B=3, T=83, S=5),
}


def reference_args(**over) -> Namespace:
    """egs/mustc/asr/conf/base.yaml + ctc.yaml flags that reach the encoder."""
    a = Namespace(input_feat_per_channel=80, input_channels=1, max_source_positions=6000,
                  encoder_embed_norm=True, encoder_no_scale_embedding=True, ctc_weight=0.3,
                  dropout=0.0, fp16=False, activation_fn="relu")
    vocab = over.pop("vocab", 10000)
    for k in ("B", "T", "S", "family"):
        over.pop(k, None)
    for k, v in over.items():
        setattr(a, k, v)
    a._src_vocab = a._tgt_vocab = vocab
    return a


def seeded_state_dict(shapes: dict, seed: int) -> dict:
    """Deterministic weights for a {name: shape} map: N(0, 0.05) matrices, LayerNorm-like vectors near 1/0,
    rounded to bf16-representable values so the bf16 product sees exactly the oracle's weights."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for name in sorted(shapes):
        shp = tuple(shapes[name])
        if name.endswith("num_batches_tracked"):
            out[name] = torch.zeros(shp, dtype=torch.long)
            continue
        t = torch.randn(shp, generator=g)
        if name.endswith("running_var"):
            t = 1.0 + 0.1 * t.abs()
        elif name.endswith("running_mean"):
            t = 0.1 * t
        elif len(shp) == 1 and name.endswith("weight"):
            t = 1.0 + 0.1 * t
        elif len(shp) == 1:
            t = 0.1 * t
        else:
            fan_in = int(np.prod(shp[1:])) if len(shp) > 1 else shp[0]
            t = t * (1.0 / np.sqrt(fan_in))
        out[name] = t.to(torch.bfloat16).to(torch.float32)
    return out


def state_checksum(sd: dict) -> str:
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def synth_batch(B: int, T: int, S: int, vocab: int, seed: int):
    """asr_test_base.get_dummy_input style (tests/speech_recognition/asr_test_base.py:70-101): randn
    features, descending ragged lengths, targets with pad=1 / eos=2."""
    g = torch.Generator().manual_seed(seed)
    feats = torch.randn((B, T, 80), generator=g)
    lens = torch.tensor(sorted([T] + [int(v) for v in torch.randint(T // 2, T, (B - 1,), generator=g)], reverse=True))
    for b in range(B):
        feats[b, lens[b]:] = 0
    tokens = torch.full((B, S + 1), 1, dtype=torch.long)
    for b in range(B):
        n = S - b
        tokens[b, :n] = torch.randint(4, vocab, (n,), generator=g)
        tokens[b, n] = 2
    return feats, lens, tokens


GRAD_KEYS = ["pae.embed_adapter.weight", "xctc_pae.embed_adapter.weight", "xctc.ctc_projection.weight", "ctc_norm2.weight",
             "xctc_norm3.bias", "layers.2.conv_module.depthwise_conv.weight", "layers.1.self_attn.linear_pos.weight",
             "downsampling1.conv.0.weight", "downsampling3.conv.0.bias", "downsampling2.norm.weight",
             "stage1.0.ffn.w_1.weight", "stage4.0.self_attn.q_proj.weight", "stage2.0.self_attn_layer_norm.bias",
             "subsample.layers.0.0.weight", "subsample.layers.1.0.bias", "embed_ln.weight", "subsample.layers.1.0.weight",
             "subsample.layers.0.1.1.weight", "subsample.layers.1.1.1.bias", "subsample.linear.bias",
             "layers.0.ffn.w_1.weight", "layers.1.ffn.w_2.bias", "layers.0.self_attn_layer_norm.bias",
             "layer_norm.weight", "ctc.ctc_projection.weight"]


def build_reference_encoder(cfg, pds: bool = False):
    """The reference's own encoder class (unmodified files through oracle/ref_shim.py) for an ``OracleConfig`` — the
    inverse of ``oracle.encoder.config_from_args``.  Used by ``bench.py --impl reference`` when the reference tree is
    available (``S2T_REFERENCE``: /root/reference here, baseline/_ref on the GPU box)."""
    from oracle import ref_shim
    ref = ref_shim.load()
    us = lambda t: "_".join(str(int(v)) for v in t)
    cs = lambda t: ",".join(str(int(v)) for v in t) if t else None
    over = dict(
        encoder_layers=cfg.layers, encoder_embed_dim=cfg.embed_dim, encoder_ffn_embed_dim=cfg.ffn_dim,
        encoder_attention_heads=cfg.heads, subsampling_type=cfg.subsampling_type, subsampling_layers=cfg.subsampling_layers,
        subsampling_filter=cfg.subsampling_filter, subsampling_kernel=cfg.subsampling_kernel,
        subsampling_stride=cfg.subsampling_stride, subsampling_norm=cfg.subsampling_norm,
        subsampling_activation=cfg.subsampling_activation, encoder_attention_type=cfg.attn_type,
        macaron_style=cfg.macaron_style, use_cnn_module=cfg.use_cnn_module, cnn_module_kernel=cfg.cnn_module_kernel,
        encoder_activation_fn=cfg.encoder_activation_fn, activation_fn=cfg.activation_fn,
        layer_padding_mask=cfg.layer_padding_mask, dropout=cfg.dropout, vocab=cfg.vocab)
    if cfg.tgt_vocab > 0 or cfg.inter_ctc_layers:
        over.update(xctc_weight=0.1 if cfg.tgt_vocab > 0 else 0.0, inter_ctc_weight=0.1 if cfg.inter_ctc_layers else 0.0,
                    inter_ctc_layers=cs(cfg.inter_ctc_layers), share_inter_ctc=cfg.share_inter_ctc,
                    inter_xctc_weight=0.05 if cfg.inter_xctc_layers else 0.0, inter_xctc_layers=cs(cfg.inter_xctc_layers),
                    ctc_pae=cfg.ctc_pae, xctc_pae=cfg.xctc_pae, pae_unnorm_input=cfg.pae_unnorm_input)
    if pds:
        n = len(cfg.pds_layers)
        over.update(pds_stages=n, pds_layers=us(cfg.pds_layers), pds_ratios=us(cfg.pds_ratios), pds_fusion=False,
                    pds_embed_dims=us(cfg.pds_dims), pds_kernel_sizes=us(cfg.pds_kernels),
                    pds_ffn_ratios=us(cfg.pds_ffn_ratios), pds_attn_heads=us(cfg.pds_heads),
                    pds_position_embed=us(cfg.pds_position_embed), pds_embed_norm=cfg.pds_embed_norm,
                    pds_ds_method="conv", pds_attn_ds_ratios=us([1] * n), pds_dropout=cfg.pds_dropout)
    args = reference_args(**over)
    task = ref.FakeTask(cfg.vocab, cfg.tgt_vocab or cfg.vocab)
    if pds:
        ref.pds.pdss2t_transformer_s_8(args)
        return ref.pds.PDSS2TTransformerEncoder(args, task, None)
    ref.s2t.s2t_transformer_s(args)
    return ref.s2t.S2TTransformerEncoder(args, task, None)


def run_reference_encoder(case: str, seed: int = 0):
    from oracle import encoder as OE
    from oracle import ref_shim
    ref = ref_shim.load()
    spec = dict(ENC_CASES[case])
    B, T, S, vocab = spec["B"], spec["T"], spec["S"], spec["vocab"]
    family = spec.get("family", "s2t")
    args = reference_args(**spec)
    task = ref.FakeTask(vocab, vocab)
    torch.manual_seed(seed)
    if family == "pds":
        ref.pds.pdss2t_transformer_s_8(args)                # arch defaults, pdss2t_transformer.py:1562
        enc = ref.pds.PDSS2TTransformerEncoder(args, task, None)
    else:
        ref.s2t.s2t_transformer_s(args)                     # arch defaults, s2t_transformer.py:2408-2415
        enc = ref.s2t.S2TTransformerEncoder(args, task, None)
    shapes = {k: v.shape for k, v in enc.state_dict().items() if not k.endswith("_float_tensor")}
    sd = seeded_state_dict(shapes, seed)
    missing = enc.load_state_dict(sd, strict=False)
    assert not missing.unexpected_keys and all(k.endswith("_float_tensor") for k in missing.missing_keys), missing
    enc.train()                                             # dropout = 0, BatchNorm uses batch statistics

    feats, lens, tokens = synth_batch(B, T, S, vocab, seed + 1)
    multi = getattr(args, "xctc_weight", 0) > 0 or bool(getattr(args, "inter_ctc_layers", None))
    tgt_tokens = synth_batch(B, T, S + 2, vocab, seed + 2)[2] if multi else None
    wkw = dict(ctc_weight=args.ctc_weight, xctc_weight=getattr(args, "xctc_weight", 0.0),
               inter_ctc_weight=getattr(args, "inter_ctc_weight", 0.0),
               inter_xctc_weight=getattr(args, "inter_xctc_weight", 0.0))
    crit = (lambda o: OE.ctc_criterion_total(o, tokens, tgt_tokens, **wkw)) if multi else \
           (lambda o: OE.ctc_criterion_loss(o["ctc_logit"][0], o["encoder_padding_mask"][0], tokens))
    out = enc(feats, lens)
    logit = out["ctc_logit"][0]
    pad = out["encoder_padding_mask"][0]
    loss = crit(out)
    loss.backward()
    ref_grads = {k: p.grad.clone() for k, p in enc.named_parameters()}

    # ---- validate the restatement against the reference on the spot --------------------------------
    cfg = OE.config_from_args(args)
    orc = OE.PDSS2TTransformerEncoder(cfg) if family == "pds" else OE.S2TTransformerEncoder(cfg)
    o_keys, r_keys = set(orc.state_dict().keys()), set(shapes.keys())
    assert o_keys == r_keys, (sorted(o_keys - r_keys)[:5], sorted(r_keys - o_keys)[:5])
    orc.load_state_dict(sd)
    orc.train()
    o_out = orc(feats, lens)
    o_loss = crit(o_out)
    o_loss.backward()
    for key in ("inter_ctc_logits", "inter_xctc_logits", "xctc_logit"):
        assert len(o_out[key]) == len(out[key]), key
        for a, b in zip(o_out[key], out[key]):
            a, b = (a[0] if isinstance(a, list) else a), (b[0] if isinstance(b, list) else b)
            assert (a - b).abs().max().item() < 1e-4, key
    err_x = (o_out["encoder_out"][0] - out["encoder_out"][0]).abs().max().item()
    err_l = (o_out["ctc_logit"][0] - logit).abs().max().item()
    assert err_x < 1e-4 and err_l < 1e-4, (err_x, err_l)
    assert abs(o_loss.item() - loss.item()) < 1e-3 * abs(loss.item())
    for k, p in orc.named_parameters():
        den = ref_grads[k].norm().item() + 1e-8
        rel = (p.grad - ref_grads[k]).norm().item() / den
        assert rel < 1e-3, (k, rel)
    print(f"[{case}] oracle == reference: |dx|={err_x:.2e} |dlogit|={err_l:.2e} loss={loss.item():.4f}")

    fix = dict(
        case=case, seed=seed, weights_sha256=state_checksum(sd),
        param_names=np.array(sorted(shapes)), param_shapes=np.array([str(tuple(shapes[k])) for k in sorted(shapes)]),
        feats=feats.numpy(), lens=lens.numpy(), tokens=tokens.numpy(),
        encoder_out=out["encoder_out"][0].detach().numpy(), ctc_logit=logit.detach().numpy(),
        padding_mask=pad.numpy(), loss=np.float64(loss.item()),
        torch_version=torch.__version__,
    )
    if multi:
        fix["tgt_tokens"] = tgt_tokens.numpy()
        fix["xctc_logit"] = out["xctc_logit"][0].detach().numpy()
        fix["inter_ctc_logit_last"] = out["inter_ctc_logits"][-1][0].detach().numpy()
        fix["inter_xctc_logit_last"] = out["inter_xctc_logits"][-1].detach().numpy()
        fix["loss_weights"] = np.array([wkw["ctc_weight"], wkw["xctc_weight"], wkw["inter_ctc_weight"],
                                        wkw["inter_xctc_weight"]])
    for k in GRAD_KEYS:
        if k in ref_grads:
            fix["grad::" + k] = ref_grads[k].numpy()
    fix["grad_norms"] = np.array([ref_grads[k].norm().item() for k in sorted(ref_grads)])
    fix["grad_names"] = np.array(sorted(ref_grads))
    return fix


CURR_CASE = dict(ENC_CASES["c5_small"], ctc_pae_ground_truth_ratio=0.3, xctc_pae_ground_truth_ratio=0.4,
                 xctc_pae_ground_truth_only_mistake=True, pae_oracle_smooth=True)


def run_reference_curriculum(seed: int = 0):
    """The ground-truth curriculum of the prediction-aware adapters (reproduction_bil_ctc_synchronous.yaml:62-65) through
    the REFERENCE encoder: first pass (no grad) -> alignment oracle (restated criterion arithmetic over the Viterbi
    restatement: criterions/ctc.py and best_alignment.cu cannot run here) -> second pass with ``ctc_alignment_oracle``,
    the masks drawn by the reference's own torch.rand under a fixed seed.  The restatement must reproduce that pass
    bit-level with the same seed; the drawn masks, oracle tokens and forced states go into the fixture so that the CUDA
    path can be given exactly the same ones."""
    from oracle import encoder as OE
    from oracle import ref_shim
    ref = ref_shim.load()
    spec = dict(CURR_CASE)
    B, T, S, vocab = spec["B"], spec["T"], spec["S"], spec["vocab"]
    args = reference_args(**spec)
    task = ref.FakeTask(vocab, vocab)
    torch.manual_seed(seed)
    ref.s2t.s2t_transformer_s(args)
    enc = ref.s2t.S2TTransformerEncoder(args, task, None)
    shapes = {k: v.shape for k, v in enc.state_dict().items() if not k.endswith("_float_tensor")}
    sd = seeded_state_dict(shapes, seed)
    enc.load_state_dict(sd, strict=False)
    enc.train()
    feats, lens, tokens = synth_batch(B, T, S, vocab, seed + 1)
    tgt_tokens = synth_batch(B, T, S + 2, vocab, seed + 2)[2]
    wkw = dict(ctc_weight=args.ctc_weight, xctc_weight=args.xctc_weight, inter_ctc_weight=args.inter_ctc_weight,
               inter_xctc_weight=args.inter_xctc_weight)
    with torch.no_grad():
        first = enc(feats, lens)
    align = OE.ground_truth_alignment(first, tokens, tgt_tokens)
    assert set(align) == {"ctc", "xctc"}
    torch.manual_seed(1234)
    out = enc(feats, lens, ctc_alignment_oracle=align)
    fe_c, fe_x = out["inter_ctc_logits"][0][2], out["inter_xctc_logits"][0][2]
    assert (fe_c >= 0).any() and (fe_x >= 0).any(), "the curriculum drew no forced frame: raise the ratios"
    results = {}
    for masked in (False, True):
        enc.zero_grad(set_to_none=True)
        o = out if not masked else None
        if masked:
            torch.manual_seed(1234)
            o = enc(feats, lens, ctc_alignment_oracle=align)
        loss = OE.ctc_criterion_total(o, tokens, tgt_tokens, masked_loss=masked, **wkw)
        loss.backward()
        results[masked] = (loss.item(), {k: p.grad.clone() for k, p in enc.named_parameters()})
    # ---- the restatement, same seed -----------------------------------------------------------------
    cfg = OE.config_from_args(args)
    orc = OE.S2TTransformerEncoder(cfg)
    orc.load_state_dict(sd)
    orc.train()
    for masked in (False, True):
        orc.zero_grad(set_to_none=True)
        torch.manual_seed(1234)
        o_out = orc(feats, lens, ctc_alignment_oracle=align)
        assert torch.equal(o_out["inter_ctc_logits"][0][2], fe_c) and torch.equal(o_out["inter_xctc_logits"][0][2], fe_x)
        for a, b in zip(o_out["inter_xctc_logits"] + o_out["inter_ctc_logits"], out["inter_xctc_logits"] + out["inter_ctc_logits"]):
            assert (a[0] - b[0]).abs().max().item() < 1e-4
        assert (o_out["encoder_out"][0] - out["encoder_out"][0]).abs().max().item() < 1e-4
        o_loss = OE.ctc_criterion_total(o_out, tokens, tgt_tokens, masked_loss=masked, **wkw)
        o_loss.backward()
        ref_loss, ref_grads = results[masked]
        assert abs(o_loss.item() - ref_loss) < 1e-3 * abs(ref_loss)
        for k, p in orc.named_parameters():
            rel = (p.grad - ref_grads[k]).norm().item() / (ref_grads[k].norm().item() + 1e-8)
            assert rel < 1e-3, (k, rel)
    # the injected-mask interface reproduces the drawn pass
    masks = {"ctc": fe_c >= 0, "xctc": fe_x >= 0}
    inj = orc(feats, lens, ctc_alignment_oracle=align, oracle_masks=masks)
    assert (inj["encoder_out"][0] - out["encoder_out"][0]).abs().max().item() < 1e-5
    print(f"[c5_curr_small] oracle == reference under the curriculum: forced frames ctc {int((fe_c >= 0).sum())}, "
          f"xctc {int((fe_x >= 0).sum())} of {fe_c.numel()}; loss {results[False][0]:.4f} (masked loss {results[True][0]:.4f})")
    fix = dict(
        seed=seed, weights_sha256=state_checksum(sd), param_names=np.array(sorted(shapes)),
        param_shapes=np.array([str(tuple(shapes[k])) for k in sorted(shapes)]),
        feats=feats.numpy(), lens=lens.numpy(), tokens=tokens.numpy(), tgt_tokens=tgt_tokens.numpy(),
        oracle_ctc=align["ctc"][0].numpy(), oracle_xctc=align["xctc"][0].numpy(),
        states_ctc=align["ctc"][1].numpy(), states_xctc=align["xctc"][1].numpy(),
        mask_ctc=masks["ctc"].numpy(), mask_xctc=masks["xctc"].numpy(),
        force_emit_ctc=fe_c.numpy(), force_emit_xctc=fe_x.numpy(),
        encoder_out=out["encoder_out"][0].detach().numpy(), padding_mask=out["encoder_padding_mask"][0].numpy(),
        loss=np.float64(results[False][0]), loss_masked=np.float64(results[True][0]),
        torch_version=torch.__version__,
    )
    # a few whole gradients (the tensors the curriculum touches most directly) and the norm of every gradient
    keep = ["xctc_pae.embed_adapter.weight", "pae.embed_adapter.weight", "xctc_norm3.weight", "ctc_norm2.weight",
            "xctc.ctc_projection.weight", "layers.0.self_attn_layer_norm.bias", "embed_ln.weight"]
    for masked, tag in ((False, "grad::"), (True, "gradm::")):
        for k in keep:
            fix[tag + k] = results[masked][1][k].numpy()
        names = sorted(results[masked][1])
        fix[tag[:-2] + "_norms"] = np.array([results[masked][1][k].norm().item() for k in names])
    fix["grad_names"] = np.array(sorted(results[False][1]))
    return fix


def run_reference_fbank(seed: int = 0):
    import torchaudio
    import torchaudio.compliance.kaldi as K
    from oracle import fbank as OF
    from oracle import ref_shim
    ref = ref_shim.load()
    g = torch.Generator().manual_seed(seed)
    n = 16000 * 2 + 321
    wave = torch.randn((1, n), generator=g) * 3000.0 + 1500.0 * torch.sin(2 * np.pi * 440.0 * torch.arange(n) / 16000.0)
    feat = K.fbank(wave, num_mel_bins=80, sample_frequency=16000).numpy()          # audio_utils.py:74-76
    ucmvn = ref.utterance_cmvn.UtteranceCMVN()(feat.copy())
    mean = feat.mean(axis=0).astype(np.float32) + 0.25
    std = feat.std(axis=0).astype(np.float32) + 0.5
    gc = ref.global_cmvn.GlobalCMVN.__new__(ref.global_cmvn.GlobalCMVN)
    gc.mean, gc.std = mean, std
    gcmvn = gc(feat.copy())
    # validate the restatement
    o = OF.fbank(wave[0].numpy())
    assert np.abs(o - feat).max() < 2e-4, np.abs(o - feat).max()
    assert np.abs(OF.utterance_cmvn(feat) - ucmvn).max() == 0
    assert np.abs(OF.global_cmvn(feat, mean, std) - gcmvn).max() == 0
    print(f"[fbank] oracle == torchaudio {torchaudio.__version__}: max-abs {np.abs(o - feat).max():.2e}; CMVN bit-exact")
    return dict(wave=wave[0].numpy(), fbank=feat, utt_cmvn=ucmvn, gmean=mean, gstd=std, global_cmvn=gcmvn,
                torchaudio_version=torchaudio.__version__)


def main():
    GOLDEN.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(GOLDEN / "fbank_cmvn.npz", **run_reference_fbank())
    np.savez_compressed(GOLDEN / "enc_c5_curr_small.npz", **run_reference_curriculum())
    for case in ENC_CASES:
        np.savez_compressed(GOLDEN / f"enc_{case}.npz", **run_reference_encoder(case))
    for f in sorted(GOLDEN.glob("*.npz")):
        print(f.name, f.stat().st_size // 1024, "KiB")


if __name__ == "__main__":
    sys.exit(main())
