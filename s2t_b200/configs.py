"""Named encoder configurations = BASELINE.json ``configs`` at their stated size.

Keyword sets for ``engine.EncoderConfig`` (the same keys ``oracle.encoder.OracleConfig`` reads), shared by
``bench.py --config`` and the full-size parity tests so that the measured and the tested model are the same.

  c2  s2t_transformer_s + CTC from raw waveform       egs/mustc/asr/conf/base.yaml + purectc.yaml (12 layers, d 256)
  c3  pdss2t_transformer_s_8 + CTC                    egs/mustc/asr/conf/pds_base_8.yaml, pdss2t_transformer.py:1496-1541
  c4  Conformer 12 layers d 256 + CTC                 egs/mustc/st/conf/conformer.yaml on s2t_transformer_s
  c5  BiL-CTC synchronous ST encoder (dual heads)     egs/mustc/st/conf/reproduction_bil_ctc_synchronous.yaml
"""
from __future__ import annotations

from typing import Dict

VOCAB = 10000

CONFIGS: Dict[str, dict] = {
    "c2": dict(vocab=VOCAB, dropout=0.1),
    "c3": dict(encoder_type="pds", vocab=VOCAB, dropout=0.1, pds_layers=(3, 3, 3, 3), pds_ratios=(2, 2, 1, 2),
               pds_dims=(256, 256, 256, 256), pds_kernels=(5, 5, 5, 5), pds_ffn_ratios=(8, 8, 8, 8),
               pds_heads=(4, 4, 4, 4), pds_embed_norm=True),
    "c4": dict(vocab=VOCAB, dropout=0.1, layers=12, macaron_style=True, use_cnn_module=True, cnn_module_kernel=15,
               attn_type="rel_pos", encoder_activation_fn="swish", layer_padding_mask=True),
    "c5": dict(vocab=VOCAB, tgt_vocab=VOCAB, dropout=0.15, embed_dim=512, ffn_dim=2048, heads=8, layers=18,
               subsampling_filter=2048, macaron_style=True, use_cnn_module=True, cnn_module_kernel=15,
               attn_type="rel_pos", encoder_activation_fn="swish", ctc_weight=0.2, xctc_weight=0.1,
               inter_ctc_layers=(6, 9, 12, 15), inter_ctc_weight=0.1, inter_xctc_layers=(6, 9, 12, 15),
               inter_xctc_weight=0.05, share_inter_ctc=True, ctc_pae="inter_league", xctc_pae="inter_league",
               pae_unnorm_input=True),
}

DESCRIPTIONS: Dict[str, str] = {
    "c2": "C2: s2t_transformer_s enc+CTC",
    "c3": "C3: pdss2t_transformer_s_8 (PDS 3_3_3_3, ratios 2_2_1_2, ffn x8) enc+CTC",
    "c4": "C4: Conformer 12L d256 (rel_pos MHA, macaron FFN, dwconv k15) enc+CTC",
    "c5": "C5: BiL-CTC s2t_transformer_m Conformer 18L d512 H8, CTC+XCTC heads, inter-CTC 6,9,12,15 + inter_league PAE",
}


def encoder_kwargs(name: str, **override) -> dict:
    if name not in CONFIGS:
        raise KeyError(f"unknown config {name!r}; choose from {sorted(CONFIGS)}")
    kw = dict(CONFIGS[name])
    kw.update(override)
    return kw
