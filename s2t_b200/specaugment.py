"""SpecAugment for a padded device batch: host-side mask draws in the reference's order, device-side apply.

Reference: ``SpecAugmentTransform`` (fairseq/data/audio/feature_transforms/specaugment.py:12-131), applied by the
dataset per utterance on the CPU after CMVN (``CompositeAudioFeatureTransform``, _train transforms
``[utterance_cmvn, specaugment]``).  Here the features never leave the GPU (fbank + CMVN kernels), so only the few
integers that describe the masks are produced on the host — with numpy's global RandomState and exactly the
reference's call order (f, f0 per frequency mask; then t, t0 per time mask; the early returns for empty / narrow
inputs and for max_time_mask_t < 1), so the same seed gives the same masks — and ``apply`` hands them to
``s2t_specaugment`` (csrc/specaugment.cu), which fills the rectangles in place (fill = constant or utterance mean).

Time warping (time_warp_W > 0, OpenCV resize on the CPU in the reference) is not part of this path: the recipes'
policies (examples/speech_to_text/data_utils.py: "lb", "ld", "sm", "ss") all use W = 0; a non-zero W is rejected.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops


@dataclass
class SpecAugmentConfig:
    """Same fields as the reference's config dict (specaugment.py:16-27)."""
    time_warp_W: int = 0
    freq_mask_N: int = 0
    freq_mask_F: int = 0
    time_mask_N: int = 0
    time_mask_T: int = 0
    time_mask_p: float = 0.0
    mask_value: Optional[float] = None

    @classmethod
    def from_config_dict(cls, config=None) -> "SpecAugmentConfig":
        c = {} if config is None else config
        return cls(c.get("time_warp_W", 0), c.get("freq_mask_N", 0), c.get("freq_mask_F", 0), c.get("time_mask_N", 0),
                   c.get("time_mask_T", 0), c.get("time_mask_p", 0.0), c.get("mask_value", None))

    def __post_init__(self):
        if self.time_warp_W:
            raise NotImplementedError("time_warp_W > 0 (OpenCV time warping) is outside the device path")
        if self.freq_mask_N > 0 and self.freq_mask_F <= 0:
            raise ValueError(f"freq_mask_F ({self.freq_mask_F}) must be larger than 0 when doing freq masking.")
        if self.time_mask_N > 0 and self.time_mask_T <= 0:
            raise ValueError(f"time_mask_T ({self.time_mask_T}) must be larger than 0 when doing time masking.")


def draw_masks(cfg: SpecAugmentConfig, num_frames: Sequence[int], num_freqs: int, rng=np.random
               ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Masks for a batch, utterance after utterance (the order a single-worker DataLoader calls the transform in):
    int32 [B, freq_mask_N, 2] = (f0, f) and [B, time_mask_N, 2] = (t0, t); rows of zeros where the reference draws
    nothing (T == 0, F < freq_mask_F, max_time_mask_t < 1)."""
    B = len(num_frames)
    fm = np.zeros((B, max(cfg.freq_mask_N, 0), 2), dtype=np.int32)
    tm = np.zeros((B, max(cfg.time_mask_N, 0), 2), dtype=np.int32)
    for b, T in enumerate(int(n) for n in num_frames):
        if T == 0 or num_freqs < cfg.freq_mask_F:                  # specaugment.py:91-95
            continue
        for i in range(cfg.freq_mask_N):                           # :115-119
            f = rng.randint(0, cfg.freq_mask_F)
            f0 = rng.randint(0, num_freqs - f)
            fm[b, i] = (f0, f)
        max_t = min(cfg.time_mask_T, math.floor(T * cfg.time_mask_p))   # :121-125
        if max_t < 1:
            continue
        for i in range(cfg.time_mask_N):                           # :127-131
            t = rng.randint(0, max_t)
            t0 = rng.randint(0, T - t)
            tm[b, i] = (t0, t)
    return torch.from_numpy(fm), torch.from_numpy(tm)


def apply_(cfg: SpecAugmentConfig, feats: torch.Tensor, num_frames: torch.Tensor, freq_masks: torch.Tensor,
           time_masks: torch.Tensor, n_freqs: Optional[int] = None, means: Optional[torch.Tensor] = None) -> torch.Tensor:
    """In place on the device batch feats [B, T, >= n_freqs] (fp32 or bf16); masks as returned by ``draw_masks``
    (moved to the device by the caller, e.g. with the batch's other small tensors)."""
    return ops.specaugment_(feats, num_frames, freq_masks if freq_masks.shape[1] else None,
                            time_masks if time_masks.shape[1] else None, cfg.mask_value, n_freqs, means)
