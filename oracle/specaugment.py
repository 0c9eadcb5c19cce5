"""CPU restatement of the reference's SpecAugment feature transform (TEST INFRASTRUCTURE — only tests/, smoke() and
bench.py's CPU legs may import this; the product path is s2t_b200/specaugment.py + csrc/specaugment.cu).

Follows fairseq/data/audio/feature_transforms/specaugment.py:80-131 (``SpecAugmentTransform.__call__``):
  * mask value = the configured number, or the mean of the whole (T, F) spectrogram when None (:88-89);
  * no-op when T == 0 (:91-92) or F < freq_mask_F (:94-95);
  * freq_mask_N times: f = randint(0, F_max), f0 = randint(0, F - f), columns [f0, f0+f) := value when f != 0 (:115-119);
  * max_t = min(time_mask_T, floor(T * time_mask_p)); nothing more when max_t < 1 (:121-125);
  * time_mask_N times: t = randint(0, max_t), t0 = randint(0, T - t), rows [t0, t0+t) := value when t != 0 (:127-131).
The draws come from numpy's GLOBAL RandomState in exactly this order, so seeding it reproduces the reference's masks.
Time warping (:97-113, needs OpenCV) is outside the path: the recipes set time_warp_W = 0
(examples/speech_to_text/data_utils.py gen_config_yaml policies "lb", "ld", "sm", "ss").

Parity pinned: tests/golden/specaugment.npz holds outputs of the reference class itself (oracle/make_golden.py loads
the reference file unmodified); tests/test_oracle_cpu.py checks this restatement against them bit for bit.
"""
from __future__ import annotations

import math
from typing import List, Optional, Tuple

import numpy as np


def draw_masks(num_frames: int, num_freqs: int, *, freq_mask_n: int = 0, freq_mask_f: int = 0, time_mask_n: int = 0,
               time_mask_t: int = 0, time_mask_p: float = 0.0, rng=np.random
               ) -> Tuple[List[Tuple[int, int]], List[Tuple[int, int]]]:
    """The (start, width) pairs the reference would draw for one utterance, in its draw order (specaugment.py:115-131).
    Pairs with width 0 are kept (they consumed random numbers) and mean "no mask"."""
    freq: List[Tuple[int, int]] = []
    time: List[Tuple[int, int]] = []
    if num_frames == 0 or num_freqs < freq_mask_f:
        return freq, time
    for _ in range(freq_mask_n):
        f = rng.randint(0, freq_mask_f)
        f0 = rng.randint(0, num_freqs - f)
        freq.append((int(f0), int(f)))
    max_t = min(time_mask_t, math.floor(num_frames * time_mask_p))
    if max_t < 1:
        return freq, time
    for _ in range(time_mask_n):
        t = rng.randint(0, max_t)
        t0 = rng.randint(0, num_frames - t)
        time.append((int(t0), int(t)))
    return freq, time


def specaugment(spectrogram: np.ndarray, *, freq_mask_n: int = 0, freq_mask_f: int = 0, time_mask_n: int = 0,
                time_mask_t: int = 0, time_mask_p: float = 0.0, mask_value: Optional[float] = 0.0,
                rng=np.random) -> np.ndarray:
    """specaugment.py:80-131 without time warping."""
    assert spectrogram.ndim == 2
    out = spectrogram.copy()
    T, F = spectrogram.shape
    value = spectrogram.mean() if mask_value is None else mask_value      # :88-89 (before the early returns)
    if T == 0 or F < freq_mask_f:
        return spectrogram
    freq, time = draw_masks(T, F, freq_mask_n=freq_mask_n, freq_mask_f=freq_mask_f, time_mask_n=time_mask_n,
                            time_mask_t=time_mask_t, time_mask_p=time_mask_p, rng=rng)
    for f0, f in freq:
        if f != 0:
            out[:, f0:f0 + f] = value
    for t0, t in time:
        if t != 0:
            out[t0:t0 + t, :] = value
    return out
