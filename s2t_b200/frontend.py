"""Host side of the fbank front end: constant tables + the drop-in ``get_fbank`` replacements.

Mirrors the reference interface ``fairseq/data/audio/audio_utils.py``:
  * ``_get_torchaudio_fbank(waveform, sample_rate, n_bins=80) -> np.ndarray`` (``:59-79``)
  * ``get_fbank(path_or_fp, n_bins=80)`` (``:82-103``; file decoding stays with the reference's
    ``get_waveform``, only the feature computation is replaced)
and adds the batched device entry point used when raw waveforms are collated to the GPU (config C2).

The mel filter matrix and the povey window are computed once on the host with the same fp32 torch
formulas torchaudio uses (``torchaudio/compliance/kaldi.py:436-511`` get_mel_banks, ``:_feature_window_function``)
— SURVEY.md §8(a1) measured that re-deriving them in another precision costs 2/3 of the error budget.
"""
from __future__ import annotations

import math
from functools import lru_cache
from typing import Optional, Tuple

import numpy as np
import torch

from . import ops

SAMPLE_RATE = 16000
FRAME_LEN, FRAME_SHIFT, FFT_LEN = 400, 160, 512


def _mel_scalar(f: float) -> float:
    return 1127.0 * math.log(1.0 + f / 700.0)


def mel_filterbank(n_mels: int = 80, low_freq: float = 20.0, high_freq: float = 0.0,
                   sample_rate: float = SAMPLE_RATE) -> torch.Tensor:
    """[n_mels, 257] fp32 triangular mel weights (Kaldi, no VTLN); last column is zero."""
    nyquist = 0.5 * sample_rate
    if high_freq <= 0.0:
        high_freq += nyquist
    bin_width = sample_rate / FFT_LEN
    lo, hi = _mel_scalar(low_freq), _mel_scalar(high_freq)
    delta = (hi - lo) / (n_mels + 1)
    idx = torch.arange(n_mels).unsqueeze(1)
    left, center, right = lo + idx * delta, lo + (idx + 1.0) * delta, lo + (idx + 2.0) * delta
    mel = (1127.0 * (1.0 + (bin_width * torch.arange(FFT_LEN // 2)) / 700.0).log()).unsqueeze(0)
    w = torch.max(torch.zeros(1), torch.min((mel - left) / (center - left), (right - mel) / (right - center)))
    return torch.nn.functional.pad(w, (0, 1)).to(torch.float32).contiguous()


def mel_ranges(mel_w: torch.Tensor) -> torch.Tensor:
    """[n_mels, 2] int32 = [lo, hi) range of non-zero bins of every filter."""
    nz = mel_w != 0
    n_bins = mel_w.shape[1]
    ar = torch.arange(n_bins).unsqueeze(0).expand_as(mel_w)
    lo = torch.where(nz, ar, torch.full_like(ar, n_bins)).min(dim=1).values
    hi = torch.where(nz, ar + 1, torch.zeros_like(ar)).max(dim=1).values
    lo = torch.minimum(lo, hi)
    return torch.stack([lo, hi], dim=1).to(torch.int32).contiguous()


def povey_window() -> torch.Tensor:
    return torch.hann_window(FRAME_LEN, periodic=False, dtype=torch.float32).pow(0.85).contiguous()


@lru_cache(maxsize=8)
def _tables(device_str: str, n_mels: int):
    dev = torch.device(device_str)
    w = mel_filterbank(n_mels)
    return w.to(dev), mel_ranges(w).to(dev), povey_window().to(dev)


def num_frames(num_samples: int) -> int:
    return 0 if num_samples < FRAME_LEN else 1 + (num_samples - FRAME_LEN) // FRAME_SHIFT


def fbank_batch(wave: torch.Tensor, num_samples: torch.Tensor, n_mels: int = 80, *,
                cmvn: Optional[str] = None, out: Optional[torch.Tensor] = None,
                out_dtype: torch.dtype = torch.float32, max_frames: Optional[int] = None,
                global_stats: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
                scratch: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Device entry point: wave fp32 [B, N] (int16 scale) on CUDA -> (features [B, T, n_mels], frames int32 [B]).

    ``cmvn`` in {None, "utterance", "global"}; padded frames are zero (collater semantics).
    """
    if not wave.is_cuda:
        raise RuntimeError("fbank_batch needs CUDA tensors (s2t_b200 has no CPU path)")
    mel_w, rng, win = _tables(str(wave.device), n_mels)
    if max_frames is None:
        max_frames = num_frames(wave.shape[1])
    if cmvn is None:
        return ops.fbank(wave, num_samples, mel_w, rng, win, max_frames, out=out, out_dtype=out_dtype)
    raw, frames = ops.fbank(wave, num_samples, mel_w, rng, win, max_frames, out=scratch)
    if cmvn == "utterance":
        if out is None:
            out = torch.empty(raw.shape, device=raw.device, dtype=out_dtype)
        ops.cmvn_utt(raw, frames, out=out)
    elif cmvn == "global":
        if out is None:
            out = torch.empty(raw.shape, device=raw.device, dtype=out_dtype)
        ops.cmvn_global(raw, frames, global_stats[0], global_stats[1], out=out)
    else:
        raise ValueError(f"unknown cmvn {cmvn!r}")
    return out, frames


def _get_torchaudio_fbank(waveform, sample_rate, n_bins: int = 80) -> Optional[np.ndarray]:
    """Drop-in for audio_utils.py:59-79: (N,) samples or the (channels, N) array ``get_waveform`` returns, int16 scale
    -> (T, n_bins) fp32 numpy.  Several channels are merged to mono by averaging, as the reference's sox ``channels 1``
    effect does (:67-72); the reference's own ``get_fbank`` hands the raw 2-D array to this hook."""
    if sample_rate != SAMPLE_RATE:
        raise ValueError("s2t_b200 fbank supports 16 kHz input (the path's sample rate)")
    arr = np.asarray(waveform.cpu() if isinstance(waveform, torch.Tensor) else waveform, dtype=np.float32)
    if arr.ndim == 2:
        arr = arr[0] if arr.shape[0] == 1 else arr.mean(axis=0)
    elif arr.ndim != 1:
        raise ValueError(f"s2t_b200 fbank: waveform must be (N,) or (channels, N), got shape {arr.shape}")
    w = torch.as_tensor(arr, dtype=torch.float32).reshape(1, -1).cuda()
    n = torch.tensor([w.shape[1]], dtype=torch.int32, device=w.device)
    feats, _ = fbank_batch(w, n, n_bins)
    return feats[0].cpu().numpy()


def get_fbank(path_or_fp, n_bins: int = 80, offset=None, size=None, get_waveform=None) -> np.ndarray:
    """Drop-in for ``get_fbank`` (audio_utils.py:82-103): decode with the reference's ``get_waveform(...,
    normalization=False)`` (int16 scale, quirk A1) — passed in, or imported from fairseq when available — and compute
    the features on the GPU.  Returns (T, n_bins) float32 numpy like the reference."""
    if get_waveform is None:
        try:
            from fairseq.data.audio.audio_utils import get_waveform  # type: ignore
        except Exception as e:  # pragma: no cover
            raise ImportError("get_fbank needs the reference's get_waveform (soundfile decoding stays with fairseq); "
                              "pass get_waveform=") from e
    sound, sample_rate = get_waveform(path_or_fp, normalization=False, offset=offset, size=size)
    sound = np.asarray(sound)
    if sound.ndim == 2:                      # (channels, N): the reference merges channels to mono (:70-72)
        sound = sound.mean(axis=0)
    return _get_torchaudio_fbank(sound, sample_rate, n_bins)


class UtteranceCMVNB200:
    """``UtteranceCMVN`` (feature_transforms/utterance_cmvn.py:8-45) on the device: (T, F) float32 numpy in / out."""

    def __init__(self, norm_means: bool = True, norm_vars: bool = True):
        self.norm_means, self.norm_vars = norm_means, norm_vars

    def __call__(self, x: np.ndarray) -> np.ndarray:
        t = torch.as_tensor(np.ascontiguousarray(x), dtype=torch.float32).unsqueeze(0).cuda()
        n = torch.tensor([t.shape[1]], dtype=torch.int32, device=t.device)
        return ops.cmvn_utt(t, n, norm_means=self.norm_means, norm_vars=self.norm_vars)[0].cpu().numpy()


class GlobalCMVNB200:
    """``GlobalCMVN`` (feature_transforms/global_cmvn.py:8-29): (x - mean) / std with ``.npz`` statistics."""

    def __init__(self, stats_npz_path: Optional[str] = None, mean=None, std=None):
        if stats_npz_path is not None:
            stats = np.load(stats_npz_path)
            mean, std = stats["mean"], stats["std"]
        self.mean = torch.as_tensor(np.asarray(mean), dtype=torch.float32).cuda()
        self.std = torch.as_tensor(np.asarray(std), dtype=torch.float32).cuda()

    def __call__(self, x: np.ndarray) -> np.ndarray:
        t = torch.as_tensor(np.ascontiguousarray(x), dtype=torch.float32).unsqueeze(0).cuda()
        n = torch.tensor([t.shape[1]], dtype=torch.int32, device=t.device)
        return ops.cmvn_global(t, n, self.mean, self.std)[0].cpu().numpy()


def patch_audio_utils() -> bool:
    """Replace the reference's fbank implementation in place (what the ``--user-dir`` module does for the data path):
    ``fairseq.data.audio.audio_utils._get_kaldi_fbank`` -> None (skip PyKaldi) and ``_get_torchaudio_fbank`` -> ours, so
    ``get_fbank`` and every caller (speech_to_text_dataset.py:204-226, data_utils.py:86-88) use the CUDA kernel."""
    try:
        from fairseq.data.audio import audio_utils  # type: ignore
    except Exception:
        return False
    audio_utils._get_kaldi_fbank = lambda *a, **k: None
    audio_utils._get_torchaudio_fbank = _get_torchaudio_fbank
    return True
