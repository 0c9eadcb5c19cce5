"""CPU oracle for the fbank front end (TEST INFRASTRUCTURE — see oracle/__init__.py).

Restates, in numpy / torch-CPU fp32, the algorithm the reference runs at
``fairseq/data/audio/audio_utils.py:59-79`` — ``torchaudio.compliance.kaldi.fbank(waveform,
num_mel_bins=80, sample_frequency=16000)`` with every other argument at its default (torchaudio is a
third-party, unpinned dependency of the reference: setup.py:206; the algorithm restated here is
``torchaudio/compliance/kaldi.py:514-646`` + ``get_mel_banks`` ``:436-511`` of torchaudio 2.11.0) —
and the reference's own CMVN transforms
``fairseq/data/audio/feature_transforms/utterance_cmvn.py:30-45`` and ``global_cmvn.py:26-29``.

Pinned by tests/test_oracle_cpu.py against torchaudio itself and against golden vectors produced by
the reference's CMVN classes (oracle/make_golden.py).
"""
from __future__ import annotations

import math

import numpy as np
import torch

SAMPLE_RATE = 16000.0
FRAME_LEN = 400      # 25 ms
FRAME_SHIFT = 160    # 10 ms
FFT_LEN = 512        # round_to_power_of_two
PREEMPH = 0.97
EPS = float(np.finfo(np.float32).eps)  # torchaudio _get_epsilon: 1.1920929e-07


def num_frames(num_samples: int) -> int:
    """snip_edges=True (kaldi.py:_get_strided): 1 + (n - window) // shift."""
    if num_samples < FRAME_LEN:
        return 0
    return 1 + (num_samples - FRAME_LEN) // FRAME_SHIFT


def povey_window() -> np.ndarray:
    """kaldi.py:_feature_window_function 'povey': hann(N, periodic=False) ** 0.85, fp32."""
    return torch.hann_window(FRAME_LEN, periodic=False, dtype=torch.float32).pow(0.85).numpy()


def _mel_scalar(f: float) -> float:
    return 1127.0 * math.log(1.0 + f / 700.0)


def mel_banks(n_mels: int = 80, low_freq: float = 20.0, high_freq: float = 0.0) -> np.ndarray:
    """kaldi.py:get_mel_banks (no VTLN) in fp32 torch ops, right-padded with one zero column -> [n_mels, 257]."""
    num_fft_bins = FFT_LEN // 2
    nyquist = 0.5 * SAMPLE_RATE
    if high_freq <= 0.0:
        high_freq += nyquist
    fft_bin_width = SAMPLE_RATE / FFT_LEN
    mel_low, mel_high = _mel_scalar(low_freq), _mel_scalar(high_freq)
    delta = (mel_high - mel_low) / (n_mels + 1)
    b = torch.arange(n_mels).unsqueeze(1)
    left = mel_low + b * delta
    center = mel_low + (b + 1.0) * delta
    right = mel_low + (b + 2.0) * delta
    mel = (1127.0 * (1.0 + (fft_bin_width * torch.arange(num_fft_bins)) / 700.0).log()).unsqueeze(0)
    up = (mel - left) / (center - left)
    down = (right - mel) / (right - center)
    bins = torch.max(torch.zeros(1), torch.min(up, down))
    bins = torch.nn.functional.pad(bins, (0, 1), mode="constant", value=0.0)
    return bins.to(torch.float32).numpy()


def fbank(wave: np.ndarray, n_mels: int = 80) -> np.ndarray:
    """wave: fp32 [N] in int16 scale (audio_utils.py:92 passes un-normalised samples) -> [m, n_mels] fp32."""
    wave = np.asarray(wave, dtype=np.float32)
    m = num_frames(wave.shape[0])
    if m == 0:
        return np.zeros((0, n_mels), dtype=np.float32)
    idx = np.arange(FRAME_LEN)[None, :] + FRAME_SHIFT * np.arange(m)[:, None]
    frames = wave[idx]                                            # (m, 400) strided view (kaldi.py:_get_strided)
    frames = frames - frames.mean(axis=1, keepdims=True, dtype=np.float32)       # remove_dc_offset
    prev = np.concatenate([frames[:, :1], frames[:, :-1]], axis=1)               # replicate-pad left
    frames = frames - np.float32(PREEMPH) * prev                                  # preemphasis
    frames = frames * povey_window()[None, :]
    padded = np.zeros((m, FFT_LEN), dtype=np.float32)
    padded[:, :FRAME_LEN] = frames
    spec = torch.fft.rfft(torch.from_numpy(padded)).abs().pow(2.0)               # power spectrum (m, 257)
    mel = torch.mm(spec, torch.from_numpy(mel_banks(n_mels)).T)
    return torch.max(mel, torch.tensor(EPS)).log().numpy().astype(np.float32)


def utterance_cmvn(x: np.ndarray, norm_means: bool = True, norm_vars: bool = True) -> np.ndarray:
    """utterance_cmvn.py:30-45, same fp32 arithmetic order."""
    x = np.asarray(x, dtype=np.float32)
    mean = x.mean(axis=0)
    square_sums = (x ** 2).sum(axis=0)
    if norm_means:
        x = np.subtract(x, mean)
    if norm_vars:
        var = square_sums / x.shape[0] - mean ** 2
        std = np.sqrt(np.maximum(var, 1e-10))
        x = np.divide(x, std)
    return x.astype(np.float32)


def global_cmvn(x: np.ndarray, mean: np.ndarray, std: np.ndarray) -> np.ndarray:
    """global_cmvn.py:26-29."""
    x = np.subtract(np.asarray(x, dtype=np.float32), mean)
    x = np.divide(x, std)
    return x.astype(np.float32)
