"""Batch collation with the padding done on the device.

The reference's ``SpeechToTextDataset.collater`` (fairseq/data/audio/speech_to_text_dataset.py:411-485) pads every
utterance to the longest one on the host (``_collate_frames``, :267-285), sorts the batch by descending frame count and
ships the padded [B, T, F] tensor.  Here the host only CONCATENATES the utterances (``pack_frames``: sum of lengths, no
padding bytes cross PCIe) and computes the same order; ``collate_frames`` (``s2t_collate_frames``) writes the zero-padded,
sorted batch straight into the engine's input buffer.  Token matrices stay on the host (``collate_tokens`` restates
fairseq/data/data_utils.py:33-67: they are a few hundred integers).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import ops


def pack_frames(frames: Sequence[torch.Tensor], pin: bool = True):
    """frames: list of [L_i, F] (features) or [L_i] (waveform) fp32 host tensors.  Returns (packed fp32 [sum L_i * F],
    offsets int64 [B] in elements, lengths int32 [B], order int64 [B]) — ``order`` is the collater's
    ``n_frames.sort(descending=True)`` permutation (:420-423), computed with the same torch call."""
    if len(frames) == 0:
        raise ValueError("pack_frames: empty batch")
    F = 1 if frames[0].dim() == 1 else int(frames[0].shape[1])
    lengths = torch.tensor([int(f.shape[0]) for f in frames], dtype=torch.long)
    offsets = torch.zeros(len(frames), dtype=torch.long)
    offsets[1:] = torch.cumsum(lengths[:-1] * F, 0)
    packed = torch.empty(int((lengths * F).sum()), dtype=torch.float32)
    if pin and torch.cuda.is_available():
        packed = packed.pin_memory()
    for f, o, n in zip(frames, offsets.tolist(), lengths.tolist()):
        packed[o:o + n * F].copy_(f.reshape(-1))
    _, order = lengths.sort(descending=True)
    return packed, offsets, lengths.to(torch.int32), order


def collate_frames(packed: torch.Tensor, offsets: torch.Tensor, lengths: torch.Tensor, order: Optional[torch.Tensor], T: int,
                   F: int, out: Optional[torch.Tensor] = None, lengths_out: Optional[torch.Tensor] = None
                   ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Device tensors in, padded batch out: out fp32 [B, T, F] ([B, T] when F == 1 and ``out`` is 2-D) in ``order``,
    frames >= length zero; lengths_out int32 [B].  ``out`` may be a (larger) plan buffer: its row stride is honoured."""
    if not packed.is_cuda:
        raise ops._lib.S2TLibraryError("collate_frames needs CUDA tensors (s2t_b200 has no CPU path)")
    B = lengths.shape[0]
    if out is None:
        out = torch.empty((B, T, F), device=packed.device, dtype=torch.float32)
    if lengths_out is None:
        lengths_out = torch.empty((B,), device=packed.device, dtype=torch.int32)
    if out.stride(0) < T * F or out.dtype != torch.float32:
        raise ValueError("collate_frames: out must be fp32 with a row stride of at least T * F")
    ops._call("s2t_collate_frames", ops._p(packed), ops._p(offsets), ops._p(lengths),
              ops._p(order.to(torch.int32)) if order is not None else None, B, T, F, ops._p(out), out.stride(0),
              ops._p(lengths_out), ops._stream())
    return out, lengths_out


def collate_tokens(values: List[torch.Tensor], pad_idx: int, eos_idx: int, left_pad: bool = False,
                   move_eos_to_beginning: bool = False) -> torch.Tensor:
    """fairseq.data.data_utils.collate_tokens (data_utils.py:33-67) for 1-D token lists (host side)."""
    size = max(v.size(0) for v in values)
    res = values[0].new_full((len(values), size), pad_idx)
    for i, v in enumerate(values):
        dst = res[i][size - len(v):] if left_pad else res[i][: len(v)]
        if move_eos_to_beginning:
            dst[0] = eos_idx
            dst[1:] = v[:-1]
        else:
            dst.copy_(v)
    return res
