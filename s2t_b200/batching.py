"""Length-bucketed data-parallel batching: the host side of SURVEY.md §8(e).

Mirrors the reference's data path for ``--max-tokens N --required-batch-size-multiple M`` training
(egs/mustc/asr/run.sh:74-75):

  * ``ordered_indices``  — utterances by descending frame count (speech_to_text_dataset.py:505-512);
  * ``batch_by_size``    — greedy packing into batches of <= max_tokens padded frames whose size is < M or a multiple
                           of M (data_utils_fast.pyx:20-103; here the native ``s2t_batch_by_size`` of the C ABI);
  * ``shard_batches``    — batch i goes to rank i mod world; ranks that run out get ``None`` and run their dummy batch
                           with zero weight so that every rank joins every gradient all-reduce
                           (iterators.py:537-568, trainer.py:1045-1058,687-691);
  * ``BucketSet``        — NOT in the reference: quantises the padded frame count of a batch to a fixed grid so that
                           the engine's pre-allocated plans and captured CUDA graphs are reused across batches
                           (engine.plan() keeps them in an LRU cache with a byte cap).

No device work happens here; the module needs the shared library only for the native packing routine and raises if it
is missing (no Python fallback).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Iterator, List, Optional, Sequence, Tuple

import numpy as np

from . import lib as _lib


def ordered_indices(n_frames: Sequence[int], shuffle_perm: Optional[np.ndarray] = None) -> np.ndarray:
    """Descending frame count; ties keep the (shuffled or original) order — ``np.lexsort`` with the frame key last,
    as speech_to_text_dataset.py:505-512."""
    first = np.arange(len(n_frames)) if shuffle_perm is None else np.asarray(shuffle_perm)
    return np.lexsort([first, -np.asarray(n_frames, dtype=np.int64)])


def batch_by_size(indices: Sequence[int], num_tokens_vec: Sequence[int], max_tokens: Optional[int] = None,
                  max_sentences: Optional[int] = None, required_batch_size_multiple: int = 1) -> List[np.ndarray]:
    """fairseq.data.data_utils.batch_by_size (data_utils.py:272-347) for the ``num_tokens_vec`` case:
    ``num_tokens_vec[i]`` is the length of ``indices[i]``.  Returns the list of index arrays."""
    idx = np.ascontiguousarray(indices, dtype=np.int64)
    vec = np.ascontiguousarray(num_tokens_vec, dtype=np.int64)
    if idx.shape != vec.shape or idx.ndim != 1:
        raise ValueError("batch_by_size: indices and num_tokens_vec must be 1-D and of equal length")
    if idx.shape[0] == 0:
        return []
    ends = np.zeros(idx.shape[0] + 1, dtype=np.int64)
    n = _lib.load().s2t_batch_by_size(vec.ctypes.data_as(C.c_void_p), int(vec.shape[0]),
                                      int(max_tokens) if max_tokens is not None else -1,
                                      int(max_sentences) if max_sentences is not None else -1,
                                      int(required_batch_size_multiple), ends.ctypes.data_as(C.c_void_p))
    if n < 0:
        # the reference asserts "Sentences lengths should not exceed max_tokens" (data_utils_fast.pyx:30-32)
        raise AssertionError(_lib.load().s2t_last_error().decode(errors="replace"))
    return np.split(idx, ends[: n - 1])


def shard_batches(batches: Sequence, num_shards: int, shard_id: int, fill_value=None) -> list:
    """ShardedIterator (iterators.py:537-568): every ``num_shards``-th batch starting at ``shard_id``, padded with
    ``fill_value`` to ceil(len / num_shards) entries."""
    if shard_id < 0 or shard_id >= num_shards:
        raise ValueError("shard_id must be between 0 and num_shards")
    sharded_len = -(-len(batches) // num_shards)
    mine = list(batches[shard_id::num_shards])
    return mine + [fill_value] * (sharded_len - len(mine))


@dataclass(frozen=True)
class BucketSet:
    """Fixed grid of padded frame counts.  Steps of ``fine`` frames up to ``fine_until``, ``mid`` up to ``mid_until``,
    ``coarse`` beyond: at most ~8 % padding above 800 frames, and a bounded number of distinct plans."""
    fine: int = 64
    fine_until: int = 1024
    mid: int = 128
    mid_until: int = 2048
    coarse: int = 256
    max_frames: int = 6000     # max_source_positions

    def frames(self, t: int) -> int:
        if t <= 0:
            raise ValueError("BucketSet.frames: empty batch")
        if t <= self.fine_until:
            step = self.fine
        elif t <= self.mid_until:
            step = self.mid
        else:
            step = self.coarse
        return min(-(-t // step) * step, max(self.max_frames, t))

    def all_frames(self) -> List[int]:
        out, t = [], 1
        while t <= self.max_frames:
            q = self.frames(t)
            if not out or out[-1] != q:
                out.append(q)
            t = q + 1
        return out

    def step_of(self, tq: int) -> int:
        return self.fine if tq <= self.fine_until else (self.mid if tq <= self.mid_until else self.coarse)

    def key(self, batch_size: int, max_frames: int, max_tokens: Optional[int] = None, bsz_mult: int = 8) -> Tuple[int, int]:
        """Plan key (B_q, T_q) of a batch.  Without ``max_tokens`` the batch size is kept (batch_by_size already makes it
        < M or a multiple of M).  With it, B_q is the LARGEST batch size max-tokens batching can produce for this frame
        bucket (every batch with T in (T_q - step, T_q] has B <= max_tokens / T), rounded to the multiple — so one plan
        (and one captured graph) serves the whole bucket; the caller pads the batch with EMPTY utterances (0 samples, 0
        targets), which contribute exactly nothing: no valid frame, CTC nll 0 (criterions/ctc.py zero_infinity path)."""
        tq = self.frames(int(max_frames))
        if max_tokens is None:
            return int(batch_size), tq
        lower = max(tq - self.step_of(tq) + 1, 1)
        cap = max(bsz_mult, (int(max_tokens) // lower) // bsz_mult * bsz_mult)
        need = -(-int(batch_size) // 64) * 64          # very short utterances: do not size the plan for max_tokens / 1
        return max(int(batch_size), min(cap, need)), tq


@dataclass
class RankBatch:
    indices: Optional[np.ndarray]      # dataset indices of this rank's batch (the dummy batch's when ``dummy``)
    dummy: bool                        # zero-weight stand-in: this rank ran out of batches (trainer.py:1045-1058)
    key: Tuple[int, int]               # (B, quantised T)


def rank_schedule(n_frames: Sequence[int], max_tokens: int, world: int, rank: int, *, bsz_mult: int = 8,
                  max_sentences: Optional[int] = None, buckets: Optional[BucketSet] = None,
                  shuffle_perm: Optional[np.ndarray] = None, pad_batch: bool = False) -> List[RankBatch]:
    """The sequence of batches rank ``rank`` of ``world`` runs for one epoch over utterances of ``n_frames`` frames:
    order -> pack -> shard, short ranks padded with their first batch as a zero-weight dummy.  ``pad_batch``: plan keys
    carry the bucket's full batch size (see BucketSet.key) — the caller pads with empty utterances."""
    buckets = buckets or BucketSet()
    nf = np.asarray(n_frames, dtype=np.int64)
    order = ordered_indices(nf, shuffle_perm)
    batches = batch_by_size(order, nf[order], max_tokens, max_sentences, bsz_mult)
    mine = shard_batches(batches, world, rank)
    out: List[RankBatch] = []
    dummy = next((b for b in mine if b is not None), batches[0] if batches else None)
    for b in mine:
        src = b if b is not None else dummy
        if src is None:
            raise RuntimeError("rank_schedule: no batch at all (fewer batches than ranks and no dummy available)")
        out.append(RankBatch(src, b is None, buckets.key(len(src), int(nf[src].max()), max_tokens if pad_batch else None,
                                                          bsz_mult)))
    return out
