"""TEST INFRASTRUCTURE — CPU restatement of the reference's length-bucketed batching for the data-parallel path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline leg may import this module.

Restated (pure Python loops, small cases) from
  * ``fairseq/data/data_utils_fast.pyx:20-103``  ``batch_by_size_vec`` — greedy packing of an ORDERED index list into
    batches of at most ``max_tokens`` padded tokens (= sentences x longest member) and ``max_sentences`` members, whose
    size is < ``bsz_mult`` or a multiple of it;
  * ``fairseq/data/audio/speech_to_text_dataset.py:505-512`` ``ordered_indices`` — descending frame count, ties in
    original order (``np.lexsort`` keys: last key primary);
  * ``fairseq/data/iterators.py:537-568`` ``ShardedIterator`` — batch i goes to rank i mod world, short ranks are padded
    with ``fill_value`` (the trainer then runs its dummy batch with zero weight, ``trainer.py:1045-1058,687-691``).

Pinned by ``oracle/make_golden_batching.py``: the reference's own Cython routine, compiled from its source where it
lies, is run on seeded length sets and its batch boundaries are stored in ``tests/golden/batch_by_size.npz``.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np


def batch_by_size_vec(indices: np.ndarray, num_tokens_vec: np.ndarray, max_tokens: int, max_sentences: int,
                      bsz_mult: int) -> List[np.ndarray]:
    """data_utils_fast.pyx:20-103.  ``num_tokens_vec[pos]`` is the length of ``indices[pos]``; a limit <= 0 is off."""
    n = int(indices.shape[0])
    if n == 0:
        return []
    assert max_tokens <= 0 or int(np.max(num_tokens_vec)) <= max_tokens, \
        f"Sentences lengths should not exceed max_tokens={max_tokens}"
    ends = np.zeros(n, dtype=np.int32)
    count = 0                    # index of the batch being built; ends[count] = end of the last complete batch
    batch_start = 0
    tail_max = 0                 # longest member of the tail [ends[count], pos]
    batch_max = 0                # longest member of the complete part [batch_start, ends[count])
    for pos in range(n):
        t = int(num_tokens_vec[pos])
        tail_max = tail_max if tail_max > t else t
        new_end = pos + 1
        new_max = batch_max if batch_max > tail_max else tail_max
        new_sent = new_end - batch_start
        new_tok = new_sent * new_max
        overflow = (max_sentences > 0 and new_sent > max_sentences) or (max_tokens > 0 and new_tok > max_tokens)
        fits_mult = new_sent < bsz_mult or new_sent % bsz_mult == 0
        if overflow:
            tail_tok = tail_max * (new_end - int(ends[count]))
            if max_tokens > 0 and tail_tok > max_tokens:      # the tail alone overflows: finalize two batches
                count += 1
                ends[count] = pos
                tail_max = t
            batch_start = int(ends[count])
            count += 1
            new_max = tail_max
        if overflow or fits_mult:
            ends[count] = new_end
            batch_max = new_max
            tail_max = 0
    if int(ends[count]) != n:
        count += 1
    return np.split(indices, ends[:count])


def ordered_indices(n_frames: Sequence[int], shuffle_perm: Optional[np.ndarray] = None) -> np.ndarray:
    """speech_to_text_dataset.py:505-512: primary key -n_frames, secondary the (shuffled or original) order."""
    first = np.arange(len(n_frames)) if shuffle_perm is None else np.asarray(shuffle_perm)
    return np.lexsort([first, [-int(v) for v in n_frames]])


def shard(batches: Sequence, num_shards: int, shard_id: int, fill_value=None) -> list:
    """iterators.py:537-568: every ``num_shards``-th batch from ``shard_id``, padded to ceil(len / num_shards)."""
    if shard_id < 0 or shard_id >= num_shards:
        raise ValueError("shard_id must be between 0 and num_shards")
    sharded_len = -(-len(batches) // num_shards)
    mine = list(batches[shard_id::num_shards])
    return mine + [fill_value] * (sharded_len - len(mine))
