"""CPU tests of the length-bucketed data-parallel batching.

The golden fixture tests/golden/batch_by_size.npz holds the batch boundaries the REFERENCE's own Cython routine
(fairseq/data/data_utils_fast.pyx, compiled from the source where it lies by oracle/make_golden_batching.py) produced for
seeded utterance-length sets.  Both the oracle restatement and the product's native routine (s2t_batch_by_size behind
s2t_b200.batching.batch_by_size — pure host code, callable without a GPU) must reproduce them bit-exactly.
"""
from pathlib import Path

import numpy as np
import pytest

GOLDEN = Path(__file__).resolve().parent / "golden" / "batch_by_size.npz"


def _cases():
    fx = np.load(GOLDEN)
    names = sorted({k.split("::")[0] for k in fx.files})
    return fx, names


@pytest.mark.parametrize("impl", ["oracle", "product"])
def test_batch_by_size_matches_reference_golden(impl):
    from oracle import batching as OB
    from s2t_b200 import batching as PB
    fx, names = _cases()
    assert len(names) >= 8
    for name in names:
        n_frames, order, ends = fx[f"{name}::n_frames"], fx[f"{name}::order"], fx[f"{name}::ends"]
        max_tokens, max_sent, mult = (int(v) for v in fx[f"{name}::params"])
        got_order = (OB if impl == "oracle" else PB).ordered_indices(n_frames)
        assert np.array_equal(got_order, order), name
        vec = n_frames[order].astype(np.int64)
        if impl == "oracle":
            batches = OB.batch_by_size_vec(order.astype(np.int64), vec, max_tokens, max_sent, mult)
        else:
            batches = PB.batch_by_size(order, vec, max_tokens if max_tokens > 0 else None,
                                       max_sent if max_sent > 0 else None, mult)
        got_ends = np.cumsum([len(b) for b in batches])
        assert np.array_equal(got_ends, ends), (name, impl)
        assert np.array_equal(np.concatenate(batches), order), name
        for b in batches:                                   # the constraints the packing promises
            if max_tokens > 0:
                assert len(b) * int(n_frames[b].max()) <= max_tokens
            if max_sent > 0:
                assert len(b) <= max_sent


def test_batch_by_size_edge_cases():
    from s2t_b200 import batching as PB
    assert PB.batch_by_size(np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.int64), 100, None, 8) == []
    with pytest.raises(AssertionError):                     # the reference asserts lengths <= max_tokens (pyx:30-32)
        PB.batch_by_size(np.arange(3), np.array([10, 200, 10]), 100, None, 1)
    one = PB.batch_by_size(np.array([5]), np.array([40000]), 40000, None, 8)
    assert len(one) == 1 and one[0].tolist() == [5]
    # property: any ordered random set is partitioned in order, every batch legal
    rng = np.random.RandomState(0)
    for trial in range(20):
        n = int(rng.randint(1, 300))
        lens = np.sort(rng.randint(1, 3000, size=n))[::-1].astype(np.int64)
        mt, mult = int(rng.randint(3000, 50000)), int(rng.choice([1, 2, 8]))
        bs = PB.batch_by_size(np.arange(n), lens, mt, None, mult)
        assert np.array_equal(np.concatenate(bs), np.arange(n))
        assert all(len(b) * int(lens[b].max()) <= mt for b in bs)


def test_shard_and_rank_schedule_cover_every_utterance_once():
    from oracle import batching as OB
    from s2t_b200.batching import BucketSet, rank_schedule, shard_batches
    assert shard_batches(list(range(7)), 4, 3, fill_value=None) == OB.shard(list(range(7)), 4, 3) == [3, None]
    assert shard_batches(list(range(8)), 4, 0) == [0, 4]
    with pytest.raises(ValueError):
        shard_batches([1], 2, 2)
    rng = np.random.RandomState(1)
    n_frames = np.clip(rng.lognormal(np.log(550), 0.6, size=700), 50, 3000).astype(np.int64)
    for world in (1, 2, 8):
        scheds = [rank_schedule(n_frames, 40000, world, r) for r in range(world)]
        assert len({len(s) for s in scheds}) == 1            # every rank runs the same number of steps (dummy padded)
        real = np.concatenate([rb.indices for s in scheds for rb in s if not rb.dummy])
        assert np.array_equal(np.sort(real), np.arange(700))
        n_dummy = sum(rb.dummy for s in scheds for rb in s)
        assert n_dummy < world
        for s in scheds:
            for rb in s:
                B, Tq = rb.key
                assert B == len(rb.indices) and Tq >= int(n_frames[rb.indices].max())
                assert B * int(n_frames[rb.indices].max()) <= 40000
    bs = BucketSet()
    grid = bs.all_frames()
    assert grid == sorted(set(grid)) and len(grid) <= 48 and grid[0] == 64
    for t in (1, 63, 64, 65, 998, 1024, 1025, 2048, 2049, 5999, 6000):
        q = bs.frames(t)
        assert q >= t and q in grid and (q - t) < 256
    assert bs.frames(998) == 1024 and bs.key(40, 998) == (40, 1024)
