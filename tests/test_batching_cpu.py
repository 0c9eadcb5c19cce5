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


def _sched_worker(rank, world, port, q):
    """One data-parallel rank on CPU (gloo): runs its length-bucketed schedule with a stand-in 'engine' whose gradient is
    a known function of the batch, ships it through BucketedAllReduce every step (dummy batches with weight 0), and
    mirrors the capture protocol of the real engine: the first time a plan key is met the warm-up pass must NOT issue
    collectives (engine.collective_free_warmup), or the ranks' all-reduce counts diverge."""
    import os
    import torch
    import torch.distributed as dist
    from s2t_b200.batching import rank_schedule
    from s2t_b200.dp import BucketedAllReduce
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.RandomState(3)
        n_frames = np.clip(rng.lognormal(np.log(550), 0.6, size=333), 50, 3000).astype(np.int64)
        sched = rank_schedule(n_frames, 40000, world, rank, pad_batch=True)
        flat = torch.zeros(64, dtype=torch.float32)
        ar = BucketedAllReduce(flat, [16, 32, 48], n_buckets=2)
        seen, n_collective_steps, sums = set(), 0, []
        for rb in sched:
            if rb.key not in seen:          # new plan: warm-up without the exchange, then "capture"
                seen.add(rb.key)
            w = 0.0 if rb.dummy else 1.0
            flat.fill_(w * float(n_frames[rb.indices].sum()))       # the step's "gradient"
            ar.reset()
            for off in (48, 32, 16, 0):
                ar.on_ready(off)
            ar.finish()
            n_collective_steps += 1
            sums.append(float(flat[0]))
        q.put((rank, len(sched), n_collective_steps, sum(rb.dummy for rb in sched), sums,
               float(sum(n_frames[rb.indices].sum() for rb in sched if not rb.dummy))))
    finally:
        dist.destroy_process_group()


def test_bucketed_schedule_world2_gloo():
    """world_size 2 on CPU: both ranks run the same number of steps (dummy-padded), every step's averaged 'gradient' is
    the mean over ranks of the REAL batches' contributions (a dummy contributes zero), every utterance is used once."""
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_sched_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    (r0, n0, c0, d0, s0, f0), (r1, n1, c1, d1, s1, f1) = res
    assert n0 == n1 == c0 == c1 and d0 + d1 <= 1
    assert s0 == s1                                             # the all-reduce gave both ranks the same averages
    rng = np.random.RandomState(3)
    n_frames = np.clip(rng.lognormal(np.log(550), 0.6, size=333), 50, 3000).astype(np.int64)
    assert f0 + f1 == float(n_frames.sum())
    assert abs(sum(s0) * 2 - float(n_frames.sum())) <= 1e-3 * float(n_frames.sum())


def test_out_frames_matches_the_reference_length_arithmetic():
    """engine.out_frames(T) — the number of output frames the REFERENCE returns for a batch padded to T frames (used to
    slice the outputs of a larger frame-bucket plan) — against the oracle encoders (pinned to the reference) run on CPU
    for the three stems: Conv1d x2, Conv2d x2 and the PDS stages with their padding quirk (A7)."""
    from types import SimpleNamespace
    import torch
    from oracle import encoder as OE
    from s2t_b200.engine import EncoderConfig, S2TEncoderEngine
    cases = [
        (dict(layers=1, ffn_dim=64, embed_dim=64, heads=1, subsampling_filter=64, vocab=0), False),
        (dict(layers=1, ffn_dim=64, embed_dim=64, heads=1, subsampling_type="conv2d", subsampling_filter=16,
              subsampling_kernel=3, subsampling_norm="batch2d", subsampling_activation="swish", vocab=0), False),
        (dict(encoder_type="pds", pds_layers=(1, 1, 1, 1), pds_dims=(64, 64, 64, 64), pds_heads=(1, 1, 1, 1),
              pds_ffn_ratios=(1, 1, 1, 1), vocab=0), True),
    ]
    from dataclasses import fields
    for kw, pds in cases:
        okw = {k: v for k, v in kw.items() if k in {f.name for f in fields(OE.OracleConfig)}}
        orc = (OE.PDSS2TTransformerEncoder if pds else OE.S2TTransformerEncoder)(OE.OracleConfig(dropout=0.0, **okw)).eval()
        fake = SimpleNamespace(cfg=EncoderConfig(dropout=0.0, **kw))
        for T in (9, 16, 63, 64, 83, 128, 250, 998):
            with torch.no_grad():
                out = orc(torch.randn(2, T, 80), torch.tensor([T, max(T - 3, 8)]))
            assert S2TEncoderEngine.out_frames(fake, T) == out["encoder_out"][0].shape[0], (kw.get("encoder_type", kw.get("subsampling_type", "conv1d")), T)
