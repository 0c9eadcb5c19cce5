"""Regenerate tests/golden/batch_by_size.npz from the REFERENCE's own Cython routine (TEST INFRASTRUCTURE).

    python -m oracle.make_golden_batching        # needs /root/reference, cython and g++

``fairseq/data/data_utils_fast.pyx`` is translated by Cython and compiled by g++ from the source where it lies
(outputs only under oracle/_ref/, git-ignored); ``batch_by_size_vec`` is then run on seeded utterance-length sets
(the recipe's max_tokens 40000 / bsz_mult 8 and edge cases: tail overflow, max_sentences, bsz_mult 1, a single
utterance, equal lengths), the restatement ``oracle/batching.py`` is asserted equal on the spot, and the reference's
batch boundaries are stored.  ``ordered_indices`` / ``ShardedIterator`` are pure Python in the reference and are
exercised through ``oracle/ref_shim``-free imports of numpy semantics only (lexsort / islice), stored as well.
"""
from __future__ import annotations

import importlib.util
import itertools
import math
import operator
import subprocess
import sys
import sysconfig
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
REF_PYX = Path("/root/reference/fairseq/data/data_utils_fast.pyx")
OUT_DIR = ROOT / "oracle" / "_ref"
GOLDEN = ROOT / "tests" / "golden" / "batch_by_size.npz"

# name -> (seed, n utterances, length distribution, max_tokens, max_sentences, bsz_mult)
CASES = {
    "recipe_lognormal": (0, 600, "lognormal", 40000, -1, 8),
    "recipe_small_cap": (1, 257, "lognormal", 6000, -1, 8),
    "max_sentences": (2, 100, "uniform", 20000, 12, 4),
    "mult1": (3, 77, "uniform", 9000, -1, 1),
    "tail_overflow": (4, 64, "spiky", 3000, -1, 8),
    "single": (5, 1, "uniform", 40000, -1, 8),
    "equal": (6, 41, "equal", 8000, -1, 8),
    "no_token_cap": (7, 50, "uniform", -1, 7, 2),
}


def build_reference_ext() -> Path:
    """cython + g++ on the reference's .pyx where it lies; the .so lands in oracle/_ref/."""
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    so = OUT_DIR / f"data_utils_fast{ext}"
    if so.exists():
        return so
    cpp = OUT_DIR / "data_utils_fast.cpp"
    subprocess.run([sys.executable, "-m", "cython", "--cplus", "-3", str(REF_PYX), "-o", str(cpp)], check=True)
    inc = [sysconfig.get_paths()["include"], np.get_include()]
    cmd = ["g++", "-O2", "-shared", "-fPIC", "-std=c++11", "-w", str(cpp), "-o", str(so)] + [f"-I{i}" for i in inc]
    subprocess.run(cmd, check=True)
    cpp.unlink()
    return so


def load_reference_ext():
    so = build_reference_ext()
    spec = importlib.util.spec_from_file_location("data_utils_fast", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def lengths(seed: int, n: int, kind: str) -> np.ndarray:
    rng = np.random.RandomState(seed)
    if kind == "lognormal":      # MuST-C-like durations, median ~5.5 s, clipped to 0.5-30 s, in frames
        secs = np.clip(rng.lognormal(mean=math.log(5.5), sigma=0.6, size=n), 0.5, 30.0)
        return (secs * 100).astype(np.int64)
    if kind == "uniform":
        return rng.randint(50, 3000, size=n).astype(np.int64)
    if kind == "spiky":          # short utterances with a few near-cap ones: exercises the tail-overflow branch
        v = rng.randint(20, 200, size=n).astype(np.int64)
        v[rng.choice(n, size=max(1, n // 8), replace=False)] = rng.randint(1400, 3000, size=max(1, n // 8))
        return v
    if kind == "equal":
        return np.full(n, 777, dtype=np.int64)
    raise ValueError(kind)


def reference_sharded(batches, num_shards, shard_id, fill_value=None):
    """fairseq/data/iterators.py:552-562 executed as written (zip_longest over islice)."""
    sharded_len = int(math.ceil(len(batches) / float(num_shards)))
    itr = map(operator.itemgetter(1),
              itertools.zip_longest(range(sharded_len), itertools.islice(batches, shard_id, len(batches), num_shards),
                                    fillvalue=fill_value))
    return list(itr)


def main():
    from oracle import batching as OB
    ref = load_reference_ext()
    out = {}
    for name, (seed, n, kind, max_tokens, max_sent, mult) in CASES.items():
        n_frames = lengths(seed, n, kind)
        order = np.lexsort([np.arange(n), [-int(v) for v in n_frames]])       # speech_to_text_dataset.py:505-512
        assert np.array_equal(order, OB.ordered_indices(n_frames))
        idx = order.astype(np.int64)
        vec = n_frames[idx].astype(np.int64)
        got = ref.batch_by_size_vec(idx, vec, max_tokens, max_sent, mult)
        mine = OB.batch_by_size_vec(idx, vec, max_tokens, max_sent, mult)
        assert len(got) == len(mine) and all(np.array_equal(a, b) for a, b in zip(got, mine)), name
        ends = np.cumsum([len(b) for b in got]).astype(np.int64)
        for world in (2, 8):
            for r in range(world):
                a = reference_sharded(list(range(len(got))), world, r, fill_value=-1)
                assert a == OB.shard(list(range(len(got))), world, r, fill_value=-1), (name, world, r)
        out[f"{name}::n_frames"] = n_frames
        out[f"{name}::order"] = order.astype(np.int64)
        out[f"{name}::ends"] = ends
        out[f"{name}::params"] = np.array([max_tokens, max_sent, mult], dtype=np.int64)
        print(f"[{name}] {n} utterances -> {len(got)} batches, restatement == reference")
    np.savez_compressed(GOLDEN, **out)
    print(f"wrote {GOLDEN}")


if __name__ == "__main__":
    main()
