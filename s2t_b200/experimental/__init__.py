"""Measured experiments that are NOT part of the default path (nothing in s2t_b200 imports this package).

* ``microbatch.MicroBatchEngine`` — utterance groups of one batch as concurrent kernel chains (update_freq semantics).
  GPU-tested for equivalence (tests/test_microbatch_gpu.py) but measured SLOWER than one chain at C2 (2.35 -> 2.65 ms with
  two groups: one-CTA-per-SM GEMMs of two chains cannot co-reside), so it stays an opt-in (``bench.py --micro N``).
"""
