"""Data-parallel gradient exchange: bucketed all-reduce overlapped with the backward pass.

Reference behaviour: ``LegacyDistributedDataParallel.all_reduce_grads``
(fairseq/distributed/legacy_distributed_data_parallel.py:76-160) copies every gradient into one flat
buffer, divides by the world size and issues ONE blocking all-reduce after backward has finished.

Here the gradients already live in one flat bf16 buffer laid out in forward order (params.ParamStore), so
the backward pass finalises it from the end.  The buffer is cut into a few contiguous buckets; the engine
reports "everything at or after offset X is final" as it goes (``on_ready``) and every bucket that became
final is handed to NCCL at once (async: it runs on the process group's own stream over NVLink/NVSwitch
while the remaining backward kernels run on the compute stream).  ``finish`` makes the compute stream wait
for the outstanding buckets — a stream dependency, not a host sync, so the whole step still captures into
one CUDA graph.  Result = mean over ranks, the same value as the reference's divide-then-sum.

The class only needs a flat tensor and ``torch.distributed``; the CPU (gloo) tests drive it directly.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def plan_buckets(total: int, cut_points: Sequence[int], n_buckets: int,
                 tail_fraction: float = 0.0) -> List[Tuple[int, int]]:
    """Split [0, total) into <= n_buckets contiguous ranges of roughly equal size whose boundaries are taken
    from ``cut_points`` (offsets at which the backward pass reports progress).  Returned in the order the
    backward pass completes them: from the end of the buffer to the start.

    ``tail_fraction`` > 0 asks for a small LAST bucket [0, c) with c the largest cut point <= tail_fraction * total:
    that bucket's all-reduce is the only one no backward kernel is left to hide, so it should carry little."""
    cuts = sorted({int(c) for c in cut_points if 0 < int(c) < total})
    n_buckets = max(1, int(n_buckets))
    lo = 0
    if tail_fraction > 0.0 and n_buckets > 1:
        small = [c for c in cuts if c <= tail_fraction * total]
        if small:
            lo = small[-1]
            n_buckets -= 1
    target = (total - lo) / n_buckets
    bounds = [total]
    hi = total
    for c in reversed(cuts):
        if len(bounds) >= n_buckets or c <= lo:
            break
        if hi - c >= target:
            bounds.append(c)
            hi = c
    if lo > 0:
        bounds.append(lo)
    bounds.append(0)
    out = []
    for i in range(len(bounds) - 1):
        if bounds[i] > bounds[i + 1]:
            out.append((bounds[i + 1], bounds[i]))
    return out


class BucketedAllReduce:
    def __init__(self, flat_grad: torch.Tensor, cut_points: Sequence[int], n_buckets: int = 4,
                 group: Optional[dist.ProcessGroup] = None, tail_fraction: float = 0.0):
        self.flat = flat_grad
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.buckets = plan_buckets(flat_grad.numel(), cut_points, n_buckets, tail_fraction)
        self.next = 0
        self.works: list = []
        self.launched: List[Tuple[int, int]] = []
        # NCCL has a native average; gloo (CPU tests) does not: sum then scale
        self._avg = bool(dist.is_initialized() and dist.get_backend(group) == "nccl")

    def reset(self) -> None:
        self.next = 0
        self.works = []
        self.launched = []

    def _launch(self, lo: int, hi: int) -> None:
        self.launched.append((lo, hi))
        if self.world == 1:
            return
        view = self.flat[lo:hi]
        if self._avg:
            self.works.append((dist.all_reduce(view, op=dist.ReduceOp.AVG, group=self.group, async_op=True), None))
        else:
            self.works.append((dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.group, async_op=True), view))

    def on_ready(self, offset: int) -> None:
        """Gradients in [offset, end) are final: ship every bucket that lies entirely inside."""
        while self.next < len(self.buckets) and self.buckets[self.next][0] >= offset:
            self._launch(*self.buckets[self.next])
            self.next += 1

    def would_ship(self, offset: int) -> bool:
        """True when ``on_ready(offset)`` would hand at least one bucket to the collective."""
        return self.next < len(self.buckets) and self.buckets[self.next][0] >= offset

    def finish(self) -> None:
        self.on_ready(0)
        for w, view in self.works:
            w.wait()                       # stream-level dependency for NCCL; blocking for gloo
            if view is not None:
                view.div_(self.world)
        self.works = []

    def bytes_on_wire_per_rank(self) -> float:
        """Ring/tree all-reduce moves 2 (n-1)/n of the payload per rank."""
        n = self.world
        return 2.0 * (n - 1) / n * self.flat.numel() * self.flat.element_size() if n > 1 else 0.0
