"""Double-buffered host -> device input pipeline for the training step.

The reference feeds the GPU from a prefetching DataLoader with pinned memory and non-blocking ``.cuda()`` copies
(fairseq/data/iterators.py: BufferedIterator; fairseq/utils.py: move_to_cuda), so the copy of batch i+1 overlaps the
compute of batch i.  The engine's step is one CUDA graph over FIXED input buffers, so the overlap needs one level of
staging: batch i+1 is copied H2D into staging tensors on a copy stream while the graph of batch i runs; ``commit``
(compute stream) waits for that copy and moves the staged batch into the plan's buffers with device-to-device copies
(a few microseconds for 5 MB), after which the staging tensors are free for batch i+2.

    feeder = InputFeeder(eng, P)
    feeder.stage(wave=w0, num_samples=n0, targets=t0, target_lengths=l0)     # pinned host tensors
    for batch in batches:
        feeder.commit()                 # batch i -> plan buffers (waits for its H2D)
        feeder.stage(**next_batch)      # H2D of batch i+1 overlaps ...
        eng.run(P)                      # ... this step
"""
from __future__ import annotations

from typing import Dict, Optional

import torch


class InputFeeder:
    def __init__(self, eng, P):
        self.eng, self.P = eng, P
        self.copy_stream = torch.cuda.Stream(device=eng.device)
        self.staged: Dict[str, torch.Tensor] = {}
        self.ev_staged = torch.cuda.Event()
        self.ev_free = torch.cuda.Event()
        self._pending = False

    def stage(self, **batch: Optional[torch.Tensor]) -> None:
        """Start the asynchronous H2D copy of one batch (same keywords as ``S2TEncoderEngine.set_inputs``; host tensors
        must be pinned for the copy to be asynchronous).  At most one batch is in flight."""
        if self._pending:
            raise RuntimeError("InputFeeder.stage: the previous batch has not been committed")
        cs = self.copy_stream
        cs.wait_event(self.ev_free)            # the last commit has finished reading the staging tensors
        with torch.cuda.stream(cs):
            for k, v in batch.items():
                if v is None:
                    self.staged.pop(k, None)
                    continue
                d = self.staged.get(k)
                if d is None or d.shape != v.shape or d.dtype != v.dtype:
                    d = torch.empty(v.shape, dtype=v.dtype, device=self.eng.device)
                    self.staged[k] = d
                d.copy_(v, non_blocking=True)
            self.ev_staged.record(cs)
        self._pending = True

    def commit(self) -> None:
        """Move the staged batch into the plan's input buffers on the current stream (device-to-device)."""
        if not self._pending:
            raise RuntimeError("InputFeeder.commit: no staged batch")
        cur = torch.cuda.current_stream()
        cur.wait_event(self.ev_staged)
        self.eng.set_inputs(self.P, **self.staged)
        self.ev_free.record(cur)
        self._pending = False
