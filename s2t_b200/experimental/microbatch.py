"""Concurrent micro-batches: one training step = several independent utterance groups on parallel CUDA streams.

Why: at the reference's recipe sizes (8 x 10 s per GPU: 2000 encoder rows) every kernel of the step is a few
microseconds long and uses a fraction of the 148 SMs, and the forward pass is one strictly serial chain of ~120
kernels — the step is bound by per-kernel latency, not by throughput.  Utterances are independent through
fbank -> encoder -> CTC (the only cross-utterance coupling is BatchNorm batch statistics in the
Conformer convolution module), so the batch is cut into ``n_micro`` groups whose forward/backward chains run side by
side on their own streams; kernels of different groups fill the SMs the other group leaves idle.

Semantics are those of the reference's ``update_freq`` gradient accumulation (fairseq/trainer.py:716-734 sums the
gradients and losses of the sub-batches before the optimizer step): every group has its own engine state (plan,
activation buffers, gradient buffer, dropout streams) over ONE shared parameter buffer; after the groups join, their
gradient buffers are summed into group 0's buffer (one fused add over the flat bf16 buffer) and the losses are added.
Models with BatchNorm (``use_cnn_module``) would change their batch statistics and are rejected.

The whole step — forks, joins, the gradient sum and the data-parallel all-reduce — is captured into one CUDA graph.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from .. import ops
from ..engine import EncoderConfig, Plan, S2TEncoderEngine


class MicroPlan:
    """The plans of the micro-batches of one (B, T, S_max) bucket plus the step-level outputs."""

    def __init__(self, subs: List[Plan], B: int, T: int):
        self.subs = subs
        self.B, self.T = B, T
        self.loss = torch.zeros_like(subs[0].loss)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.graph_kind = ""

    @property
    def out_len(self) -> torch.Tensor:
        return torch.cat([p.out_len for p in self.subs])


class MicroBatchEngine:
    def __init__(self, cfg: EncoderConfig, n_micro: int = 2, device: Optional[torch.device] = None):
        if n_micro < 1:
            raise ValueError("n_micro must be >= 1")
        if cfg.use_cnn_module and n_micro > 1:
            raise ValueError("micro-batches would change BatchNorm batch statistics (convolution module); use n_micro=1")
        self.cfg = cfg
        self.n_micro = n_micro
        self.engines = [S2TEncoderEngine(cfg, device) for _ in range(n_micro)]
        e0 = self.engines[0]
        for i, e in enumerate(self.engines[1:], 1):
            e.store.flat = e0.store.flat           # ONE parameter buffer; gradients stay per engine until the join
            e.SEED = e0.SEED + 7919 * i            # independent dropout masks per group
        if n_micro > 1:
            # every group gets its share of the SMs (a GEMM CTA needs a whole SM: unbounded GEMMs of two groups would
            # serialise): forward GEMMs, backward chain GEMMs, backward weight-gradient GEMMs
            import os
            sh = os.environ.get("S2T_MICRO_SHARES", "")
            if sh:
                f, c_, w = (int(v) for v in sh.split(","))
            else:
                f, c_, w = 148 // n_micro, 100 // n_micro, 48 // n_micro
            for e in self.engines:
                e.fwd_share = f
                e.sm_shares = (c_, w, -1)
        self.device = e0.device
        self.store = e0.store
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(n_micro - 1)]
        self.plans: Dict[tuple, MicroPlan] = {}
        self.dp = None
        self.grad_ready_hook = None     # (interface parity with S2TEncoderEngine; the groups report no partial progress)

    # ---- parameters (engine 0 owns the shared buffer) ---------------------------------------------------------
    def init_parameters(self, seed: int = 0) -> None:
        self.engines[0].init_parameters(seed)

    def load_state_dict(self, sd) -> None:
        self.engines[0].load_state_dict(sd)

    def state_dict(self):
        return self.engines[0].state_dict()

    def grads(self):
        return self.engines[0].grads()

    def num_params(self) -> int:
        return self.engines[0].num_params()

    def train(self, mode: bool = True):
        for e in self.engines:
            e.train(mode)
        return self

    def eval(self):
        return self.train(False)

    # ---- plans / inputs -----------------------------------------------------------------------------------------
    def plan(self, B: int, T: int, smax: int = 0, wave_len: int = 0) -> MicroPlan:
        if B % self.n_micro:
            raise ValueError(f"batch {B} is not a multiple of n_micro={self.n_micro}")
        key = (B, T, smax, wave_len)
        if key not in self.plans:
            b = B // self.n_micro
            self.plans[key] = MicroPlan([e.plan(b, T, smax, wave_len) for e in self.engines], B, T)
        return self.plans[key]

    def set_inputs(self, MP: MicroPlan, **kw) -> None:
        """Same keywords as ``S2TEncoderEngine.set_inputs``; every tensor is cut along the batch dimension into
        contiguous groups (host tensors pinned: the copies are asynchronous on the current stream)."""
        b = MP.B // self.n_micro
        for i, (e, P) in enumerate(zip(self.engines, MP.subs)):
            part = {k: (v[i * b:(i + 1) * b] if v is not None else None) for k, v in kw.items()}
            e.set_inputs(P, **part)

    # ---- the step ---------------------------------------------------------------------------------------------------
    def _fan_out(self, MP: MicroPlan, fn_name: str) -> None:
        main = torch.cuda.current_stream()
        for s, e, P in zip(self.streams, self.engines[1:], MP.subs[1:]):
            s.wait_stream(main)
            with torch.cuda.stream(s):
                getattr(e, fn_name)(P)
        getattr(self.engines[0], fn_name)(MP.subs[0])
        for s in self.streams:
            main.wait_stream(s)

    def forward(self, MP: MicroPlan) -> None:
        self._fan_out(MP, "forward")

    def train_step(self, MP: MicroPlan) -> None:
        """fwd+bwd of every group side by side; summed loss in MP.loss, summed gradients in store.grad."""
        if self.dp is not None:
            self.dp.reset()
        self._fan_out(MP, "train_step")
        g0 = self.engines[0].store.grad
        for e in self.engines[1:]:
            ops.add_(g0, e.store.grad)
        if self.n_micro == 1:
            MP.loss.copy_(MP.subs[0].loss)
        else:
            torch.add(MP.subs[0].loss, MP.subs[1].loss, out=MP.loss)
            for P in MP.subs[2:]:
                MP.loss.add_(P.loss)
        if self.dp is not None:
            self.dp.finish()          # on_ready(0): ships every bucket, then the compute stream waits for them

    def enable_data_parallel(self, n_buckets: int = 4, group=None) -> None:
        """Gradient all-reduce after the groups' gradients have been summed (bucketed so that NCCL pipelines the
        buckets; there is no backward left to overlap with once the sum exists)."""
        from ..dp import BucketedAllReduce
        S = self.store
        total = S.grad.numel()
        cuts = [total * k // n_buckets // 64 * 64 for k in range(1, n_buckets)]
        self.dp = BucketedAllReduce(S.grad, cuts, n_buckets, group)

    def capture(self, MP: MicroPlan, kind: str = "train") -> None:
        fn = self.train_step if kind == "train" else self.forward
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn(MP)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn(MP)
        MP.graph, MP.graph_kind = g, kind

    def run(self, MP: MicroPlan, kind: str = "train") -> None:
        if MP.graph is not None and MP.graph_kind == kind:
            MP.graph.replay()
        elif kind == "train":
            self.train_step(MP)
        else:
            self.forward(MP)
