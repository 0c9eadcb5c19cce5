"""Fused optimizer step over the engine's flat parameter buffer.

The reference trainer runs, per update, ``optimizer.multiply_grads(c)`` -> ``clip_grad_norm(clip_norm)`` ->
``optimizer.step()`` over a few hundred parameter tensors, with fp32 master copies under ``--fp16``
(fairseq/trainer.py:729-759, fairseq/utils.py:328-369, fairseq/optim/adam.py:145-231, fairseq/optim/fp16_optimizer.py).
The engine keeps ONE flat bf16 parameter buffer and ONE flat gradient buffer, so the whole update is two passes over the
gradients: ``s2t_grad_norm`` (global L2 norm into a device scalar) and ``s2t_fused_adam`` (scale, clip, Adam on fp32
master weights and moments, bf16 parameters written back) — no host sync, graph-capturable (lr and the step count may
live in device scalars).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops


class FusedAdamB200:
    """Adam with the reference's arithmetic (fairseq/optim/adam.py: bias-corrected step size, eps added to sqrt(v), weight
    decay applied to the parameter before the Adam update) on fp32 master weights of a flat bf16 parameter buffer."""

    def __init__(self, flat_param: torch.Tensor, flat_grad: torch.Tensor, lr: float = 1e-3,
                 betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
        if not flat_param.is_cuda:
            raise ops._lib.S2TLibraryError("FusedAdamB200 needs CUDA tensors (s2t_b200 has no CPU path)")
        assert flat_param.dtype == torch.bfloat16 and flat_grad.dtype == torch.bfloat16 and flat_param.numel() == flat_grad.numel()
        self.param, self.grad = flat_param, flat_grad
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        n, dev = flat_param.numel(), flat_param.device
        self.master = flat_param.detach().float().clone()
        self.exp_avg = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(n, device=dev, dtype=torch.float32)
        self.step_count = 0
        self._partials = torch.empty(ops._lib.load().s2t_grad_norm_partials(n), device=dev, dtype=torch.float32)
        self.total_norm = torch.zeros(1, device=dev, dtype=torch.float32)

    def grad_norm(self, grad_mult: float = 1.0) -> torch.Tensor:
        """total_norm = grad_mult * ||grad||_2 as a device scalar (utils.clip_grad_norm_'s norm after multiply_grads)."""
        ops._call("s2t_grad_norm", ops._p(self.grad), self.grad.numel(), float(grad_mult), ops._p(self._partials),
                  ops._p(self.total_norm), ops._stream())
        return self.total_norm

    def step(self, grad_mult: float = 1.0, max_norm: float = 0.0, lr: Optional[float] = None,
             lr_dev: Optional[torch.Tensor] = None, step_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
        """multiply_grads(grad_mult) -> clip_grad_norm(max_norm) -> Adam step -> bf16 parameters.  Returns the (device)
        gradient norm before clipping, as the trainer logs it (``gnorm``)."""
        self.step_count += 1
        self.grad_norm(grad_mult)
        b1, b2 = self.betas
        ops._call("s2t_fused_adam", ops._p(self.param), ops._p(self.master), ops._p(self.grad), ops._p(self.exp_avg),
                  ops._p(self.exp_avg_sq), self.param.numel(), float(self.lr if lr is None else lr), float(b1), float(b2),
                  float(self.eps), float(self.weight_decay), int(self.step_count), float(grad_mult), ops._p(self.total_norm),
                  float(max_norm), ops._p(lr_dev), ops._p(step_dev), ops._stream())
        return self.total_norm

    def state_dict(self):
        return {"step": self.step_count, "master": self.master, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq}

    def load_state_dict(self, sd) -> None:
        self.step_count = int(sd["step"])
        for k in ("master", "exp_avg", "exp_avg_sq"):
            getattr(self, k).copy_(sd[k])
        self.param.copy_(self.master.to(torch.bfloat16))
