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

    def check_finite(self) -> float:
        """Host-side half of the overflow guard (one device->host read of the norm): the kernel has already skipped the
        update if the gradient norm was NaN / Inf; this undoes the host step count and raises FloatingPointError like
        the reference trainer (fairseq/trainer.py:751-753).  Returns the norm otherwise."""
        tn = float(self.total_norm.item())
        if not (tn == tn and abs(tn) != float("inf")):
            self.step_count -= 1
            raise FloatingPointError("gradients are Nan/Inf")
        return tn

    # ---- state --------------------------------------------------------------------------------------------------------
    def state_dict(self):
        """Flat state (the layout of the engine's parameter buffer): the cheapest form to save and restore between runs of
        THIS library.  ``reference_state_dict`` gives the reference's optimizer format instead."""
        return {"step": self.step_count, "master": self.master, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq}

    def load_state_dict(self, sd) -> None:
        self.step_count = int(sd["step"])
        for k in ("master", "exp_avg", "exp_avg_sq"):
            getattr(self, k).copy_(sd[k])
        self.param.copy_(self.master.to(torch.bfloat16))

    def reference_state_dict(self, engine, names=None):
        """The optimizer state in the reference's checkpoint format: torch's ``Optimizer.state_dict()`` of its Adam
        (fairseq/optim/adam.py:145-231: per parameter ``step`` / ``exp_avg`` / ``exp_avg_sq`` in fp32, the parameter's own
        shape) as ``FairseqOptimizer.state_dict`` returns it (last_optimizer_state of a checkpoint saved WITHOUT --fp16
        flattening, i.e. one state entry per parameter).  ``names``: the reference model's parameter order
        (``[n for n, _ in model.named_parameters()]``); default = the engine's own order."""
        return pack_reference_state(engine.export_flat(self.exp_avg), engine.export_flat(self.exp_avg_sq), self.step_count,
                                    names or engine.store.names(), lr=self.lr, betas=self.betas, eps=self.eps,
                                    weight_decay=self.weight_decay)

    def load_reference_state_dict(self, engine, sd, names=None) -> None:
        """Inverse of ``reference_state_dict``: moments and step count from a reference optimizer state; the fp32 master
        weights restart from the engine's current bf16 parameters (the reference keeps no separate master copy in this
        format; under --fp16 it rebuilds it from the model the same way, fp16_optimizer.py)."""
        ea, es, step = unpack_reference_state(sd, names or engine.store.names())
        engine.import_flat(ea, self.exp_avg)
        engine.import_flat(es, self.exp_avg_sq)
        self.step_count = step
        self.master.copy_(self.param.float())


def pack_reference_state(exp_avg, exp_avg_sq, step, names, *, lr, betas, eps, weight_decay):
    """name -> tensor dicts of the two moments -> torch ``Optimizer.state_dict()`` layout (state keyed by the position of
    the parameter in ``names``, one param group carrying the reference Adam's hyper-parameters)."""
    names = list(names)
    missing = [n for n in names if n not in exp_avg]
    if missing:
        raise KeyError(f"no optimizer state for {missing[:3]}")
    state = {i: {"step": int(step), "exp_avg": exp_avg[n].float().cpu(), "exp_avg_sq": exp_avg_sq[n].float().cpu()}
             for i, n in enumerate(names)}
    group = {"lr": float(lr), "betas": (float(betas[0]), float(betas[1])), "eps": float(eps),
             "weight_decay": float(weight_decay), "amsgrad": False, "params": list(range(len(names)))}
    return {"state": state, "param_groups": [group]}


def unpack_reference_state(sd, names):
    """torch ``Optimizer.state_dict()`` layout -> (exp_avg by name, exp_avg_sq by name, step).  Every parameter must carry
    the same step count (one flat update has one)."""
    names = list(names)
    order = [i for g in sd["param_groups"] for i in g["params"]]
    if len(order) != len(names):
        raise ValueError(f"optimizer state holds {len(order)} parameters, the model has {len(names)}")
    ea, es, steps = {}, {}, set()
    for n, i in zip(names, order):
        st = sd["state"][i]
        ea[n], es[n] = st["exp_avg"], st["exp_avg_sq"]
        steps.add(int(st["step"]))
    if len(steps) != 1:
        raise ValueError(f"parameters disagree on the step count: {sorted(steps)}")
    return ea, es, steps.pop()
