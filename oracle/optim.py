"""TEST INFRASTRUCTURE — the reference's update arithmetic for the fused-optimizer parity tests.

``reference_objects()`` extracts the inner ``Adam`` optimizer class (fairseq/optim/adam.py:104-231) and
``clip_grad_norm_`` (fairseq/utils.py:328-369) from the reference's files BY SOURCE (ast) and executes exactly those
definitions (the modules import the whole fairseq package and cannot be imported here).  ``update`` below restates the
trainer's per-update sequence on an fp32 master copy of bf16 parameters (fp16_optimizer.py: grads -> fp32, multiply,
clip, step, copy back); tests/golden/fused_adam.npz (oracle/make_golden_optim.py) pins it to the extracted objects.
"""
from __future__ import annotations

import ast
import math
import os
import warnings
from typing import Dict

import torch

REF = os.environ.get("S2T_REFERENCE", "/root/reference")


def reference_objects() -> Dict[str, object]:
    out = {}
    src = open(os.path.join(REF, "fairseq/optim/adam.py")).read()
    cls = next(n for n in ast.parse(src).body if isinstance(n, ast.ClassDef) and n.name == "Adam")
    ns = {"torch": torch, "math": math}
    exec(compile(ast.Module(body=[cls], type_ignores=[]), "fairseq/optim/adam.py", "exec"), ns)
    out["Adam"] = ns["Adam"]
    src = open(os.path.join(REF, "fairseq/utils.py")).read()
    fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "clip_grad_norm_")
    ns = {"torch": torch, "warnings": warnings, "multi_tensor_l2norm_available": False}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "fairseq/utils.py", "exec"), ns)
    out["clip_grad_norm_"] = ns["clip_grad_norm_"]
    return out


def update(master: torch.Tensor, exp_avg: torch.Tensor, exp_avg_sq: torch.Tensor, grad_bf16: torch.Tensor, step: int, *,
           lr: float, betas, eps: float, weight_decay: float, grad_mult: float, max_norm: float):
    """One update in place on fp32 (master, exp_avg, exp_avg_sq) from a bf16 gradient.  Returns (total_norm, bf16 params)."""
    g = grad_bf16.float() * grad_mult                                     # multiply_grads on the fp32 copy
    total_norm = torch.norm(g, p=2, dtype=torch.float32)                  # utils.py:340
    if max_norm > 0:
        g = g * (max_norm / (total_norm + 1e-6)).clamp_(max=1)            # :364-368
    b1, b2 = betas
    exp_avg.mul_(b1).add_(g, alpha=1 - b1)                                # adam.py:204-205
    exp_avg_sq.mul_(b2).addcmul_(g, g, value=1 - b2)
    denom = exp_avg_sq.sqrt().add_(eps)                                   # :212
    step_size = lr * math.sqrt(1 - b2 ** step) / (1 - b1 ** step)         # :214-216
    if weight_decay != 0:
        master.add_(master, alpha=-weight_decay * lr)                     # :218-221
    master.addcdiv_(exp_avg, denom, value=-step_size)                     # :223
    return total_norm, master.to(torch.bfloat16)
