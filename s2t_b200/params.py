"""Flat bf16 parameter / gradient storage with the reference's parameter names.

The reference keeps one ``nn.Parameter`` per tensor and flattens the gradients into one buffer only
for the all-reduce (``fairseq/distributed/legacy_distributed_data_parallel.py:82-113``).  Here the flat
layout is the primary one: every parameter is a 16-byte aligned view into ONE bf16 buffer, every
gradient the same view into a second buffer, laid out in forward order so that the backward pass
completes contiguous ranges from the end — the ranges the bucketed all-reduce ships while the rest of
the backward is still running (s2t_b200/dp.py).

Names and shapes are exactly the reference ``state_dict`` keys, so checkpoints move in
both directions; tensors that the kernels want in another layout (Conv1d weights as [Cout][k][Cin]) are
converted at this boundary only.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Iterable, List, Tuple

import torch

BF16 = torch.bfloat16
_ALIGN = 64  # elements (128 B): keeps every view TMA-legal and vector friendly


class ParamStore:
    def __init__(self, device: torch.device):
        self.device = device
        self._specs: "OrderedDict[str, Tuple[Tuple[int, ...], int]]" = OrderedDict()
        self._size = 0
        self.flat: torch.Tensor = None  # type: ignore
        self.grad: torch.Tensor = None  # type: ignore
        self._contig_groups: List[List[str]] = []

    # ---- declaration -------------------------------------------------------------------------
    def declare(self, name: str, shape: Iterable[int], pack: bool = False) -> None:
        """Reserve a tensor.  ``pack=True`` places it directly after the previous one (no alignment gap;
        used to make q/k/v projections one fused [3C][C] operand) — the caller guarantees its size is a
        multiple of 8 elements."""
        shape = tuple(int(s) for s in shape)
        n = 1
        for s in shape:
            n *= s
        if name in self._specs:
            raise KeyError(f"duplicate parameter {name}")
        if not pack:
            self._size = (self._size + _ALIGN - 1) // _ALIGN * _ALIGN
        self._specs[name] = (shape, self._size)
        self._size += n

    def finalize(self) -> None:
        total = (self._size + _ALIGN - 1) // _ALIGN * _ALIGN
        self.flat = torch.zeros((total,), device=self.device, dtype=BF16)
        self.grad = torch.zeros((total,), device=self.device, dtype=BF16)

    # ---- views -------------------------------------------------------------------------------
    def names(self) -> List[str]:
        return list(self._specs)

    def shape(self, name: str) -> Tuple[int, ...]:
        return self._specs[name][0]

    def offset(self, name: str) -> int:
        return self._specs[name][1]

    def numel(self, name: str) -> int:
        n = 1
        for s in self._specs[name][0]:
            n *= s
        return n

    def p(self, name: str) -> torch.Tensor:
        shape, off = self._specs[name]
        return self.flat[off:off + self.numel(name)].view(shape)

    def g(self, name: str) -> torch.Tensor:
        shape, off = self._specs[name]
        return self.grad[off:off + self.numel(name)].view(shape)

    def fused(self, first: str, rows: int, cols: int, grad: bool = False) -> torch.Tensor:
        """[rows][cols] view starting at ``first`` spanning the packed tensors that follow it."""
        off = self._specs[first][1]
        buf = self.grad if grad else self.flat
        n = rows * cols if cols else rows
        v = buf[off:off + n]
        return v.view(rows, cols) if cols else v

    def __contains__(self, name: str) -> bool:
        return name in self._specs

    @property
    def total(self) -> int:
        return int(self.flat.numel())

    def num_params(self) -> int:
        return sum(self.numel(n) for n in self._specs)
