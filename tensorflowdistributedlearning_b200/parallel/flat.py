"""Flat parameter / gradient / moving-statistic storage with gradient buckets.

All trainable tensors of a tower live in ONE fp32 buffer (the master weights), their gradients in a
second buffer of identical layout, BatchNorm moving statistics in a third.  ``nn.Parameter.data`` /
``.grad`` and the module buffers become views, so autograd accumulates straight into the bucket the
communication kernel reads, one fused kernel updates a whole bucket, and the cross-tower traffic of
SURVEY 2.5 (X1 gradients, X2 initial mirror, X3 moving statistics) is three contiguous ranges.

Layout is in *reverse* registration order: the gradients that backward finishes first sit at the
lowest addresses, so bucket k is the contiguous range that becomes ready k-th.

The reference packs its 208 gradient tensors into one NCCL message after backward
(``num_packs = 1``, Test.ipynb:200); buckets here exist so the reduction can overlap backward.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, List, Optional

import torch
import torch.nn as nn

ALIGN = 8  # elements; keeps every tensor 32-byte (fp32) / 16-byte (bf16) aligned inside the flat buffers
N_SCALARS = 8  # per-step scalars that ride behind the moving statistics: slot 0 = loss, 1.. = metric sums (X4/X5)


@dataclass
class Bucket:
    index: int
    offset: int          # first element
    numel: int           # padded element count (multiple of ALIGN)
    params: List[nn.Parameter]


def _round_up(n: int, a: int = ALIGN) -> int:
    return (n + a - 1) // a * a


class FlatParams:
    def __init__(self, model: nn.Module, bucket_cap_mb: float = 32.0, grad_dtype: torch.dtype = torch.float32,
                 alloc: Optional[Callable[[int, torch.dtype, str], torch.Tensor]] = None):
        """``alloc(numel, dtype, tag)`` returns a zeroed flat tensor (tag in {'param','grad','stats'}); the P2P
        runtime passes an allocator backed by the symmetric heap.  Call after ``model.to(device)``."""
        if grad_dtype != torch.float32:
            # autograd requires .grad to match the fp32 master dtype; the bf16 wire format is exercised by the
            # kernel-level sweep (scripts/allreduce_sweep.py), not through nn.Parameter.grad.
            raise ValueError("FlatParams keeps fp32 gradients (autograd: .grad dtype must match the parameter)")
        self.model = model
        params = [p for p in model.parameters() if p.requires_grad]
        if not params:
            raise ValueError("model has no trainable parameters")
        device = params[0].device
        self.device = device
        alloc = alloc or (lambda n, dt, tag: torch.zeros(n, dtype=dt, device=device))
        order = list(reversed(params))
        self.offsets: Dict[int, int] = {}
        off = 0
        for p in order:
            self.offsets[id(p)] = off
            off += _round_up(p.numel())
        self.total = off
        self.param_flat = alloc(self.total, torch.float32, "param")
        self.grad_flat = alloc(self.total, grad_dtype, "grad")
        self.params = order
        for p in order:
            o = self.offsets[id(p)]
            view = self.param_flat[o:o + p.numel()].view(p.shape)
            view.copy_(p.data)
            p.data = view
            p.grad = self.grad_flat[o:o + p.numel()].view(p.shape)
        # buckets (contiguous, in readiness order)
        cap = max(int(bucket_cap_mb * (1 << 20) / 4), 1)
        self.buckets: List[Bucket] = []
        cur: List[nn.Parameter] = []
        start = 0
        size = 0
        for p in order:
            n = _round_up(p.numel())
            if cur and size + n > cap:
                self.buckets.append(Bucket(len(self.buckets), start, size, cur))
                start += size
                cur, size = [], 0
            cur.append(p)
            size += n
        if cur:
            self.buckets.append(Bucket(len(self.buckets), start, size, cur))
        self.bucket_of: Dict[int, int] = {id(p): b.index for b in self.buckets for p in b.params}
        # moving statistics (all floating-point buffers)
        self.stat_refs = []
        soff = 0
        for m in model.modules():
            for name, b in list(m._buffers.items()):
                if b is not None and b.is_floating_point():
                    self.stat_refs.append((m, name, soff, b.numel(), b.shape))
                    soff += _round_up(b.numel())
        self.stats_total = soff
        # the logged loss (tower mean, X5) and the per-step metric sums (X4) travel in the same buffer, so one packed
        # one-shot reduction per step covers moving statistics and scalars
        self.scalar_offset = soff
        self.stats_flat = alloc(self.stats_total + N_SCALARS, torch.float32, "stats")
        self.scalars = self.stats_flat[self.scalar_offset:self.scalar_offset + N_SCALARS]
        for m, name, o, n, shape in self.stat_refs:
            view = self.stats_flat[o:o + n].view(shape)
            view.copy_(m._buffers[name])
            m._buffers[name] = view

    # ---- helpers ----
    def zero_grad(self) -> None:
        self.grad_flat.zero_()

    def num_params(self) -> int:
        return sum(p.numel() for p in self.params)

    def check_views(self) -> None:
        """Assert nobody replaced .data/.grad (e.g. ``optimizer.zero_grad(set_to_none=True)``)."""
        base = self.grad_flat.untyped_storage().data_ptr()
        for p in self.params:
            if p.grad is None or p.grad.untyped_storage().data_ptr() != base:
                raise RuntimeError("a parameter's .grad no longer aliases the flat gradient buffer")
