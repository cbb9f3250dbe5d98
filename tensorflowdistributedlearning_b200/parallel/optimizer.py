"""Synchronous multi-tower optimizer: tower-mean of gradients + Adam / momentum-SGD in fused kernels.

Semantics are MirroredStrategy's: every tower computes gradients of its own per-tower
mean loss, gradients are averaged over towers, and every mirror applies the identical update, so
mirrors stay bit-identical.  Adam is TensorFlow's variant (A19) with the exponential-decay schedule
of model.py:457-459 evaluated on the device (so CUDA graphs replay with a moving learning rate).

Gradient paths (``comm``):
  * ``"p2p"``   -- product path: per-bucket one-shot / two-shot kernels over NVLink peer memory, fused
                   with the 1/N scale and the optimizer (csrc/p2p_allreduce.cu).  No NCCL call.
  * ``"nccl"``  -- baseline: ``dist.all_reduce`` of the flat gradient + the fused local optimizer kernel.
  * ``"gloo"``  -- CPU towers (plumbing tests): ``dist.all_reduce`` + torch-op optimizer.
  * single tower: one fused optimizer kernel over the whole flat buffer.
Moving BatchNorm statistics are averaged over towers every step (X3) through the same machinery.
"""
from __future__ import annotations

import os

import math
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import reference as R
from .flat import FlatParams
from .tower import TowerGroup


class DistributedOptimizer:
    def __init__(self, model: torch.nn.Module, group: TowerGroup, kind: str = "adam", lr: float = 1e-3,
                 momentum: float = 0.9, weight_decay: float = 0.0, beta1: float = 0.9, beta2: float = 0.999,
                 eps: float = 1e-8, decay_steps: float = 0.0, decay_rate: float = 1.0, comm: Optional[str] = None,
                 bucket_cap_mb: float = 32.0, overlap: bool = True, sync_bn_stats: bool = True,
                 comm_max_blocks: int = 64, async_wgrad: bool = True, grad_wire: Optional[str] = None):
        if kind not in ("adam", "sgdm"):
            raise ValueError("kind must be 'adam' or 'sgdm'")
        self.group, self.kind = group, kind
        self.lr, self.momentum, self.weight_decay = lr, momentum, weight_decay
        self.beta1, self.beta2, self.eps = beta1, beta2, eps
        self.decay_steps, self.decay_rate = decay_steps, decay_rate
        self.sync_bn_stats = sync_bn_stats
        # CTAs per bucket kernel: they share the SMs with the backward pass they overlap (TDL_COMM_MAX_BLOCKS overrides)
        self.comm_max_blocks = int(os.environ.get("TDL_COMM_MAX_BLOCKS", comm_max_blocks))
        dev = group.device
        self.on_gpu = dev.type == "cuda"
        if comm is None:
            comm = ("p2p" if self.on_gpu else "gloo") if group.distributed else "local"
        if not group.distributed:
            comm = "local"
        self.comm = comm
        # how gradients cross NVLink on the P2P path: "bf16" (default: the fp32 buckets are packed to bf16 inside the
        # bucket kernel and reduced with fp32 accumulation -- half the reduce-scatter bytes) or "fp32"
        self.grad_wire = grad_wire or os.environ.get("TDL_GRAD_WIRE", "bf16")
        if self.grad_wire not in ("bf16", "fp32"):
            raise ValueError("grad_wire must be 'bf16' or 'fp32'")
        self._p2p = None
        self._p2p_stats = None
        alloc = None
        self._symm = {}
        self.nvls = False
        if comm == "nvls":  # p2p with the reduction inside NVSwitch (multimem); falls back to plain p2p if unavailable
            comm, self.nvls = "p2p", True
            self.comm = "p2p"
        if comm == "p2p":
            from .p2p import SymmetricBuffer

            def alloc(n, dtype, tag):  # noqa: E306
                want_mc = self.nvls and tag in ("param", "grad")
                try:
                    buf = SymmetricBuffer(n * torch.empty((), dtype=dtype).element_size(), multicast=want_mc)
                except Exception as e:  # noqa: BLE001
                    if not want_mc:
                        raise
                    if group.is_primary:
                        print(f"WARNING: NVLS symmetric memory unavailable ({type(e).__name__}: {e}); using two-shot P2P")
                    self.nvls = False
                    buf = SymmetricBuffer(n * torch.empty((), dtype=dtype).element_size())
                self._symm[tag] = buf
                return buf.tensor(dtype)

        self.flat = FlatParams(model, bucket_cap_mb=bucket_cap_mb, alloc=alloc)
        n = self.flat.total
        self.m1 = torch.zeros(n, dtype=torch.float32, device=dev)
        self.m2 = torch.zeros(n, dtype=torch.float32, device=dev) if kind == "adam" else None
        self.global_step = torch.zeros(1, dtype=torch.int64, device=dev)
        self.lr_state = torch.zeros(2, dtype=torch.float32, device=dev)
        # X2: tower 0's initial values everywhere (parameters over the P2P kernel when the symmetric heap exists,
        # the small statistics buffer and every non-P2P configuration through the process group)
        if comm == "p2p":
            from .p2p import P2PComm

            self._p2p = P2PComm(self._symm["grad"], self._symm["param"])
            self.nvls = self.nvls and self._p2p.has_multicast
            if self.grad_wire == "bf16" and not self.nvls:
                from .p2p import SymmetricBuffer

                self._symm["wire"] = SymmetricBuffer(self.flat.total * 2)
                self._p2p.set_wire(self._symm["wire"])
            else:
                self.grad_wire = "fp32"
            self._p2p_stats = P2PComm(self._symm["stats"], None)
            self._stats_staging = torch.zeros_like(self.flat.stats_flat)
            self._p2p.broadcast_params_(self.flat.total)
            torch.cuda.synchronize()
        else:
            self.grad_wire = "fp32"  # only the P2P kernels have a wire format of their own
            group.broadcast_(self.flat.param_flat)
        group.broadcast_(self.flat.stats_flat)
        # overlap machinery
        self.overlap = bool(overlap and comm == "p2p")
        self._pending = [0] * len(self.flat.buckets)
        self._launched = [False] * len(self.flat.buckets)
        self._side = torch.cuda.Stream(priority=-1) if self.overlap else None
        self._hooks = []
        if self.on_gpu:
            from ..ops import functional as TF

            TF.enable_direct_grad(True)  # native backward kernels accumulate straight into the flat gradient buckets
            TF.enable_async_wgrad(async_wgrad)
        if self.overlap:
            from ..ops import functional as TF

            for p in self.flat.params:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad_ready))
                TF.register_grad_ready(p, self._on_grad_ready)
        self._lr_ready = False

    # ---- schedule ----
    def _advance_schedule(self) -> None:
        if self.on_gpu:
            from ..ops import native_ops as O

            O.lr_schedule_step(self.global_step, self.lr_state, self.lr, self.decay_rate, self.decay_steps,
                               self.beta1, self.beta2)
        else:
            step = int(self.global_step.item())
            lr = R.exponential_decay(self.lr, step, self.decay_steps, self.decay_rate)
            t = step + 1
            self.lr_state[0] = lr
            self.lr_state[1] = lr * math.sqrt(1 - self.beta2 ** t) / (1 - self.beta1 ** t)
            self.global_step += 1

    def begin_step(self) -> None:
        """Call before backward when overlapping: advances the schedule so bucket kernels can run mid-backward."""
        self._advance_schedule()
        self._lr_ready = True
        if self.overlap:
            for b in self.flat.buckets:
                self._pending[b.index] = len(b.params)
                self._launched[b.index] = False
            self._side.wait_stream(torch.cuda.current_stream())

    # ---- bucket kernels ----
    def _launch_bucket(self, b) -> None:
        a, bb, c = ((self.momentum, self.weight_decay, 0.0) if self.kind == "sgdm"
                    else (self.beta1, self.beta2, self.eps))
        if self.grad_wire == "bf16":
            self._p2p.fused_step_wire(b.offset, b.numel, kind=1 if self.kind == "sgdm" else 2, zero_grad=True,
                                      m1=self.m1, m2=self.m2, lr_state=self.lr_state, a=a, b=bb, c=c, mode=0,
                                      max_blocks=self.comm_max_blocks)
            return
        self._p2p.fused_step(b.offset, b.numel, kind=1 if self.kind == "sgdm" else 2, grad_is_bf16=False,
                             zero_grad=True, m1=self.m1, m2=self.m2, lr_state=self.lr_state, a=a, b=bb, c=c,
                             mode=3 if (self.nvls and b.numel * 4 > (1 << 19)) else 0,
                             # NVLS: 32 CTAs are the optimum at every size (more only queue up inside the switch)
                             max_blocks=min(self.comm_max_blocks, 32) if self.nvls else self.comm_max_blocks)

    def _on_grad_ready(self, p) -> None:
        i = self.flat.bucket_of[id(p)]
        self._pending[i] -= 1
        if self._pending[i] == 0 and not self._launched[i]:
            # buckets must launch in the same order on every rank: only launch when all earlier ones did
            self._flush_ready()

    def _flush_ready(self) -> None:
        for b in self.flat.buckets:
            if self._launched[b.index]:
                continue
            if self._pending[b.index] != 0:
                break
            self._side.wait_stream(torch.cuda.current_stream())
            if self.on_gpu:
                from ..ops import functional as TF

                TF.join_side_streams(self._side)  # weight gradients may still be in flight on the wgrad stream
            with torch.cuda.stream(self._side):
                self._launch_bucket(b)
            self._launched[b.index] = True

    # ---- the step ----
    def step(self) -> None:
        """Average gradients over towers and apply the update; leaves gradients zeroed for the next step."""
        if not self._lr_ready:
            self._advance_schedule()
        self._lr_ready = False
        if self.on_gpu:
            from ..ops import functional as TF

            TF.join_side_streams()  # side-stream weight gradients must have landed in the flat buckets
        f = self.flat
        if self.comm == "p2p":
            if self.overlap:
                for b in f.buckets:  # anything autograd did not touch (unused params) still has to go out
                    self._pending[b.index] = 0
                self._flush_ready()
                torch.cuda.current_stream().wait_stream(self._side)
            else:
                for b in f.buckets:
                    self._launch_bucket(b)
            # one packed one-shot: moving statistics (X3) + loss / metric scalars (X4, X5)
            lo = 0 if self.sync_bn_stats else f.scalar_offset
            self._p2p_stats.all_reduce_mean_(self._stats_staging, lo, f.stats_flat.numel() - lo)
            return
        if self.comm in ("nccl", "gloo"):
            dist.all_reduce(f.grad_flat)
            packed = f.stats_flat if self.sync_bn_stats else f.scalars
            dist.all_reduce(packed)
            packed.div_(self.group.world)
        scale = 1.0 / self.group.world
        if self.on_gpu:
            from ..ops import native_ops as O

            if self.kind == "sgdm":
                O.sgd_momentum_flat(f.param_flat, f.grad_flat, self.m1, None, self.lr_state, self.momentum,
                                    self.weight_decay, scale, True)
            else:
                O.adam_flat(f.param_flat, f.grad_flat, self.m1, self.m2, None, self.lr_state, self.beta1, self.beta2,
                            self.eps, scale, True)
        else:
            g = f.grad_flat * scale
            with torch.no_grad():
                if self.kind == "sgdm":
                    g = g + self.weight_decay * f.param_flat
                    self.m1.mul_(self.momentum).add_(g)
                    f.param_flat.sub_(self.lr_state[0] * self.m1)
                else:
                    self.m1.mul_(self.beta1).add_((1 - self.beta1) * g)
                    self.m2.mul_(self.beta2).add_((1 - self.beta2) * g * g)
                    f.param_flat.sub_(self.lr_state[1] * self.m1 / (self.m2.sqrt() + self.eps))
                f.grad_flat.zero_()

    # ---- checkpoint state ----
    TWO_SHOT_MIN_BYTES = 1 << 19  # buckets above this take the reduce-scatter path (csrc/p2p_allreduce.cu, mode 0)

    def owned_ranges(self):
        """Element ranges of the flat buffers whose optimizer slots THIS tower keeps up to date.

        One-shot buckets (and every non-P2P path) update the full state on every tower.  Two-shot / NVLS buckets run
        the optimizer on the owned slice only -- rank r's share of the bucket's 4-element vectors, the same split
        ``two_shot_kernel`` computes -- so the other (W-1)/W of ``m1`` / ``m2`` on a tower is stale."""
        if self.comm != "p2p" or not self.group.distributed:
            return [(0, self.flat.total)]
        w, r = self.group.world, self.group.rank
        vec = 8 if self.grad_wire == "bf16" else 4  # elements per work unit of the bucket kernels
        out = []
        for b in self.flat.buckets:
            if b.numel * 4 <= self.TWO_SHOT_MIN_BYTES:
                out.append((b.offset, b.offset + b.numel))
                continue
            nv = b.numel // vec
            per = (nv + w - 1) // w
            lo, hi = min(per * r, nv), min(per * r + per, nv)
            if hi > lo:
                out.append((b.offset + lo * vec, b.offset + hi * vec))
            if r == w - 1 and b.numel % vec:
                out.append((b.offset + nv * vec, b.offset + b.numel))
        return out

    def consolidate_state(self) -> None:
        """Make ``m1`` / ``m2`` complete on every tower (collective: every tower calls it).  Needed before tower 0
        writes a checkpoint (X7: the reference's Saver reads the primary copy, which there holds the full Adam slots,
        model.py:117-121): with the sharded two-shot update each tower only owns a slice of every large bucket."""
        if self.comm != "p2p" or not self.group.distributed:
            return
        one_shot = {(b.offset, b.offset + b.numel) for b in self.flat.buckets
                    if b.numel * 4 <= self.TWO_SHOT_MIN_BYTES}
        for slot in (self.m1, self.m2):
            if slot is None:
                continue
            full = torch.zeros_like(slot)
            for lo, hi in self.owned_ranges():
                if (lo, hi) in one_shot and self.group.rank != 0:
                    continue  # replicated state: tower 0's copy stands for all
                full[lo:hi] = slot[lo:hi]
            self.group.all_reduce_sum_(full)  # disjoint supports: the sum is the concatenation of the owned slices
            slot.copy_(full)

    def state_dict(self):
        return {"global_step": int(self.global_step.item()), "m1": self.m1, "m2": self.m2, "kind": self.kind}

    def peer_timeout(self):
        """Who a P2P kernel gave up waiting for (None if nobody): see ``P2PComm.timed_out``."""
        for c in (self._p2p, self._p2p_stats):
            who = c.timed_out() if c is not None else None
            if who is not None:
                return who
        return None

    def current_lr(self) -> float:
        return float(self.lr_state[0].item())

    def close(self) -> None:
        from ..ops import functional as TF

        for h in self._hooks:
            h.remove()
        for p in self.flat.params:
            TF.unregister_grad_ready(p)
        for c in (self._p2p, self._p2p_stats):
            if c is not None:
                c.close()
        for b in self._symm.values():
            b.close()
        self._symm = {}
