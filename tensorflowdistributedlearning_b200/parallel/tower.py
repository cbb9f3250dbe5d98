"""Tower group: one process per GPU ("tower" == rank), CPU towers over gloo for plumbing tests.

Replaces ``tf.contrib.distribute.MirroredStrategy(devices=gpus[:n_gpus])`` (reference: model.py:114-119).
The reference replicates the graph inside one process; here every tower is its own process bound to
one GPU, bootstrapped by torchrun-style environment variables.  torch.distributed (NCCL / gloo) is
used for rendezvous, the initial parameter mirror, scalar reductions and as the *baseline* gradient
path; the product gradient path is the P2P kernel of ``parallel/p2p.py``.
"""
from __future__ import annotations

import datetime
import os
from typing import List, Optional

import torch
import torch.distributed as dist


class TowerGroup:
    def __init__(self, device: Optional[str] = None, backend: Optional[str] = None, timeout_s: Optional[int] = None):
        # collective timeout = the failure detector of the host-side path: a tower that dies or hangs makes its
        # peers raise after this many seconds instead of waiting forever (TDL_COLLECTIVE_TIMEOUT_S overrides)
        if timeout_s is None:
            timeout_s = int(os.environ.get("TDL_COLLECTIVE_TIMEOUT_S", "600"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", str(self.rank)))
        if device is None:
            device = "cuda" if torch.cuda.is_available() else "cpu"
        if device.startswith("cuda"):
            torch.cuda.set_device(self.local_rank % max(torch.cuda.device_count(), 1))
            self.device = torch.device("cuda", torch.cuda.current_device())
        else:
            self.device = torch.device("cpu")
        self.backend = backend or ("nccl" if self.device.type == "cuda" else "gloo")
        self._own_pg = False
        if self.world > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            kw = {}
            if self.backend == "nccl":
                kw["device_id"] = self.device
            dist.init_process_group(self.backend, rank=self.rank, world_size=self.world,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kw)
            self._own_pg = True

    # ---- identity ----
    @property
    def is_primary(self) -> bool:
        return self.rank == 0

    @property
    def distributed(self) -> bool:
        return self.world > 1

    # ---- collectives (bootstrap / baseline / scalar traffic) ----
    def barrier(self) -> None:
        if self.distributed:
            dist.barrier()

    def broadcast_(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        """X2: mirror tower-0's initial values into every tower."""
        if self.distributed:
            dist.broadcast(t, src)
        return t

    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        if self.distributed:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def all_reduce_mean_(self, t: torch.Tensor) -> torch.Tensor:
        if self.distributed:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            t.div_(self.world)
        return t

    def max_over_ranks(self, value: float) -> float:
        """Timing rule: every multi-GPU number is the max over ranks."""
        if not self.distributed:
            return float(value)
        t = torch.tensor([float(value)], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather_objects(self, obj) -> List:
        if not self.distributed:
            return [obj]
        out = [None] * self.world
        dist.all_gather_object(out, obj)
        return out

    def shutdown(self) -> None:
        if self._own_pg and dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:  # noqa: BLE001
                pass
            self._own_pg = False


def get_available_gpus() -> List[str]:
    """reference: utils.py:6-8 -- names of the local GPU devices."""
    return [f"/device:GPU:{i}" for i in range(torch.cuda.device_count())] if torch.cuda.is_available() else []
