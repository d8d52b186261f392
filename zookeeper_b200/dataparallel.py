"""Host-side data-parallel plumbing: one process per GPU, ``torch.distributed`` (NCCL on the GPU
box, gloo in the CPU tests).  The batch shards by sample; the only exchange on the data path is a
sum all-reduce of the flat gradient arena (and, with SyncBN, of per-channel BatchNorm sums).
The reference has no multi-device code; N-GPU == 1-GPU at equal global batch
is the behaviour to match (section 8e)."""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.distributed as dist


class DataParallel:
    def __init__(self, process_group=None, enabled: bool = True):
        self.group = process_group
        active = enabled and dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(process_group) if active else 1
        self.rank = dist.get_rank(process_group) if active else 0

    # ---- sharding -------------------------------------------------------------------------
    def shard(self, global_batch_indices: np.ndarray) -> np.ndarray:
        """This rank's samples of a global batch (round-robin, so every rank gets B/world)."""
        n = len(global_batch_indices)
        if n % self.world:
            raise ValueError(f"global batch of {n} is not divisible by world size {self.world}")
        return global_batch_indices[self.rank::self.world]

    def global_count(self, local_rows: int, sync: bool) -> float:
        return float(local_rows * (self.world if sync else 1))

    @property
    def loss_scale(self) -> float:
        """Each rank scales its loss gradient by 1/(local batch * world): the SUM all-reduce then
        yields the gradient of the mean loss over the global batch."""
        return 1.0 / self.world

    # ---- collectives ------------------------------------------------------------------------
    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def all_reduce_sum_async_(self, t: torch.Tensor):
        """Start a sum all-reduce ordered behind the work already enqueued on the CURRENT stream and
        return its handle (``None`` on one replica); ``handle.wait()`` orders the then-current stream
        behind the collective without blocking the host."""
        if self.world == 1:
            return None
        return dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def reduce_gradients_(self, flat_grads: torch.Tensor) -> torch.Tensor:
        return self.all_reduce_sum_(flat_grads)

    def reduce_bn_sums_(self, *sums: torch.Tensor) -> None:
        for s in sums:
            self.all_reduce_sum_(s)
