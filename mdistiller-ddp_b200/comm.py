"""Collectives of the sharded step: reduce-scatter of a gradient bucket, all-gather of P.

Replaces torch DDP's bucketed all-reduce (created at tools/train.py:86) on this path:
per step two ``ncclReduceScatter`` (KD bucket, task bucket) and one ``ncclAllGather``
(parameters), student tensors only, issued through ``torch.distributed`` over NCCL /
NVLink.  Both run in place on the flat buffers (NCCL's in-place convention: the rank's
output/input is the slice ``[rank*shard, (rank+1)*shard)`` of the full buffer).
"""

from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist


class ShardComm:
    """NCCL implementation.  ``world_size == 1`` degenerates to no-ops."""

    def __init__(self, group: Optional[dist.ProcessGroup] = None):
        self.group = group
        if dist.is_available() and dist.is_initialized():
            self.world_size = dist.get_world_size(group)
            self.rank = dist.get_rank(group)
        else:
            self.world_size, self.rank = 1, 0

    # -- helpers -----------------------------------------------------------------
    def _shard(self, full: torch.Tensor) -> torch.Tensor:
        n = full.numel() // self.world_size
        return full[self.rank * n:(self.rank + 1) * n]

    # -- collectives -------------------------------------------------------------
    def reduce_scatter_sum(self, full: torch.Tensor, async_op: bool = False):
        """Sum over ranks; the rank's shard of ``full`` holds the result afterwards."""
        if self.world_size == 1:
            return None
        return dist.reduce_scatter_tensor(self._shard(full), full, op=dist.ReduceOp.SUM,
                                          group=self.group, async_op=async_op)

    def all_gather(self, full: torch.Tensor, async_op: bool = False):
        """Every rank's shard of ``full`` is distributed into everybody's ``full``."""
        if self.world_size == 1:
            return None
        return dist.all_gather_into_tensor(full, self._shard(full), group=self.group, async_op=async_op)

    def broadcast(self, full: torch.Tensor, src: int = 0) -> None:
        if self.world_size > 1:
            dist.broadcast(full, src=src, group=self.group)

    def all_reduce_sum(self, t: torch.Tensor, async_op: bool = False):
        if self.world_size == 1:
            return None
        return dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)

    def all_reduce_max(self, t: torch.Tensor):
        if self.world_size > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
