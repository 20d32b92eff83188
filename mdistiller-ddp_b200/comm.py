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


class SymmetricShardComm(ShardComm):
    """Flat buffers in NVLink symmetric memory, so the step can run as ONE fused kernel
    (`mdb200_dot_step_fused`: in-switch gradient sum -> DOT update -> multicast parameter store)
    instead of reduce-scatter / update / all-gather launches.

    ``fuse_kd=False`` (default) keeps the KD bucket on an asynchronous NCCL reduce-scatter issued
    from ``step_kd()`` so that it overlaps the second backward; the fused kernel then pulls only
    the task bucket through the switch.  ``fuse_kd=True`` pulls both.
    ``use_multicast=False`` forces plain peer loads/stores even where NVLS is available.
    """

    fused = True

    def __init__(self, group: Optional[dist.ProcessGroup] = None, fuse_kd: bool = False,
                 use_multicast: bool = True, barrier_timeout_ms: int = 20000, in_kernel_sync: bool = True):
        super().__init__(group)
        if self.world_size > 16:
            raise ValueError("the fused kernel addresses at most 16 peers")
        self.fuse_kd = fuse_kd
        self.use_multicast = use_multicast
        self.barrier_timeout_ms = barrier_timeout_ms
        # True: the fused kernel orders itself against the other ranks through a sync pad in symmetric memory
        # (no barrier launches around it); False: two symmetric-memory barrier launches bracket it
        self.in_kernel_sync = in_kernel_sync
        self._handles = {}
        self._pad = None
        self._pad_ptrs = None
        self._epoch = 0

    def next_sync(self):
        """(pad address on every rank, this rank, epoch) for the next fused launch -- every rank must call this
        the same number of times -- or None when the launch is bracketed by barrier launches instead."""
        if not self.in_kernel_sync:
            return None
        if self._pad is None:
            import torch.distributed._symmetric_memory as symm_mem
            dev = torch.device("cuda", torch.cuda.current_device())
            self._pad = symm_mem.empty(64, dtype=torch.int32, device=dev)
            hdl = symm_mem.rendezvous(self._pad, self.group if self.group is not None else dist.group.WORLD)
            self._pad.zero_()
            self._pad_ptrs = [int(a) for a in hdl.buffer_ptrs]
            self._pad_handle = hdl
            hdl.barrier(channel=0, timeout_ms=self.barrier_timeout_ms)     # every pad is zero before anyone signals
        # (the kernel can also count the launches itself -- epoch argument 0, pad[34] -- which is what a CUDA-graph
        #  replay of the launch would need; the host-counted form is the one verified on 8 GPUs and is what ships)
        self._epoch += 1
        return self._pad_ptrs, self.rank, self._epoch

    def probe(self) -> None:
        """Collective: allocate and rendezvous a tiny symmetric buffer, so that a box without NVLink
        symmetric memory fails HERE (where the caller can still fall back to NCCL) and not half-way through
        adopting the parameters."""
        dev = torch.device("cuda", torch.cuda.current_device())
        t = self.alloc_flat(1024, dev)
        if self.use_multicast and not self.addresses(t)[0]:
            raise RuntimeError("no NVSwitch multicast address for symmetric memory on this box")
        self._handles.pop(t.data_ptr(), None)

    def alloc_flat(self, n: int, device: torch.device) -> torch.Tensor:
        """Collective: every rank must allocate the same sequence of buffers."""
        import torch.distributed._symmetric_memory as symm_mem
        t = symm_mem.empty(n, dtype=torch.float32, device=device)
        hdl = symm_mem.rendezvous(t, self.group if self.group is not None else dist.group.WORLD)
        t.zero_()
        self._handles[t.data_ptr()] = hdl
        return t

    def handle(self, t: torch.Tensor):
        return self._handles[t.data_ptr()]

    def addresses(self, t: torch.Tensor):
        """(multicast address or 0, [base address on every rank]) of a symmetric flat buffer."""
        h = self.handle(t)
        mc = int(h.multicast_ptr) if self.use_multicast else 0
        return mc, [int(a) for a in h.buffer_ptrs]

    def barrier(self, t: torch.Tensor, channel: int = 0) -> None:
        """Device-side barrier across ranks on the current stream (signal pad of `t`'s allocation)."""
        self.handle(t).barrier(channel=channel, timeout_ms=self.barrier_timeout_ms)
