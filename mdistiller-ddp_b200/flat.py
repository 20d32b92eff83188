"""Flat fp32 parameter space and its data-parallel sharding.

Layout in HBM: all learnable student tensors, in
``get_learnable_parameters()`` order, live back to back in ONE buffer ``P``; each tensor
starts on a 16-byte boundary (4 elements) and the total is padded to a multiple of
``4 * world_size`` so that every rank's shard is itself 16-byte aligned.  Gradient
buckets have the same layout, so a parameter's gradient is the same slice of a bucket.
Momentum lives only for the rank's own shard (ZeRO-1 style).  Padding elements are zero
and stay zero under every update rule used here.
"""

from __future__ import annotations

from typing import List, Sequence, Tuple

import torch

ALIGN = 4  # elements (16 bytes)


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class FlatSpace:
    def __init__(self, params: Sequence[torch.Tensor], world_size: int = 1, rank: int = 0, flat_alloc=None):
        params = list(params)
        if not params:
            raise ValueError("no parameters")
        dev, dt = params[0].device, params[0].dtype
        if dt != torch.float32:
            raise TypeError("the flat update kernels are fp32 only")
        for p in params:
            if p.device != dev or p.dtype != dt:
                raise ValueError("all parameters must share one device and dtype")
        self.params: List[torch.Tensor] = params
        self.device = dev
        self.world_size, self.rank = int(world_size), int(rank)
        self.offsets: List[int] = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off = _round_up(off + p.numel(), ALIGN)
        self.used = off
        self.total = _round_up(max(off, ALIGN), ALIGN * self.world_size)
        self.shard_len = self.total // self.world_size
        self.shard_lo = self.rank * self.shard_len
        self.shard_hi = self.shard_lo + self.shard_len
        # tensor i owns [offsets[i], ends[i]) including its alignment padding
        self.ends: List[int] = self.offsets[1:] + [self.total]
        # full-length buffers (P and the gradient buckets) may come from a special allocator
        # (NVLink symmetric memory for the fused step); everything else is ordinary device memory
        self._flat_alloc = flat_alloc
        self.flat = self.new_flat_buffer()
        self._adopt()

    def _adopt(self) -> None:
        """Move every parameter's storage into the flat buffer (values preserved)."""
        with torch.no_grad():
            for p, v in zip(self.params, self.views(self.flat)):
                v.copy_(p.data)
                p.data = v

    def new_flat_buffer(self) -> torch.Tensor:
        """Zeroed full-length buffer from the flat allocator (collective if that one is)."""
        if self._flat_alloc is not None:
            return self._flat_alloc(self.total, self.device)
        return torch.zeros(self.total, dtype=torch.float32, device=self.device)

    def new_buffer(self) -> torch.Tensor:
        return torch.zeros(self.total, dtype=torch.float32, device=self.device)

    def new_shard_buffer(self) -> torch.Tensor:
        return torch.zeros(self.shard_len, dtype=torch.float32, device=self.device)

    def views(self, buf: torch.Tensor) -> List[torch.Tensor]:
        return [buf[o:o + p.numel()].view(p.shape) for o, p in zip(self.offsets, self.params)]

    def chunk_table(self, chunk_elems: int = 16384) -> torch.Tensor:
        """int64 [n_chunks, 4] = {tensor index, count, source offset, bucket offset}: how
        `mdb200_gather_flat` tiles the per-parameter gradients into a bucket (64 KB per thread block)."""
        rows = []
        for i, (o, p) in enumerate(zip(self.offsets, self.params)):
            n = p.numel()
            for s in range(0, n, chunk_elems):
                rows.append((i, min(chunk_elems, n - s), s, o + s))
        return torch.tensor(rows, dtype=torch.int64, device=self.device).reshape(-1, 4).contiguous()

    def shard(self, buf: torch.Tensor) -> torch.Tensor:
        return buf[self.shard_lo:self.shard_hi]

    def clip_to_shard(self, lo: int, hi: int) -> Tuple[int, int]:
        """Intersection of the global range [lo, hi) with this rank's shard (may be empty)."""
        return max(lo, self.shard_lo), min(hi, self.shard_hi)
