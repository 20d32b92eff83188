#!/usr/bin/env python
"""Small / mid shapes of the loss kernel, 20 launches each (warm caches): where the fixed cost goes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdistiller_ddp_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
for (B, C, T, beta, ce) in ((64, 100, 4.0, 8.0, ops.CE_SPLIT), (256, 1000, 1.0, 0.5, ops.CE_SPLIT),
                            (4096, 1000, 1.0, 0.5, ops.CE_SPLIT), (16384, 1000, 1.0, 0.5, ops.CE_NONE),
                            (16384, 100, 4.0, 8.0, ops.CE_NONE), (65536, 100, 4.0, 8.0, ops.CE_NONE)):
    zs, zt = torch.randn(B, C, device=dev) * 3, torch.randn(B, C, device=dev) * 3
    y = torch.randint(0, C, (B,), device=dev)
    gk, gc, lo = torch.empty_like(zs), torch.empty_like(zs), torch.empty(2, device=dev)
    for _ in range(20):
        ops.distill_loss(zs, zt, y, kind=ops.LOSS_DKD, alpha=1.0, beta=beta, temperature=T, ce_mode=ce,
                         grad_kd=gk, grad_ce=gc, loss_out=lo)
    torch.cuda.synchronize()
print("ok")
