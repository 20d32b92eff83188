#!/usr/bin/env python
"""Launch each hot kernel a few times at BASELINE.json's sweep sizes -- the target of
`ncu --set full` captures (see profiles/README.md for the command lines)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdistiller_ddp_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3


def loss(B, C, T, beta, ce):
    zs, zt = torch.randn(B, C, device=dev) * 3, torch.randn(B, C, device=dev) * 3
    y = torch.randint(0, C, (B,), device=dev)
    gk, gc, lo = torch.empty_like(zs), torch.empty_like(zs), torch.empty(2, device=dev)
    for _ in range(reps):
        ops.distill_loss(zs, zt, y, kind=ops.LOSS_DKD, alpha=1.0, beta=beta, temperature=T, ce_mode=ce,
                         grad_kd=gk, grad_ce=gc, loss_out=lo)
    torch.cuda.synchronize()


def dot(n):
    p, gt, gk = (torch.randn(n, device=dev) * 0.1 for _ in range(3))
    vt, vk = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    for _ in range(reps):
        ops.dot_step(p, gt, gk, vt, vk, lr=0.2, momentum=0.81, momentum_kd=0.99, weight_decay=1e-4)
    for _ in range(reps):
        ops.sgd_step(p, gt, vt, lr=0.05, momentum=0.9, weight_decay=5e-4)
    torch.cuda.synchronize()


def gather(shapes):
    from mdistiller_ddp_b200.flat import FlatSpace
    params = [torch.nn.Parameter(torch.randn(*s, device=dev)) for s in shapes]
    sp = FlatSpace(params)
    table, bucket = sp.chunk_table(), torch.zeros(sp.total, device=dev)
    grads = [torch.randn_like(p) for p in params]
    for _ in range(reps):
        ops.gather_flat(bucket, [g.data_ptr() for g in grads], table)
    torch.cuda.synchronize()


def adam(n):
    p, g, m = (torch.randn(n, device=dev) * 0.1 for _ in range(3))
    v = torch.rand(n, device=dev) * 1e-3
    for _ in range(reps):
        ops.adam_step(p, g, m, v, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-4, step=10)
    torch.cuda.synchronize()


def kd(B, C, T):
    zs, zt = torch.randn(B, C, device=dev) * 3, torch.randn(B, C, device=dev) * 3
    gk, lo = torch.empty_like(zs), torch.empty(2, device=dev)
    for _ in range(reps):
        ops.distill_loss(zs, zt, None, kind=ops.LOSS_KD, temperature=T, grad_kd=gk, loss_out=lo)
    torch.cuda.synchronize()


loss(65536, 1000, 1.0, 0.5, ops.CE_NONE)
loss(65536, 1000, 1.0, 0.5, ops.CE_SPLIT)
loss(65536, 100, 4.0, 8.0, ops.CE_NONE)       # lean staged kernel <4,7,100>
kd(65536, 100, 4.0)
loss(256, 1000, 1.0, 0.5, ops.CE_SPLIT)
loss(16384, 1000, 1.0, 0.5, ops.CE_NONE)      # lean staged kernel <32,8,1000>
dot(25_557_032)
dot(11_689_512)
adam(25_557_032)
import torchvision.models as tvm
gather([tuple(p.shape) for p in tvm.resnet18(weights=None).parameters()])
print("ok")
