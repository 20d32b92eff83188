"""A compact pass over every kernel variant with ragged shapes -- run under `compute-sanitizer --tool memcheck`
(or racecheck); no timing, small sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mdistiller_ddp_b200 import ops
from mdistiller_ddp_b200.flat import FlatSpace
dev = "cuda"
torch.manual_seed(0)
n_launch = 0
for C in (8, 37, 100, 112, 128, 200, 512, 1000, 2048, 4099):
    for B in (1, 5, 67):
        zs, zt = torch.randn(B, C, device=dev) * 3, torch.randn(B, C, device=dev) * 3
        y = torch.randint(0, C, (B,), device=dev)
        for kind in (ops.LOSS_DKD, ops.LOSS_KD):
            for T in (1.0, 4.0):
                for ce in (ops.CE_NONE, ops.CE_SPLIT, ops.CE_COMBINED):
                    lp, g, gc = ops.distill_loss(zs, zt, y, kind=kind, alpha=1.0, beta=2.0, temperature=T, ce_mode=ce)
                    n_launch += 1
    assert torch.isfinite(lp).all() and torch.isfinite(g).all()
# strided / misaligned logits
w = torch.randn(9, 140, device=dev)
for lo in (0, 1, 4):
    ops.distill_loss(w[:, lo:lo + 100], w[:, lo + 8:lo + 108], torch.randint(0, 100, (9,), device=dev), kind=ops.LOSS_DKD,
                     alpha=1.0, beta=8.0, temperature=4.0, ce_mode=ops.CE_SPLIT)
for n in (1, 3, 4, 5, 1023, 4099, 100_003):
    p, gt, gk, vt, vk = (torch.randn(n + 1, device=dev) for _ in range(5))
    for off in (0, 1):
        a = slice(off, off + n)
        for mem in (ops.DOT_BOTH, ops.DOT_TASK_ONLY, ops.DOT_KD_ONLY):
            for first in (False, True):
                ops.dot_step(p[a], gt[a], gk[a], vt[a], vk[a], lr=0.1, momentum=0.8, momentum_kd=0.9, weight_decay=1e-4,
                             first_task=first, first_kd=first, membership=mem)
        ops.dot_step(p[a], gt[a], gk[a], None, vk[a], lr=0.1, momentum=0.0, momentum_kd=0.9)
        ops.sgd_step(p[a], gt[a], vt[a], lr=0.1, momentum=0.9, weight_decay=1e-4)
        ops.sgd_step(p[a], gt[a], None, lr=0.1, momentum=0.0)
        vk.abs_()
        ops.adam_step(p[a], gt[a], vt[a], vk[a], lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True, step=3)
        s = ops.sum_squares(gt[a], scale=0.5)
        c = ops.clip_coefficient(s, 1.0)
        ops.sgd_step(p[a], gt[a], vt[a], lr=0.1, momentum=0.9, grad_scale_dev=c)
        ops.scale_by_device_scalar(gt[a].contiguous(), c[0])
params = [torch.nn.Parameter(torch.randn(*s, device=dev)) for s in [(7, 3), (5,), (1,), (33000,), (4, 4, 3, 3)]]
sp = FlatSpace(params)
bucket = torch.zeros(sp.total, device=dev)
grads = [torch.randn_like(p) for p in params]; grads[1] = None
ops.gather_flat(bucket, [0 if g is None else g.data_ptr() for g in grads], sp.chunk_table(chunk_elems=1000))
torch.cuda.synchronize()
print("sanitize pass ok, loss launches", n_launch)
