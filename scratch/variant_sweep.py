import os, sys, statistics, subprocess, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mdistiller_ddp_b200 import ops
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev); drain = torch.empty(64 << 20, dtype=torch.float32, device=dev)
def timed(fn, reps=12):
    ts = []
    for i in range(reps + 3):
        flush.fill_(i & 0xFF); drain.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        if i >= 3: ts.append(a.elapsed_time(b) * 1e3)
    return statistics.median(ts), min(ts)
for (B, C, T, beta, ce) in [(65536, 100, 4.0, 8.0, 0), (65536, 100, 4.0, 8.0, 1), (16384, 100, 4.0, 8.0, 0), (65536, 1000, 1.0, 0.5, 0), (16384, 1000, 1.0, 0.5, 0), (4096, 1000, 1.0, 0.5, 1)]:
    zs, zt = torch.randn(B, C, device=dev) * 3, torch.randn(B, C, device=dev) * 3
    y = torch.randint(0, C, (B,), device=dev)
    gk, gc, lo = torch.empty_like(zs), torch.empty_like(zs), torch.empty(2, device=dev)
    med, best = timed(lambda: ops.distill_loss(zs, zt, y, kind=ops.LOSS_DKD, alpha=1.0, beta=beta, temperature=T, ce_mode=ce, grad_kd=gk, grad_ce=gc, loss_out=lo))
    nb = (16 if ce else 12) * B * C
    print(f"variant={os.environ.get('MDB200_LOSS_STAGED','0')} {B}x{C} ce={ce}: {med:7.1f} us (best {best:6.1f})  {nb/med/1e3:7.0f} GB/s", flush=True)
