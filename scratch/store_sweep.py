import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mdistiller_ddp_b200 import ops
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev); drain = torch.empty(64 << 20, dtype=torch.float32, device=dev)
def timed(fn, reps=15):
    ts = []
    for i in range(reps + 3):
        flush.fill_(i & 0xFF); drain.sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        if i >= 3: ts.append(a.elapsed_time(b) * 1e3)
    return statistics.median(ts), min(ts)
tag = os.environ.get("MDB200_LIB", "default").split("/")[-1]
B, C = 65536, 1000
zs, zt = torch.randn(B, C, device=dev) * 3, torch.randn(B, C, device=dev) * 3
y = torch.randint(0, C, (B,), device=dev)
gk, gc, lo = torch.empty_like(zs), torch.empty_like(zs), torch.empty(2, device=dev)
for ce in (0, 1):
    med, best = timed(lambda: ops.distill_loss(zs, zt, y, kind=ops.LOSS_DKD, alpha=1.0, beta=0.5, temperature=1.0, ce_mode=ce, grad_kd=gk, grad_ce=gc, loss_out=lo))
    print(f"{tag:12s} loss ce={ce}: {med:7.1f} us (best {best:6.1f})")
del zs, zt, gk, gc
for n in (11_689_512, 25_557_032):
    p, gt, g2 = (torch.randn(n, device=dev) * 0.1 for _ in range(3)); vt, vk = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    med, best = timed(lambda: ops.dot_step(p, gt, g2, vt, vk, lr=0.2, momentum=0.81, momentum_kd=0.99, weight_decay=1e-4))
    print(f"{tag:12s} dot n={n}: {med:7.1f} us (best {best:6.1f})")
