"""Host-side cost of the tail per step (CIFAR KD+DOT, B=64): where the Python time goes."""
import os, sys, time, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mdistiller_ddp_b200 as b200
from mdistiller_ddp_b200 import workloads as W, ops
torch.backends.cudnn.benchmark = True
wl = sys.argv[1] if len(sys.argv) > 1 else "cifar_res32x4_res8x4_kd_dot"
teacher, student, hw, classes = W.build_pair(wl)
cfg = W.default_cfg(wl)
d = b200.distiller_dict[cfg.DISTILLER.TYPE](student, teacher, cfg).cuda().train()
tr = b200.StandaloneTrainer(d, cfg)
B = W.WORKLOADS[wl][7]
image, target, index = W.synthetic_batch(B, hw, classes, seed=0, device="cuda")
acc = collections.defaultdict(float)
def wrap(obj, name, key):
    f = getattr(obj, name)
    def g(*a, **k):
        t = time.perf_counter(); r = f(*a, **k); acc[key] += time.perf_counter() - t; return r
    setattr(obj, name, g)
opt = tr.optimizer
wrap(opt, "zero_grad", "zero_grad"); wrap(opt, "step_kd", "step_kd"); wrap(opt, "step", "step")
wrap(ops, "distill_loss", "loss_call")
for _ in range(20): tr.tail_step(image, target, 1)
torch.cuda.synchronize(); acc.clear()
N = 100
t0 = time.perf_counter()
for _ in range(N): tr.tail_step(image, target, 1)
t_host = time.perf_counter() - t0
torch.cuda.synchronize(); t_all = time.perf_counter() - t0
print(f"{wl}: {1e3*t_all/N:.3f} ms/step wall, host issue time {1e3*t_host/N:.3f} ms/step")
for k, v in acc.items(): print(f"  {k:10s} {1e6*v/N:8.1f} us/step (host)")
meters = b200.new_train_meters()
t0 = time.perf_counter()
for _ in range(N): tr.report(torch.randn(B, classes, device="cuda"), target, {"loss_ce": torch.zeros((), device="cuda"), "loss_kd": torch.zeros((), device="cuda")}, 1, meters)
print(f"  report()   {1e6*(time.perf_counter()-t0)/N:8.1f} us/step (host, includes its sync)")
