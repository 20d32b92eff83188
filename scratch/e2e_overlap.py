import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mdistiller_ddp_b200 as b200
from mdistiller_ddp_b200 import workloads as W
torch.backends.cudnn.benchmark = True
wl = "imagenet_r34_r18_dkd_dot"
teacher, student, hw, classes = W.build_pair(wl); cfg = W.default_cfg(wl)
tr = b200.StandaloneTrainer(b200.DKD(student, teacher, cfg).cuda().train(), cfg)
host = W.synthetic_batch(256, hw, classes, seed=1, pin=True)
dev = tuple(t.cuda() for t in host)
m = b200.new_train_meters()
def run(fn, n=8):
    for _ in range(3): fn()
    tr.flush_metrics(m); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    tr.flush_metrics(m); torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / n
print("device-resident tail_step     ms/step", round(run(lambda: tr.tail_step(dev[0], dev[1], 1)), 2))
tr.metrics_lag = 0
print("train_iter lag0               ms/step", round(run(lambda: tr.train_iter(host, 1, m)), 2))
tr.metrics_lag = 1
print("train_iter lag1               ms/step", round(run(lambda: tr.train_iter(host, 1, m)), 2))
tr.metrics_lag = 3
print("train_iter lag3               ms/step", round(run(lambda: tr.train_iter(host, 1, m)), 2))
# pure copy time
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): x = host[0].cuda(non_blocking=True)
torch.cuda.synchronize(); print("H2D 154MB ms", round(1e3 * (time.perf_counter() - t0) / 5, 2))
# tail_step + report only (no H2D)
tr.metrics_lag = 1
def no_h2d():
    p, l = tr.tail_step(dev[0], dev[1], 1); tr.report(p, dev[1], l, 1, m)
print("tail_step+report lag1 (no H2D) ms/step", round(run(no_h2d), 2))
