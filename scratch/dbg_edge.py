import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, 'tests')
import numpy as np, torch, _cases
from mdistiller_ddp_b200 import ops
G = np.load('tests/golden/loss_golden.npz')
for c in _cases.LOSS_CASES:
    zs, zt, y = [t.cuda() for t in _cases.make_logits(c)]
    for kind, key in ((ops.LOSS_DKD, 'dkd'), (ops.LOSS_KD, 'kd')):
        lp, g, gc = ops.distill_loss(zs, zt, y, kind=kind, alpha=c.alpha, beta=c.beta, temperature=c.T, ce_mode=ops.CE_SPLIT)
        r64, r32 = float(G[f'{c.name}/{key}64']), float(G[f'{c.name}/{key}32'])
        ce64 = float(G[f'{c.name}/ce64'])
        print(f"{c.name:22s} {key:3s} ours={lp[0].item():.9g} r64={r64:.9g} rel={abs(lp[0].item()-r64)/abs(r64):.2e} ref32rel={abs(r32-r64)/abs(r64):.2e} | ce rel={abs(lp[1].item()-ce64)/ce64:.2e}")
