"""Where do the patched and the reference run of tools/train.py differ?  (GPU box; prints per-tensor errors and the
run-to-run noise floor of the reference itself.)"""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import oracle
from oracle import ref_harness

def errs(a, b, min_dim=0):
    out = []
    for k, v in b["student"].items():
        if v.dtype.is_floating_point and v.dim() >= min_dim:
            out.append((oracle.max_rel_err(a["student"][k], v), k, float(v.abs().max())))
    return sorted(out, reverse=True)

yaml_path = sys.argv[1] if len(sys.argv) > 1 else "configs/cifar100/dot/res32x4_res8x4.yaml"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
tmp = tempfile.mkdtemp()
runs = {}
for name, patched in (("ref1", False), ("ref2", False), ("ours1", True), ("ours2", True)):
    runs[name] = ref_harness.run_train(yaml_path, patched=patched, out_dir=os.path.join(tmp, name), iters=iters, batch=64)
for a, b in (("ref2", "ref1"), ("ours2", "ours1"), ("ours1", "ref1")):
    e = errs(runs[a], runs[b])
    print(f"== {a} vs {b}: worst {e[0][0]:.3e}")
    for x in e[:4]:
        print("   %.3e  %-40s max|ref|=%.3e" % x)
    e2 = errs(runs[a], runs[b], 2)
    print("   weights (dim>=2): worst %.3e %s; median %.3e" % (e2[0][0], e2[0][1], e2[len(e2) // 2][0]))
    e1 = [x for x in e if "running" in x[1]]
    print("   running stats: worst %.3e %s" % (e1[0][0], e1[0][1]))
