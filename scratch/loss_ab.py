"""A/B timing of the loss kernel variants on the sweep shapes (cold L2), one process per variant via env."""
import os, sys, json, statistics, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    from mdistiller_ddp_b200 import ops
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    drain = torch.empty(64 << 20, dtype=torch.float32, device=dev)
    def timed(fn, reps=20):
        ts = []
        for i in range(reps + 3):
            flush.fill_(i & 0xFF); drain.sum()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); b.synchronize()
            if i >= 3: ts.append(a.elapsed_time(b) * 1e3)
        return statistics.median(ts), min(ts)
    res = {}
    # launch floor: a 1-thread kernel timed the same way
    ss = torch.zeros(1, dtype=torch.float64, device=dev)
    res["floor_1thread_kernel"] = timed(lambda: ops.clip_coefficient(ss, 1.0))
    for (B, C, T, beta, ce) in ((65536, 100, 4.0, 8.0, ops.CE_NONE), (65536, 100, 4.0, 8.0, ops.CE_SPLIT), (16384, 100, 4.0, 8.0, ops.CE_NONE),
                                (4096, 100, 4.0, 8.0, ops.CE_NONE), (64, 100, 4.0, 8.0, ops.CE_SPLIT),
                                (16384, 1000, 1.0, 0.5, ops.CE_NONE), (4096, 1000, 1.0, 0.5, ops.CE_NONE), (256, 1000, 1.0, 0.5, ops.CE_SPLIT),
                                (32768, 1000, 1.0, 0.5, ops.CE_NONE), (65536, 1000, 1.0, 0.5, ops.CE_NONE)):
        zs, zt = torch.randn(B, C, device=dev) * 3, torch.randn(B, C, device=dev) * 3
        y = torch.randint(0, C, (B,), device=dev)
        gk, gc, lo = torch.empty_like(zs), torch.empty_like(zs), torch.empty(2, device=dev)
        for kind, kn in ((ops.LOSS_DKD, "dkd"), (ops.LOSS_KD, "kd")):
            if kn == "kd" and (B, C) not in ((65536, 100), (16384, 1000)): continue
            med, best = timed(lambda: ops.distill_loss(zs, zt, y, kind=kind, alpha=1.0, beta=beta, temperature=T, ce_mode=ce, grad_kd=gk, grad_ce=gc, loss_out=lo))
            nbytes = (16 if ce == ops.CE_SPLIT else 12) * B * C + 8 * B + 8
            res[f"{kn}{'+ce' if ce else ''}_{B}x{C}"] = (round(med, 2), round(best, 2), round(nbytes / med / 8e6, 3))
        del zs, zt, gk, gc
    print(json.dumps(res))
else:
    for name, env in (("r1_staged", {"MDB200_LOSS_LEAN": "0"}), ("lean_4x7_b2", {"MDB200_LOSS_LEAN": "1"}), ("lean_4x7_b3", {"MDB200_LOSS_LEAN": "2"}),
                      ("lean_8x4_b4", {"MDB200_LOSS_LEAN": "3"}), ("lean_8x4_b3", {"MDB200_LOSS_LEAN": "4"})):
        out = subprocess.run([sys.executable, __file__, "child"], env=dict(os.environ, **env), capture_output=True, text=True)
        line = [l for l in out.stdout.splitlines() if l.startswith("{")]
        print(name, line[0] if line else out.stderr[-800:])
