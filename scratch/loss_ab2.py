"""A/B: lean kernel with fixed stride (MDB200_LOSS_LEAN=1) vs dynamic tiles (=2); cold-L2 single launches + parity check."""
import os, sys, json, statistics, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    import oracle
    from mdistiller_ddp_b200 import ops
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    drain = torch.empty(64 << 20, dtype=torch.float32, device=dev)
    def timed(fn, reps=30):
        ts = []
        for i in range(reps + 3):
            flush.fill_(i & 0xFF); drain.sum()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); b.synchronize()
            if i >= 3: ts.append(a.elapsed_time(b) * 1e3)
        return round(statistics.median(ts), 2), round(min(ts), 2)
    res = {}
    # parity first (small enough for the CPU oracle), incl. ragged row counts
    for (B, C, T, beta) in ((2500, 100, 4.0, 8.0), (777, 1000, 1.0, 0.5), (40000, 100, 4.0, 8.0)):
        g = torch.Generator().manual_seed(B)
        zs, zt = torch.randn(B, C, generator=g) * 3, torch.randn(B, C, generator=g) * 3
        y = torch.randint(0, C, (B,), generator=g)
        l_ref, g_ref = oracle.dkd_closed_form(zs.numpy(), zt.numpy(), y.numpy(), 1.0, beta, T)
        for rep in range(3):
            lp, gk, gc = ops.distill_loss(zs.cuda(), zt.cuda(), y.cuda(), kind=ops.LOSS_DKD, alpha=1.0, beta=beta, temperature=T, ce_mode=ops.CE_SPLIT if T == 1.0 else ops.CE_NONE)
            assert abs(lp[0].item() - l_ref) <= 1e-5 * abs(l_ref), (B, C, rep, lp[0].item(), l_ref)
            assert oracle.max_rel_err(gk, g_ref) <= 1e-5, (B, C, rep)
    res["parity"] = "ok"
    for (B, C, T, beta, ce, kind, tag) in ((65536, 100, 4.0, 8.0, ops.CE_NONE, ops.LOSS_DKD, "dkd_65536x100"), (65536, 100, 4.0, 8.0, ops.CE_NONE, ops.LOSS_KD, "kd_65536x100"),
                                           (65536, 100, 4.0, 8.0, ops.CE_SPLIT, ops.LOSS_KD, "kd+ce_65536x100"), (16384, 100, 4.0, 8.0, ops.CE_NONE, ops.LOSS_DKD, "dkd_16384x100"),
                                           (16384, 1000, 1.0, 0.5, ops.CE_NONE, ops.LOSS_DKD, "dkd_16384x1000"), (32768, 1000, 1.0, 0.5, ops.CE_NONE, ops.LOSS_DKD, "dkd_32768x1000"),
                                           (4096, 1000, 1.0, 0.5, ops.CE_NONE, ops.LOSS_DKD, "dkd_4096x1000"), (256, 1000, 1.0, 0.5, ops.CE_SPLIT, ops.LOSS_DKD, "dkd+ce_256x1000")):
        zs, zt = torch.randn(B, C, device=dev) * 3, torch.randn(B, C, device=dev) * 3
        y = torch.randint(0, C, (B,), device=dev)
        gk, gc, lo = torch.empty_like(zs), torch.empty_like(zs), torch.empty(2, device=dev)
        res[tag] = timed(lambda: ops.distill_loss(zs, zt, y, kind=kind, alpha=1.0, beta=beta, temperature=T, ce_mode=ce, grad_kd=gk, grad_ce=gc, loss_out=lo))
        del zs, zt, gk, gc
    print(json.dumps(res))
else:
    for name, env in (("lean_fixed", {"MDB200_LOSS_LEAN": "1"}), ("lean_dynamic", {"MDB200_LOSS_LEAN": "2"}), ("lean_fixed_again", {"MDB200_LOSS_LEAN": "1"}), ("lean_dynamic_again", {"MDB200_LOSS_LEAN": "2"})):
        out = subprocess.run([sys.executable, __file__, "child"], env=dict(os.environ, **env), capture_output=True, text=True)
        line = [l for l in out.stdout.splitlines() if l.startswith("{")]
        print(name, line[0] if line else out.stderr[-1500:])
