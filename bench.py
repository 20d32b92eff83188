#!/usr/bin/env python
"""Benchmark of the distillation tail: train images/sec + kernel HBM rooflines.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

Contract (one JSON line on rank 0):
  metric  train images/sec on BASELINE.json's headline config (DKD+DOT ResNet34->ResNet18,
          synthetic ImageNet 224x224 / 1000 classes, 256 images per GPU), whole job over N GPUs.
  value   inputs already resident in HBM; K steps between barrier+synchronize, CUDA events,
          max over ranks.
  e2e     the same steps through the public call (`trainer.train_iter(batch)`): pinned host batch
          -> H2D inside the timed region, loss/top-k metrics read back D2H every step.
  roofline  the DOT update kernel timed live inside the timed region (CUDA events on its stream),
          algorithmic bytes 32*N_shard per launch vs the measured HBM copy peak.
  kernels   isolated large-shape sweep of both kernels (L2 flushed between launches).
  cpu_baseline  the reference's own path on the host cores, bounded sample of the same config: the reference's
          distiller / optimizer classes and backbones from the staged tree (oracle/make_ref.sh; kind "reference"),
          or the oracle restatement when none is staged (kind "port").
  dp_parity  (N > 1) the shipped data-parallel step against the oracle on rank-averaged gradients, before the
          timed region; the run exits non-zero above rel 1e-5.
`--impl reference` prints the CPU arm alone (rank 0; other ranks exit 0), same workload, batch, steps and warm-up.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "train_images_per_sec"
UNIT = "images/s"
FALLBACK_HBM_GBS = 6650.0


def workload_label(name: str) -> str:
    """The same `config.workload` string in both arms (ours and --impl reference)."""
    labels = {
        "imagenet_r34_r18_dkd_dot": "DKD+DOT ResNet34->ResNet18, synthetic ImageNet 224x224x3, 1000 classes, random-init",
        "imagenet_r34_r18_kd_dot": "KD+DOT ResNet34->ResNet18, synthetic ImageNet 224x224x3, 1000 classes, random-init",
        "cifar_res32x4_res8x4_kd_dot": "KD+DOT ResNet32x4->ResNet8x4, synthetic CIFAR-100 32x32x3, 100 classes, random-init",
        "cifar_res32x4_res8x4_dkd_base": "DKD (base trainer, SGD) ResNet32x4->ResNet8x4, synthetic CIFAR-100 32x32x3, random-init",
        "cifar_wrn40_2_wrn16_2_dkd_base": "DKD (base trainer, SGD) WRN-40-2->WRN-16-2, synthetic CIFAR-100 32x32x3, random-init",
        "cifar_vgg13_vgg8_dkd_base": "DKD (base trainer, SGD) VGG13->VGG8, synthetic CIFAR-100 32x32x3, random-init",
    }
    return f"{name}: {labels.get(name, name)}"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="imagenet_r34_r18_dkd_dot")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (0 = the workload's)")
    ap.add_argument("--cpu-batch", type=int, default=0,
                    help="images per CPU step of the reference arm (0 = the workload's per-GPU batch, i.e. the same config)")
    ap.add_argument("--comm", default=None,
                    choices=["auto", "nccl", "fused", "fused_kd", "fused_p2p", "fused_kd_p2p"],
                    help="data-parallel step (N > 1); default = whatever the package ships with "
                         "(TailStepMixin.comm_mode, normally 'auto': trainer.resolve_comm_mode, NCCL launches if "
                         "NVLink symmetric memory cannot be set up)")
    ap.add_argument("--sweep", default="grid", choices=["grid", "short"],
                    help="kernel sweep: the full SURVEY 8(d) grid or the eight headline shapes")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


def profiled_traffic(kernel: str, n_elements: int):
    """dram bytes per launch from the committed ncu capture -- only if it was taken at this launch size."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            rec = json.load(f).get(kernel)
        return rec["dram_bytes"] if rec and rec.get("n_elements") == n_elements else None
    except Exception:
        return None


# ------------------------------------------------------------------------------- CPU arm

def _reference_step(workload: str, batch: int, device: str):
    """The reference's OWN objects on its OWN backbones (staged tree, see oracle/make_ref.sh): `DKD` / `KD`
    distiller (distillers/DKD.py:54-84, KD.py:16-39), `DistillationOrientedTrainer` (engine/dot.py:58) or
    `optim.SGD` (trainer.py:86-91), driven in the order of DOT.train_iter (trainer.py:412-444) /
    BaseTrainer.train_iter (:257-321) -- minus their hard-coded `.cuda()` calls, which is why the trainer classes
    themselves cannot run on the host."""
    from oracle import ref_loader
    from mdistiller_ddp_b200 import workloads as W
    DKD, KD = ref_loader.distiller_classes()
    teacher, student, hw, classes = W.build_pair(workload, seed=0, reference="require")
    cfg = W.default_cfg(workload)
    distiller = (DKD if cfg.DISTILLER.TYPE == "DKD" else KD)(student, teacher, cfg).to(device).train()
    params = distiller.get_learnable_parameters()
    m, d = cfg.SOLVER.SGD.MOMENTUM, cfg.SOLVER.DOT.DELTA
    dot = cfg.SOLVER.TRAINER == "dot"
    if dot:
        opt = ref_loader.dot_module().DistillationOrientedTrainer(params, lr=cfg.SOLVER.LR, momentum=m - d,
                                                                  momentum_kd=m + d, weight_decay=cfg.SOLVER.WEIGHT_DECAY)
    else:
        opt = torch.optim.SGD(params, lr=cfg.SOLVER.LR, momentum=m, weight_decay=cfg.SOLVER.WEIGHT_DECAY)
    image, target, _ = W.synthetic_batch(batch, hw, classes, seed=1234, device=device)

    def step():
        preds, losses = distiller(image=image, target=target, epoch=1)
        if dot:
            opt.zero_grad(set_to_none=True)
            losses["loss_kd"].backward(retain_graph=True)
            opt.step_kd()
            opt.zero_grad(set_to_none=True)
            losses["loss_ce"].backward()
            opt.step()
        else:
            opt.zero_grad()
            sum(l.mean() for l in losses.values()).backward()
            opt.step()
        return sum(float(v.detach()) for v in losses.values())     # the reference reads its losses back every iteration too
    return step, hw


def _port_step(workload: str, batch: int, device: str):
    """Fallback when no reference tree is staged: the oracle's restatement of the same sequence on stand-in nets."""
    import oracle
    from mdistiller_ddp_b200 import workloads as W
    teacher, student, hw, classes = W.build_pair(workload, seed=0, reference="never")
    teacher.to(device).eval()
    student.to(device).train()
    cfg = W.default_cfg(workload)
    params = [p for _, p in student.named_parameters()]
    m, d = cfg.SOLVER.SGD.MOMENTUM, cfg.SOLVER.DOT.DELTA
    opt = oracle.DotListOracle([p.data for p in params], lr=cfg.SOLVER.LR, momentum=m - d, momentum_kd=m + d,
                               weight_decay=cfg.SOLVER.WEIGHT_DECAY)
    kind = cfg.DISTILLER.TYPE
    base_trainer = cfg.SOLVER.TRAINER == "base"
    sgd = torch.optim.SGD(params, lr=cfg.SOLVER.LR, momentum=m, weight_decay=cfg.SOLVER.WEIGHT_DECAY)
    image, target, _ = W.synthetic_batch(batch, hw, classes, seed=1234, device=device)

    def step():
        zs, _ = student(image)
        with torch.no_grad():
            zt, _ = teacher(image)
        if kind == "DKD":
            l_ce = cfg.DKD.CE_WEIGHT * oracle.ce_loss(zs, target)
            l_kd = min(1 / cfg.DKD.WARMUP, 1.0) * oracle.dkd_loss(zs, zt, target, cfg.DKD.ALPHA, cfg.DKD.BETA, cfg.DKD.T)
        else:
            l_ce = cfg.KD.LOSS.CE_WEIGHT * oracle.ce_loss(zs, target)
            l_kd = cfg.KD.LOSS.KD_WEIGHT * oracle.kd_loss(zs, zt, cfg.KD.TEMPERATURE)
        if base_trainer:
            sgd.zero_grad()
            (l_kd + l_ce).backward()
            sgd.step()
        else:
            g_kd = torch.autograd.grad(l_kd, params, retain_graph=True)
            opt.step_kd(list(g_kd))
            g_ce = torch.autograd.grad(l_ce, params)
            opt.step(list(g_ce))
        return l_kd.item() + l_ce.item()
    return step, hw


def cpu_arm(workload: str, batch: int, steps: int, warmup: int, device: str = "cpu"):
    """The reference's per-step path -- backbones forward, dkd/kd + CE loss, two backwards, per-tensor DOT loop --
    timed on `device`.  device="cpu" is the CPU baseline / --impl reference arm on all host cores; a CUDA device
    gives the "reference's torch code on the same GPU" comparator (never the product path).  kind "reference" =
    the reference's own classes from the staged tree; "port" = the oracle restatement when none is staged."""
    from mdistiller_ddp_b200 import refhost
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if refhost.available():
        step, hw = _reference_step(workload, batch, device)
        kind, what = "reference", "the reference's own distiller / optimizer classes and backbones (staged tree)"
    else:
        step, hw = _port_step(workload, batch, device)
        kind, what = "port", "oracle restatement on stand-in nets (no reference tree staged)"
    on_gpu = str(device) != "cpu"
    for _ in range(warmup):
        step()
    if on_gpu:
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    if on_gpu:
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"value": batch * steps / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{steps} steps x {batch} images of {workload} ({hw}x{hw} fp32; {what}) on "
                      f"{'the host CPU' if not on_gpu else device}, torch {torch.__version__} intra-op threads={cores}",
            "ms_per_step": 1e3 * dt / steps}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from mdistiller_ddp_b200 import workloads as W
    steps, warmup = max(1, args.steps), max(0, args.warmup)        # exactly what was asked for, like our arm
    batch = args.cpu_batch or args.batch or W.WORKLOADS[args.workload][7]
    r = cpu_arm(args.workload, batch, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_label(args.workload), "per_gpu_batch": batch, "per_step_images": batch,
                   "device": "host CPU, all cores; " + r["sample"]},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------- clocks

class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()

    def summary(self, t0: float, t1: float):
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 9] or [r for _, r in self.rows if len(r) >= 9]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = [float(r[1]) for r in rows]
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": float(rows[0][2]), "reasons": sorted(reasons),
                "samples": len(rows), "power_w_max": max(float(r[3]) for r in rows)}


# ------------------------------------------------------------------------------- GPU arm

def kernel_sweep(dev, grid=True):
    """Both kernels alone at BASELINE.json's sweep sizes; L2 flushed between launches (see `timed`).  Shapes below
    a few MB are launch-latency bound: read their `us` (one launch each), not their GB/s."""
    from mdistiller_ddp_b200 import ops
    peak, _ = hbm_peak()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    drain = torch.empty(64 << 20, dtype=torch.float32, device=dev)
    out = {}

    def floor_us():
        """What the cold single-launch protocol below reads for a ONE-THREAD kernel (mdb200_clip_coefficient): the
        two dependent-launch gaps around it plus event resolution -- a constant that every `us` in this table carries."""
        ss, cf = torch.zeros(1, dtype=torch.float64, device=dev), torch.zeros(1, device=dev)
        return timed(lambda: ops.clip_coefficient(ss, 1.0, out=cf), reps=20)[0] * 1e6

    def streamed(make_call, nbytes):
        """Sustained per-launch time: launches back to back inside ONE event pair, each on its own buffer set, the
        sets together > 3x the 126 MB L2 (inputs larger than L2 -- nothing a launch touches is cache-resident),
        so neither the event/launch floor nor L2 hits flatter the figure."""
        n_sets = max(2, min(12, -(-3 * (126 << 20) // nbytes)))
        calls = [make_call() for _ in range(n_sets)]
        for c in calls:
            c()
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _rep in range(2):
                for c in calls:
                    c()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b) * 1e-3 / (2 * n_sets))
        return statistics.median(ts), n_sets

    def timed(fn, reps=10):
        ts = []
        for i in range(reps + 2):
            # cold L2: overwrite it (256 MB of writes), then stream 256 MB of reads through it so the
            # dirty lines are written back BEFORE the timed launch instead of during it
            flush.fill_(i & 0xFF)
            drain.sum()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            if i >= 2:
                ts.append(a.elapsed_time(b) * 1e-3)
        return statistics.median(ts), min(ts)

    head = [
        (65536, 1000, ops.LOSS_DKD, 1.0, 0.5, ops.CE_NONE, "dkd_65536x1000"),
        (65536, 1000, ops.LOSS_DKD, 1.0, 0.5, ops.CE_SPLIT, "dkd+ce_split_65536x1000"),
        (65536, 1000, ops.LOSS_KD, 1.0, 1.0, ops.CE_NONE, "kd_65536x1000"),
        (16384, 1000, ops.LOSS_DKD, 1.0, 0.5, ops.CE_NONE, "dkd_16384x1000"),
        (65536, 100, ops.LOSS_DKD, 4.0, 8.0, ops.CE_NONE, "dkd_65536x100"),
        (65536, 100, ops.LOSS_DKD, 4.0, 8.0, ops.CE_SPLIT, "dkd+ce_split_65536x100_T4"),
        (256, 1000, ops.LOSS_DKD, 1.0, 0.5, ops.CE_SPLIT, "dkd+ce_split_256x1000"),
        (64, 100, ops.LOSS_DKD, 4.0, 8.0, ops.CE_SPLIT, "dkd+ce_split_64x100_T4")]
    shapes = list(head)
    if grid:
        # B in {64 ... 65536} x C in {100, 1000}; DKD hyper-sets (alpha, beta, T) = (1,8,4), (1,6,4),
        # (1,0.5,1); KD at T in {4, 1}.  Tags: <kind>_<B>x<C>_b<beta>_T<T>.
        seen = {(B, C, k, T, be, ce) for (B, C, k, T, be, ce, _) in head}
        for C in (100, 1000):
            for B in (64, 256, 1024, 4096, 16384, 65536):
                for (beta, T) in ((8.0, 4.0), (6.0, 4.0), (0.5, 1.0)):
                    if (B, C, ops.LOSS_DKD, T, beta, ops.CE_NONE) not in seen:
                        shapes.append((B, C, ops.LOSS_DKD, T, beta, ops.CE_NONE, f"grid_dkd_{B}x{C}_b{beta:g}_T{T:g}"))
                for T in (4.0, 1.0):
                    if (B, C, ops.LOSS_KD, T, 1.0, ops.CE_NONE) not in seen:
                        shapes.append((B, C, ops.LOSS_KD, T, 1.0, ops.CE_NONE, f"grid_kd_{B}x{C}_T{T:g}"))
    bufs = {}
    for (B, C, kind, T, beta, ce, tag) in shapes:
        if (B, C) not in bufs:
            bufs.clear()
            g = torch.Generator(device=dev).manual_seed(0)
            bufs[(B, C)] = (torch.randn(B, C, device=dev, generator=g) * 3, torch.randn(B, C, device=dev, generator=g) * 3,
                            torch.randint(0, C, (B,), device=dev, generator=g),
                            torch.empty(B, C, device=dev), torch.empty(B, C, device=dev), torch.empty(2, device=dev))
        zs, zt, y, gk, gc, lo = bufs[(B, C)]
        med, best = timed(lambda: ops.distill_loss(zs, zt, y, kind=kind, alpha=1.0, beta=beta, temperature=T,
                                                   ce_mode=ce, grad_kd=gk, grad_ce=gc, loss_out=lo),
                          reps=10 if not tag.startswith("grid_") else 6)
        nbytes = (16 if ce == ops.CE_SPLIT else 12) * B * C + 8 * B + 8
        out[tag] = {"us": med * 1e6, "GB/s": nbytes / med / 1e9, "frac_of_measured_peak": nbytes / med / 1e9 / peak,
                    "frac_of_8TBs": nbytes / med / 8e12, "bytes": nbytes, "launches": 1}
        if not tag.startswith("grid_") and nbytes >= (16 << 20):
            def make_call(B=B, C=C, kind=kind, T=T, beta=beta, ce=ce):
                g = torch.Generator(device=dev).manual_seed(1)
                a = (torch.randn(B, C, device=dev, generator=g) * 3, torch.randn(B, C, device=dev, generator=g) * 3,
                     torch.randint(0, C, (B,), device=dev, generator=g), torch.empty(B, C, device=dev),
                     torch.empty(B, C, device=dev) if ce == ops.CE_SPLIT else None, torch.empty(2, device=dev))
                return lambda: ops.distill_loss(a[0], a[1], a[2], kind=kind, alpha=1.0, beta=beta, temperature=T,
                                                ce_mode=ce, grad_kd=a[3], grad_ce=a[4], loss_out=a[5])
            bufs.clear()
            st, n_sets = streamed(make_call, nbytes)
            out[tag].update({"us_stream": st * 1e6, "stream_sets": n_sets, "frac_of_measured_peak_stream": nbytes / st / 1e9 / peak,
                             "frac_of_8TBs_stream": nbytes / st / 8e12})
    bufs.clear()
    out["_method"] = {
        "us": "one cold launch between two events; before it L2 is overwritten (256 MB fill) and drained (256 MB read)",
        "us_stream": "sustained: back-to-back launches in one event pair, every launch on its own buffer set, sets "
                     "together > 3x L2 (inputs larger than L2)",
        "single_launch_floor_us": floor_us()}
    for n, tag in ((11_689_512, "dot_resnet18_11.7M"), (25_557_032, "dot_resnet50_25.6M")):
        p, gt, gk2 = (torch.randn(n, device=dev) * 0.1 for _ in range(3))
        vt, vk = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        med, best = timed(lambda: ops.dot_step(p, gt, gk2, vt, vk, lr=0.2, momentum=0.81, momentum_kd=0.99,
                                               weight_decay=1e-4))
        out[tag] = {"us": med * 1e6, "GB/s": 32 * n / med / 1e9, "frac_of_measured_peak": 32 * n / med / 1e9 / peak,
                    "frac_of_8TBs": 32 * n / med / 8e12, "bytes": 32 * n}

        def make_dot(n=n):
            a = [torch.randn(n, device=dev) * 0.1 for _ in range(5)]
            return lambda: ops.dot_step(a[0], a[1], a[2], a[3], a[4], lr=0.2, momentum=0.81, momentum_kd=0.99, weight_decay=1e-4)
        st, n_sets = streamed(make_dot, 32 * n)
        out[tag].update({"us_stream": st * 1e6, "stream_sets": n_sets, "frac_of_measured_peak_stream": 32 * n / st / 1e9 / peak,
                         "frac_of_8TBs_stream": 32 * n / st / 8e12})
        med, best = timed(lambda: ops.sgd_step(p, gt, vt, lr=0.05, momentum=0.9, weight_decay=5e-4))
        out[tag.replace("dot_", "sgd_")] = {"us": med * 1e6, "GB/s": 20 * n / med / 1e9,
                                            "frac_of_measured_peak": 20 * n / med / 1e9 / peak, "bytes": 20 * n}
        vk.fill_(1e-3)
        med, best = timed(lambda: ops.adam_step(p, gt, vt, vk, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8,
                                                weight_decay=1e-4, step=10))
        out[tag.replace("dot_", "adam_")] = {"us": med * 1e6, "GB/s": 28 * n / med / 1e9,
                                             "frac_of_measured_peak": 28 * n / med / 1e9 / peak, "bytes": 28 * n}
    return out


def torch_on_gpu_baseline(dev, workload=None, batch=0):
    """What the kernels displace ON THE SAME GPU: the reference's torch
    formulation -- the oracle restatement of dkd_loss + autograd, and the per-tensor DOT loop -- run
    on CUDA tensors.  A baseline leg like cpu_baseline: never part of the product path."""
    import oracle
    import torchvision.models as tvm
    from mdistiller_ddp_b200 import refhost
    out = {}
    if refhost.available():            # the reference's own functions (staged tree) ...
        from oracle import ref_loader
        ref_dkd, _ = ref_loader.loss_functions()
        RefDot = ref_loader.dot_module().DistillationOrientedTrainer
        out["source"] = "reference functions (dkd_loss DKD.py:8, F.cross_entropy, DistillationOrientedTrainer dot.py:58)"
    else:                              # ... or the oracle's restatement of them
        ref_dkd, RefDot = oracle.dkd_loss, None
        out["source"] = "oracle restatement (no reference tree staged)"
    import torch.nn.functional as F

    def timed(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b) * 1e-3)
        return statistics.median(ts)

    for B, C, T, beta in ((256, 1000, 1.0, 0.5), (65536, 1000, 1.0, 0.5)):
        zs = (torch.randn(B, C, device=dev) * 3).requires_grad_(True)
        zt = torch.randn(B, C, device=dev) * 3
        y = torch.randint(0, C, (B,), device=dev)

        def ref_loss():
            zs.grad = None
            (ref_dkd(zs, zt, y, 1.0, beta, T) + F.cross_entropy(zs, y)).backward()
        t = timed(ref_loss)
        out[f"torch_dkd+ce_fwd_bwd_{B}x{C}"] = {"us": t * 1e6, "algorithmic_GB/s": (12 * B * C) / t / 1e9}
        del zs, zt
    params = [p.detach().clone().to(dev) for p in tvm.resnet18(weights=None).parameters()]
    g1 = [torch.randn_like(p) * 0.1 for p in params]
    g2 = [torch.randn_like(p) * 0.1 for p in params]
    if RefDot is not None:
        params = [p.requires_grad_(True) for p in params]
        opt = RefDot(params, lr=0.2, momentum=0.81, momentum_kd=0.99, weight_decay=1e-4)

        def ref_dot():
            for p, g in zip(params, g1):
                p.grad = g
            opt.step_kd()
            for p, g in zip(params, g2):
                p.grad = g
            opt.step()
    else:
        opt = oracle.DotListOracle(params, lr=0.2, momentum=0.81, momentum_kd=0.99, weight_decay=1e-4)

        def ref_dot():
            opt.step_kd(g1)
            opt.step(g2)
    t = timed(ref_dot, reps=10)
    n = sum(p.numel() for p in params)
    out["torch_dot_step_resnet18_62_tensors"] = {"us": t * 1e6, "algorithmic_GB/s": 32 * n / t / 1e9}
    del params, opt, g1, g2
    # the whole training step in the reference's torch formulation (same nets, same batch) on this GPU
    if workload is not None:
        r = cpu_arm(workload, batch, steps=10, warmup=5, device=dev)
        out["torch_train_step"] = {"images/s": r["value"], "ms_per_step": r["ms_per_step"], "per_gpu_batch": batch,
                                   "what": r["sample"] + "; losses read back each step (as trainer.py:439-444 does)"}
    return out


def dp_parity(trainer_cls, kind, world, rank, dev):
    """Numerics of the data-parallel step AS SHIPPED (same comm plumbing the timed trainer gets), N > 1: 12 steps of
    tests/_cases.DOT_CASES[6] (a res8x4-like list of 32 tensors, 1.23 M elements) with per-rank-scaled gradients whose
    mean over ranks is the single-process gradient, against the CPU oracle fed the rank-averaged gradients
    (mdistiller/engine/dot.py:15-55,165-170 for DOT; torch.optim.SGD for the base trainer) --
    parameters AND momenta (through sync_checkpoint_state + rank 0's state_dict).  All ranks must end bit-equal."""
    import torch.distributed as dist
    import oracle
    import mdistiller_ddp_b200 as b200
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _cases
    case = _cases.DOT_CASES[6]
    host = trainer_cls.__new__(trainer_cls)            # only _make_comm is used: the package's own comm policy
    host.trainer_kind = kind
    comm = host._make_comm()
    params = [torch.nn.Parameter((p.clone() + (0.5 if rank else 0.0)).to(dev)) for p in _cases.make_params(case)]
    if kind == "dot":
        opt = b200.DistillationOrientedTrainer(params, lr=case.lr, momentum=case.momentum - case.delta,
                                               momentum_kd=case.momentum + case.delta, weight_decay=case.wd, comm=comm)
    else:
        opt = b200.FlatSGD(params, lr=case.lr, momentum=case.momentum, weight_decay=case.wd, comm=comm)
    shapes = [p.shape for p in params]
    f_kd = 1.0 + 0.25 * (rank - (world - 1) / 2)
    f_task = 1.0 - 0.5 * (rank - (world - 1) / 2)
    for step in range(12):
        if step == 6:
            opt.param_groups[0]["lr"] = case.lr * 0.1
        kd, task = _cases.make_grads(case, step, shapes)
        if kind == "dot":
            opt.zero_grad()
            for p, g in zip(params, kd):
                p.grad = (g * f_kd).to(dev)
            opt.step_kd()
        opt.zero_grad()
        for p, g in zip(params, task):
            p.grad = (g * f_task).to(dev)
        opt.step()
    flat = opt.flat_params.clone()
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    bit_equal = all(torch.equal(g, flat) for g in gathered)
    opt.sync_checkpoint_state()
    worst = 0.0
    if rank == 0:
        sd = opt.state_dict()["state"]
        ref_p = [p.clone() for p in _cases.make_params(case)]
        if kind == "dot":
            ref = oracle.DotListOracle(ref_p, lr=case.lr, momentum=case.momentum - case.delta,
                                       momentum_kd=case.momentum + case.delta, weight_decay=case.wd)
        else:
            ref_p = [torch.nn.Parameter(p) for p in ref_p]
            ref = torch.optim.SGD(ref_p, lr=case.lr, momentum=case.momentum, weight_decay=case.wd)
        for step in range(12):
            kd, task = _cases.make_grads(case, step, shapes)
            if kind == "dot":
                if step == 6:
                    ref.hp["lr"] = case.lr * 0.1
                ref.step_kd(kd)
                ref.step(task)
            else:
                if step == 6:
                    ref.param_groups[0]["lr"] = case.lr * 0.1
                for p, g in zip(ref_p, task):
                    p.grad = g.clone()
                ref.step()
        for i, (p, q) in enumerate(zip(params, ref_p)):
            worst = max(worst, oracle.max_rel_err(p.detach().cpu(), q.detach()))
            if kind == "dot":
                worst = max(worst, oracle.max_rel_err(sd[i]["momentum_buffer"].cpu(), ref.buf_task[i]),
                            oracle.max_rel_err(sd[i]["momentum_kd_buffer"].cpu(), ref.buf_kd[i]))
            else:
                worst = max(worst, oracle.max_rel_err(sd[i]["momentum_buffer"].cpu(), ref.state[q]["momentum_buffer"]))
    t = torch.tensor([worst, 0.0 if bit_equal else 1.0], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    worst, diverged = t.tolist()
    fused = getattr(opt, "last_fused", None)
    del opt, params
    return {"mode": host.comm_mode_used, "fused_step": fused, "max_rel_err": worst, "ranks_bit_equal": diverged == 0.0,
            "steps": 12, "case": case.name, "tolerance": 1e-5,
            "oracle": "single-process reference arithmetic on rank-averaged gradients (params + momenta)"}


def run_b200_arm(args):
    import torch.distributed as dist
    import mdistiller_ddp_b200 as b200
    from mdistiller_ddp_b200 import ops, workloads as W

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.backends.cudnn.benchmark = True          # as tools/train.py:9

    per_gpu = args.batch or W.WORKLOADS[args.workload][7]
    teacher, student, hw, classes = W.build_pair(args.workload, seed=0)
    cfg = W.default_cfg(args.workload)
    distiller = b200.distiller_dict[cfg.DISTILLER.TYPE](student, teacher, cfg).to(dev).train()
    if args.comm is not None:                  # otherwise: exactly what the package ships with
        b200.TailStepMixin.comm_mode = args.comm
    parity = None
    if world > 1:
        parity = dp_parity(b200.StandaloneTrainer, cfg.SOLVER.TRAINER, world, rank, dev)
        if not (parity["max_rel_err"] <= 1e-5 and parity["ranks_bit_equal"]):
            if rank == 0:
                print(json.dumps({"dp_parity": parity, "error": "data-parallel step disagrees with the oracle"}), flush=True)
            dist.destroy_process_group()
            raise SystemExit(3)
    trainer = b200.StandaloneTrainer(distiller, cfg)
    args.comm = trainer.comm_mode_used
    comm_note = ""
    if getattr(trainer, "comm_fallback_reason", None):
        comm_note = f" (fused step unavailable: {trainer.comm_fallback_reason}; using NCCL launches)"
    n_params = sum(p.numel() for p in distiller.get_learnable_parameters())
    shard_elems = trainer.optimizer._space.shard_len

    host = W.synthetic_batch(per_gpu, hw, classes, seed=1234 + rank, pin=True)
    resident = tuple(t.to(dev) for t in host)
    epoch = 1

    # live timing of our two kernels inside the timed region (events on the launching stream)
    probes = {"dot": [], "loss": []}
    probing = {"on": False}
    orig_dot, orig_loss, orig_fused, orig_sgd = ops.dot_step, ops.distill_loss, ops.dot_step_fused, ops.sgd_step

    def wrap(name, fn):
        def inner(*a, **k):
            if not probing["on"] or torch.cuda.is_current_stream_capturing():
                return fn(*a, **k)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*a, **k)
            e1.record()
            probes[name].append((e0, e1))
            return r
        return inner

    probes["opt_step"] = []
    opt = trainer.optimizer
    opt.step = wrap("opt_step", opt.step)
    ops.dot_step = wrap("dot", orig_dot)
    ops.sgd_step = wrap("dot", ops.sgd_step)
    ops.dot_step_fused = wrap("dot", orig_fused)
    ops.sgd_step_fused = wrap("dot", ops.sgd_step_fused)
    ops.distill_loss = wrap("loss", orig_loss)
    # losses.py / optim.py call through the `ops` module attribute, so the wrappers are seen

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    def device_step():
        return trainer.tail_step(resident[0], resident[1], epoch)

    for _ in range(max(args.warmup, 3)):
        device_step()
    # launch-bound workloads switch to a CUDA-graph replay of the step after a few eager steps (graphstep.py):
    # let that settle before anything is timed
    runner = trainer._graph_runner
    settle = 0
    while runner.enabled and runner.graph is None and settle < 8:
        device_step()
        settle += 1
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    probing["on"] = True
    launches0 = ops.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        device_step()
    e1.record()
    barrier()
    t_wall1 = time.perf_counter()
    probing["on"] = False
    launches = ops.launch_count - launches0
    secs = e0.elapsed_time(e1) * 1e-3
    clocks = sampler.summary(t_wall0, t_wall1)

    # end-to-end: public call with host (pinned) batch, H2D + metrics D2H inside the timed region
    # (the trainer's default metrics_lag = 1: the loss/top-k read-back of step i is awaited during step i+1,
    #  which lets the H2D copy of the next batch overlap the current step; every step's result is still read
    #  inside the timed region, the last one by flush_metrics)
    meters = b200.new_train_meters()
    for _ in range(2):
        trainer.train_iter(host, epoch, meters)
    trainer.flush_metrics(meters)
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(args.steps):
        trainer.train_iter(host, epoch, meters)
    trainer.flush_metrics(meters)
    f1.record()
    barrier()
    assert meters["top1"].count == (2 + args.steps) * per_gpu * world, "a step's metrics were not read back"
    sampler.stop()
    e2e_secs = f0.elapsed_time(f1) * 1e-3

    t = torch.tensor([secs, e2e_secs], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    secs, e2e_secs = t.tolist()

    graph_summary = dict(runner.summary(), settle_steps=settle)
    if runner.graph is not None:
        # the timed steps were graph replays, whose nodes cannot carry event probes: time our kernels live in a few
        # EAGER steps of the same workload right after the timed regions (same buffers, same position in the step)
        runner.enabled, runner.graph = False, None
        probing["on"] = True
        for _ in range(min(10, max(3, args.steps))):
            device_step()
        torch.cuda.synchronize()
        probing["on"] = False
    dot_t = [a.elapsed_time(b) * 1e-3 for a, b in probes["dot"]]
    loss_t = [a.elapsed_time(b) * 1e-3 for a, b in probes["loss"]]
    probe_note = "timed in-step right after backward: part of the working set is L2-resident" + (
        "" if runner.replays == 0 else "; the timed steps were CUDA-graph replays, the probes come from eager steps of the same workload run right after them")
    if not dot_t or not loss_t:
        # the step was replayed from a CUDA graph, whose nodes cannot carry event probes: time the same two kernels on
        # same-sized buffers right after the timed region instead (warm L2)
        probe_note = "no in-step probes were recorded: kernels timed standalone after the timed region, warm L2"
        n = shard_elems
        bufs = [torch.randn(n, device=dev) * 0.1 for _ in range(5)]
        zs, zt = torch.randn(per_gpu, classes, device=dev), torch.randn(per_gpu, classes, device=dev)
        yy = torch.randint(0, classes, (per_gpu,), device=dev)
        dot_t, loss_t = [], []
        for i in range(12):
            a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            a.record()
            if cfg.SOLVER.TRAINER == "dot":
                orig_dot(bufs[0], bufs[1], bufs[2], bufs[3], bufs[4], lr=0.05, momentum=0.825, momentum_kd=0.975, weight_decay=5e-4)
            else:
                orig_sgd(bufs[0], bufs[1], bufs[3], lr=0.05, momentum=0.9, weight_decay=5e-4)
            b.record()
            orig_loss(zs, zt, yy, kind=ops.LOSS_DKD if cfg.DISTILLER.TYPE == "DKD" else ops.LOSS_KD, alpha=1.0, beta=8.0,
                      temperature=4.0, ce_mode=ops.CE_SPLIT if cfg.SOLVER.TRAINER == "dot" else ops.CE_COMBINED)
            c.record()
            c.synchronize()
            if i >= 2:
                dot_t.append(a.elapsed_time(b) * 1e-3)
                loss_t.append(b.elapsed_time(c) * 1e-3)
    peak, peak_src = hbm_peak()
    dot_avg = sum(dot_t) / len(dot_t)
    is_dot = cfg.SOLVER.TRAINER == "dot"
    dot_bytes = (32 if is_dot else 20) * shard_elems
    loss_avg = sum(loss_t) / len(loss_t)
    step_t = [a.elapsed_time(b) * 1e-3 for a, b in probes["opt_step"]]
    loss_bytes = (16 if is_dot else 12) * per_gpu * classes + 8 * per_gpu + 8

    fused_info = getattr(trainer.optimizer, "last_fused", None)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    h2d = sum(t.numel() * t.element_size() for t in host[:2])
    if fused_info is None:
        roofline = {"kernel": "flat_step_vec4<true,true> (mdb200_dot_step)" if is_dot
                    else "flat_step_vec4<true,false> (mdb200_sgd_step)", "bound": "hbm",
                    "achieved": dot_bytes / dot_avg / 1e9, "peak": peak, "unit": "GB/s",
                    "frac": dot_bytes / dot_avg / 1e9 / peak, "traffic": profiled_traffic("dot_step", shard_elems),
                    "peak_source": peak_src, "bytes_per_launch": dot_bytes, "us_per_launch": dot_avg * 1e6,
                    "launches_timed": len(dot_t), "note": probe_note}
    else:
        # fused reduce + update + broadcast kernel: the bytes that must cross NVLink bound it, not HBM
        S, Wn = 4 * shard_elems, world
        nb = 2 if fused_info["kd_in_kernel"] else 1
        if fused_info.get("in_kernel_sync"):
            fused_info = dict(fused_info, note="us_per_launch includes the in-kernel wait for the slowest rank")
        if fused_info["multicast"]:     # every replica of a bucket shard travels to the switch once; P comes back W times
            out_b, in_b = nb * Wn * S + S, nb * S + Wn * S
        else:
            out_b, in_b = nb * (Wn - 1) * S + (Wn - 1) * S, nb * (Wn - 1) * S + (Wn - 1) * S
        link = max(out_b, in_b)
        roofline = {"kernel": "dot_fused_shard (mdb200_dot_step_fused)" if is_dot else "dot_fused_shard<HAS_KD=false> (mdb200_sgd_step_fused)",
                    "bound": "nvlink",
                    "achieved": link / dot_avg / 1e9, "peak": 770.0, "unit": "GB/s", "frac": link / dot_avg / 1e9 / 770.0,
                    "traffic": None, "peak_source": "B200_PROFILING.md: measured peer copy 770 GB/s per direction per GPU",
                    "bytes_per_launch": link, "us_per_launch": dot_avg * 1e6, "launches_timed": len(dot_t),
                    "note": f"bytes = busier NVLink direction per GPU (out {out_b}, in {in_b}); HBM-side algorithmic "
                            f"bytes are {dot_bytes}; mode {fused_info}"}
    line = {
        "metric": METRIC, "value": world * per_gpu * args.steps / secs, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3) + settle, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_label(args.workload),
                   "per_gpu_batch": per_gpu, "global_batch": per_gpu * world, "student_params": n_params,
                   "backbones": ("the reference's own modules (staged tree)" if W.backbone_source == "reference"
                                 else "architecture-equivalent stand-ins (no reference tree staged)"),
                   "parallelism": (f"dp{world}: reduce-scatter x2 -> DOT on 1/{world} shard -> all-gather (NCCL)" if args.comm == "nccl" or world == 1
                                   else f"dp{world}: fused NVLink kernel ({args.comm}): gradient sum over ranks -> DOT on 1/{world} shard -> parameters to every rank") + comm_note,
                   "backbone_math": "fp32 storage, cuDNN TF32 convolutions allowed (torch default, as the reference)",
                   "l2": (f"inputs_exceed_l2 ({h2d / 1e6:.0f} MB batch, activations of a step far beyond the 126 MB L2)"
                          if h2d > (126 << 20) else
                          f"step working set (batch {h2d / 1e6:.1f} MB + activations + 5 x {4 * n_params / 1e6:.0f} MB flat buffers) "
                          "is streamed once per step; isolated-kernel numbers in kernels{} flush L2 before every launch")},
        "e2e": {"value": world * per_gpu * args.steps / e2e_secs, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 20, "ms_per_step": 1e3 * e2e_secs / args.steps,
                "api": "StandaloneTrainer.train_iter((image, target, index), epoch, meters) -- the mirror of "
                       f"DOT.train_iter (trainer.py:412-461); metrics_lag={trainer.metrics_lag} (the package default), flushed inside the timed region"},
        "gpu_launches": launches,
        "optimizer_step": {"us": 1e6 * sum(step_t) / max(1, len(step_t)),
                           "what": "optimizer.step(): task-bucket collective + DOT kernel + parameter write-back, "
                                   "CUDA events on the compute stream", "comm": args.comm if world > 1 else "none",
                           "fused": getattr(trainer.optimizer, "last_fused", None)},
        "clocks": clocks,
        "cuda_graph": graph_summary,
        "dp_parity": parity,
        "roofline": roofline,
        "loss_kernel_in_step": {"kernel": "loss_rows_in_registers (mdb200_distill_loss)", "us_per_launch": loss_avg * 1e6,
                                "bytes_per_launch": loss_bytes, "GB/s": loss_bytes / loss_avg / 1e9,
                                "note": "256x1000 is 4 MB: launch-latency bound, see kernels{} for the large-shape sweep"},
    }
    if not args.no_sweep:
        line["kernels"] = kernel_sweep(dev, grid=(args.sweep == "grid"))
        if world == 1:
            line["torch_on_b200_baseline"] = torch_on_gpu_baseline(dev, args.workload, per_gpu)
    if world == 1 and not args.no_cpu:
        # bounded sample of the SAME config (same per-step batch): about 10-30 s of host work
        cpu_steps = 3 if args.workload.startswith("imagenet") else 20
        line["cpu_baseline"] = {k: v for k, v in cpu_arm(args.workload, args.cpu_batch or per_gpu, cpu_steps, 1).items()
                                if k != "ms_per_step"}
    else:
        line["cpu_baseline"] = None
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
