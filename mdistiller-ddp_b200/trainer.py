"""Per-iteration training step for the DOT and base trainers, B200 path.

Mirrors the hot section of the reference engine and nothing else:
  DOT.init_optimizer / DOT.train_iter      mdistiller/engine/trainer.py:376-391, 412-461
  BaseTrainer.init_optimizer / train_iter  mdistiller/engine/trainer.py:83-110, 257-321
Epoch loop, logging, validation and checkpoints stay the reference's (out of scope);
``install.py`` grafts these methods onto the reference's ``BaseTrainer`` so an unchanged
``tools/train.py`` picks them up through ``trainer_dict``.

Differences from the reference, all documented reference defects:
  D1/D2  m_task = SGD.MOMENTUM - DOT.DELTA, m_kd = SGD.MOMENTUM + DOT.DELTA, AMP read from cfg.
  D3     both gradient sets are averaged across ranks (the reference syncs only the KD one).
  D4     only student parameters are communicated (DDP's reducer is switched off).
"""

from __future__ import annotations

import os
import time
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import ops
from .comm import ShardComm, SymmetricShardComm
from .amp import DeviceLossScaler
from .graphstep import GraphedStep
from .optim import DistillationOrientedTrainer, FlatAdam, FlatSGD


def _unwrap(distiller):
    return distiller.module if hasattr(distiller, "module") else distiller


def silence_ddp_reducer(distiller) -> None:
    """An already DDP-wrapped distiller (tools/train.py:86) must not all-reduce on its own:
    the flat buckets are reduce-scattered by the optimizer instead."""
    if hasattr(distiller, "require_backward_grad_sync"):
        distiller.require_backward_grad_sync = False


class _Meter:
    def __init__(self):
        self.val = self.avg = self.sum = self.count = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count


def new_train_meters() -> Dict:
    return {"training_time": _Meter(), "data_time": _Meter(), "losses": dict(), "top1": _Meter(), "top5": _Meter()}


class TailStepMixin:
    """`init_optimizer` + `train_iter` for both trainers; host class supplies
    ``self.distiller``, ``self.cfg`` and (after init) ``self.optimizer``."""

    trainer_kind = "dot"   # "dot" | "base"
    # how the data-parallel step talks: "auto" (default) = the measured policy of resolve_comm_mode(), falling
    # back to NCCL launches when NVLink symmetric memory cannot be set up; "nccl" = reduce-scatter / update /
    # all-gather launches; "fused" = one kernel over NVLink symmetric memory (KD bucket still on an overlapped
    # NCCL reduce-scatter); "fused_kd" = both buckets in the kernel; "*_p2p" = no NVSwitch multicast
    comm_mode = os.environ.get("MDB200_COMM", "auto")
    comm_mode_used = "none"       # what init_optimizer actually set up (bench.py reports it)

    # ---------------------------------------------------------------- optimizer
    def _make_comm(self):
        """The data-parallel plumbing this trainer ships with -- bench.py measures exactly this."""
        if not (dist.is_available() and dist.is_initialized()):
            self.comm_mode_used = "none"
            return None
        world = dist.get_world_size()
        mode = resolve_comm_mode(self.comm_mode, world)
        # the fused NVLink step exists for DOT and for the base trainer's momentum SGD; "auto" picks it for DOT
        # only: measured on 8 B200s (profiles/r2_n8_*), it wins on the DOT trainer (3 collectives -> 1 kernel:
        # +18 % images/s on KD+DOT ResNet32x4->ResNet8x4) and loses to two NCCL launches on the base trainer
        fusable = self.trainer_kind == "dot" or (self.comm_mode != "auto" and
                                                 self.__dict__.get("_base_solver", "SGD") == "SGD")
        if mode != "nccl" and world > 1 and fusable:
            try:
                comm = SymmetricShardComm(fuse_kd=mode.startswith("fused_kd"), use_multicast=not mode.endswith("p2p"),
                                          in_kernel_sync=os.environ.get("MDB200_FUSED_SYNC", "pad") != "barrier")
                comm.probe()                    # allocation + rendezvous work on this box?
                self.comm_mode_used = mode
                return comm
            except Exception as e:              # no symmetric memory / multicast here
                if self.comm_mode != "auto":
                    raise
                self.comm_fallback_reason = f"{type(e).__name__}: {str(e)[:160]}"
        self.comm_mode_used = "nccl" if world > 1 else "none"
        return ShardComm()

    def init_optimizer(self, cfg):
        core = _unwrap(self.distiller)
        silence_ddp_reducer(self.distiller)
        self._base_solver = cfg.SOLVER.TYPE          # the fused NVLink step exists for DOT and for momentum SGD
        comm = self._make_comm()
        params = core.get_learnable_parameters()
        if self.trainer_kind == "dot":
            if cfg.EXPERIMENT.AMP:
                raise NotImplementedError("AMP is not supported for DOT.")
            if cfg.SOLVER.TYPE != "SGD":
                raise NotImplementedError(cfg.SOLVER.TYPE)
            return _trust(DistillationOrientedTrainer(
                params, lr=cfg.SOLVER.LR,
                momentum=cfg.SOLVER.SGD.MOMENTUM - cfg.SOLVER.DOT.DELTA,
                momentum_kd=cfg.SOLVER.SGD.MOMENTUM + cfg.SOLVER.DOT.DELTA,
                weight_decay=cfg.SOLVER.WEIGHT_DECAY, comm=comm))
        # AMP (trainer.py:79-80, 268, 304-313): autocast forward; the loss scale lives on the device and its
        # unscale / overflow check / update are folded into the flat step (amp.DeviceLossScaler)
        self.use_amp = bool(cfg.EXPERIMENT.AMP)
        # (kept on the optimizer: the reference's BaseTrainer.__init__ assigns its own GradScaler to self.amp_scaler
        #  right after this call, trainer.py:79-80 -- it is simply never used on this path)
        loss_scaler = DeviceLossScaler(params[0].device) if self.use_amp else None
        core.combined_tail = True                 # one gradient buffer: the base trainer sums the losses anyway
        kind = cfg.SOLVER.TYPE
        if kind == "SGD":
            opt = FlatSGD(params, lr=cfg.SOLVER.LR, momentum=cfg.SOLVER.SGD.MOMENTUM,
                          weight_decay=cfg.SOLVER.WEIGHT_DECAY, comm=comm)
        elif kind in ("Adam", "AdamW"):
            opt = FlatAdam(params, lr=cfg.SOLVER.LR, betas=tuple(cfg.SOLVER.ADAM.BETAS),
                           eps=cfg.SOLVER.ADAM.EPSILON, weight_decay=cfg.SOLVER.WEIGHT_DECAY,
                           decoupled=(kind == "AdamW"), comm=comm)
        else:
            raise NotImplementedError(kind)
        opt.max_grad_norm = float(getattr(cfg.SOLVER, "GRAD_CLIP", 0.0) or 0.0)   # clip inside step(), on the shard
        opt.loss_scaler = loss_scaler
        return _trust(opt)

    # ---------------------------------------------------------------- one iteration
    # metrics_lag = 0: the progress line of an iteration contains that iteration (the reference's
    # behaviour, one host sync per step).  metrics_lag = k > 0 (default 1, MDB200_METRICS_LAG): the
    # device-to-host read of step i is only waited for at step i+k, so the host can already enqueue the
    # next batch's H2D copy and kernels while the GPU finishes step i; the progress line then trails by k
    # iterations, and flush_metrics() (called automatically after an epoch's last iteration when the host
    # set iters_per_epoch) settles the rest before anybody reads the epoch averages.
    metrics_lag = int(os.environ.get("MDB200_METRICS_LAG", "1"))
    # set by a host that knows the epoch length (install.py: len(train_loader)): after the epoch's last
    # iteration every rank settles the lagged metrics and takes the collective checkpoint snapshot, so the
    # reference's epoch loop can log exact averages and checkpoint on the master rank alone
    iters_per_epoch = 0

    def train_iter(self, data, epoch, train_meters):
        msg = self._train_iter(data, epoch, train_meters)
        if self.iters_per_epoch:
            if self.__dict__.get("_epoch_seen") != epoch:
                self._epoch_seen, self._iter_in_epoch = epoch, 0
            self._iter_in_epoch += 1
            if self._iter_in_epoch == self.iters_per_epoch:
                self.flush_metrics(train_meters)
                self.optimizer.sync_checkpoint_state()
                msg = self._progress_line(epoch, train_meters)
        return msg

    def _train_iter(self, data, epoch, train_meters):
        t0 = time.time()
        image, target, index = data
        train_meters["data_time"].update(time.time() - t0)
        # H2D on a side stream: with lagged metrics the copy of batch i+1 overlaps the kernels of step i
        copy_stream = self.__dict__.get("_copy_stream")
        if copy_stream is None:
            copy_stream = self._copy_stream = torch.cuda.Stream()
        main = torch.cuda.current_stream()
        with torch.cuda.stream(copy_stream):
            image = image.float().cuda(non_blocking=True)
            target = target.cuda(non_blocking=True)
        main.wait_stream(copy_stream)
        image.record_stream(main)
        target.record_stream(main)
        preds, losses = self.tail_step(image, target, epoch)
        train_meters["training_time"].update(time.time() - t0)
        return self.report(preds, target, losses, epoch, train_meters)

    def tail_step(self, image, target, epoch):
        """forward -> backward(s) -> optimizer; returns (logits, {'loss_ce','loss_kd'}) on device.  Launch-bound
        steady-state steps are replayed from a CUDA graph (graphstep.py), everything else runs eagerly."""
        runner = self.__dict__.get("_graph_runner")
        if runner is None:
            runner = self._graph_runner = GraphedStep(self)
        return runner.step(image, target, epoch)

    def _eager_tail_step(self, image, target, epoch, graphed: bool = False):
        opt = self.optimizer
        core = _unwrap(self.distiller)
        # inside a graph capture the DDP wrapper (whose reducer is switched off anyway) is bypassed: its forward
        # hooks are host logic that must not be frozen into the graph
        model = core if graphed else self.distiller
        if self.trainer_kind == "dot":
            preds, losses = model(image=image, target=target, epoch=epoch)
            tail = getattr(core, "last_tail", None)
            opt.zero_grad(set_to_none=True)
            if tail is not None and not tail.combined:
                # seed both backwards with the gradients the fused launch already produced
                torch.autograd.backward([tail.logits], [tail.grad_kd], retain_graph=True)
                opt.step_kd()
                opt.zero_grad(set_to_none=True)
                torch.autograd.backward([tail.logits], [tail.grad_ce])
            else:
                losses["loss_kd"].backward(retain_graph=True)
                opt.step_kd()
                opt.zero_grad(set_to_none=True)
                losses["loss_ce"].backward()
            opt.step()
        elif getattr(self, "use_amp", False):
            opt.zero_grad()
            with torch.autocast("cuda"):
                preds, losses = model(image=image, target=target, epoch=epoch)
            tail = getattr(core, "last_tail", None)
            scaler = opt.loss_scaler
            if tail is not None and tail.combined:
                # GradScaler.scale(loss).backward(): seed the backward with (dloss/dlogits) x scale, in the logits' dtype
                seed = scaler.scale_gradient(tail.grad_kd).to(tail.logits.dtype)
                torch.autograd.backward([tail.logits], [seed])
            else:
                (sum(losses.values()) * scaler.scale.squeeze(0)).backward()
            opt.step()        # unscale, overflow check, clipping and the scaler's update happen inside, on the device
        else:
            opt.zero_grad()
            preds, losses = model(image=image, target=target, epoch=epoch)
            tail = getattr(core, "last_tail", None)
            if tail is not None and tail.combined:
                torch.autograd.backward([tail.logits], [tail.grad_kd])
            else:
                sum(losses.values()).backward()
            opt.step()        # SOLVER.GRAD_CLIP (trainer.py:316-320) happens inside, after the reduce-scatter
        if tail is not None:
            core.last_tail = None
        return preds, losses

    # ---------------------------------------------------------------- metrics (one collective, one D2H)
    @torch.no_grad()
    def report(self, preds, target, losses, epoch, train_meters):
        """Cross-rank loss means and top-1/top-5 for the progress line (trainer.py:436-461):
        local counts are packed into one tensor, summed with ONE all-reduce and read back with
        ONE device-to-host copy (the reference does 2 all-gathers + 4 all-reduces + 4 syncs)."""
        pack = self._metrics_pack(preds.detach(), target, losses["loss_ce"].detach(), losses["loss_kd"].detach())
        world = 1
        if dist.is_available() and dist.is_initialized():
            dist.all_reduce(pack)
            world = dist.get_world_size()
        pending = self.__dict__.setdefault("_pending_metrics", [])
        host = torch.empty(5, dtype=torch.float32, pin_memory=True)
        host.copy_(pack, non_blocking=True)
        done = torch.cuda.Event()
        done.record()
        pending.append((done, host, world))
        while len(pending) > self.metrics_lag:
            self._consume_metrics(pending.pop(0), train_meters)
        return self._progress_line(epoch, train_meters)

    @staticmethod
    def _metrics_pack(preds, target, loss_ce, loss_kd):
        """{loss_ce, loss_kd, top-1 hits, top-5 hits, rows} on the device, one launch (fp16 logits of an
        autocast forward are widened first)."""
        return ops.metrics_record(preds.float(), target, loss_ce.float(), loss_kd.float())

    def _consume_metrics(self, entry, train_meters) -> None:
        done, host, world = entry
        done.synchronize()
        ce, kd, c1, c5, n = host.tolist()
        vals = {"loss_ce": ce / world, "loss_kd": kd / world, "_total": (ce + kd) / world}
        n = int(n)
        for name, v in vals.items():
            train_meters["losses"].setdefault(name, _Meter()).update(v, n)
        train_meters["top1"].update(100.0 * c1 / n, n)
        train_meters["top5"].update(100.0 * c5 / n, n)

    def flush_metrics(self, train_meters) -> None:
        """Wait for and account every outstanding iteration (needed only with metrics_lag > 0)."""
        pending = self.__dict__.get("_pending_metrics", [])
        while pending:
            self._consume_metrics(pending.pop(0), train_meters)

    @staticmethod
    def _progress_line(epoch, train_meters) -> str:
        total = train_meters["losses"]["_total"].avg if "_total" in train_meters["losses"] else float("nan")
        return "Epoch:{}| Time(data):{:.3f}| Time(train):{:.3f}| Loss:{:.4f}| Top-1:{:.3f}| Top-5:{:.3f}".format(
            epoch, train_meters["data_time"].avg, train_meters["training_time"].avg, total,
            train_meters["top1"].avg, train_meters["top5"].avg)


def resolve_comm_mode(mode: str, world: int) -> str:
    """"auto" -> the step measured best on one HGX B200 box (profiles/): the fused NVLink kernel with plain peer
    loads/stores up to 4 GPUs, NVSwitch multicast with both buckets in the kernel above; anything else is
    taken literally."""
    if mode != "auto":
        return mode
    if world <= 1:
        return "nccl"
    return "fused_p2p" if world <= 4 else "fused_kd"


def _trust(opt):
    """tail_step owns the gradient protocol: on CUDA let autograd keep its own gradient tensors and gather
    them into the bucket with one launch; elsewhere (CPU test doubles) trust the bucket views."""
    if opt.flat_params.is_cuda and os.environ.get("MDB200_NO_GATHER") != "1":
        opt.enable_gather()
    else:
        opt.trust_bucket_grads = True
    return opt


class StandaloneTrainer(TailStepMixin):
    """Minimal host for the mixin when the reference engine is not around (bench, smoke, tests)."""

    def __init__(self, distiller, cfg, kind: Optional[str] = None):
        self.distiller = distiller
        self.cfg = cfg
        self.trainer_kind = kind or cfg.SOLVER.TRAINER
        self.grad_clip = getattr(cfg.SOLVER, "GRAD_CLIP", 0.0)
        self.optimizer = self.init_optimizer(cfg)
