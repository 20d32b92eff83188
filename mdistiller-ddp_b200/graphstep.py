"""Whole training step as ONE CUDA graph for the launch-bound configurations.

On the CIFAR-sized workloads (BASELINE.json configs[1], configs[2]) a step is ~250 small kernels
and the host needs 4-5 ms to enqueue what the GPU runs in far less: the GPU idles.  Once the step
has reached its steady state (every momentum buffer exists, shapes fixed), ``GraphedStep`` captures
forward -> fused loss -> backward(s) -> gradient gather -> flat optimizer step into a
``torch.cuda.CUDAGraph`` and replays it with one launch per iteration.

What is baked into a captured graph are kernel ARGUMENTS: learning rate, momenta, weight decay,
the DKD warm-up factor ``min(epoch / WARMUP, 1)``.  They form the graph's key together with the
batch shape; when the key changes (the reference's ``adjust_learning_rate`` at an epoch boundary,
a new epoch during warm-up, a ragged last batch) the step is captured again, and a key that keeps
changing (a per-iteration schedule) switches the mechanism off.  Anything that cannot be replayed
-- Adam's host-side step count, AMP's first-step logic, a data-parallel world (the collectives and
the sync-pad epoch are launch arguments) -- stays on the eager path.

``MDB200_GRAPH`` = ``auto`` (default: capture when the eager step is host-bound), ``1`` (always try),
``0`` (never).
"""

from __future__ import annotations

import os
import time
from typing import Optional

import torch


def graph_mode() -> str:
    return os.environ.get("MDB200_GRAPH", "auto")


class GraphedStep:
    """Owns the capture / replay decision for one trainer (see the module docstring)."""

    WARM_STEPS = 4            # eager steps before anything is measured or captured
    MAX_RECAPTURES = 8        # key changes in a row without a single replay -> give up

    def __init__(self, trainer):
        self.trainer = trainer
        self.mode = graph_mode()
        self.enabled = self.mode != "0"
        self.seen = 0
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.key = None
        self.static_in = None
        self.static_out = None
        self.replays = 0
        self.captures = 0
        self.misses_in_a_row = 0
        self.why_off = None
        self._host_s, self._gpu_ev = [], []

    # ------------------------------------------------------------------ eligibility
    def _eligible(self) -> bool:
        tr, opt = self.trainer, self.trainer.optimizer
        if opt.world_size != 1:
            return self._off("data-parallel world: collectives and the sync-pad epoch are launch arguments")
        if getattr(opt, "loss_scaler", None) is not None:
            return self._off("AMP")
        if type(opt).__name__ == "FlatAdam":
            return self._off("Adam's bias correction changes every step")
        if not getattr(opt, "gather_grads", False):
            return self._off("gradient gather mode is off")
        return True

    def _off(self, why: str) -> bool:
        self.enabled, self.why_off = False, why
        self.graph = None
        return False

    def _steady(self) -> bool:
        opt = self.trainer.optimizer
        if hasattr(opt, "_steady"):
            return bool(opt._steady)
        return all(getattr(opt, "_has_v", [True])) or opt.param_groups[0]["momentum"] == 0

    def _key(self, image, target, epoch):
        g = self.trainer.optimizer.param_groups[0]
        hp = tuple((k, g[k]) for k in sorted(g) if k != "params" and not isinstance(g[k], torch.Tensor))
        return (tuple(image.shape), image.dtype, tuple(target.shape), int(epoch), hp)

    # ------------------------------------------------------------------ the step
    def step(self, image, target, epoch):
        """One training step on device-resident inputs: replay, capture-then-replay, or eager."""
        tr = self.trainer
        self.seen += 1
        if not self.enabled or self.seen <= self.WARM_STEPS or not self._steady():
            return self._eager(image, target, epoch, measure=self.enabled and self.seen > 2)
        if self.graph is None and self.captures == 0:
            if not self._eligible():
                return tr._eager_tail_step(image, target, epoch)
            if self.mode == "auto" and not self._host_bound():
                self._off("the eager step is not host-bound")
                return tr._eager_tail_step(image, target, epoch)
        key = self._key(image, target, epoch)
        if self.graph is None or key != self.key:
            self.misses_in_a_row += 1
            if self.misses_in_a_row > self.MAX_RECAPTURES:
                self._off("launch arguments change every step (per-iteration schedule)")
                return tr._eager_tail_step(image, target, epoch)
            if not self._capture(image, target, epoch, key):
                return tr._eager_tail_step(image, target, epoch)
        else:
            self.misses_in_a_row = 0
        self.static_in[0].copy_(image, non_blocking=True)
        self.static_in[1].copy_(target, non_blocking=True)
        self.graph.replay()
        self.replays += 1
        from . import ops
        ops.launch_count += self.launches_per_step
        tr.optimizer.note_replayed_step()
        return self.static_out

    def _eager(self, image, target, epoch, measure: bool):
        if not measure:
            return self.trainer._eager_tail_step(image, target, epoch)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        out = self.trainer._eager_tail_step(image, target, epoch)
        e1.record()
        self._host_s.append(time.perf_counter() - t0)
        self._gpu_ev.append((e0, e1))
        return out

    def _host_bound(self) -> bool:
        """Did enqueueing the eager step take (nearly) as long as the step took on the GPU timeline?  Then the
        GPU was waiting for the host and one graph launch per step is the cure."""
        if not self._gpu_ev:
            return False
        e0, e1 = self._gpu_ev[-1]
        e1.synchronize()
        gpu_s = e0.elapsed_time(e1) * 1e-3
        self.host_fraction = self._host_s[-1] / max(gpu_s, 1e-9)
        self._host_s, self._gpu_ev = [], []
        return self.host_fraction >= 0.8

    # ------------------------------------------------------------------ capture
    def _capture(self, image, target, epoch, key) -> bool:
        """Record the step for this key.  Nothing runs during capture; the caller replays right afterwards, so the
        captured step IS this iteration's step.  A capture that fails leaves the eager path in charge for good."""
        tr, opt = self.trainer, self.trainer.optimizer
        self.graph, self.key = None, None
        if self.static_in is None or tuple(self.static_in[0].shape) != tuple(image.shape):
            self.static_in = (torch.empty_like(image), torch.empty_like(target))
        self.static_in[0].copy_(image)
        self.static_in[1].copy_(target)
        from . import ops
        saved = opt.host_state()
        launches0 = ops.launch_count
        graph = torch.cuda.CUDAGraph()
        try:
            torch.cuda.synchronize()
            with torch.cuda.graph(graph):
                preds, losses = tr._eager_tail_step(self.static_in[0], self.static_in[1], epoch, graphed=True)
        except Exception as e:              # an op that cannot be captured (host sync, allocation by a library, ...)
            torch.cuda.synchronize()
            opt.restore_host_state(saved)
            self._off(f"capture failed: {type(e).__name__}: {str(e)[:160]}")
            return False
        opt.restore_host_state(saved)       # the replay below is what performs (and counts) this step
        self.launches_per_step = ops.launch_count - launches0      # our kernels inside the graph, per replay
        ops.launch_count = launches0                               # recording launched nothing
        self.graph, self.key = graph, key
        self.static_out = (preds, {k: v for k, v in losses.items()})
        self.captures += 1
        return True

    def summary(self) -> dict:
        return {"mode": self.mode, "enabled": self.enabled, "captures": self.captures, "replays": self.replays,
                "why_off": self.why_off, "host_over_gpu_time": getattr(self, "host_fraction", None)}
