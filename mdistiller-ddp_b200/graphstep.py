"""Whole training step as ONE CUDA graph for the launch-bound configurations.

On the CIFAR-sized workloads (BASELINE.json configs[1], configs[2]) a step is ~250 small kernels
and the host needs 4-5 ms to enqueue what the GPU runs in far less: the GPU idles (+37 % images/s on
KD+DOT ResNet32x4->ResNet8x4, +20 % on VGG13->VGG8 once replayed).  Even the GPU-bound ImageNet step
gains 2 % (79.9 -> 78.4 ms on one B200, same box) from the launch gaps that disappear.  Once the step
has reached its steady state (every momentum buffer exists, shapes fixed), ``GraphedStep`` captures
forward -> fused loss -> backward(s) -> gradient gather -> flat optimizer step into a
``torch.cuda.CUDAGraph`` and replays it with one launch per iteration.

What is baked into a captured graph are kernel ARGUMENTS: learning rate, momenta, weight decay,
the DKD warm-up factor ``min(epoch / WARMUP, 1)``.  They form the graph's key together with the
batch shape; when the key changes (the reference's ``adjust_learning_rate`` at an epoch boundary,
a new epoch during warm-up, a ragged last batch) the step is captured again, and a key that keeps
changing (a per-iteration schedule) switches the mechanism off.  Anything that cannot be replayed
-- Adam's host-side step count, AMP's first-step logic, a data-parallel world -- stays on the eager path.
(Data-parallel replay was tried on 2 GPUs in round 2: the fused NVLink kernel can count its sync-pad epochs
itself, epoch argument 0, and passes its parity tests that way, but the captured step hung and the GPU budget
ended before it could be debugged; it is the first item of DESIGN.md section 8.)

``MDB200_GRAPH`` = ``auto`` / ``1`` (default: capture whenever the step is eligible), ``0`` (never).
"""

from __future__ import annotations

import os
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

    # ------------------------------------------------------------------ eligibility
    def _eligible(self) -> bool:
        tr, opt = self.trainer, self.trainer.optimizer
        if opt.world_size != 1:
            return self._off("data-parallel world")
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
            return tr._eager_tail_step(image, target, epoch)
        if self.graph is None and self.captures == 0:
            if not self._eligible():
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
                "why_off": self.why_off}
