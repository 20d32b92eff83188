"""Flat-buffer optimizers: DOT's dual-momentum SGD and plain momentum SGD.

``DistillationOrientedTrainer`` keeps the reference's constructor, ``step_kd()`` /
``step()`` / ``zero_grad()`` protocol, ``param_groups[0]["lr"]`` mutability and
``state_dict()`` layout (mdistiller/engine/dot.py:58-174), but stores parameters,
the two gradient sets and the two momentum buffers as flat fp32 buffers so one
sm_100a launch updates everything (434 launches per step for ResNet18 in the
reference).  With ``torch.distributed`` initialised it is also the data-parallel
step: reduce-scatter(KD bucket) -> reduce-scatter(task bucket) -> update on the rank's
shard -> all-gather(parameters), replacing DDP's all-reduce + the optimizer loop
(tools/train.py:86, trainer.py:427-432).

How the two gradient sets get into flat buckets without extra copies: ``zero_grad``
zeroes the bucket of the coming backward and points every ``p.grad`` at its slice, so
autograd accumulates straight into the bucket.  A parameter counts as "has a gradient"
for a pass (dot.py:112,144 test ``p.grad is not None``) if autograd touched it in that
pass (post-accumulate hook) or the caller assigned a gradient tensor of its own, which
is then copied into the bucket.
"""

from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch.optim.optimizer import Optimizer, required

from . import ops
from .comm import ShardComm
from .flat import FlatSpace

_NONE, _TASK, _KD, _BOTH = 0, 1, 2, 3
_MEMBERSHIP = {_BOTH: ops.DOT_BOTH, _TASK: ops.DOT_TASK_ONLY, _KD: ops.DOT_KD_ONLY}


class _FlatOptimizer(Optimizer):
    """Shared plumbing: flat space, gradient buckets, membership tracking, sharding."""

    _bucket_names: Tuple[str, ...] = ()

    def __init__(self, params, defaults, comm: Optional[ShardComm], kernels):
        super().__init__(params, defaults)
        assert len(self.param_groups) == 1, "Only implement for one-group params."
        self._k = kernels if kernels is not None else ops
        self._comm = comm if comm is not None else ShardComm()
        plist = self.param_groups[0]["params"]
        self._space = FlatSpace(plist, self._comm.world_size, self._comm.rank,
                                flat_alloc=getattr(self._comm, "alloc_flat", None))
        self._comm.broadcast(self._space.flat, src=0)       # DDP's constructor does the same
        self._buckets: Dict[str, torch.Tensor] = {}
        self._views: Dict[str, List[torch.Tensor]] = {}
        for name in self._bucket_names:
            b = self._space.new_flat_buffer()
            self._buckets[name] = b
            self._views[name] = self._space.views(b)
        self._phase = self._bucket_names[0]
        # set by a trainer that only ever lets autograd write the gradients: skips the per-parameter
        # "did the caller assign a tensor of its own?" check in step()/step_kd()
        self.trust_bucket_grads = False
        # gather mode (CUDA, trusted callers): zero_grad() sets p.grad = None, autograd hands over its own
        # gradient tensors, and ONE multi-tensor launch copies them into the bucket -- instead of one add_
        # launch per parameter per backward when autograd accumulates into bucket views
        self.gather_grads = False
        self._chunk_table = None
        self._all_true = [True] * len(plist)
        self._touched = set()
        self._index = {id(p): i for i, p in enumerate(plist)}
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in plist if p.requires_grad]
        self._run_cache: Dict[tuple, list] = {}
        self._n_steps = 0                       # completed step() calls; dates the checkpoint snapshot
        self._snapshot: Optional[Dict[str, torch.Tensor]] = None
        self._snapshot_step = -1

    # -- gradient capture ----------------------------------------------------------
    def _on_grad(self, p: torch.Tensor) -> None:
        self._touched.add(self._index[id(p)])

    def enable_gather(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self._chunk_table = self._space.chunk_table()
        self.trust_bucket_grads = self.gather_grads = True

    def _arm(self, name: str) -> None:
        """Zero bucket `name` and make it the accumulation target of the next backward."""
        if self.gather_grads:
            for p in self._space.params:
                p.grad = None          # nothing to zero: the gather overwrites (or zero-fills) every slice
            return
        self._buckets[name].zero_()
        for p, v in zip(self._space.params, self._views[name]):
            p.grad = v
        self._touched.clear()

    def zero_grad(self, set_to_none: bool = True) -> None:  # noqa: D401 - same signature as torch's
        # `set_to_none` is accepted for drop-in compatibility: the reference protocol only needs
        # "no stale gradient survives", which a zeroed bucket guarantees.
        self._arm(self._phase)

    def _collect(self, name: str) -> List[bool]:
        """Which parameters have a gradient for this pass; foreign tensors are copied in."""
        if self.gather_grads:
            return self._gather(name)
        if self.trust_bucket_grads:          # caller guarantees p.grad are still our views (our own trainers do)
            touched = self._touched
            if len(touched) == len(self._space.params):
                return self._all_true
            return [i in touched for i in range(len(self._space.params))]
        views = self._views[name]
        has = []
        for i, (p, v) in enumerate(zip(self._space.params, views)):
            g = p.grad
            if g is None:
                has.append(False)
            elif g.data_ptr() == v.data_ptr() and g.shape == v.shape:
                has.append(i in self._touched)
            else:
                v.copy_(g)
                has.append(True)
        return has

    def _gather(self, name: str) -> List[bool]:
        """Multi-tensor gather of whatever autograd left in p.grad into bucket `name`."""
        ptrs, has, odd = [], [], []
        dev = self._space.device
        for i, p in enumerate(self._space.params):
            g = p.grad
            if g is None:
                ptrs.append(0)
                has.append(False)
            elif g.is_contiguous() and g.dtype == torch.float32 and g.device == dev:
                ptrs.append(g.data_ptr())
                has.append(True)
            else:                          # strided / other dtype: zero-filled by the kernel, then copied
                ptrs.append(0)
                has.append(True)
                odd.append(i)
        self._k.gather_flat(self._buckets[name], ptrs, self._chunk_table)
        for i in odd:
            self._views[name][i].copy_(self._space.params[i].grad)
        return self._all_true if all(has) else has

    # -- runs of tensors sharing one update rule --------------------------------------
    def _runs(self, modes: Sequence[tuple]) -> list:
        """Merge consecutive tensors with equal `modes[i]` into (lo, hi, mode) element ranges,
        clipped to this rank's shard."""
        key = tuple(modes)
        cached = self._run_cache.get(key)
        if cached is not None:
            return cached
        sp = self._space
        runs = []
        i, n = 0, len(modes)
        while i < n:
            j = i
            while j + 1 < n and modes[j + 1] == modes[i]:
                j += 1
            lo, hi = sp.clip_to_shard(sp.offsets[i], sp.ends[j])
            if hi > lo and modes[i][0] != _NONE:
                runs.append((lo, hi, modes[i]))
            i = j + 1
        if len(self._run_cache) > 64:
            self._run_cache.clear()
        self._run_cache[key] = runs
        return runs

    def _clip_coefficient(self, bucket: torch.Tensor, max_norm: float) -> Optional[torch.Tensor]:
        """Device scalar min(1, max_norm / (||g|| + 1e-6)) over ALL ranks' averaged gradient --
        nn.utils.clip_grad_norm_ (trainer.py:316-320) without a host sync.  `bucket` holds this
        rank's reduce-scattered SUM in its shard; the 1/world average is applied inside the norm."""
        if not max_norm or max_norm <= 0:
            return None
        sp = self._space
        sumsq = self._k.sum_squares(sp.shard(bucket), scale=1.0 / self._comm.world_size)
        self._comm.all_reduce_sum(sumsq)
        return self._k.clip_coefficient(sumsq, float(max_norm))

    # mixed precision: set by the trainer (a DeviceLossScaler); the bucket then holds SCALED gradients
    loss_scaler = None

    def _step_multiplier(self, bucket: torch.Tensor, max_norm: float) -> Optional[torch.Tensor]:
        """Device scalar that multiplies the shard's gradient inside the step kernel: the clipping coefficient,
        and under AMP 1/scale (x clipping) -- or a negative value, which makes the kernel skip the update
        (the scaler's overflow verdict; identical on every rank because it comes from the all-reduced norm)."""
        if self.loss_scaler is None:
            return self._clip_coefficient(bucket, max_norm)
        sumsq = self._k.sum_squares(self._space.shard(bucket), scale=1.0 / self._comm.world_size)
        self._comm.all_reduce_sum(sumsq)
        return self.loss_scaler.verdict(sumsq, float(max_norm or 0.0))

    # -- CUDA-graph replay of a whole step (graphstep.py): the Python side of step() does not run on a replay
    _HOST_STATE = ("_n_steps", "_snapshot", "_snapshot_step", "_phase", "_kd_members", "_kd_work", "_kd_prereduced")

    def host_state(self) -> dict:
        return {k: getattr(self, k) for k in self._HOST_STATE if hasattr(self, k)}

    def restore_host_state(self, saved: dict) -> None:
        for k, v in saved.items():
            setattr(self, k, v)

    def note_replayed_step(self) -> None:
        """A captured step has just been replayed: account for it like step() does."""
        self._n_steps += 1
        self._snapshot = None

    @property
    def flat_params(self) -> torch.Tensor:
        return self._space.flat

    @property
    def world_size(self) -> int:
        return self._comm.world_size

    # -- checkpoint format -------------------------------------------------------------
    def _full_momentum(self, shard_buf: torch.Tensor) -> torch.Tensor:
        """Gather a sharded momentum buffer into a full-length flat tensor."""
        full = self._space.new_buffer()
        self._space.shard(full).copy_(shard_buf)
        self._comm.all_gather(full)
        return full

    def _state_pairs(self) -> Sequence[Tuple[str, torch.Tensor, List[bool]]]:
        """(checkpoint key, sharded buffer, per-parameter "exists" flags) of every state tensor."""
        raise NotImplementedError

    def sync_checkpoint_state(self) -> None:
        """COLLECTIVE -- call on every rank.  Gathers the sharded optimizer state into a full-length
        snapshot so that a later ``state_dict()`` on any ONE rank needs no communication.

        Why it exists: the reference asks only the master rank for ``optimizer.state_dict()``
        (mdistiller/engine/trainer.py:230-236, under ``if IS_MASTER``), which is fine for a replicated
        torch optimizer but would leave rank 0 alone inside an all-gather when the momenta are sharded.
        The hosted trainers call this after the last iteration of every epoch (trainer.py here), so the
        reference's epoch loop checkpoints unchanged.  The snapshot is dropped by the next ``step()``."""
        if self._comm.world_size == 1:
            return
        self._snapshot = {key: self._full_momentum(buf) for key, buf, has in self._state_pairs() if any(has)}
        self._snapshot_step = self._n_steps

    def _export_state(self, pairs: Sequence[Tuple[str, torch.Tensor, List[bool]]]) -> None:
        sharded = self._comm.world_size > 1
        if sharded and (self._snapshot is None or self._snapshot_step != self._n_steps):
            raise RuntimeError(
                "optimizer state is sharded over ranks: call optimizer.sync_checkpoint_state() on EVERY rank "
                "(after the last step) before state_dict(); state_dict() itself never communicates, so it is "
                "safe on one rank, as the reference's master-only checkpointing needs")
        self.state.clear()
        for key, shard_buf, has in pairs:
            if not any(has):
                continue
            full = self._snapshot[key] if sharded else shard_buf     # world 1: the shard IS the full buffer
            for p, v, h in zip(self._space.params, self._space.views(full), has):
                if h:
                    self.state[p][key] = v.clone()

    def _import_state(self, pairs: Sequence[Tuple[str, torch.Tensor, List[bool]]]) -> None:
        sp = self._space
        for key, shard_buf, has in pairs:
            full = sp.new_buffer()
            for i, (p, v) in enumerate(zip(sp.params, sp.views(full))):
                t = self.state.get(p, {}).get(key)
                has[i] = t is not None
                if t is not None:
                    v.copy_(t.to(v.dtype))
            shard_buf.copy_(sp.shard(full))
        self.state.clear()


class DistillationOrientedTrainer(_FlatOptimizer):
    r"""Distillation-Oriented Trainer (drop-in for mdistiller/engine/dot.py:58-174).

    Usage is the reference's::

        optimizer.zero_grad(set_to_none=True)
        loss_kd.backward(retain_graph=True)
        optimizer.step_kd()      # KD gradients captured (and, multi-GPU, reduce-scattered)
        optimizer.zero_grad(set_to_none=True)
        loss_task.backward()
        optimizer.step()         # both momenta and the parameters updated in one launch
    """

    _bucket_names = ("kd", "task")

    def __init__(self, params, lr=required, momentum=0, momentum_kd=0, dampening=0, weight_decay=0, *,
                 comm: Optional[ShardComm] = None, kernels=None):
        if lr is not required and lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if momentum < 0.0:
            raise ValueError("Invalid momentum value: {}".format(momentum))
        if momentum_kd < 0.0:
            raise ValueError("Invalid momentum kd value: {}".format(momentum_kd))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        defaults = dict(lr=lr, momentum=momentum, momentum_kd=momentum_kd, dampening=dampening,
                        weight_decay=weight_decay)
        super().__init__(params, defaults, comm, kernels)
        n = len(self._space.params)
        self._v_task = self._space.new_shard_buffer()
        self._v_kd = self._space.new_shard_buffer()
        self._has_vt = [False] * n     # 'momentum_buffer' exists (dot.py:165-167)
        self._has_vk = [False] * n     # 'momentum_kd_buffer' exists (dot.py:168-170)
        self._kd_members: Optional[List[bool]] = None
        self._kd_work = None
        self._kd_prereduced = True
        self._steady, self._steady_runs = False, None
        self.last_fused = None          # set by every fused step: {"multicast": bool, "kd_in_kernel": bool}

    @torch.no_grad()
    def step_kd(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        assert len(self.param_groups) == 1, "Only implement for one-group params."
        self._kd_members = self._collect("kd")
        # the KD bucket is complete: start its reduce-scatter now so it overlaps backward #2
        # (unless the fused step kernel is going to pull it through the switch itself)
        self._kd_prereduced = not (getattr(self._comm, "fused", False) and self._comm.fuse_kd)
        if self._kd_prereduced:
            self._kd_work = self._comm.reduce_scatter_sum(self._buckets["kd"], async_op=True)
        self._phase = "task"
        return loss

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        assert len(self.param_groups) == 1, "Only implement for one-group params."
        group = self.param_groups[0]
        task_members = self._collect("task")
        kd_members = self._kd_members if self._kd_members is not None else [False] * len(task_members)
        sp = self._space
        momentum = group["momentum"]
        steady = (self._steady and task_members is self._all_true and kd_members is self._all_true)
        modes: list = []
        if steady:      # every parameter in both lists and every buffer already exists: one cached run
            runs = self._steady_runs
        else:
            for i, (t, k) in enumerate(zip(task_members, kd_members)):
                m = (_BOTH if k else _TASK) if t else (_KD if k else _NONE)
                modes.append((m, t and not self._has_vt[i], k and not self._has_vk[i]))
            runs = self._runs(modes)
        hp = dict(lr=group["lr"], momentum=momentum, momentum_kd=group["momentum_kd"],
                  weight_decay=group["weight_decay"], dampening=group["dampening"],
                  grad_scale=1.0 / self._comm.world_size)
        # The fused kernel is a rendezvous of ALL ranks, so the decision must come out the same everywhere:
        # it is taken from the global membership (every tensor in both lists, identical first-step flags),
        # never from `runs`, which is clipped to this rank's shard and may look uniform on one rank only.
        uniform = steady or (len(modes) > 0 and modes[0][0] == _BOTH and all(m == modes[0] for m in modes))
        fused = (getattr(self._comm, "fused", False) and self._comm.world_size > 1 and uniform
                 and self._kd_members is not None)
        if fused:
            self._fused_step(modes[0] if not steady else (_BOTH, False, False), hp)
        else:
            if getattr(self, "_kd_prereduced", True) is False:   # KD bucket was left for a fused step that cannot run
                self._kd_work = self._comm.reduce_scatter_sum(self._buckets["kd"], async_op=True)
            work = self._comm.reduce_scatter_sum(self._buckets["task"], async_op=True)
            for w in (self._kd_work, work):
                if w is not None:
                    w.wait()
            self._kd_work = None
            for lo, hi, (m, first_t, first_k) in runs:
                a, b = lo - sp.shard_lo, hi - sp.shard_lo
                self._k.dot_step(
                    sp.flat[lo:hi],
                    self._buckets["task"][lo:hi] if m != _KD else None,
                    self._buckets["kd"][lo:hi] if m != _TASK else None,
                    self._v_task[a:b] if (m != _KD and momentum != 0) else None,
                    self._v_kd[a:b] if m != _TASK else None,
                    first_task=first_t, first_kd=first_k, membership=_MEMBERSHIP[m], **hp)
            self._comm.all_gather(sp.flat)

        if not steady:
            for i, (t, k) in enumerate(zip(task_members, kd_members)):
                if t and momentum != 0:
                    self._has_vt[i] = True
                if k:
                    self._has_vk[i] = True
            if all(self._has_vk) and (momentum == 0 or all(self._has_vt)):
                self._steady = True
                self._steady_runs = self._runs([(_BOTH, False, False)] * len(self._has_vk))
        self._kd_members = None
        self._phase = "kd"
        self._n_steps += 1
        self._snapshot = None
        return loss

    def _fused_step(self, mode, hp) -> None:
        """gradient sum over ranks -> update -> parameters to every rank, one launch (dot_fused.cu)."""
        comm, sp = self._comm, self._space
        _, first_t, first_k = mode
        if self._kd_work is not None:
            self._kd_work.wait()
            self._kd_work = None
        p, gt, gk = sp.flat, self._buckets["task"], self._buckets["kd"]
        (p_mc, p_peers), (gt_mc, gt_peers), (gk_mc, gk_peers) = (comm.addresses(t) for t in (p, gt, gk))
        use_mc = bool(p_mc and gt_mc and gk_mc)
        sync = comm.next_sync()
        self.last_fused = {"multicast": use_mc, "kd_in_kernel": not self._kd_prereduced, "in_kernel_sync": sync is not None}
        if sync is None:
            comm.barrier(p, channel=0)      # every rank's second backward (and KD reduce-scatter) has landed
        self._k.dot_step_fused(
            p, gt, gk, self._v_task if hp["momentum"] != 0 else None, self._v_kd,
            multicast=(p_mc, gt_mc, gk_mc) if use_mc else None,
            peers=None if use_mc else (p_peers, gt_peers, gk_peers),
            kd_is_reduced=self._kd_prereduced, shard_lo=sp.shard_lo, shard_len=sp.shard_len,
            world=comm.world_size, first_task=first_t, first_kd=first_k, sync=sync, **hp)
        if sync is None:
            comm.barrier(p, channel=1)      # nobody reads P before every shard owner has written it

    # -- reference checkpoint layout: state[p]['momentum_buffer'|'momentum_kd_buffer'] -----
    def _state_pairs(self):
        return (("momentum_buffer", self._v_task, self._has_vt), ("momentum_kd_buffer", self._v_kd, self._has_vk))

    def state_dict(self):
        self._export_state(self._state_pairs())
        try:
            return super().state_dict()
        finally:
            self.state.clear()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._import_state(self._state_pairs())
        self._run_cache.clear()
        self._steady = False
        self._snapshot = None

    def momentum_views(self):
        """Per-parameter views of the (unsharded) momentum buffers -- single-process use."""
        if self._comm.world_size != 1:
            raise RuntimeError("momentum is sharded; use state_dict()")
        return self._space.views(self._v_task), self._space.views(self._v_kd)


class FlatSGD(_FlatOptimizer):
    """torch.optim.SGD(momentum, weight_decay, dampening; no Nesterov) on the flat space --
    the optimizer BaseTrainer.init_optimizer builds for SOLVER.TYPE == 'SGD'
    (mdistiller/engine/trainer.py:86-91), plus the sharded gradient sync."""

    _bucket_names = ("grad",)

    def __init__(self, params, lr=required, momentum=0, dampening=0, weight_decay=0, *,
                 comm: Optional[ShardComm] = None, kernels=None):
        if lr is not required and lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if momentum < 0.0:
            raise ValueError("Invalid momentum value: {}".format(momentum))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay)
        super().__init__(params, defaults, comm, kernels)
        self._v = self._space.new_shard_buffer()
        self._has_v = [False] * len(self._space.params)
        self.max_grad_norm = 0.0      # SOLVER.GRAD_CLIP; > 0 clips the global gradient norm inside step()
        self.last_fused = None        # set by every fused step, like the DOT optimizer's

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        group = self.param_groups[0]
        members = self._collect("grad")
        momentum = group["momentum"]
        modes = [((_TASK if t else _NONE), t and not self._has_v[i], False) for i, t in enumerate(members)]
        sp = self._space
        comm = self._comm
        # one fused kernel over NVLink (gradient sum over ranks -> SGD on the shard -> parameters to every rank)
        # when every parameter has a gradient with the same first-step flag -- a GLOBAL property, so all ranks
        # agree -- and no clipping norm has to be formed from the reduced gradient first
        fused = (getattr(comm, "fused", False) and comm.world_size > 1 and not (self.max_grad_norm and self.max_grad_norm > 0)
                 and self.loss_scaler is None and modes[0][0] == _TASK and all(m == modes[0] for m in modes))
        if fused:
            p, g = sp.flat, self._buckets["grad"]
            (p_mc, p_peers), (g_mc, g_peers) = comm.addresses(p), comm.addresses(g)
            use_mc = bool(p_mc and g_mc)
            sync = comm.next_sync()
            self.last_fused = {"multicast": use_mc, "kd_in_kernel": False, "in_kernel_sync": sync is not None}
            if sync is None:
                comm.barrier(p, channel=0)
            self._k.sgd_step_fused(p, g, self._v if momentum != 0 else None,
                                   multicast=(p_mc, g_mc) if use_mc else None, peers=None if use_mc else (p_peers, g_peers),
                                   shard_lo=sp.shard_lo, shard_len=sp.shard_len, world=comm.world_size, lr=group["lr"],
                                   momentum=momentum, weight_decay=group["weight_decay"], dampening=group["dampening"],
                                   grad_scale=1.0 / comm.world_size, first_step=modes[0][1], sync=sync)
            if sync is None:
                comm.barrier(p, channel=1)
        else:
            work = comm.reduce_scatter_sum(self._buckets["grad"], async_op=True)
            if work is not None:
                work.wait()
            clip = self._step_multiplier(self._buckets["grad"], self.max_grad_norm)
            for lo, hi, (_, first, _) in self._runs(modes):
                a, b = lo - sp.shard_lo, hi - sp.shard_lo
                self._k.sgd_step(sp.flat[lo:hi], self._buckets["grad"][lo:hi],
                                 self._v[a:b] if momentum != 0 else None,
                                 lr=group["lr"], momentum=momentum, weight_decay=group["weight_decay"],
                                 dampening=group["dampening"], grad_scale=1.0 / comm.world_size,
                                 first_step=first, grad_scale_dev=clip)
            comm.all_gather(sp.flat)
        if momentum != 0 and self.loss_scaler is None:
            for i, t in enumerate(members):
                if t:
                    self._has_v[i] = True
        elif momentum != 0 and not all(self._has_v):
            # under AMP the very first update may be skipped on the device; "buffer exists" is only certain once a
            # step has gone through, so the first-step flag is kept until the host has seen one clean verdict
            if float(self.loss_scaler.coef.item()) > 0:          # a host read, on the first step(s) only
                for i, t in enumerate(members):
                    if t:
                        self._has_v[i] = True
        self._n_steps += 1
        self._snapshot = None
        return loss

    def _state_pairs(self):
        return (("momentum_buffer", self._v, self._has_v),)

    def state_dict(self):
        self._export_state(self._state_pairs())
        try:
            return super().state_dict()
        finally:
            self.state.clear()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._import_state(self._state_pairs())
        self._run_cache.clear()
        self._snapshot = None


class FlatAdam(_FlatOptimizer):
    """torch.optim.Adam / AdamW (no amsgrad) on the flat space -- what BaseTrainer.init_optimizer
    builds for SOLVER.TYPE "Adam" / "AdamW" (mdistiller/engine/trainer.py:92-107), plus the sharded
    gradient sync and optional global-norm clipping.  State layout in checkpoints follows torch:
    ``step``, ``exp_avg``, ``exp_avg_sq`` per parameter."""

    _bucket_names = ("grad",)

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, *, decoupled=False,
                 comm: Optional[ShardComm] = None, kernels=None):
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if eps < 0.0:
            raise ValueError("Invalid epsilon value: {}".format(eps))
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError("Invalid beta parameter at index 0: {}".format(betas[0]))
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError("Invalid beta parameter at index 1: {}".format(betas[1]))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults, comm, kernels)
        self.decoupled = bool(decoupled)
        self._m = self._space.new_shard_buffer()
        self._v = self._space.new_shard_buffer()
        self._steps = [0] * len(self._space.params)     # per parameter, like torch's state['step']
        self._amp_pending = None
        self.max_grad_norm = 0.0

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        group = self.param_groups[0]
        members = self._collect("grad")
        work = self._comm.reduce_scatter_sum(self._buckets["grad"], async_op=True)
        if work is not None:
            work.wait()
        # the previous step's overflow verdict, copied to pinned memory back then: a skipped update must not count as
        # an Adam step (GradScaler would not have called optimizer.step() at all)
        self._settle_amp()
        clip = self._step_multiplier(self._buckets["grad"], self.max_grad_norm)
        if self.loss_scaler is not None:
            host = torch.empty(1, dtype=torch.float32, pin_memory=True)
            host.copy_(clip, non_blocking=True)
            done = torch.cuda.Event()
            done.record()
            self._amp_pending = (done, host, [i for i, t in enumerate(members) if t])
        for i, t in enumerate(members):
            if t:
                self._steps[i] += 1
        modes = [((_TASK if t else _NONE), self._steps[i], False) for i, t in enumerate(members)]
        sp = self._space
        b1, b2 = group["betas"]
        for lo, hi, (_, step, _) in self._runs(modes):
            a, b = lo - sp.shard_lo, hi - sp.shard_lo
            self._k.adam_step(sp.flat[lo:hi], self._buckets["grad"][lo:hi], self._m[a:b], self._v[a:b],
                              lr=group["lr"], beta1=b1, beta2=b2, eps=group["eps"],
                              weight_decay=group["weight_decay"], decoupled=self.decoupled, step=step,
                              grad_scale=1.0 / self._comm.world_size, grad_scale_dev=clip)
        self._comm.all_gather(sp.flat)
        self._n_steps += 1
        self._snapshot = None
        return loss

    def _settle_amp(self):
        if self._amp_pending is not None:
            done, host, who = self._amp_pending
            done.synchronize()
            if float(host.item()) < 0:
                for i in who:
                    self._steps[i] -= 1
            self._amp_pending = None

    def _state_pairs(self):
        self._settle_amp()
        has = [n > 0 for n in self._steps]
        return (("exp_avg", self._m, has), ("exp_avg_sq", self._v, list(has)))

    def state_dict(self):
        self._export_state(self._state_pairs())
        for p, n in zip(self._space.params, self._steps):
            if n > 0:
                self.state[p]["step"] = torch.tensor(float(n))
        try:
            return super().state_dict()
        finally:
            self.state.clear()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._steps = [int(self.state.get(p, {}).get("step", 0)) for p in self._space.params]
        has = [False] * len(self._steps)
        self._import_state((("exp_avg", self._m, has), ("exp_avg_sq", self._v, list(has))))
        self._run_cache.clear()
        self._snapshot = None
