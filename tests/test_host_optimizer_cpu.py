"""Host logic of the flat optimizers on CPU, kernels replaced by the oracle test double.

Covers: flat layout/views, membership runs (dot.py:36-39,49-54 rules), first-step flags,
lr mutation through param_groups, state_dict layout round trip, ValueError contract, and
the world_size-2 sharded step over gloo (reduce-scatter -> shard update -> all-gather).
"""

import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _cases
import oracle
from _kernels_cpu import OracleKernels

import mdistiller_ddp_b200 as b200
from mdistiller_ddp_b200.comm import ShardComm


def _flat(ts, like):
    return torch.cat([(torch.zeros_like(l) if t is None else t.detach()).reshape(-1) for t, l in zip(ts, like)]).numpy()


def _drive_dot(case, opt, params, steps=None):
    shapes = [p.shape for p in params]
    for step in range(case.steps if steps is None else steps):
        if step == case.lr_drop_at:
            opt.param_groups[0]["lr"] = case.lr * 0.1
        kd, task = _cases.make_grads(case, step, shapes)
        opt.zero_grad(set_to_none=True)
        for p, g in zip(params, kd):
            p.grad = None if g is None else g.clone()
        opt.step_kd()
        opt.zero_grad(set_to_none=True)
        for p, g in zip(params, task):
            p.grad = None if g is None else g.clone()
        opt.step()


@pytest.mark.parametrize("case", [c for c in _cases.DOT_CASES if c.kind == "dot" and len(c.shapes) < 20],
                         ids=lambda c: c.name)
def test_flat_dot_optimizer_reproduces_reference_trajectory(case, dot_golden):
    """Foreign-gradient protocol (p.grad assigned by the caller, None = no gradient)."""
    params = [torch.nn.Parameter(p.clone()) for p in _cases.make_params(case)]
    k = OracleKernels()
    opt = b200.DistillationOrientedTrainer(params, lr=case.lr, momentum=case.momentum - case.delta,
                                           momentum_kd=case.momentum + case.delta, dampening=case.dampening,
                                           weight_decay=case.wd, kernels=k)
    _drive_dot(case, opt, params)
    got = _flat(params, params)
    ref = dot_golden[f"{case.name}/p32"]
    assert oracle.max_rel_err(got, ref) <= 1e-6
    vt, vk = opt.momentum_views()
    assert oracle.max_rel_err(_flat(vt, params), dot_golden[f"{case.name}/vt32"]) <= 1e-6
    assert oracle.max_rel_err(_flat(vk, params), dot_golden[f"{case.name}/vk32"]) <= 1e-6
    if not (case.kd_missing or case.task_missing or case.kd_from_step):
        assert all(c[1] >= sum(p.numel() for p in params) for c in k.calls), "uniform membership => one launch/step"
        assert len(k.calls) == case.steps


def test_autograd_accumulates_into_the_flat_bucket_and_membership_follows_usage():
    torch.manual_seed(0)
    a = torch.nn.Parameter(torch.randn(8, 4))
    b = torch.nn.Parameter(torch.randn(4))
    unused = torch.nn.Parameter(torch.randn(12))
    k = OracleKernels()
    opt = b200.DistillationOrientedTrainer([a, unused, b], lr=0.1, momentum=0.8, momentum_kd=0.95,
                                           weight_decay=0.01, kernels=k)
    flat = opt.flat_params
    assert a.data_ptr() == flat.data_ptr() and unused.data_ptr() == flat[32:].data_ptr()
    x = torch.randn(3, 8)
    before = unused.detach().clone()
    for _ in range(3):
        out = (x @ a + b)
        loss_kd, loss_ce = out.pow(2).mean(), out.abs().mean()
        opt.zero_grad(set_to_none=True)
        assert a.grad.data_ptr() == opt._buckets["kd"].data_ptr()
        loss_kd.backward(retain_graph=True)
        assert a.grad.data_ptr() == opt._buckets["kd"].data_ptr(), "autograd must accumulate in place"
        opt.step_kd()
        opt.zero_grad(set_to_none=True)
        loss_ce.backward()
        opt.step()
    assert torch.equal(unused.detach(), before), "a parameter no backward touched is skipped (dot.py:112,144)"
    assert any(c[1] == 32 for c in k.calls) and any(c[1] == 4 for c in k.calls)


def test_hyperparameter_validation_matches_reference():
    p = [torch.nn.Parameter(torch.zeros(4))]
    for kw in (dict(lr=-1.0), dict(lr=0.1, momentum=-0.1), dict(lr=0.1, momentum_kd=-0.1),
               dict(lr=0.1, weight_decay=-1e-4)):
        with pytest.raises(ValueError):
            b200.DistillationOrientedTrainer(p, kernels=OracleKernels(), **kw)
    with pytest.raises(ValueError):
        b200.FlatSGD(p, lr=-0.1, kernels=OracleKernels())


def test_state_dict_has_the_reference_layout_and_round_trips():
    case = _cases.DOT_CASES[0]
    params = [torch.nn.Parameter(p.clone()) for p in _cases.make_params(case)]
    opt = b200.DistillationOrientedTrainer(params, lr=case.lr, momentum=0.825, momentum_kd=0.975,
                                           weight_decay=case.wd, kernels=OracleKernels())
    _drive_dot(case, opt, params, steps=5)
    sd = opt.state_dict()
    assert set(sd["param_groups"][0]) >= {"lr", "momentum", "momentum_kd", "dampening", "weight_decay", "params"}
    assert set(sd["state"][0]) == {"momentum_buffer", "momentum_kd_buffer"}      # dot.py:165-170
    assert sd["state"][3]["momentum_buffer"].shape == params[3].shape
    # load into a fresh optimizer and continue: identical trajectory
    params2 = [torch.nn.Parameter(p.detach().clone()) for p in params]
    opt2 = b200.DistillationOrientedTrainer(params2, lr=1.0, momentum=0.1, momentum_kd=0.1, kernels=OracleKernels())
    opt2.load_state_dict(sd)
    assert opt2.param_groups[0]["lr"] == case.lr and opt2.param_groups[0]["momentum_kd"] == 0.975
    shapes = [p.shape for p in params]
    for o, ps in ((opt, params), (opt2, params2)):
        kd, task = _cases.make_grads(case, 5, shapes)
        o.zero_grad()
        for p, g in zip(ps, kd):
            p.grad = g.clone()
        o.step_kd()
        o.zero_grad()
        for p, g in zip(ps, task):
            p.grad = g.clone()
        o.step()
    for p, q in zip(params, params2):
        assert torch.equal(p, q)


def test_reference_optimizer_state_dict_loads_into_ours():
    """Checkpoint interoperability: a state_dict written by the reference class."""
    from oracle import ref_loader
    if not ref_loader.available():
        pytest.skip("reference tree not mounted")
    case = _cases.DOT_CASES[0]
    dotmod = ref_loader.dot_module()
    ref_params = [p.clone().requires_grad_(True) for p in _cases.make_params(case)]
    ref_opt = dotmod.DistillationOrientedTrainer(ref_params, lr=case.lr, momentum=0.825, momentum_kd=0.975,
                                                 weight_decay=case.wd)
    ours = [torch.nn.Parameter(p.detach().clone()) for p in ref_params]
    our_opt = b200.DistillationOrientedTrainer(ours, lr=case.lr, momentum=0.825, momentum_kd=0.975,
                                               weight_decay=case.wd, kernels=OracleKernels())
    shapes = [p.shape for p in ours]

    def one(o, ps, step):
        kd, task = _cases.make_grads(case, step, shapes)
        o.zero_grad(set_to_none=True)
        for p, g in zip(ps, kd):
            p.grad = g.clone()
        o.step_kd()
        o.zero_grad(set_to_none=True)
        for p, g in zip(ps, task):
            p.grad = g.clone()
        o.step()

    for s in range(3):
        one(ref_opt, ref_params, s)
    with torch.no_grad():
        for p, q in zip(ours, ref_params):
            p.copy_(q)
    our_opt.load_state_dict(ref_opt.state_dict())
    one(ref_opt, ref_params, 3)
    one(our_opt, ours, 3)
    for p, q in zip(ours, ref_params):
        assert torch.equal(p.detach(), q.detach())
    # and back: ours -> reference
    ref_opt.load_state_dict(our_opt.state_dict())
    one(ref_opt, ref_params, 4)
    one(our_opt, ours, 4)
    for p, q in zip(ours, ref_params):
        assert torch.equal(p.detach(), q.detach())


def _make_base_optimizer(case, params, **kw):
    if case.kind == "sgd":
        opt = b200.FlatSGD(params, lr=case.lr, momentum=case.momentum, weight_decay=case.wd,
                           dampening=case.dampening, **kw)
    else:
        opt = b200.FlatAdam(params, lr=case.lr, betas=_cases.ADAM_BETAS, eps=_cases.ADAM_EPS, weight_decay=case.wd,
                            decoupled=(case.kind == "adamw"), **kw)
    opt.max_grad_norm = case.clip
    return opt


@pytest.mark.parametrize("case", [c for c in _cases.DOT_CASES if c.kind != "dot" and len(c.shapes) < 20],
                         ids=lambda c: c.name)
def test_flat_base_optimizers_reproduce_torch(case, dot_golden):
    """SGD / Adam / AdamW (+ global-norm clipping) as the base trainer drives them (trainer.py:86-107,316-321)."""
    params = [torch.nn.Parameter(p.clone()) for p in _cases.make_params(case)]
    opt = _make_base_optimizer(case, params, kernels=OracleKernels())
    shapes = [p.shape for p in params]
    for step in range(case.steps):
        if step == case.lr_drop_at:
            opt.param_groups[0]["lr"] = case.lr * 0.1
        _, task = _cases.make_grads(case, step, shapes)
        opt.zero_grad()
        for p, g in zip(params, task):
            p.grad = None if g is None else g.clone()
        opt.step()
    assert oracle.max_rel_err(_flat(params, params), dot_golden[f"{case.name}/p32"]) <= 2e-6
    if case.kind != "sgd":       # checkpoint layout follows torch.optim.Adam: step / exp_avg / exp_avg_sq
        sd = opt.state_dict()
        assert set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"} and float(sd["state"][0]["step"]) == case.steps
        assert 4 not in sd["state"] or 4 not in case.task_missing
        ref = torch.optim.AdamW(params, lr=1.0) if case.kind == "adamw" else torch.optim.Adam(params, lr=1.0)
        ref.load_state_dict(sd)                      # torch's own class accepts our checkpoint
        assert ref.param_groups[0]["lr"] == pytest.approx(case.lr * 0.1)
        opt2 = _make_base_optimizer(case, [torch.nn.Parameter(p.detach().clone()) for p in params], kernels=OracleKernels())
        opt2.load_state_dict(sd)
        assert opt2._steps == opt._steps and torch.equal(opt2._m, opt._m) and torch.equal(opt2._v, opt._v)


# ---------------------------------------------------------------------------- world_size 2, gloo

class GlooShardComm(ShardComm):
    """gloo has no reduce-scatter: emulate the two tensor collectives with all_reduce /
    all_gather so the sharding logic above them is what gets tested."""

    def reduce_scatter_sum(self, full, async_op=False):
        dist.all_reduce(full, group=self.group)
        return None

    def all_gather(self, full, async_op=False):
        n = full.numel() // self.world_size
        parts = [torch.empty(n, dtype=full.dtype) for _ in range(self.world_size)]
        dist.all_gather(parts, full[self.rank * n:(self.rank + 1) * n].clone(), group=self.group)
        full.copy_(torch.cat(parts))
        return None


def _worker(rank, world, port, case_idx, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        case = _cases.DOT_CASES[case_idx]
        params = [torch.nn.Parameter(p.clone() + (0.5 if rank else 0.0)) for p in _cases.make_params(case)]
        opt = b200.DistillationOrientedTrainer(params, lr=case.lr, momentum=case.momentum - case.delta,
                                               momentum_kd=case.momentum + case.delta, weight_decay=case.wd,
                                               comm=GlooShardComm(), kernels=OracleKernels())
        shapes = [p.shape for p in params]
        for step in range(12):
            if step == 6:
                opt.param_groups[0]["lr"] = case.lr * 0.1
            kd, task = _cases.make_grads(case, step, shapes)
            # rank-dependent gradients whose mean over ranks is the single-process gradient
            sign = 1.0 if rank == 0 else -1.0
            opt.zero_grad()
            for i, (p, g) in enumerate(zip(params, kd)):
                p.grad = None if g is None else g * (1.0 + 0.25 * sign)
            opt.step_kd()
            opt.zero_grad()
            for p, g in zip(params, task):
                p.grad = None if g is None else g * (1.0 - 0.5 * sign)
            opt.step()
        # the reference checkpoints on the master rank only (trainer.py:230-236): without the collective
        # snapshot that must be a clear error, never a one-sided collective (= a hang)
        with pytest.raises(RuntimeError, match="sync_checkpoint_state"):
            opt.state_dict()
        opt.sync_checkpoint_state()          # every rank (the hosted trainer does it after the epoch's last step)
        if rank == 0:                        # ... and then rank 0 ALONE asks for the checkpoint
            sd = opt.state_dict()
            torch.save({"p": [p.detach().clone() for p in params], "sd": sd,
                        "shard": (opt._space.shard_lo, opt._space.shard_hi, opt._space.total)}, out)
        # every rank must hold identical parameters after the all-gather
        flat = opt.flat_params.clone()
        other = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(other, flat)
        assert all(torch.equal(o, flat) for o in other)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case_idx", [0, 2])
def test_sharded_step_world2_matches_single_process_on_averaged_gradients(case_idx, tmp_path):
    """the multi-rank oracle is the single-process reference fed rank-averaged
    gradients.  Rank 0's initial parameters win (broadcast), momentum is sharded."""
    port = 29500 + (os.getpid() % 2000) + case_idx
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, port, case_idx, out), nprocs=2, join=True)
    got = torch.load(out, weights_only=False)
    case = _cases.DOT_CASES[case_idx]
    params = [p.clone() for p in _cases.make_params(case)]
    ref = oracle.DotListOracle(params, lr=case.lr, momentum=case.momentum - case.delta,
                               momentum_kd=case.momentum + case.delta, weight_decay=case.wd)
    shapes = [p.shape for p in params]
    for step in range(12):
        if step == 6:
            ref.hp["lr"] = case.lr * 0.1
        kd, task = _cases.make_grads(case, step, shapes)
        ref.step_kd(kd)
        ref.step(task)
    for p, q in zip(got["p"], params):
        assert oracle.max_rel_err(p, q) <= 2e-6
    lo, hi, total = got["shard"]
    assert hi - lo == total // 2 and total % 8 == 0
    # gathered checkpoint has full-size per-parameter buffers
    st = got["sd"]["state"]
    for i, q in enumerate(params):
        if ref.buf_kd[i] is not None:
            assert oracle.max_rel_err(st[i]["momentum_kd_buffer"], ref.buf_kd[i]) <= 2e-6
        if ref.buf_task[i] is not None:
            assert oracle.max_rel_err(st[i]["momentum_buffer"], ref.buf_task[i]) <= 2e-6


class _RecordingFusedComm(ShardComm):
    """Pretends to be rank `rank` of 2 with the fused NVLink step available; records which path ran."""
    fused, fuse_kd = True, False

    def __init__(self, rank):
        self.group, self.world_size, self.rank = None, 2, rank
        self.path = []

    def reduce_scatter_sum(self, full, async_op=False):
        self.path.append("reduce_scatter")
        return None

    def all_gather(self, full, async_op=False):
        self.path.append("all_gather")
        return None

    def broadcast(self, full, src=0):
        pass

    def addresses(self, t):
        return 0, [t.data_ptr(), t.data_ptr()]

    def barrier(self, t, channel=0):
        self.path.append("barrier")

    def next_sync(self):
        return None


class _FusedRecorder(OracleKernels):
    def dot_step_fused(self, *a, **k):
        self.calls.append(("fused",))


def test_fused_or_not_is_decided_from_global_membership_so_every_rank_agrees():
    """A parameter without a KD gradient sits in rank 1's shard only.  Rank 0's own shard looks uniform, but
    it must still take the unfused path, because rank 1 cannot enter the fused rendezvous."""
    took = []
    for rank in (0, 1):
        params = [torch.nn.Parameter(torch.randn(64)), torch.nn.Parameter(torch.randn(8))]
        comm, k = _RecordingFusedComm(rank), _FusedRecorder()
        opt = b200.DistillationOrientedTrainer(params, lr=0.1, momentum=0.8, momentum_kd=0.95, comm=comm, kernels=k)
        assert (opt._space.shard_lo, opt._space.shard_hi) == ((0, 36) if rank == 0 else (36, 72))
        for has_kd_on_small in (False, True, True):     # 2nd pass: first-step flags still differ; 3rd: uniform
            comm.path.clear()
            k.calls.clear()
            opt.zero_grad()
            params[0].grad = torch.randn(64)
            params[1].grad = torch.randn(8) if has_kd_on_small else None
            opt.step_kd()
            opt.zero_grad()
            params[0].grad, params[1].grad = torch.randn(64), torch.randn(8)
            opt.step()
            took.append((rank, has_kd_on_small, ("fused",) in k.calls))
    assert [t[2] for t in took] == [False, False, True, False, False, True], took


def _base_worker(rank, world, port, case_idx, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        case = _cases.DOT_CASES[case_idx]
        params = [torch.nn.Parameter(p.clone()) for p in _cases.make_params(case)]
        opt = _make_base_optimizer(case, params, comm=GlooShardComm(), kernels=OracleKernels())
        shapes = [p.shape for p in params]
        sign = 1.0 if rank == 0 else -1.0
        for step in range(15):
            _, task = _cases.make_grads(case, step, shapes)
            opt.zero_grad()
            for p, g in zip(params, task):
                p.grad = None if g is None else g * (1.0 + 0.5 * sign)     # mean over the two ranks = g
            opt.step()
        if rank == 0:
            torch.save([p.detach().clone() for p in params], out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["sgd_clip_small", "adamw_clip_partial"])
def test_sharded_base_optimizers_with_clipping_world2(name, tmp_path):
    """The clipping norm is global: per-shard sums of squares of the rank-averaged gradient, all-reduced."""
    case_idx = [c.name for c in _cases.DOT_CASES].index(name)
    case = _cases.DOT_CASES[case_idx]
    out = str(tmp_path / "r0.pt")
    mp.spawn(_base_worker, args=(2, 29300 + (os.getpid() % 500) + case_idx, case_idx, out), nprocs=2, join=True)
    got = torch.load(out, weights_only=False)
    params = [torch.nn.Parameter(p.clone()) for p in _cases.make_params(case)]
    ref = _make_base_optimizer(case, params, kernels=OracleKernels())
    shapes = [p.shape for p in params]
    for step in range(15):
        _, task = _cases.make_grads(case, step, shapes)
        ref.zero_grad()
        for p, g in zip(params, task):
            p.grad = None if g is None else g.clone()
        ref.step()
    for p, q in zip(got, params):
        assert oracle.max_rel_err(p, q) <= 2e-6


def test_flat_space_layout_and_chunk_table_invariants():
    """Every tensor starts on a 16-byte boundary, shards are equal and 16-byte aligned, and the gather
    chunk table tiles every tensor exactly once into its slice of the bucket."""
    import random
    from mdistiller_ddp_b200.flat import FlatSpace
    rng = random.Random(0)
    for world in (1, 2, 4, 8):
        shapes = [tuple(rng.randint(1, 9) for _ in range(rng.randint(1, 3))) for _ in range(rng.randint(1, 12))]
        shapes.append((40000,))
        params = [torch.nn.Parameter(torch.randn(*s)) for s in shapes]
        before = [p.detach().clone() for p in params]
        sp = FlatSpace(params, world_size=world, rank=world - 1)
        assert sp.total % (4 * world) == 0 and sp.shard_len * world == sp.total and sp.shard_hi == sp.total
        assert all(o % 4 == 0 for o in sp.offsets) and sp.shard_lo % 4 == 0
        for p, b, o in zip(params, before, sp.offsets):
            assert torch.equal(p, b) and p.data_ptr() == sp.flat[o:].data_ptr()      # values kept, storage moved
        table = sp.chunk_table(chunk_elems=1000).tolist()
        covered = torch.zeros(sp.total, dtype=torch.int32)
        for t, n, src, dst in table:
            assert 0 < n <= 1000 and dst == sp.offsets[t] + src and src + n <= params[t].numel()
            covered[dst:dst + n] += 1
        for o, e, p in zip(sp.offsets, sp.ends, params):
            assert bool((covered[o:o + p.numel()] == 1).all()) and bool((covered[o + p.numel():e] == 0).all())
        lo, hi = sp.clip_to_shard(0, sp.total)
        assert (lo, hi) == (sp.shard_lo, sp.shard_hi)
