"""Sharded step over NCCL on real GPUs (needs >= 2 devices; skipped on a 1-GPU box).

Oracle for N > 1: the single-process reference fed rank-averaged gradients.
"""

import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _cases
import oracle

pytestmark = pytest.mark.gpu


def _need_gpus(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs, have {torch.cuda.device_count()}")


def _make_comm(mode):
    from mdistiller_ddp_b200.comm import ShardComm, SymmetricShardComm
    if mode == "nccl":
        return ShardComm()
    mode, _, sync = mode.partition("+")        # "+barrier": bracket the kernel with barrier launches instead of the sync pad
    return SymmetricShardComm(fuse_kd=(mode in ("fused_kd", "fused_kd_p2p")), use_multicast=not mode.endswith("p2p"),
                              barrier_timeout_ms=15000, in_kernel_sync=(sync != "barrier"))


def _opt_worker(rank, world, port, case_idx, out, mode="nccl"):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import mdistiller_ddp_b200 as b200
        case = _cases.DOT_CASES[case_idx]
        params = [torch.nn.Parameter((p.clone() + (0.5 if rank else 0.0)).cuda()) for p in _cases.make_params(case)]
        opt = b200.DistillationOrientedTrainer(params, lr=case.lr, momentum=case.momentum - case.delta,
                                               momentum_kd=case.momentum + case.delta, weight_decay=case.wd,
                                               comm=_make_comm(mode))
        assert opt.world_size == world
        shapes = [p.shape for p in params]
        # per-rank gradient factors with mean 1 over ranks
        f_kd = 1.0 + 0.25 * (rank - (world - 1) / 2)
        f_task = 1.0 - 0.5 * (rank - (world - 1) / 2)
        for step in range(12):
            if step == 6:
                opt.param_groups[0]["lr"] = case.lr * 0.1
            kd, task = _cases.make_grads(case, step, shapes)
            opt.zero_grad()
            for p, g in zip(params, kd):
                p.grad = None if g is None else (g * f_kd).cuda()
            opt.step_kd()
            opt.zero_grad()
            for p, g in zip(params, task):
                p.grad = None if g is None else (g * f_task).cuda()
            opt.step()
        flat = opt.flat_params.clone()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        assert all(torch.equal(g, flat) for g in gathered), "ranks diverged after the all-gather"
        opt.sync_checkpoint_state()      # collective; afterwards the master alone checkpoints, as the reference does
        if rank == 0:
            sd = opt.state_dict()
            torch.save({"p": [p.detach().cpu() for p in params],
                        "state": {k: {n: t.cpu() for n, t in v.items()} for k, v in sd["state"].items()}}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,case_idx,mode", [
    (2, 0, "nccl"), (2, 2, "nccl"), (4, 6, "nccl"), (8, 6, "nccl"),
    # one fused kernel over NVLink symmetric memory: NVLS multicast, both buckets in-switch, plain peer access
    (2, 0, "fused"), (2, 6, "fused_kd"), (2, 6, "fused_p2p"), (2, 2, "fused"),
    (2, 6, "fused_p2p+barrier"), (2, 6, "fused_kd+barrier"),
    (4, 6, "fused_kd"), (4, 0, "fused_p2p"),
    (8, 6, "fused"), (8, 6, "fused_kd"), (8, 6, "fused_kd_p2p"), (8, 6, "fused_kd+barrier"),
])
def test_sharded_dot_step_matches_oracle_on_rank_averaged_gradients(world, case_idx, mode, tmp_path):
    _need_gpus(world)
    out = str(tmp_path / "r0.pt")
    port = 29600 + (os.getpid() % 1500) + world * 3 + case_idx + 17 * len(mode)
    mp.spawn(_opt_worker, args=(world, port, case_idx, out, mode), nprocs=world, join=True)
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
        assert oracle.max_rel_err(p, q) <= 1e-5
    for i in range(len(params)):
        if ref.buf_kd[i] is not None:
            assert oracle.max_rel_err(got["state"][i]["momentum_kd_buffer"], ref.buf_kd[i]) <= 1e-5
        if ref.buf_task[i] is not None:
            assert oracle.max_rel_err(got["state"][i]["momentum_buffer"], ref.buf_task[i]) <= 1e-5


def _sgd_worker(rank, world, port, case_idx, out, mode):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import mdistiller_ddp_b200 as b200
        case = _cases.DOT_CASES[case_idx]
        params = [torch.nn.Parameter((p.clone() + (0.5 if rank else 0.0)).cuda()) for p in _cases.make_params(case)]
        opt = b200.FlatSGD(params, lr=case.lr, momentum=case.momentum, weight_decay=case.wd, comm=_make_comm(mode))
        shapes = [p.shape for p in params]
        f = 1.0 - 0.5 * (rank - (world - 1) / 2)          # per-rank gradient factor, mean 1 over ranks
        fused_steps = 0
        for step in range(12):
            if step == 6:
                opt.param_groups[0]["lr"] = case.lr * 0.1
            _, task = _cases.make_grads(case, step, shapes)
            opt.zero_grad()
            for p, g in zip(params, task):
                p.grad = (g * f).cuda()
            opt.last_fused = None
            opt.step()
            fused_steps += opt.last_fused is not None
        assert fused_steps == (0 if mode == "nccl" else 12), "the fused kernel must have served every step"
        flat = opt.flat_params.clone()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        assert all(torch.equal(g, flat) for g in gathered), "ranks diverged"
        opt.sync_checkpoint_state()
        if rank == 0:
            sd = opt.state_dict()
            torch.save({"p": [p.detach().cpu() for p in params],
                        "v": [sd["state"][i]["momentum_buffer"].cpu() for i in range(len(params))]}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,mode", [(2, "nccl"), (2, "fused_p2p"), (2, "fused"), (2, "fused_p2p+barrier"),
                                        (4, "fused_p2p"), (8, "fused"), (8, "fused_p2p")])
def test_sharded_sgd_step_matches_torch_sgd_on_rank_averaged_gradients(world, mode, tmp_path):
    """Base trainer's data-parallel step (BaseTrainer + DDP all-reduce + optim.SGD, trainer.py:86-91,303-321): NCCL
    launches and the fused NVLink kernel against torch.optim.SGD fed the rank-averaged gradient."""
    _need_gpus(world)
    case_idx = 6
    out = str(tmp_path / "r0.pt")
    port = 29100 + (os.getpid() % 300) + world * 7 + len(mode)
    mp.spawn(_sgd_worker, args=(world, port, case_idx, out, mode), nprocs=world, join=True)
    got = torch.load(out, weights_only=False)
    case = _cases.DOT_CASES[case_idx]
    params = [torch.nn.Parameter(p.clone()) for p in _cases.make_params(case)]
    ref = torch.optim.SGD(params, lr=case.lr, momentum=case.momentum, weight_decay=case.wd)
    shapes = [p.shape for p in params]
    for step in range(12):
        if step == 6:
            ref.param_groups[0]["lr"] = case.lr * 0.1
        _, task = _cases.make_grads(case, step, shapes)
        for p, g in zip(params, task):
            p.grad = g.clone()
        ref.step()
    for p, q, v in zip(got["p"], params, got["v"]):
        assert oracle.max_rel_err(p, q.detach()) <= 1e-5
        assert oracle.max_rel_err(v, ref.state[q]["momentum_buffer"]) <= 1e-5


def _trainer_worker(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), LOCAL_RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import mdistiller_ddp_b200 as b200
        from mdistiller_ddp_b200 import workloads as W
        from torch.nn.parallel import DistributedDataParallel as DDP
        torch.manual_seed(0)
        teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
        cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot")
        distiller = b200.DKD(student, teacher, cfg).cuda().train()
        wrapped = DDP(distiller, device_ids=[rank], find_unused_parameters=True)   # as tools/train.py:86
        tr = b200.StandaloneTrainer(wrapped, cfg, kind="dot")
        meters = b200.new_train_meters()
        for step in range(3):
            batch = W.synthetic_batch(16, 32, 100, seed=100 * step + rank)
            msg = tr.train_iter(batch, 2, meters)
        tr.flush_metrics(meters)
        flat = tr.optimizer.flat_params
        other = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(other, flat.clone())
        assert all(torch.equal(o, flat) for o in other)
        assert torch.isfinite(flat).all() and meters["top1"].count == 3 * 16 * world, msg
    finally:
        dist.destroy_process_group()


def test_trainer_under_ddp_wrapper_keeps_ranks_in_sync():
    _need_gpus(2)
    mp.spawn(_trainer_worker, args=(2, 29900 + os.getpid() % 90), nprocs=2, join=True)


def test_one_process_driving_two_gpus_gets_per_device_kernel_setup():
    """The dynamic shared-memory opt-in and the occupancy look-ups of the staged loss kernels are per
    device: a process that used cuda:0 first must still be able to launch them on cuda:1 (both loss
    entry points; 100- and 1000-class staged shapes need > 48 KB of shared memory)."""
    _need_gpus(2)
    import ctypes
    from mdistiller_ddp_b200 import _lib, ops
    lib = _lib.load()
    for dev in (0, 1, 0):
        with torch.cuda.device(dev):
            for (B, C, T, beta) in ((2048, 100, 4.0, 8.0), (512, 1000, 1.0, 0.5), (300, 128, 2.0, 6.0)):
                g = torch.Generator().manual_seed(B + C)
                zs, zt = torch.randn(B, C, generator=g) * 3, torch.randn(B, C, generator=g) * 3
                y = torch.randint(0, C, (B,), generator=g)
                l_ref, g_ref = oracle.dkd_closed_form(zs.numpy(), zt.numpy(), y.numpy(), 1.0, beta, T)
                zs, zt, y = zs.cuda(dev), zt.cuda(dev), y.cuda(dev)
                lp, gk, _ = ops.distill_loss(zs, zt, y, kind=ops.LOSS_DKD, alpha=1.0, beta=beta, temperature=T)
                assert abs(lp[0].item() - l_ref) <= 1e-5 * abs(l_ref) and oracle.max_rel_err(gk, g_ref) <= 1e-5
                gk2, loss2 = torch.empty_like(zs), torch.empty(2, device=zs.device)
                rc = lib.mdb200_distill_loss_static_scratch(
                    zs.data_ptr(), zt.data_ptr(), y.data_ptr(), B, C, C, C, ops.LOSS_DKD, 1.0, beta, T, 1.0,
                    ops.CE_NONE, 1.0, gk2.data_ptr(), None, C, loss2.data_ptr(),
                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
                assert rc == 0, lib.mdb200_error_string(rc)
                torch.cuda.synchronize()
                assert torch.equal(gk2, gk) and torch.equal(loss2[:1], lp[:1])

