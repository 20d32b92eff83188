"""Parity of the DOT / SGD flat update kernels and of the optimizer built on them -- B200 only.

Bar (BASELINE.json north_star): parameters after 100 DOT steps within rel 1e-5 of the
reference, rel = max|a-b| / max|b| per tensor.  Calls go through the C-ABI.
"""

import numpy as np
import pytest
import torch

import _cases
import oracle

pytestmark = pytest.mark.gpu

TOL = 1e-5


@pytest.fixture(scope="module")
def b200():
    import mdistiller_ddp_b200 as m
    return m


def _flat(ts, like):
    return torch.cat([(torch.zeros_like(l) if t is None else t.detach()).reshape(-1)
                      for t, l in zip(ts, like)]).double().cpu().numpy()


def _ref(G, name, key, arr):
    if f"{name}/{key}" in G:
        return arr, G[f"{name}/{key}"]
    return arr[_cases.sample_index(arr.size)], G[f"{name}/{key}_s"]


@pytest.mark.parametrize("case", [c for c in _cases.DOT_CASES if c.kind == "dot"], ids=lambda c: c.name)
def test_100_step_trajectory_matches_reference(case, dot_golden, b200):
    params = [torch.nn.Parameter(p.clone().cuda()) for p in _cases.make_params(case)]
    opt = b200.DistillationOrientedTrainer(params, lr=case.lr, momentum=case.momentum - case.delta,
                                           momentum_kd=case.momentum + case.delta, dampening=case.dampening,
                                           weight_decay=case.wd)
    shapes = [p.shape for p in params]
    for step in range(case.steps):
        if step == case.lr_drop_at:
            opt.param_groups[0]["lr"] = case.lr * 0.1       # what adjust_learning_rate does (engine/utils.py:91-92)
        kd, task = _cases.make_grads(case, step, shapes)
        opt.zero_grad(set_to_none=True)
        for p, g in zip(params, kd):
            p.grad = None if g is None else g.cuda()
        opt.step_kd()
        opt.zero_grad(set_to_none=True)
        for p, g in zip(params, task):
            p.grad = None if g is None else g.cuda()
        opt.step()
    vt, vk = opt.momentum_views()
    for key, arr in (("p", _flat(params, params)), ("vt", _flat(vt, params)), ("vk", _flat(vk, params))):
        for tag in ("64", "32"):
            got, ref = _ref(dot_golden, case.name, key + tag, arr)
            assert oracle.max_rel_err(got, ref) <= TOL, (case.name, key, tag)


def _dot_steps(case, opt, params, steps):
    shapes = [p.shape for p in params]
    for step in steps:
        if step == case.lr_drop_at:
            opt.param_groups[0]["lr"] = case.lr * 0.1
        kd, task = _cases.make_grads(case, step, shapes)
        opt.zero_grad(set_to_none=True)
        for p, g in zip(params, kd):
            p.grad = None if g is None else g.cuda()
        opt.step_kd()
        opt.zero_grad(set_to_none=True)
        for p, g in zip(params, task):
            p.grad = None if g is None else g.cuda()
        opt.step()


@pytest.mark.parametrize("name", ["dot_cifar_small", "dot_membership"])
def test_checkpoint_on_cuda_has_the_reference_layout_and_momenta(name, dot_golden, b200):
    """SURVEY 8 f-4 on the GPU: state_dict() after the fixture's 100 steps holds the reference's
    'momentum_buffer' / 'momentum_kd_buffer' per parameter (dot.py:165-170) with the reference's values,
    and a fresh optimizer that loads it continues bit-identically."""
    case = next(c for c in _cases.DOT_CASES if c.name == name)
    mk = lambda ps: b200.DistillationOrientedTrainer(ps, lr=case.lr, momentum=case.momentum - case.delta,
                                                     momentum_kd=case.momentum + case.delta,
                                                     dampening=case.dampening, weight_decay=case.wd)
    params = [torch.nn.Parameter(p.clone().cuda()) for p in _cases.make_params(case)]
    opt = mk(params)
    _dot_steps(case, opt, params, range(case.steps))
    sd = opt.state_dict()
    assert all(t.is_cuda for st in sd["state"].values() for t in st.values())
    for key, gkey in (("momentum_buffer", "vt64"), ("momentum_kd_buffer", "vk64")):
        bufs = [sd["state"].get(i, {}).get(key) for i in range(len(params))]
        missing = case.task_missing if key == "momentum_buffer" else case.kd_missing
        assert [b is None for b in bufs] == [i in missing for i in range(len(params))]
        got, ref = _ref(dot_golden, case.name, gkey, _flat(bufs, params))
        assert oracle.max_rel_err(got, ref) <= TOL
    params2 = [torch.nn.Parameter(p.detach().clone()) for p in params]
    opt2 = mk(params2)
    opt2.load_state_dict(sd)
    assert opt2.param_groups[0]["lr"] == opt.param_groups[0]["lr"]
    for o, ps in ((opt, params), (opt2, params2)):
        _dot_steps(case, o, ps, range(case.steps, case.steps + 3))
    for p, q in zip(params, params2):
        assert torch.equal(p, q)
    sd1, sd2 = opt.state_dict(), opt2.state_dict()
    for i in sd1["state"]:
        for k in sd1["state"][i]:
            assert torch.equal(sd1["state"][i][k], sd2["state"][i][k])


@pytest.mark.parametrize("case", [c for c in _cases.DOT_CASES if c.kind != "dot"], ids=lambda c: c.name)
def test_base_optimizer_trajectories_match_torch(case, dot_golden, b200):
    """SGD / Adam / AdamW with and without global-norm clipping vs torch.optim + clip_grad_norm_
    (what BaseTrainer runs, trainer.py:86-107, 316-321)."""
    params = [torch.nn.Parameter(p.clone().cuda()) for p in _cases.make_params(case)]
    if case.kind == "sgd":
        opt = b200.FlatSGD(params, lr=case.lr, momentum=case.momentum, weight_decay=case.wd, dampening=case.dampening)
    else:
        opt = b200.FlatAdam(params, lr=case.lr, betas=_cases.ADAM_BETAS, eps=_cases.ADAM_EPS, weight_decay=case.wd,
                            decoupled=(case.kind == "adamw"))
    opt.max_grad_norm = case.clip
    shapes = [p.shape for p in params]
    for step in range(case.steps):
        if step == case.lr_drop_at:
            opt.param_groups[0]["lr"] = case.lr * 0.1
        _, task = _cases.make_grads(case, step, shapes)
        opt.zero_grad()
        for p, g in zip(params, task):
            p.grad = None if g is None else g.cuda()
        opt.step()
    got, ref = _ref(dot_golden, case.name, "p64", _flat(params, params))
    assert oracle.max_rel_err(got, ref) <= TOL
    if case.kind != "sgd":
        sp = opt._space
        for key, buf in (("vt64", opt._m), ("vk64", opt._v)):       # exp_avg, exp_avg_sq
            got, ref = _ref(dot_golden, case.name, key, _flat(sp.views(buf), params))
            assert oracle.max_rel_err(got, ref) <= TOL


def test_sum_squares_and_clip_coefficient(b200):
    ops = b200.ops
    for n in (5, 4099, 11_689_512):
        x = torch.randn(n, device="cuda") * 0.3
        want = (x.double() * 0.125).pow(2).sum().item()
        got = ops.sum_squares(x, scale=0.125)
        assert abs(got.item() - want) <= 1e-6 * want
        assert torch.equal(got, ops.sum_squares(x, scale=0.125))          # fixed summation order
        for max_norm in (0.01, 1e6):
            c = ops.clip_coefficient(got, max_norm).item()
            assert abs(c - min(1.0, max_norm / (want ** 0.5 + 1e-6))) <= 1e-6


@pytest.mark.parametrize("n", [11_689_512, 25_557_032, 1_000_003, 5])
def test_one_step_equals_flat_oracle_at_benchmark_sizes(n, b200):
    """ResNet18 / ResNet50 element counts (BASELINE.json configs[4]) plus ragged sizes; later
    steps and the first step; exact up to one rounding of the fused forms."""
    ops = b200.ops
    g = torch.Generator().manual_seed(n)
    p = torch.randn(n, generator=g) * 0.1
    gt, gk = torch.randn(n, generator=g) * 0.1, torch.randn(n, generator=g) * 0.1
    vt, vk = torch.randn(n, generator=g) * 0.1, torch.randn(n, generator=g) * 0.1
    hp = dict(lr=0.2, m_task=0.81, m_kd=0.99, weight_decay=1e-4, dampening=0.0)
    for first in (False, True):
        P, VT, VK = p.clone(), vt.clone(), vk.clone()
        oracle.dot_flat_step(P, gt, gk, VT, VK, first_step=first, **hp)
        dp, dvt, dvk = p.cuda(), vt.cuda(), vk.cuda()
        ops.dot_step(dp, gt.cuda(), gk.cuda(), dvt, dvk, lr=0.2, momentum=0.81, momentum_kd=0.99,
                     weight_decay=1e-4, first_task=first, first_kd=first)
        for a, b in ((dp, P), (dvt, VT), (dvk, VK)):
            assert oracle.max_rel_err(a, b) <= 2e-7
    # unaligned base pointer (scalar path) and the reduce-scatter scale
    P, VT, VK = p[1:].clone(), vt[1:].clone(), vk[1:].clone()
    oracle.dot_flat_step(P, gt[1:], gk[1:], VT, VK, grad_scale=0.125, **hp)
    dp, dvt, dvk = p.cuda(), vt.cuda(), vk.cuda()
    ops.dot_step(dp[1:], gt.cuda()[1:], gk.cuda()[1:], dvt[1:], dvk[1:], lr=0.2, momentum=0.81, momentum_kd=0.99,
                 weight_decay=1e-4, grad_scale=0.125)
    assert oracle.max_rel_err(dp[1:], P) <= 2e-7 and dp[0].item() == p[0].item()
    # single-momentum SGD mode
    P, V = p.clone(), vt.clone()
    oracle.sgd_flat_step(P, gt, V, lr=0.05, momentum=0.9, weight_decay=5e-4)
    dp, dv = p.cuda(), vt.cuda()
    ops.sgd_step(dp, gt.cuda(), dv, lr=0.05, momentum=0.9, weight_decay=5e-4)
    assert oracle.max_rel_err(dp, P) <= 2e-7 and oracle.max_rel_err(dv, V) <= 2e-7


def test_update_is_linear_in_lr_and_idempotent_at_zero(b200):
    """Size-independent properties at the ResNet-50 size: lr = 0 leaves p untouched but advances
    the momenta; a zero gradient with zero momentum and no decay is the identity."""
    ops = b200.ops
    n = 25_557_032
    p = torch.randn(n, device="cuda") * 0.1
    g1, g2 = torch.randn(n, device="cuda"), torch.randn(n, device="cuda")
    vt, vk = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    p0 = p.clone()
    ops.dot_step(p, g1, g2, vt, vk, lr=0.0, momentum=0.8, momentum_kd=0.9, weight_decay=0.0)
    assert torch.equal(p, p0) and torch.equal(vt, g1) and torch.equal(vk, g2)
    z = torch.zeros(n, device="cuda")
    vt.zero_(); vk.zero_()
    ops.dot_step(p, z, z, vt, vk, lr=0.3, momentum=0.8, momentum_kd=0.9, weight_decay=0.0)
    assert torch.equal(p, p0) and vt.abs().max().item() == 0.0


def test_trainer_step_matches_cpu_oracle_end_to_end(b200):
    """forward -> fused tail -> two backwards -> DOT step on a tiny net, against the same net on the
    CPU driven by the oracle loss + oracle optimizer (what the reference's DOT.train_iter computes)."""
    from mdistiller_ddp_b200 import workloads as W
    import copy
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    cpu_t, cpu_s = copy.deepcopy(teacher), copy.deepcopy(student)
    cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot")
    cfg.DISTILLER.TYPE = "DKD"
    dist_gpu = b200.DKD(student, teacher, cfg).cuda().train()
    tr = b200.StandaloneTrainer(dist_gpu, cfg, kind="dot")
    cpu_t.eval(); cpu_s.train()
    cpu_params = [p for _, p in cpu_s.named_parameters()]
    ref_opt = oracle.DotListOracle([p.data for p in cpu_params], lr=cfg.SOLVER.LR,
                                   momentum=cfg.SOLVER.SGD.MOMENTUM - cfg.SOLVER.DOT.DELTA,
                                   momentum_kd=cfg.SOLVER.SGD.MOMENTUM + cfg.SOLVER.DOT.DELTA,
                                   weight_decay=cfg.SOLVER.WEIGHT_DECAY)
    for step in range(4):
        image, target, _ = W.synthetic_batch(32, 32, 100, seed=step)
        epoch = 5
        preds, losses = tr.tail_step(image.cuda(), target.cuda(), epoch)
        zs, _ = cpu_s(image)
        with torch.no_grad():
            zt, _ = cpu_t(image)
        l_ce = cfg.DKD.CE_WEIGHT * oracle.ce_loss(zs, target)
        l_kd = min(epoch / cfg.DKD.WARMUP, 1.0) * oracle.dkd_loss(zs, zt, target, cfg.DKD.ALPHA, cfg.DKD.BETA, cfg.DKD.T)
        g_kd = torch.autograd.grad(l_kd, cpu_params, retain_graph=True, allow_unused=True)
        g_ce = torch.autograd.grad(l_ce, cpu_params, allow_unused=True)
        ref_opt.step_kd(list(g_kd))
        ref_opt.step(list(g_ce))
        assert abs(losses["loss_kd"].item() - l_kd.item()) <= 1e-4 * abs(l_kd.item())
        assert abs(losses["loss_ce"].item() - l_ce.item()) <= 1e-4 * abs(l_ce.item())
    for (name, p), q in zip(student.named_parameters(), cpu_params):
        assert oracle.max_rel_err(p, q) <= 1e-4, name     # cuDNN vs CPU conv rounding dominates here


@pytest.mark.parametrize("kind", ["DKD", "KD"])
def test_base_trainer_step_matches_cpu_oracle(kind, b200):
    """Non-DOT configs (BaseTrainer.train_iter, trainer.py:257-321): summed loss, one backward,
    torch.optim.SGD -- here: combined-gradient tail + FlatSGD."""
    from mdistiller_ddp_b200 import workloads as W
    import copy
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(1)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    cpu_t, cpu_s = copy.deepcopy(teacher).eval(), copy.deepcopy(student).train()
    cfg = W.default_cfg("cifar_res32x4_res8x4_dkd_base")
    cfg.DISTILLER.TYPE = kind
    dist_gpu = b200.distiller_dict[kind](student, teacher, cfg).cuda().train()
    tr = b200.StandaloneTrainer(dist_gpu, cfg, kind="base")
    assert dist_gpu.combined_tail
    ref_opt = torch.optim.SGD(cpu_s.parameters(), lr=cfg.SOLVER.LR, momentum=cfg.SOLVER.SGD.MOMENTUM,
                              weight_decay=cfg.SOLVER.WEIGHT_DECAY)
    for step in range(4):
        image, target, _ = W.synthetic_batch(32, 32, 100, seed=10 + step)
        epoch = 7
        preds, losses = tr.tail_step(image.cuda(), target.cuda(), epoch)
        ref_opt.zero_grad()
        zs, _ = cpu_s(image)
        with torch.no_grad():
            zt, _ = cpu_t(image)
        if kind == "DKD":
            l_ce = cfg.DKD.CE_WEIGHT * oracle.ce_loss(zs, target)
            l_kd = min(epoch / cfg.DKD.WARMUP, 1.0) * oracle.dkd_loss(zs, zt, target, cfg.DKD.ALPHA, cfg.DKD.BETA, cfg.DKD.T)
        else:
            l_ce = cfg.KD.LOSS.CE_WEIGHT * oracle.ce_loss(zs, target)
            l_kd = cfg.KD.LOSS.KD_WEIGHT * oracle.kd_loss(zs, zt, cfg.KD.TEMPERATURE)
        (l_ce + l_kd).backward()
        ref_opt.step()
        assert abs(losses["loss_kd"].item() - l_kd.item()) <= 1e-4 * abs(l_kd.item())
        assert abs(losses["loss_ce"].item() - l_ce.item()) <= 1e-4 * abs(l_ce.item())
    for (name, p), q in zip(student.named_parameters(), cpu_s.parameters()):
        assert oracle.max_rel_err(p, q) <= 1e-4, name


def test_kd_dot_trainer_step_matches_cpu_oracle(b200):
    """configs[1]: KD + DOT (KD.forward_train, KD.py:25-39, T = 4, CE 0.1 / KD 0.9)."""
    from mdistiller_ddp_b200 import workloads as W
    import copy
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(2)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    cpu_t, cpu_s = copy.deepcopy(teacher).eval(), copy.deepcopy(student).train()
    cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot")
    tr = b200.StandaloneTrainer(b200.KD(student, teacher, cfg).cuda().train(), cfg, kind="dot")
    cpu_params = [p for _, p in cpu_s.named_parameters()]
    m, d = cfg.SOLVER.SGD.MOMENTUM, cfg.SOLVER.DOT.DELTA
    ref = oracle.DotListOracle([p.data for p in cpu_params], lr=cfg.SOLVER.LR, momentum=m - d, momentum_kd=m + d,
                               weight_decay=cfg.SOLVER.WEIGHT_DECAY)
    for step in range(4):
        image, target, _ = W.synthetic_batch(32, 32, 100, seed=20 + step)
        tr.tail_step(image.cuda(), target.cuda(), 1)
        zs, _ = cpu_s(image)
        with torch.no_grad():
            zt, _ = cpu_t(image)
        l_ce = cfg.KD.LOSS.CE_WEIGHT * oracle.ce_loss(zs, target)
        l_kd = cfg.KD.LOSS.KD_WEIGHT * oracle.kd_loss(zs, zt, cfg.KD.TEMPERATURE)
        g_kd = torch.autograd.grad(l_kd, cpu_params, retain_graph=True)
        g_ce = torch.autograd.grad(l_ce, cpu_params)
        ref.step_kd(list(g_kd))
        ref.step(list(g_ce))
    for (name, p), q in zip(student.named_parameters(), cpu_params):
        assert oracle.max_rel_err(p, q) <= 1e-4, name


def test_generic_autograd_protocol_equals_seeded_backward(b200):
    """The reference call pattern (loss_kd.backward(retain_graph=True); step_kd(); loss_ce.backward();
    step()) through autograd gives the same parameters as the trainer's seeded-backward shortcut."""
    from mdistiller_ddp_b200 import workloads as W
    import copy
    torch.manual_seed(3)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    t2, s2 = copy.deepcopy(teacher), copy.deepcopy(student)
    cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot")
    fast = b200.StandaloneTrainer(b200.DKD(student, teacher, cfg).cuda().train(), cfg, kind="dot")
    slow_d = b200.DKD(s2, t2, cfg).cuda().train()
    slow = b200.StandaloneTrainer(slow_d, cfg, kind="dot")
    for step in range(3):
        image, target, _ = W.synthetic_batch(16, 32, 100, seed=30 + step)
        image, target = image.cuda(), target.cuda()
        fast.tail_step(image, target, 4)
        opt = slow.optimizer
        preds, losses = slow_d(image=image, target=target, epoch=4)
        opt.zero_grad(set_to_none=True)
        losses["loss_kd"].backward(retain_graph=True)
        opt.step_kd()
        opt.zero_grad(set_to_none=True)
        losses["loss_ce"].backward()
        opt.step()
    for p, q in zip(student.parameters(), s2.parameters()):
        assert oracle.max_rel_err(p, q) <= 1e-6      # identical arithmetic; cuDNN backward is not bitwise repeatable


def test_base_trainer_amp_branch(b200):
    """EXPERIMENT.AMP with the base trainer (trainer.py:268, 304-313): autocast forward, loss scale kept on the
    device and folded into the flat step.  fp16 backbones make exact parity meaningless here (the scaler's
    arithmetic is pinned to torch's GradScaler in the next test); check the mechanics: finite losses close to the
    fp32 run, parameters move, and an overflowing scale skips the step and backs off."""
    from mdistiller_ddp_b200 import workloads as W
    import copy
    torch.manual_seed(4)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    t32, s32 = copy.deepcopy(teacher), copy.deepcopy(student)
    cfg = W.default_cfg("cifar_res32x4_res8x4_dkd_base")
    cfg.EXPERIMENT.AMP = True
    cfg.SOLVER.GRAD_CLIP = 5.0
    amp_tr = b200.StandaloneTrainer(b200.DKD(student, teacher, cfg).cuda().train(), cfg, kind="base")
    assert amp_tr.use_amp and amp_tr.optimizer.loss_scaler is not None and amp_tr.optimizer.max_grad_norm == 5.0
    cfg32 = W.default_cfg("cifar_res32x4_res8x4_dkd_base")
    cfg32.SOLVER.GRAD_CLIP = 5.0
    ref_tr = b200.StandaloneTrainer(b200.DKD(s32, t32, cfg32).cuda().train(), cfg32, kind="base")
    before = amp_tr.optimizer.flat_params.clone()
    for step in range(3):
        image, target, _ = W.synthetic_batch(32, 32, 100, seed=40 + step)
        _, l16 = amp_tr.tail_step(image.cuda(), target.cuda(), 10)
        _, l32 = ref_tr.tail_step(image.cuda(), target.cuda(), 10)
        for k in ("loss_ce", "loss_kd"):
            assert torch.isfinite(l16[k]) and abs(l16[k].item() - l32[k].item()) <= 2e-2 * abs(l32[k].item()), k
    after = amp_tr.optimizer.flat_params.clone()
    assert not torch.equal(before, after)
    assert oracle.max_rel_err(after, ref_tr.optimizer.flat_params) <= 5e-3
    # overflow: a scale far beyond fp16 range -> inf gradients -> GradScaler skips our step and backs off
    amp_tr.optimizer.loss_scaler = b200.DeviceLossScaler(after.device, init_scale=2.0 ** 120)
    image, target, _ = W.synthetic_batch(32, 32, 100, seed=50)
    amp_tr.tail_step(image.cuda(), target.cuda(), 10)
    assert torch.equal(amp_tr.optimizer.flat_params, after)
    assert amp_tr.optimizer.loss_scaler.get_scale() == 2.0 ** 119


@pytest.mark.parametrize("kind,clip", [("sgd", 0.0), ("sgd", 0.5), ("adamw", 0.5)])
def test_device_loss_scaler_equals_torch_gradscaler(kind, clip, b200):
    """Scaled gradients with an overflow at steps 4 and 5: FlatSGD / FlatAdam + DeviceLossScaler (unscale, inf check,
    clipping and scale update inside the step, on the device) against torch.optim + torch.amp.GradScaler driven the way
    BaseTrainer.backward_loss drives them (trainer.py:304-313).  Same scale after every step, same parameters."""
    case = next(c for c in _cases.DOT_CASES if c.name == "sgd_cifar_small")
    mk = lambda: [torch.nn.Parameter(p.clone().cuda()) for p in _cases.make_params(case)]
    ours_p, ref_p = mk(), mk()
    if kind == "sgd":
        ours = b200.FlatSGD(ours_p, lr=case.lr, momentum=case.momentum, weight_decay=case.wd)
        ref = torch.optim.SGD(ref_p, lr=case.lr, momentum=case.momentum, weight_decay=case.wd)
    else:
        ours = b200.FlatAdam(ours_p, lr=1e-3, betas=_cases.ADAM_BETAS, eps=_cases.ADAM_EPS, weight_decay=5e-2, decoupled=True)
        ref = torch.optim.AdamW(ref_p, lr=1e-3, betas=_cases.ADAM_BETAS, eps=_cases.ADAM_EPS, weight_decay=5e-2)
    ours.max_grad_norm = clip
    ours.loss_scaler = b200.DeviceLossScaler(ours_p[0].device, init_scale=1024.0, growth_interval=3)
    scaler = torch.amp.GradScaler("cuda", init_scale=1024.0, growth_interval=3)
    scaler.scale(torch.zeros(1, device="cuda"))                      # materialises its scale tensor
    shapes = [p.shape for p in ours_p]
    for step in range(14):
        _, task = _cases.make_grads(case, step, shapes)
        scale = scaler.get_scale()
        assert ours.loss_scaler.get_scale() == scale, step
        grads = [(g * scale).cuda() for g in task]
        if step in (4, 5):
            grads[2][0] = float("inf") if step == 4 else float("nan")
        ours.zero_grad()
        ref.zero_grad()
        for p, q, g in zip(ours_p, ref_p, grads):
            p.grad, q.grad = g.clone(), g.clone()
        ours.step()
        if clip > 0:
            scaler.unscale_(ref)
            torch.nn.utils.clip_grad_norm_(ref_p, clip)
        scaler.step(ref)
        scaler.update()
    # x2 after steps 0-2, two back-offs (steps 4, 5), x2 after steps 6-8 and 9-11
    assert ours.loss_scaler.get_scale() == scaler.get_scale() == 2048.0
    for p, q in zip(ours_p, ref_p):
        assert oracle.max_rel_err(p, q) <= TOL
    if kind != "sgd":
        sd = ours.state_dict()["state"]
        assert float(sd[0]["step"]) == 12.0 == float(ref.state[ref_p[0]]["step"])     # the two skipped updates do not count


def test_lagged_metrics_account_every_iteration(b200):
    """metrics_lag > 0 defers the device-to-host read; after flush_metrics the meters equal the eager ones."""
    from mdistiller_ddp_b200 import workloads as W
    import copy
    torch.manual_seed(5)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    t2, s2 = copy.deepcopy(teacher), copy.deepcopy(student)
    cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot")
    eager = b200.StandaloneTrainer(b200.KD(student, teacher, cfg).cuda().train(), cfg)
    lagged = b200.StandaloneTrainer(b200.KD(s2, t2, cfg).cuda().train(), cfg)
    eager.metrics_lag, lagged.metrics_lag = 0, 2
    m1, m2 = b200.new_train_meters(), b200.new_train_meters()
    for step in range(5):
        batch = W.synthetic_batch(16, 32, 100, seed=60 + step, pin=True)
        eager.train_iter(batch, 1, m1)
        lagged.train_iter(batch, 1, m2)
        assert m1["top1"].count == 16 * (step + 1) and m2["top1"].count == 16 * max(0, step - 1)
    lagged.flush_metrics(m2)
    assert m2["top1"].count == m1["top1"].count == 80
    for k in ("loss_ce", "loss_kd", "_total"):
        assert abs(m1["losses"][k].avg - m2["losses"][k].avg) <= 1e-5 * abs(m1["losses"][k].avg)
    assert m1["top5"].avg == m2["top5"].avg


def test_epoch_end_settles_metrics_and_checkpoint_snapshot(b200):
    """A host that tells the mixin the epoch length (install.py does: len(train_loader)) gets exact meters and a
    checkpointable optimizer right after the epoch's last train_iter -- what the reference's train_epoch reads
    next (trainer.py:213-236) -- although metrics are read back one iteration late by default."""
    from mdistiller_ddp_b200 import workloads as W
    torch.manual_seed(6)
    cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot")
    tr = b200.StandaloneTrainer(b200.KD(W.TinyNet(100, 16), W.TinyNet(100, 24), cfg).cuda().train(), cfg)
    assert tr.metrics_lag == 1
    tr.iters_per_epoch = 3
    for epoch in (1, 2):
        meters = b200.new_train_meters()
        for it in range(3):
            tr.train_iter(W.synthetic_batch(8, 32, 100, seed=epoch * 10 + it, pin=True), epoch, meters)
            assert meters["top1"].count == (8 * it if it < 2 else 24)
        sd = tr.optimizer.state_dict()
        assert len(sd["state"]) == len(tr.optimizer.param_groups[0]["params"])


@pytest.mark.parametrize("kind,distiller", [("dot", "KD"), ("base", "DKD")])
def test_graph_replayed_steps_equal_eager_steps(kind, distiller, b200, monkeypatch):
    """Launch-bound steps are replayed from a CUDA graph (graphstep.py).  Same nets, same batches: the trainer that
    replays must end where the eager one ends, across a learning-rate change and an epoch change (both are launch
    arguments baked into the graph, so each forces a fresh capture), and its metrics must account every iteration."""
    from mdistiller_ddp_b200 import workloads as W, graphstep
    import copy
    torch.manual_seed(7)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    t2, s2 = copy.deepcopy(teacher), copy.deepcopy(student)
    cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot" if kind == "dot" else "cifar_res32x4_res8x4_dkd_base")
    cls = b200.KD if distiller == "KD" else b200.DKD
    monkeypatch.setenv("MDB200_GRAPH", "0")
    eager = b200.StandaloneTrainer(cls(student, teacher, cfg).cuda().train(), cfg, kind=kind)
    eager.tail_step(*[t.cuda() for t in W.synthetic_batch(16, 32, 100, seed=0)[:2]], 1)
    monkeypatch.setenv("MDB200_GRAPH", "1")
    graphed = b200.StandaloneTrainer(cls(s2, t2, cfg).cuda().train(), cfg, kind=kind)
    graphed.tail_step(*[t.cuda() for t in W.synthetic_batch(16, 32, 100, seed=0)[:2]], 1)
    assert not eager._graph_runner.enabled and graphed._graph_runner.enabled
    m1, m2 = b200.new_train_meters(), b200.new_train_meters()
    for step in range(1, 16):
        epoch = 1 if step < 11 else 2
        if step == 8:
            for tr in (eager, graphed):
                tr.optimizer.param_groups[0]["lr"] *= 0.1          # what adjust_learning_rate does
        batch = W.synthetic_batch(16, 32, 100, seed=step, pin=True)
        eager.train_iter(batch, epoch, m1)
        graphed.train_iter(batch, epoch, m2)
    eager.flush_metrics(m1)
    graphed.flush_metrics(m2)
    g = graphed._graph_runner.summary()
    assert g["why_off"] is None and g["captures"] == 3 and g["replays"] >= 10, g
    assert oracle.max_rel_err(graphed.optimizer.flat_params, eager.optimizer.flat_params) <= 1e-6
    assert m1["top1"].count == m2["top1"].count == 15 * 16
    for k in ("loss_ce", "loss_kd"):
        assert abs(m1["losses"][k].avg - m2["losses"][k].avg) <= 1e-6 * abs(m1["losses"][k].avg)
    sd1, sd2 = eager.optimizer.state_dict()["state"], graphed.optimizer.state_dict()["state"]
    for i in sd1:
        for key in sd1[i]:
            assert oracle.max_rel_err(sd2[i][key], sd1[i][key]) <= 1e-6


def test_multi_tensor_gather_matches_copies(b200):
    """mdb200_gather_flat: ragged tensor sizes, a missing gradient (zero-filled slice), padding untouched."""
    from mdistiller_ddp_b200.flat import FlatSpace
    torch.manual_seed(6)
    shapes = [(64, 3, 7, 7), (64,), (5,), (1000, 512), (3, 3), (1,), (40000,)]
    params = [torch.nn.Parameter(torch.randn(*s, device="cuda")) for s in shapes]
    sp = FlatSpace(params)
    table = sp.chunk_table()
    assert table.shape[1] == 4 and int(table[:, 1].sum()) == sum(p.numel() for p in params)
    bucket = torch.full((sp.total,), 7.0, device="cuda")
    for o, e, p in zip(sp.offsets, sp.ends, params):            # padding slots start (and must stay) zero
        bucket[o + p.numel():e] = 0.0
    grads = [torch.randn_like(p) for p in params]
    grads[2] = None
    b200.ops.gather_flat(bucket, [0 if g is None else g.data_ptr() for g in grads], table)
    for g, v, p in zip(grads, sp.views(bucket), params):
        assert torch.equal(v, torch.zeros_like(p) if g is None else g)
    for o, e, p in zip(sp.offsets, sp.ends, params):
        assert bucket[o + p.numel():e].abs().sum().item() == 0.0


def test_gather_mode_equals_bucket_view_mode(b200, monkeypatch):
    """The trainer's gather protocol (p.grad = None, one gather launch per backward) and the
    accumulate-into-bucket-views protocol give the same parameters."""
    from mdistiller_ddp_b200 import workloads as W
    import copy
    torch.manual_seed(7)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    t2, s2 = copy.deepcopy(teacher), copy.deepcopy(student)
    cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot")
    gather = b200.StandaloneTrainer(b200.DKD(student, teacher, cfg).cuda().train(), cfg, kind="dot")
    assert gather.optimizer.gather_grads
    monkeypatch.setenv("MDB200_NO_GATHER", "1")
    views = b200.StandaloneTrainer(b200.DKD(s2, t2, cfg).cuda().train(), cfg, kind="dot")
    assert not views.optimizer.gather_grads and views.optimizer.trust_bucket_grads
    for step in range(3):
        image, target, _ = W.synthetic_batch(16, 32, 100, seed=70 + step)
        gather.tail_step(image.cuda(), target.cuda(), 3)
        views.tail_step(image.cuda(), target.cuda(), 3)
    for p, q in zip(student.parameters(), s2.parameters()):
        assert oracle.max_rel_err(p, q) <= 1e-6


def test_tail_kernels_replay_from_a_cuda_graph(b200):
    """The workspace-taking entry points are pure stream work (no allocation, no host sync), so the
    loss + update pair can be captured once and replayed: same numbers as eager launches."""
    ops = b200.ops
    B, C, n = 256, 1000, 1_000_000
    g = torch.Generator(device="cuda").manual_seed(8)
    zs = torch.randn(B, C, device="cuda", generator=g) * 3
    zt = torch.randn(B, C, device="cuda", generator=g) * 3
    y = torch.randint(0, C, (B,), device="cuda", generator=g)
    p0 = torch.randn(n, device="cuda", generator=g) * 0.1
    gt, gk = torch.randn(n, device="cuda", generator=g), torch.randn(n, device="cuda", generator=g)
    hp = dict(lr=0.05, momentum=0.825, momentum_kd=0.975, weight_decay=5e-4)

    def run_eager(steps):
        p, vt, vk = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
        for _ in range(steps):
            lp, gkd, gce = ops.distill_loss(zs, zt, y, kind=ops.LOSS_DKD, alpha=1.0, beta=0.5, temperature=1.0,
                                            ce_mode=ops.CE_SPLIT)
            ops.dot_step(p, gt, gk, vt, vk, **hp)
        return lp.clone(), gkd.clone(), p

    want_l, want_g, want_p = run_eager(3)
    p, vt, vk = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    gkd, gce, lp = torch.empty_like(zs), torch.empty_like(zs), torch.empty(2, device="cuda")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                       # warm-up on the capture stream (workspace, lazy init)
        ops.distill_loss(zs, zt, y, kind=ops.LOSS_DKD, alpha=1.0, beta=0.5, temperature=1.0, ce_mode=ops.CE_SPLIT,
                         grad_kd=gkd, grad_ce=gce, loss_out=lp)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        ops.distill_loss(zs, zt, y, kind=ops.LOSS_DKD, alpha=1.0, beta=0.5, temperature=1.0, ce_mode=ops.CE_SPLIT,
                         grad_kd=gkd, grad_ce=gce, loss_out=lp)
        ops.dot_step(p, gt, gk, vt, vk, **hp)
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(lp, want_l) and torch.equal(gkd, want_g) and torch.equal(p, want_p)


def test_metrics_record_matches_topk_accuracy(b200):
    """top-1 / top-5 hit counts == the reference's accuracy() (engine/utils.py:125-136: topk + eq).  With exact
    ties torch.topk's choice among equal values is unspecified, so there the count must lie between the
    strict (all ties lose) and the inclusive (all ties win) count; the kernel ranks ties by class index."""
    ops = b200.ops
    torch.manual_seed(9)
    l_ce, l_kd = torch.tensor(1.25, device="cuda"), torch.tensor(-0.5, device="cuda")
    for B, C in ((64, 100), (256, 1000), (7, 5), (33, 3), (4096, 1000)):
        z = torch.randn(B, C, device="cuda")
        y = torch.randint(0, C, (B,), device="cuda")
        rec = ops.metrics_record(z, y, l_ce, l_kd).cpu()
        k = min(5, C)
        hit = z.topk(k, dim=1).indices.eq(y.view(-1, 1))
        assert rec[0].item() == 1.25 and rec[1].item() == -0.5 and rec[4].item() == B
        assert rec[2].item() == hit[:, 0].sum().item() and rec[3].item() == hit.sum().item(), (B, C)
        # ties
        zt = z.clone()
        zt[:, : C // 2] = zt[:, : C // 2].round()
        zt[0] = 0.0
        rec = ops.metrics_record(zt, y, None, None).cpu()
        zg = zt.gather(1, y.view(-1, 1))
        strict, incl = (zt > zg).sum(1), (zt >= zg).sum(1) - 1
        for col, kk in ((2, 1), (3, k)):
            lo, hi = (incl < kk).sum().item(), (strict < kk).sum().item()
            assert lo <= rec[col].item() <= hi, (B, C, kk)
        # the documented tie rule: rank = #greater + #equal with a lower class index
        idx = torch.arange(C, device="cuda").view(1, -1)
        rank = strict + ((zt == zg) & (idx < y.view(-1, 1))).sum(1)
        assert rec[2].item() == (rank < 1).sum().item() and rec[3].item() == (rank < k).sum().item()
