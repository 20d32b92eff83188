"""Pin the CPU oracle to the reference's own outputs (committed fixtures) -- CPU only.

Tolerances: the fp64 restatement/closed forms must reproduce the fp64 reference
to 1e-12 (same maths, different op order); the fp32 restatement must reproduce
the fp32 reference to 2e-6 max-rel (it follows the same op order, so in practice
it is bit-identical); the published bar for the CUDA path is 1e-5.
"""

import numpy as np
import pytest
import torch

import _cases
import oracle
from oracle import ref_loader

TOL64 = 1e-12
TOL32 = 2e-6


def _grad_ref(G, name, key, shape):
    """Full reference gradient or (index, samples)."""
    if f"{name}/g{key}" in G:
        return None, G[f"{name}/g{key}"]
    idx = _cases.sample_index(shape[0] * shape[1])
    return idx, G[f"{name}/g{key}_s"]


def _cmp_grad(got, G, name, key, shape, tol):
    idx, ref = _grad_ref(G, name, key, shape)
    got = np.asarray(got, dtype=np.float64)
    if idx is not None:
        got = got.reshape(-1)[idx]
    scale = G[f"{name}/g{key[:-2]}_max64"]
    assert np.abs(got - ref).max() / scale <= tol, (name, key)


@pytest.mark.parametrize("case", _cases.LOSS_CASES, ids=lambda c: c.name)
def test_loss_restatement_matches_reference(case, loss_golden):
    G = loss_golden
    zs, zt, y = _cases.make_logits(case)
    np.testing.assert_allclose(_cases.input_checksum(zs, zt, y), G[f"{case.name}/chk"], rtol=0, atol=1e-9)
    shape = (case.B, case.C)
    for kind in ("dkd", "kd"):
        for tag, dt, tol in (("64", torch.float64, TOL64), ("32", torch.float32, TOL32)):
            r = oracle.loss_and_grad(kind, zs, zt, y, alpha=case.alpha, beta=case.beta,
                                     temperature=case.T, dtype=dt)
            ref_l = float(G[f"{case.name}/{kind}{tag}"])
            assert abs(r["loss_kd"].item() - ref_l) <= tol * max(1.0, abs(ref_l))
            _cmp_grad(r["grad_kd"].numpy(), G, case.name, f"{kind}{tag}", shape, tol)
            ref_ce = float(G[f"{case.name}/ce{tag}"])
            assert abs(r["loss_ce"].item() - ref_ce) <= tol * max(1.0, abs(ref_ce))
            _cmp_grad(r["grad_ce"].numpy(), G, case.name, f"ce{tag}", shape, tol)


@pytest.mark.parametrize("case", _cases.LOSS_CASES, ids=lambda c: c.name)
def test_closed_forms_match_reference_fp64(case, loss_golden):
    G = loss_golden
    zs, zt, y = _cases.make_logits(case)
    shape = (case.B, case.C)
    l, g = oracle.dkd_closed_form(zs.numpy(), zt.numpy(), y.numpy(), case.alpha, case.beta, case.T)
    assert abs(l - float(G[f"{case.name}/dkd64"])) <= 1e-11 * max(1.0, abs(l))
    _cmp_grad(g, G, case.name, "dkd64", shape, 1e-11)
    l, g = oracle.kd_closed_form(zs.numpy(), zt.numpy(), case.T)
    assert abs(l - float(G[f"{case.name}/kd64"])) <= 1e-11 * max(1.0, abs(l))
    _cmp_grad(g, G, case.name, "kd64", shape, 1e-11)
    l, g = oracle.ce_closed_form(zs.numpy(), y.numpy())
    assert abs(l - float(G[f"{case.name}/ce64"])) <= 1e-11 * max(1.0, abs(l))
    _cmp_grad(g, G, case.name, "ce64", shape, 1e-11)


def test_survey_known_answers(loss_golden):
    """SURVEY.md Appendix C.1 rows, typed in by hand from the survey."""
    rows = {
        "c1_cifar_64x100": (71.1602410562, 8.8813987360, 9.0704182181, 25.8535309301),
        "c1_cifar_128x100": (54.1394492459, 8.9988500748, 8.6148377078, 19.3945846562),
        "c1_imnet_256x1000": (3.7768977593, 7.5215915538, 11.2056976294, 0.9556128137),
        "c1_imnet_1024x1000": (3.8942872776, 7.7751890591, 11.1608108121, 0.9582475390),
    }
    for case in _cases.LOSS_CASES[:4]:
        zs, zt, y = _cases.make_logits(case)
        dkd, kd, ce, gsum = rows[case.name]
        l, g = oracle.dkd_closed_form(zs.numpy(), zt.numpy(), y.numpy(), case.alpha, case.beta, case.T)
        assert abs(l - dkd) < 5e-10 and abs(np.abs(g).sum() - gsum) < 5e-10
        assert abs(oracle.kd_closed_form(zs.numpy(), zt.numpy(), case.T)[0] - kd) < 5e-10
        assert abs(oracle.ce_closed_form(zs.numpy(), y.numpy())[0] - ce) < 5e-10


def _run_oracle_lists(case, dtype):
    params = [p.to(dtype).clone() for p in _cases.make_params(case)]
    shapes = [p.shape for p in params]
    if case.kind == "dot":
        opt = oracle.DotListOracle(params, lr=case.lr, momentum=case.momentum - case.delta,
                                   momentum_kd=case.momentum + case.delta,
                                   dampening=case.dampening, weight_decay=case.wd)
        for step in range(case.steps):
            if step == case.lr_drop_at:
                opt.hp["lr"] = case.lr * 0.1
            kd, task = _cases.make_grads(case, step, shapes)
            opt.step_kd([None if g is None else g.to(dtype) for g in kd])
            opt.step([None if g is None else g.to(dtype) for g in task])
        return params, opt.buf_task, opt.buf_kd
    # base-trainer optimizers, tensor by tensor like torch does (parameters without a gradient are skipped)
    first = [None] * len(params)
    second = [None] * len(params)
    steps = [0] * len(params)
    lr = case.lr
    for step in range(case.steps):
        if step == case.lr_drop_at:
            lr = case.lr * 0.1
        _, task = _cases.make_grads(case, step, shapes)
        task = [None if g is None else g.to(dtype) for g in task]
        coef = oracle.clip_coefficient(task, case.clip) if case.clip > 0 else 1.0
        for i, (p, g) in enumerate(zip(params, task)):
            if g is None:
                continue
            steps[i] += 1
            if case.kind == "sgd":
                fresh = first[i] is None
                if fresh:
                    first[i] = torch.zeros_like(p)
                oracle.sgd_flat_step(p, g, first[i], lr=lr, momentum=case.momentum, weight_decay=case.wd,
                                     dampening=case.dampening, first_step=fresh, grad_scale=coef)
            else:
                if first[i] is None:
                    first[i], second[i] = torch.zeros_like(p), torch.zeros_like(p)
                oracle.adam_flat_step(p, g, first[i], second[i], lr=lr, beta1=_cases.ADAM_BETAS[0],
                                      beta2=_cases.ADAM_BETAS[1], eps=_cases.ADAM_EPS, weight_decay=case.wd,
                                      decoupled=(case.kind == "adamw"), step=steps[i], grad_scale=coef)
    return params, first, second


def _flat(ts, like):
    return torch.cat([(torch.zeros_like(l) if t is None else t).reshape(-1) for t, l in zip(ts, like)]).numpy()


@pytest.mark.parametrize("case", _cases.DOT_CASES, ids=lambda c: c.name)
def test_dot_restatement_matches_reference(case, dot_golden):
    G = dot_golden
    np.testing.assert_allclose(_cases.dot_input_checksum(case), G[f"{case.name}/chk"], rtol=0, atol=1e-9)
    for tag, dt, tol in (("64", torch.float64, 1e-12), ("32", torch.float32, 1e-6)):
        params, bt, bk = _run_oracle_lists(case, dt)
        got = {"p": _flat(params, params), "vt": _flat(bt, params), "vk": _flat(bk, params)}
        n = got["p"].size
        for k, arr in got.items():
            if f"{case.name}/{k}{tag}" in G:
                ref = G[f"{case.name}/{k}{tag}"]
            else:
                ref = G[f"{case.name}/{k}{tag}_s"]
                arr = arr[_cases.sample_index(n)]
            assert oracle.max_rel_err(arr, ref) <= tol, (case.name, k, tag)
        assert abs(got["p"].astype(np.float64).sum() - float(G[f"{case.name}/p{tag}_sum"])) <= \
            (1e-9 if tag == "64" else 2e-6) * float(G[f"{case.name}/p{tag}_abssum"])


@pytest.mark.parametrize("case", [c for c in _cases.DOT_CASES
                                  if c.kind == "dot" and not (c.kd_missing or c.task_missing or c.kd_from_step)
                                  and c.momentum - c.delta != 0],
                         ids=lambda c: c.name)
def test_flat_recurrence_equals_list_form(case):
    """The flat Appendix-A.3 recurrence (what the kernel computes) == dot() per tensor, bit for bit."""
    params = [p.clone() for p in _cases.make_params(case)]
    shapes = [p.shape for p in params]
    flat = torch.cat([p.reshape(-1) for p in params]).clone()
    vt, vk = torch.zeros_like(flat), torch.zeros_like(flat)
    opt = oracle.DotListOracle(params, lr=case.lr, momentum=case.momentum - case.delta,
                               momentum_kd=case.momentum + case.delta, dampening=case.dampening,
                               weight_decay=case.wd)
    lr = case.lr
    for step in range(min(case.steps, 30)):
        if step == 10:
            lr = case.lr * 0.1
            opt.hp["lr"] = lr
        kd, task = _cases.make_grads(case, step, shapes)
        opt.step_kd(kd)
        opt.step(task)
        oracle.dot_flat_step(flat, torch.cat([t.reshape(-1) for t in task]),
                             torch.cat([t.reshape(-1) for t in kd]), vt, vk, lr=lr,
                             m_task=case.momentum - case.delta, m_kd=case.momentum + case.delta,
                             weight_decay=case.wd, dampening=case.dampening, first_step=(step == 0))
    assert torch.equal(flat, torch.cat([p.reshape(-1) for p in params]))


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not mounted (GPU box)")
def test_live_reference_agrees_with_oracle():
    """Where /root/reference is mounted, re-run it live on a fresh seed (not in the fixtures)."""
    ref_dkd, ref_kd = ref_loader.loss_functions()
    case = _cases.LossCase("live", 777, 96, 100, 3.0, 1.0, 8.0, 4.0)
    zs, zt, y = _cases.make_logits(case)
    a = zs.clone().requires_grad_(True)
    l = ref_dkd(a, zt, y, case.alpha, case.beta, case.T)
    (g,) = torch.autograd.grad(l, a)
    r = oracle.loss_and_grad("dkd", zs, zt, y, alpha=case.alpha, beta=case.beta, temperature=case.T)
    assert torch.equal(r["loss_kd"], l.detach()) and torch.equal(r["grad_kd"], g)
    a = zs.clone().requires_grad_(True)
    l = ref_kd(a, zt, case.T)
    (g,) = torch.autograd.grad(l, a)
    r = oracle.loss_and_grad("kd", zs, zt, y, temperature=case.T)
    assert torch.equal(r["loss_kd"], l.detach()) and torch.equal(r["grad_kd"], g)


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not mounted (GPU box)")
@pytest.mark.parametrize("kind", ["DKD", "KD"])
def test_forward_train_composition_matches_reference_distillers(kind):
    """a3/a5: the way the GPU tests compose the oracle (CE_WEIGHT*CE, warm-up/KD_WEIGHT * distillation
    term) is exactly what the reference's DKD.forward_train / KD.forward_train return (DKD.py:65-84,
    KD.py:25-39), on the same small nets."""
    import copy
    import sys
    sys.path.insert(0, __import__("os").path.dirname(__file__))
    import _ref_shims
    _ref_shims.install_shims()
    RefDKD, RefKD = ref_loader.distiller_classes()
    from mdistiller_ddp_b200 import workloads as W
    torch.manual_seed(11)
    teacher, student = W.TinyNet(100, 24), W.TinyNet(100, 16)
    cfg = W.default_cfg("cifar_res32x4_res8x4_kd_dot")
    ref = (RefDKD if kind == "DKD" else RefKD)(copy.deepcopy(student), copy.deepcopy(teacher), cfg).train()
    image, target, _ = W.synthetic_batch(16, 32, 100, seed=3)
    for epoch in (1, 7, 40):
        logits, losses = ref.forward_train(image=image, target=target, epoch=epoch)
        student.train()
        teacher.eval()
        zs, _ = student(image)
        with torch.no_grad():
            zt, _ = teacher(image)
        if kind == "DKD":
            l_ce = cfg.DKD.CE_WEIGHT * oracle.ce_loss(zs, target)
            l_kd = min(epoch / cfg.DKD.WARMUP, 1.0) * oracle.dkd_loss(zs, zt, target, cfg.DKD.ALPHA, cfg.DKD.BETA, cfg.DKD.T)
        else:
            l_ce = cfg.KD.LOSS.CE_WEIGHT * oracle.ce_loss(zs, target)
            l_kd = cfg.KD.LOSS.KD_WEIGHT * oracle.kd_loss(zs, zt, cfg.KD.TEMPERATURE)
        assert torch.allclose(losses["loss_ce"], l_ce, rtol=1e-6) and torch.allclose(losses["loss_kd"], l_kd, rtol=1e-6)
        assert set(losses) == {"loss_ce", "loss_kd"} and torch.allclose(logits, zs, rtol=1e-6, atol=1e-7)
