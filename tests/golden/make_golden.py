#!/usr/bin/env python
"""Generate the committed fixtures in tests/golden/ from the UNMODIFIED reference.

Runs only in the build container (needs /root/reference).  It calls the
reference's own ``dkd_loss`` (DKD.py:8), ``kd_loss`` (KD.py:8),
``F.cross_entropy`` (DKD.py:71) and ``DistillationOrientedTrainer``
(engine/dot.py:58) on seeded CPU inputs and stores inputs/outputs so that the
oracle and the CUDA path can be checked where the reference is not mounted.

    python tests/golden/make_golden.py            # rewrites *.npz next to this file

Input recipe (shared with tests/_cases.py -- keep in sync):
    torch.manual_seed(seed); zs = randn(B,C)*sigma; zt = randn(B,C)*sigma;
    y = randint(0,C,(B,)); then zs[i,y[i]] += boost, zt[i,y[i]] += boost.
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_loader  # noqa: E402
import _cases  # noqa: E402


def loss_fixture():
    ref_dkd, ref_kd = ref_loader.loss_functions()
    out = {}
    names = []
    for case in _cases.LOSS_CASES:
        zs, zt, y = _cases.make_logits(case)
        rec = {}
        for tag, dt in (("32", torch.float32), ("64", torch.float64)):
            a = zs.to(dt).clone().requires_grad_(True)
            b = zt.to(dt)
            l = ref_dkd(a, b, y, case.alpha, case.beta, case.T)
            (g,) = torch.autograd.grad(l, a)
            rec[f"dkd{tag}"], rec[f"gdkd{tag}"] = l.item(), g.numpy()
            a = zs.to(dt).clone().requires_grad_(True)
            l = ref_kd(a, b, case.T)
            (g,) = torch.autograd.grad(l, a)
            rec[f"kd{tag}"], rec[f"gkd{tag}"] = l.item(), g.numpy()
            a = zs.to(dt).clone().requires_grad_(True)
            l = F.cross_entropy(a, y)
            (g,) = torch.autograd.grad(l, a)
            rec[f"ce{tag}"], rec[f"gce{tag}"] = l.item(), g.numpy()
        n = case.name
        names.append(n)
        small = case.B * case.C <= _cases.FULL_TENSOR_LIMIT
        out[f"{n}/chk"] = _cases.input_checksum(zs, zt, y)
        for k in ("dkd", "kd", "ce"):
            out[f"{n}/{k}32"] = np.float64(rec[f"{k}32"])
            out[f"{n}/{k}64"] = np.float64(rec[f"{k}64"])
            g32, g64 = rec[f"g{k}32"], rec[f"g{k}64"]
            out[f"{n}/g{k}_abssum64"] = np.float64(np.abs(g64).sum())
            out[f"{n}/g{k}_max64"] = np.float64(np.abs(g64).max())
            if small:
                out[f"{n}/g{k}32"] = g32
                out[f"{n}/g{k}64"] = g64
            else:
                idx = _cases.sample_index(case.B * case.C)
                out[f"{n}/g{k}32_s"] = g32.reshape(-1)[idx]
                out[f"{n}/g{k}64_s"] = g64.reshape(-1)[idx]
        if small:
            out[f"{n}/zs"], out[f"{n}/zt"], out[f"{n}/y"] = zs.numpy(), zt.numpy(), y.numpy()
        print(f"loss {n:28s} dkd32={rec['dkd32']:.7f} dkd64={rec['dkd64']:.10f} "
              f"kd64={rec['kd64']:.10f} ce64={rec['ce64']:.10f}")
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "loss_golden.npz"), **out)


def _run_ref_dot(case, dtype):
    dotmod = ref_loader.dot_module()
    params = [p.to(dtype).clone().requires_grad_(True) for p in _cases.make_params(case)]
    opt = dotmod.DistillationOrientedTrainer(
        params, lr=case.lr, momentum=case.momentum - case.delta,
        momentum_kd=case.momentum + case.delta, dampening=case.dampening,
        weight_decay=case.wd)
    for step in range(case.steps):
        if step == case.lr_drop_at:
            opt.param_groups[0]["lr"] = case.lr * 0.1
        g_kd, g_task = _cases.make_grads(case, step, [p.shape for p in params])
        opt.zero_grad(set_to_none=True)
        for p, g in zip(params, g_kd):
            p.grad = None if g is None else g.to(dtype).clone()
        opt.step_kd()
        opt.zero_grad(set_to_none=True)
        for p, g in zip(params, g_task):
            p.grad = None if g is None else g.to(dtype).clone()
        opt.step()
    bufs_t = [opt.state[p].get("momentum_buffer") for p in params]
    bufs_k = [opt.state[p].get("momentum_kd_buffer") for p in params]
    return params, bufs_t, bufs_k


def _run_ref_sgd(case, dtype):
    """torch.optim.{SGD,Adam,AdamW} (+ clip_grad_norm_) exactly as BaseTrainer drives them
    (mdistiller/engine/trainer.py:86-107, 316-321)."""
    params = [p.to(dtype).clone().requires_grad_(True) for p in _cases.make_params(case)]
    if case.kind == "sgd":
        opt = torch.optim.SGD(params, lr=case.lr, momentum=case.momentum, weight_decay=case.wd,
                              dampening=case.dampening)
    else:
        cls = torch.optim.Adam if case.kind == "adam" else torch.optim.AdamW
        opt = cls(params, lr=case.lr, betas=_cases.ADAM_BETAS, eps=_cases.ADAM_EPS, weight_decay=case.wd)
    for step in range(case.steps):
        if step == case.lr_drop_at:
            opt.param_groups[0]["lr"] = case.lr * 0.1
        _, g_task = _cases.make_grads(case, step, [p.shape for p in params])
        for p, g in zip(params, g_task):
            p.grad = None if g is None else g.to(dtype).clone()
        if case.clip > 0:
            torch.nn.utils.clip_grad_norm_(params, case.clip)
        opt.step()
    key1 = "momentum_buffer" if case.kind == "sgd" else "exp_avg"
    second = [opt.state[p].get("exp_avg_sq") for p in params]
    return params, [opt.state[p].get(key1) for p in params], second


def _flat(ts, like):
    return torch.cat([(torch.zeros_like(l) if t is None else t.detach()).reshape(-1)
                      for t, l in zip(ts, like)]).numpy()


def dot_fixture():
    out = {}
    names = []
    for case in _cases.DOT_CASES:
        n = case.name
        names.append(n)
        for tag, dt in (("32", torch.float32), ("64", torch.float64)):
            if case.kind == "dot":
                params, bt, bk = _run_ref_dot(case, dt)
            else:
                params, bt, bk = _run_ref_sgd(case, dt)      # (momentum | exp_avg, exp_avg_sq)
            flat_p = _flat(params, params)
            flat_t = _flat(bt, params)
            flat_k = _flat(bk, params)
            if flat_p.size <= _cases.FULL_TENSOR_LIMIT:
                out[f"{n}/p{tag}"], out[f"{n}/vt{tag}"], out[f"{n}/vk{tag}"] = flat_p, flat_t, flat_k
            else:
                idx = _cases.sample_index(flat_p.size)
                out[f"{n}/p{tag}_s"], out[f"{n}/vt{tag}_s"], out[f"{n}/vk{tag}_s"] = \
                    flat_p[idx], flat_t[idx], flat_k[idx]
            out[f"{n}/p{tag}_sum"] = np.float64(flat_p.astype(np.float64).sum())
            out[f"{n}/p{tag}_abssum"] = np.float64(np.abs(flat_p.astype(np.float64)).sum())
            out[f"{n}/p{tag}_max"] = np.float64(np.abs(flat_p).max())
        out[f"{n}/chk"] = _cases.dot_input_checksum(case)
        print(f"dot  {n:28s} N={flat_p.size:8d} sum={out[f'{n}/p64_sum']:.9f} "
              f"abssum={out[f'{n}/p64_abssum']:.6f}")
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "dot_golden.npz"), **out)


if __name__ == "__main__":
    torch.set_num_threads(8)
    loss_fixture()
    dot_fixture()
