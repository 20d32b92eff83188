"""CPU restatement of mdistiller's per-step distillation tail.

TEST INFRASTRUCTURE -- see ``oracle/__init__.py`` for who may import this.

Two independent statements of every quantity live here:

* a torch-CPU restatement that walks the same arithmetic steps as the
  reference (so fp32 rounding behaves like the reference's), differentiated by
  autograd exactly as the reference is; and
* numpy float64 closed forms with analytic gradients,
  used to cross-check the restatement and as the high-precision truth.

Reference lines followed (paths relative to the reference root):
  dkd_loss            mdistiller/distillers/DKD.py:8-32  (masks :35-51)
  kd_loss             mdistiller/distillers/KD.py:8-13
  cross entropy       mdistiller/distillers/DKD.py:71, KD.py:31
  dot()               mdistiller/engine/dot.py:15-55
  optimizer protocol  mdistiller/engine/dot.py:96-174
  plain SGD           torch.optim.SGD as built at mdistiller/engine/trainer.py:86-91
"""

from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------
# logit losses, torch restatement (differentiable)
# --------------------------------------------------------------------------

def _class_masks(logits: torch.Tensor, target: torch.Tensor):
    """Boolean one-hot of the label and its complement (DKD.py:35-44)."""
    cols = torch.arange(logits.shape[1], device=logits.device)
    is_gt = cols.unsqueeze(0) == target.reshape(-1, 1)
    return is_gt, ~is_gt


def _two_bin(prob: torch.Tensor, is_gt: torch.Tensor, not_gt: torch.Tensor):
    """[p_gt, sum of the rest] per row (DKD.py:47-51)."""
    hit = (prob * is_gt).sum(dim=1, keepdim=True)
    rest = (prob * not_gt).sum(dim=1, keepdim=True)
    return torch.cat([hit, rest], dim=1)


def _kl_sum(log_q: torch.Tensor, p: torch.Tensor) -> torch.Tensor:
    """sum p*(log p - log q) with 0*log 0 = 0, i.e. F.kl_div(..., reduction='sum')."""
    return (torch.xlogy(p, p) - p * log_q).sum()


def dkd_loss(logits_student, logits_teacher, target, alpha, beta, temperature):
    """alpha*TCKD + beta*NCKD, same op order as DKD.py:8-32."""
    rows = target.shape[0]
    is_gt, not_gt = _class_masks(logits_student, target)
    p_s = F.softmax(logits_student / temperature, dim=1)
    p_t = F.softmax(logits_teacher / temperature, dim=1)
    b_s = _two_bin(p_s, is_gt, not_gt)
    b_t = _two_bin(p_t, is_gt, not_gt)
    tckd = _kl_sum(torch.log(b_s), b_t) * (temperature ** 2) / rows
    knock = 1000.0 * is_gt
    q_t = F.softmax(logits_teacher / temperature - knock, dim=1)
    log_q_s = F.log_softmax(logits_student / temperature - knock, dim=1)
    nckd = _kl_sum(log_q_s, q_t) * (temperature ** 2) / rows
    return alpha * tckd + beta * nckd


def kd_loss(logits_student, logits_teacher, temperature):
    """T^2 * mean_rows KL(p_t || p_s), as KD.py:8-13."""
    log_p_s = F.log_softmax(logits_student / temperature, dim=1)
    p_t = F.softmax(logits_teacher / temperature, dim=1)
    per_row = (torch.xlogy(p_t, p_t) - p_t * log_p_s).sum(1)
    out = per_row.mean()
    out = out * temperature ** 2
    return out


def ce_loss(logits_student, target):
    return F.cross_entropy(logits_student, target)


def loss_and_grad(kind: str, z_s: torch.Tensor, z_t: torch.Tensor, target: torch.Tensor,
                  *, alpha: float = 1.0, beta: float = 8.0, temperature: float = 4.0,
                  kd_scale: float = 1.0, ce_scale: float = 1.0,
                  dtype: torch.dtype = torch.float32):
    """Run the restatement under autograd.

    Returns dict(loss_kd, grad_kd, loss_ce, grad_ce); the scales are the outer
    multipliers the distillers apply (warm-up factor / KD_WEIGHT / CE_WEIGHT,
    DKD.py:71-72, KD.py:31-34).
    """
    zs = z_s.detach().to(dtype).clone().requires_grad_(True)
    zt = z_t.detach().to(dtype)
    if kind == "dkd":
        l_kd = kd_scale * dkd_loss(zs, zt, target, alpha, beta, temperature)
    elif kind == "kd":
        l_kd = kd_scale * kd_loss(zs, zt, temperature)
    else:
        raise ValueError(kind)
    (g_kd,) = torch.autograd.grad(l_kd, zs)
    l_ce = ce_scale * ce_loss(zs, target)
    (g_ce,) = torch.autograd.grad(l_ce, zs)
    return {"loss_kd": l_kd.detach(), "grad_kd": g_kd, "loss_ce": l_ce.detach(), "grad_ce": g_ce}


# --------------------------------------------------------------------------
# logit losses, numpy float64 closed forms (Appendix A.1 / A.2)
# --------------------------------------------------------------------------

def _row_stats(u: np.ndarray, y: np.ndarray):
    """Shifted exponentials split into the label term and the rest."""
    m = u.max(axis=1, keepdims=True)
    e = np.exp(u - m)
    rows = np.arange(u.shape[0])
    e_gt = e[rows, y].copy()
    e_rest = e.copy()
    e_rest[rows, y] = 0.0
    s_rest = e_rest.sum(axis=1)
    return m[:, 0], e_rest, e_gt, s_rest


def dkd_closed_form(z_s, z_t, y, alpha, beta, temperature):
    """(loss, dloss/dz_s) in float64 from the Appendix A.1 formulas."""
    zs = np.asarray(z_s, dtype=np.float64)
    zt = np.asarray(z_t, dtype=np.float64)
    y = np.asarray(y, dtype=np.int64)
    T = float(temperature)
    B, _ = zs.shape
    rows = np.arange(B)
    us, ut = zs / T, zt / T
    ms, es, egs, ss = _row_stats(us, y)
    mt, et, egt, st = _row_stats(ut, y)
    pi_s, pi_t = egs / (egs + ss), egt / (egt + st)
    om_s, om_t = ss / (egs + ss), st / (egt + st)
    qs, qt = es / ss[:, None], et / st[:, None]          # label column is exactly 0
    tckd = pi_t * (np.log(pi_t) - np.log(pi_s)) + om_t * (np.log(om_t) - np.log(om_s))
    lse_s, lse_t = ms + np.log(ss), mt + np.log(st)
    cross = (qt * (ut - us)).sum(axis=1)
    nckd = cross - (lse_t - lse_s)
    loss = (T * T / B) * (alpha * tckd + beta * nckd).sum()
    d = pi_s - pi_t
    g = (T / B) * (-alpha * d[:, None] * qs + beta * (qs - qt))
    g[rows, y] = (T / B) * alpha * d
    return loss, g


def kd_closed_form(z_s, z_t, temperature):
    zs = np.asarray(z_s, dtype=np.float64)
    zt = np.asarray(z_t, dtype=np.float64)
    T = float(temperature)
    B = zs.shape[0]

    def log_softmax(u):
        m = u.max(axis=1, keepdims=True)
        return u - m - np.log(np.exp(u - m).sum(axis=1, keepdims=True))

    lps, lpt = log_softmax(zs / T), log_softmax(zt / T)
    pt = np.exp(lpt)
    loss = (T * T / B) * (pt * (lpt - lps)).sum()
    g = (T / B) * (np.exp(lps) - pt)
    return loss, g


def ce_closed_form(z_s, y):
    zs = np.asarray(z_s, dtype=np.float64)
    y = np.asarray(y, dtype=np.int64)
    B = zs.shape[0]
    m = zs.max(axis=1, keepdims=True)
    lp = zs - m - np.log(np.exp(zs - m).sum(axis=1, keepdims=True))
    rows = np.arange(B)
    loss = -lp[rows, y].mean()
    g = np.exp(lp)
    g[rows, y] -= 1.0
    return loss, g / B


# --------------------------------------------------------------------------
# DOT dual-momentum step: list form (membership semantics of dot.py:15-55)
# --------------------------------------------------------------------------

def _listed(t: torch.Tensor, pool: Sequence[torch.Tensor]) -> bool:
    """Identity membership, dot.py:9-13."""
    return any(t is other for other in pool)


def dot_lists_step(task_params: List[torch.Tensor], task_grads: List[torch.Tensor],
                   task_bufs: List[Optional[torch.Tensor]],
                   kd_grads: List[torch.Tensor], kd_bufs: List[Optional[torch.Tensor]],
                   kd_params: List[torch.Tensor], *, weight_decay: float, momentum: float,
                   momentum_kd: float, lr: float, dampening: float) -> None:
    """In-place restatement of dot() (dot.py:15-55): task pass, then KD pass."""
    blend = (momentum_kd + momentum) / 2.0
    for i, p in enumerate(task_params):
        step_dir = task_grads[i]
        if weight_decay != 0:
            step_dir = step_dir.add(p, alpha=weight_decay)        # pre-update p, :29-30
        if momentum != 0:
            v = task_bufs[i]
            if v is None:                                          # first step: plain copy, :33-35
                v = step_dir.detach().clone()
                task_bufs[i] = v
            else:
                coeff = momentum if _listed(p, kd_params) else blend
                v.mul_(coeff).add_(step_dir, alpha=1 - dampening)  # two roundings, :36-39
            step_dir = v
        p.add_(step_dir, alpha=-lr)                                # :42
    for i, p in enumerate(kd_params):
        g = kd_grads[i]
        v = kd_bufs[i]
        if v is None:                                              # :46-48, no decay on first step
            v = g.detach().clone()
            kd_bufs[i] = v
        elif _listed(p, task_params):
            v.mul_(momentum_kd).add_(g, alpha=1 - dampening)       # :49-50, no decay
        else:
            if weight_decay != 0:
                g = g.add(p, alpha=weight_decay)                   # :52-53
            v.mul_(blend).add_(g, alpha=1 - dampening)
        p.add_(v, alpha=-lr)                                       # :55


class DotListOracle:
    """Stateful mirror of the optimizer protocol at dot.py:96-174 over plain tensors.

    ``step_kd(grads)`` stashes the KD gradients (``None`` = parameter had no
    gradient on that backward), ``step(grads)`` runs both passes and stores the
    momentum buffers per parameter like ``state[p]['momentum_buffer']`` /
    ``['momentum_kd_buffer']`` do.
    """

    def __init__(self, params: Sequence[torch.Tensor], lr: float, momentum: float = 0.0,
                 momentum_kd: float = 0.0, dampening: float = 0.0, weight_decay: float = 0.0):
        if lr < 0.0 or momentum < 0.0 or momentum_kd < 0.0 or weight_decay < 0.0:
            raise ValueError("negative hyper-parameter")
        self.params = list(params)
        self.hp = dict(lr=lr, momentum=momentum, momentum_kd=momentum_kd,
                       dampening=dampening, weight_decay=weight_decay)
        self.buf_task: List[Optional[torch.Tensor]] = [None] * len(self.params)
        self.buf_kd: List[Optional[torch.Tensor]] = [None] * len(self.params)
        self._kd_idx: List[int] = []
        self._kd_grads: List[torch.Tensor] = []

    def step_kd(self, grads: Sequence[Optional[torch.Tensor]]) -> None:
        self._kd_idx = [i for i, g in enumerate(grads) if g is not None]
        self._kd_grads = [grads[i] for i in self._kd_idx]

    def step(self, grads: Sequence[Optional[torch.Tensor]]) -> None:
        t_idx = [i for i, g in enumerate(grads) if g is not None]
        t_bufs = [self.buf_task[i] for i in t_idx]
        k_bufs = [self.buf_kd[i] for i in self._kd_idx]
        dot_lists_step([self.params[i] for i in t_idx], [grads[i] for i in t_idx], t_bufs,
                       self._kd_grads, k_bufs, [self.params[i] for i in self._kd_idx],
                       weight_decay=self.hp["weight_decay"], momentum=self.hp["momentum"],
                       momentum_kd=self.hp["momentum_kd"], lr=self.hp["lr"],
                       dampening=self.hp["dampening"])
        for i, b in zip(t_idx, t_bufs):
            self.buf_task[i] = b
        for i, b in zip(self._kd_idx, k_bufs):
            self.buf_kd[i] = b
        self._kd_idx, self._kd_grads = [], []


# --------------------------------------------------------------------------
# flat recurrences (Appendix A.3): what one kernel launch must compute
# --------------------------------------------------------------------------

def dot_flat_step(p, g_task, g_kd, v_task, v_kd, *, lr, m_task, m_kd, weight_decay,
                  dampening=0.0, first_step=False, grad_scale=1.0) -> None:
    """Every element is in both gradient sets (the KD/DKD student case).

    ``grad_scale`` multiplies both gradients first (1/world_size after a
    sum-reduce-scatter); 1.0 reproduces dot.py exactly.  In place on torch
    tensors of any float dtype.
    """
    gt = g_task if grad_scale == 1.0 else g_task * grad_scale
    gk = g_kd if grad_scale == 1.0 else g_kd * grad_scale
    d = gt.add(p, alpha=weight_decay) if weight_decay != 0 else gt
    if m_task != 0:
        if first_step:
            v_task.copy_(d)
        else:
            v_task.mul_(m_task).add_(d, alpha=1 - dampening)
        p.add_(v_task, alpha=-lr)
    else:
        p.add_(d, alpha=-lr)
    if first_step:
        v_kd.copy_(gk)
    else:
        v_kd.mul_(m_kd).add_(gk, alpha=1 - dampening)
    p.add_(v_kd, alpha=-lr)


def sgd_flat_step(p, g, v, *, lr, momentum, weight_decay, dampening=0.0,
                  first_step=False, grad_scale=1.0) -> None:
    """torch.optim.SGD (no Nesterov) on flat buffers -- the base-trainer step."""
    gg = g if grad_scale == 1.0 else g * grad_scale
    d = gg.add(p, alpha=weight_decay) if weight_decay != 0 else gg
    if momentum != 0:
        if first_step:
            v.copy_(d)
        else:
            v.mul_(momentum).add_(d, alpha=1 - dampening)
        d = v
    p.add_(d, alpha=-lr)


def adam_flat_step(p, g, m, v, *, lr, beta1, beta2, eps, weight_decay=0.0, decoupled=False, step=1,
                   grad_scale=1.0) -> None:
    """torch.optim.Adam / AdamW single-tensor update (no amsgrad), in place, `step` 1-based --
    the optimizers BaseTrainer.init_optimizer builds at mdistiller/engine/trainer.py:92-107."""
    gg = g if grad_scale == 1.0 else g * grad_scale
    if weight_decay != 0:
        if decoupled:
            p.mul_(1 - lr * weight_decay)
        else:
            gg = gg.add(p, alpha=weight_decay)
    m.lerp_(gg, 1 - beta1)
    v.mul_(beta2).addcmul_(gg, gg, value=1 - beta2)
    bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
    denom = (v.sqrt() / (bc2 ** 0.5)).add_(eps)
    p.addcdiv_(m, denom, value=-(lr / bc1))


def clip_coefficient(grads: Sequence[torch.Tensor], max_norm: float) -> float:
    """nn.utils.clip_grad_norm_'s factor (mdistiller/engine/trainer.py:316-320): min(1, c / (||g||_2 + 1e-6))."""
    total = torch.sqrt(sum((t.double() ** 2).sum() for t in grads if t is not None)).item()
    return min(1.0, max_norm / (total + 1e-6))


def max_rel_err(a, b) -> float:
    """The parity metric fixed in SURVEY.md section 7: max|a-b| / max|b| per tensor."""
    a = np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a, dtype=np.float64)
    b = np.asarray(b.detach().cpu() if isinstance(b, torch.Tensor) else b, dtype=np.float64)
    denom = np.abs(b).max() if b.size else 0.0
    if denom == 0.0:
        return float(np.abs(a - b).max()) if a.size else 0.0
    return float(np.abs(a - b).max() / denom)
