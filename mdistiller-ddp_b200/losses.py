"""Autograd-facing logit losses backed by the fused sm_100a kernel.

Public signatures match the reference so existing callers keep working
(second caller: detection/model/rcnn.py:30):

    dkd_loss(logits_student, logits_teacher, target, alpha, beta, temperature)   # DKD.py:8
    kd_loss(logits_student, logits_teacher, temperature)                          # KD.py:8

Both return a 0-dim tensor differentiable w.r.t. ``logits_student`` (the teacher runs
under ``no_grad`` in the reference, DKD.py:67-68, so it gets no gradient).  The forward
launch already produces d loss / d logits_student; backward only multiplies by the
upstream scalar, on the device, without a host sync.

``distill_tail`` is the fused form the distillers use: distillation term + weighted
cross entropy in the same launch, returning both losses.
"""

from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops


class TailResult:
    """What one fused launch produced; lets a trainer seed backward at the logits directly."""

    __slots__ = ("logits", "loss_pair", "grad_kd", "grad_ce", "combined")

    def __init__(self, logits, loss_pair, grad_kd, grad_ce, combined):
        self.logits = logits          # student logits (graph-attached)
        self.loss_pair = loss_pair    # [kd, ce] on the device
        self.grad_kd = grad_kd        # d(kd term)/dlogits  (+ CE part when combined)
        self.grad_ce = grad_ce        # d(ce term)/dlogits, None when combined or absent
        self.combined = combined


class _LogitLoss(torch.autograd.Function):
    """loss = kernel(z_s, z_t, y); backward = stashed gradient * upstream scalar."""

    @staticmethod
    def forward(ctx, z_s, z_t, target, kind, alpha, beta, temperature):
        loss_pair, grad_kd, _ = ops.distill_loss(
            z_s.detach(), z_t.detach(), target, kind=kind, alpha=float(alpha), beta=float(beta),
            temperature=float(temperature), ce_mode=ops.CE_NONE)
        ctx.save_for_backward(grad_kd)
        return loss_pair[0]

    @staticmethod
    def backward(ctx, upstream):
        (grad_kd,) = ctx.saved_tensors
        g = ops.scale_by_device_scalar(grad_kd, upstream.to(torch.float32).contiguous())
        return g, None, None, None, None, None, None


class _DistillTail(torch.autograd.Function):
    """(loss_kd, loss_ce) from one launch, each with its own gradient buffer."""

    @staticmethod
    def forward(ctx, z_s, z_t, target, kind, alpha, beta, temperature, kd_scale, ce_scale, stash):
        loss_pair, grad_kd, grad_ce = ops.distill_loss(
            z_s.detach(), z_t.detach(), target, kind=kind, alpha=float(alpha), beta=float(beta),
            temperature=float(temperature), kd_scale=float(kd_scale), ce_mode=ops.CE_SPLIT,
            ce_scale=float(ce_scale))
        ctx.save_for_backward(grad_kd, grad_ce)
        ctx.set_materialize_grads(False)      # a loss that takes no part in a backward arrives as None, not zeros
        stash.extend((loss_pair, grad_kd, grad_ce))
        return loss_pair[0], loss_pair[1]

    @staticmethod
    def backward(ctx, up_kd, up_ce):
        grad_kd, grad_ce = ctx.saved_tensors
        g = None
        if up_kd is not None:
            g = ops.scale_by_device_scalar(grad_kd, up_kd.to(torch.float32).contiguous())
        if up_ce is not None:
            g_ce = ops.scale_by_device_scalar(grad_ce, up_ce.to(torch.float32).contiguous())
            g = g_ce if g is None else g.add_(g_ce)
        return g, None, None, None, None, None, None, None, None, None


def dkd_loss(logits_student, logits_teacher, target, alpha, beta, temperature):
    """Decoupled KD loss, drop-in for mdistiller/distillers/DKD.py:8-32."""
    return _LogitLoss.apply(logits_student, logits_teacher, target, ops.LOSS_DKD, alpha, beta, temperature)


def kd_loss(logits_student, logits_teacher, temperature):
    """Hinton KD loss, drop-in for mdistiller/distillers/KD.py:8-13."""
    return _LogitLoss.apply(logits_student, logits_teacher, None, ops.LOSS_KD, 1.0, 1.0, temperature)


def distill_tail(logits_student: torch.Tensor, logits_teacher: torch.Tensor, target: torch.Tensor, *,
                 kind: int, alpha: float = 1.0, beta: float = 1.0, temperature: float = 1.0,
                 kd_scale: float = 1.0, ce_scale: float = 1.0, combined: bool = False
                 ) -> Tuple[torch.Tensor, torch.Tensor, Optional[TailResult]]:
    """Fused ``ce_scale*CE + kd_scale*{DKD|KD}``: returns (loss_kd, loss_ce, TailResult).

    ``combined=False``: the two returned losses are autograd-connected through
    ``_DistillTail`` (any mix of backward calls works, as ``trainer.py:427-432`` does),
    and the TailResult also exposes both gradient buffers.
    ``combined=True``: one gradient buffer holding the sum (12 B/element instead of 16);
    the returned losses are detached -- the caller backpropagates with
    ``TailResult.logits.backward(TailResult.grad_kd)`` (what our base trainer does).
    """
    if combined:
        loss_pair, grad_kd, _ = ops.distill_loss(
            logits_student.detach(), logits_teacher.detach(), target, kind=kind, alpha=float(alpha),
            beta=float(beta), temperature=float(temperature), kd_scale=float(kd_scale),
            ce_mode=ops.CE_COMBINED, ce_scale=float(ce_scale))
        res = TailResult(logits_student, loss_pair, grad_kd, None, True)
        return loss_pair[0], loss_pair[1], res
    stash: list = []
    loss_kd, loss_ce = _DistillTail.apply(logits_student, logits_teacher, target, kind, alpha, beta,
                                          temperature, kd_scale, ce_scale, stash)
    res = TailResult(logits_student, stash[0], stash[1], stash[2], False)
    return loss_kd, loss_ce, res
