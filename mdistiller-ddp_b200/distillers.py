"""DKD / KD distillers with the fused logit tail.

Same constructor and ``forward_train(image, target, **kwargs) -> (logits_student,
{"loss_ce", "loss_kd"})`` contract as the reference classes they replace
(mdistiller/distillers/DKD.py:54-84, KD.py:16-39, base class _base.py:6-44), so they can
be registered under ``distiller_dict["DKD"]`` / ``["KD"]`` and driven by an unchanged
``tools/train.py``.  The backbones are whatever modules the caller passes in; they must
return ``(logits, feature_dict)`` like every reference model
(mdistiller/models/cifar/resnet.py:216-238).
"""

from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import ops
from .losses import TailResult, distill_tail


class Distiller(nn.Module):
    """Student + frozen-in-eval teacher pair (interface of distillers/_base.py:6-44)."""

    def __init__(self, student: nn.Module, teacher: nn.Module):
        super().__init__()
        self.student = student
        self.teacher = teacher
        # set by a trainer that backpropagates from the logits itself (see losses.distill_tail)
        self.combined_tail = False
        self.last_tail: Optional[TailResult] = None

    @property
    def module(self):          # lets code written for a DDP-wrapped distiller run unwrapped
        return self

    def train(self, mode: bool = True):
        if not isinstance(mode, bool):
            raise ValueError("training mode is expected to be boolean")
        self.training = mode
        for child in self.children():
            child.train(mode)
        self.teacher.eval()    # the teacher never leaves eval mode (_base.py:16-24)
        return self

    def get_learnable_parameters(self):
        return [p for _, p in self.student.named_parameters()]

    def get_extra_parameters(self):
        return 0

    def forward_train(self, **kwargs):
        raise NotImplementedError()

    def forward_test(self, image):
        return self.student(image)[0]

    def forward(self, **kwargs):
        if self.training:
            return self.forward_train(**kwargs)
        return self.forward_test(kwargs["image"])

    # shared by DKD and KD
    def _logits(self, image):
        logits_student, _ = self.student(image)
        with torch.no_grad():
            logits_teacher, _ = self.teacher(image)
        return logits_student, logits_teacher

    def _tail(self, logits_student, logits_teacher, target, **kw):
        # the tail computes in fp32 like the reference's softmax / cross_entropy do under autocast
        # (trainer.py:268 wraps only the forward); fp32 logits pass through untouched
        logits_fp32, logits_teacher = logits_student.float(), logits_teacher.float()
        loss_kd, loss_ce, res = distill_tail(logits_fp32, logits_teacher, target,
                                             combined=self.combined_tail, **kw)
        self.last_tail = res
        return logits_student, {"loss_ce": loss_ce, "loss_kd": loss_kd}


class DKD(Distiller):
    """Decoupled Knowledge Distillation (reference: distillers/DKD.py:54-84)."""

    def __init__(self, student, teacher, cfg):
        super().__init__(student, teacher)
        self.ce_loss_weight = cfg.DKD.CE_WEIGHT
        self.alpha = cfg.DKD.ALPHA
        self.beta = cfg.DKD.BETA
        self.temperature = cfg.DKD.T
        self.warmup = cfg.DKD.WARMUP

    def forward_train(self, image, target, **kwargs):
        logits_student, logits_teacher = self._logits(image)
        ramp = min(kwargs["epoch"] / self.warmup, 1.0)          # DKD.py:72
        return self._tail(logits_student, logits_teacher, target, kind=ops.LOSS_DKD,
                          alpha=self.alpha, beta=self.beta, temperature=self.temperature,
                          kd_scale=ramp, ce_scale=self.ce_loss_weight)


class KD(Distiller):
    """Hinton et al. logit distillation (reference: distillers/KD.py:16-39)."""

    def __init__(self, student, teacher, cfg):
        super().__init__(student, teacher)
        self.temperature = cfg.KD.TEMPERATURE
        self.ce_loss_weight = cfg.KD.LOSS.CE_WEIGHT
        self.kd_loss_weight = cfg.KD.LOSS.KD_WEIGHT

    def forward_train(self, image, target, **kwargs):
        logits_student, logits_teacher = self._logits(image)
        return self._tail(logits_student, logits_teacher, target, kind=ops.LOSS_KD,
                          temperature=self.temperature, kd_scale=self.kd_loss_weight,
                          ce_scale=self.ce_loss_weight)
