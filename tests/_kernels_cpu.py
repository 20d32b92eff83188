"""Test double for the kernel entry points, backed by the CPU oracle.

Lets the HOST logic (flat buckets, membership runs, sharding, collectives, checkpoint
format) be exercised on a machine without a GPU, including world_size-2 gloo runs.  It
is injected explicitly by tests (`kernels=OracleKernels()`); the package itself never
imports it and has no CPU path.
"""

import torch

import oracle

DOT_BOTH, DOT_TASK_ONLY, DOT_KD_ONLY = 0, 1, 2


class OracleKernels:
    def __init__(self):
        self.calls = []

    def dot_step(self, p, g_task, g_kd, v_task, v_kd, *, lr, momentum, momentum_kd, weight_decay, dampening,
                 grad_scale, first_task, first_kd, membership, n=None):
        self.calls.append(("dot", p.numel(), membership, bool(first_task), bool(first_kd)))
        blend = (momentum_kd + momentum) / 2.0
        if membership == DOT_BOTH:
            oracle.dot_flat_step(p, g_task, g_kd, v_task if v_task is not None else torch.zeros_like(p), v_kd,
                                 lr=lr, m_task=momentum, m_kd=momentum_kd, weight_decay=weight_decay,
                                 dampening=dampening, first_step=False, grad_scale=grad_scale) \
                if not (first_task or first_kd) else self._dot_first(p, g_task, g_kd, v_task, v_kd, lr, momentum,
                                                                      momentum_kd, weight_decay, dampening,
                                                                      grad_scale, first_task, first_kd)
        elif membership == DOT_TASK_ONLY:
            oracle.sgd_flat_step(p, g_task, v_task, lr=lr, momentum=blend if momentum != 0 else 0.0,
                                 weight_decay=weight_decay, dampening=dampening, first_step=first_task,
                                 grad_scale=grad_scale)
        else:
            g = g_kd * grad_scale if grad_scale != 1.0 else g_kd
            if first_kd:
                v_kd.copy_(g)
            else:
                d = g.add(p, alpha=weight_decay) if weight_decay != 0 else g
                v_kd.mul_(blend).add_(d, alpha=1 - dampening)
            p.add_(v_kd, alpha=-lr)

    @staticmethod
    def _dot_first(p, g_task, g_kd, v_task, v_kd, lr, m_task, m_kd, wd, damp, gscale, first_task, first_kd):
        gt = g_task * gscale if gscale != 1.0 else g_task
        gk = g_kd * gscale if gscale != 1.0 else g_kd
        d = gt.add(p, alpha=wd) if wd != 0 else gt
        if m_task != 0:
            if first_task:
                v_task.copy_(d)
            else:
                v_task.mul_(m_task).add_(d, alpha=1 - damp)
            p.add_(v_task, alpha=-lr)
        else:
            p.add_(d, alpha=-lr)
        if first_kd:
            v_kd.copy_(gk)
        else:
            v_kd.mul_(m_kd).add_(gk, alpha=1 - damp)
        p.add_(v_kd, alpha=-lr)

    def sgd_step(self, p, g, v, *, lr, momentum, weight_decay, dampening, grad_scale, first_step, n=None,
                 grad_scale_dev=None):
        self.calls.append(("sgd", p.numel(), bool(first_step)))
        gs = grad_scale * (float(grad_scale_dev.item()) if grad_scale_dev is not None else 1.0)
        oracle.sgd_flat_step(p, g, v if v is not None else torch.zeros_like(p), lr=lr, momentum=momentum,
                             weight_decay=weight_decay, dampening=dampening, first_step=first_step,
                             grad_scale=gs)

    def adam_step(self, p, g, m, v, *, lr, beta1, beta2, eps, weight_decay, decoupled, step, grad_scale,
                  grad_scale_dev=None, n=None):
        self.calls.append(("adam", p.numel(), int(step)))
        gs = grad_scale * (float(grad_scale_dev.item()) if grad_scale_dev is not None else 1.0)
        oracle.adam_flat_step(p, g, m, v, lr=lr, beta1=beta1, beta2=beta2, eps=eps, weight_decay=weight_decay,
                              decoupled=decoupled, step=step, grad_scale=gs)

    def sum_squares(self, x, scale=1.0, out=None):
        return ((x.double() * scale) ** 2).sum().reshape(1)

    def clip_coefficient(self, total_sumsq, max_norm, out=None):
        norm = float(total_sumsq.sqrt().item())
        return torch.tensor([min(1.0, max_norm / (norm + 1e-6))], dtype=torch.float32)
