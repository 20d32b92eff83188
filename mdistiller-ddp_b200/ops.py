"""Tensor-level wrappers over the C-ABI: pointers out of torch tensors, nothing else.

torch is used here for device memory and the current stream only.  Every function
insists on CUDA fp32 tensors and raises otherwise -- there is no CPU path.
"""

from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import (CE_COMBINED, CE_NONE, CE_SPLIT, DOT_BOTH, DOT_KD_ONLY, DOT_TASK_ONLY,  # noqa: F401
                   LOSS_DKD, LOSS_KD)

_workspaces = {}
launch_count = 0   # kernels of ours launched by this process (bench.py reports it)


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _need_cuda_f32(t: torch.Tensor, name: str) -> None:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: the distillation tail has no CPU fallback")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32, got {t.dtype}")


def loss_workspace(device: torch.device) -> torch.Tensor:
    """Per (device, stream) reduction workspace, zeroed once (the kernel re-zeroes its ticket)."""
    key = (device.index if device.index is not None else torch.cuda.current_device(),
           torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None:
        nbytes = _lib.load().mdb200_loss_workspace_bytes()
        ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def distill_loss(z_student: torch.Tensor, z_teacher: torch.Tensor, target: Optional[torch.Tensor], *,
                 kind: int, alpha: float = 1.0, beta: float = 1.0, temperature: float = 1.0,
                 kd_scale: float = 1.0, ce_mode: int = CE_NONE, ce_scale: float = 1.0,
                 grad_kd: Optional[torch.Tensor] = None, grad_ce: Optional[torch.Tensor] = None,
                 loss_out: Optional[torch.Tensor] = None
                 ) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
    """One launch: returns (loss_pair[2] = {kd, ce}, grad_kd, grad_ce or None)."""
    global launch_count
    lib = _lib.load()
    _need_cuda_f32(z_student, "logits_student")
    _need_cuda_f32(z_teacher, "logits_teacher")
    if z_student.dim() != 2 or z_student.shape != z_teacher.shape:
        raise ValueError(f"logits must be matching [rows, classes]; got {tuple(z_student.shape)} "
                         f"and {tuple(z_teacher.shape)}")
    if z_student.stride(1) != 1:
        z_student = z_student.contiguous()
    if z_teacher.stride(1) != 1:
        z_teacher = z_teacher.contiguous()
    rows, classes = z_student.shape
    tptr = None
    if target is not None:
        if not target.is_cuda or target.dtype != torch.int64:
            raise TypeError("target must be a CUDA int64 tensor")
        target = target.reshape(-1).contiguous()
        if target.numel() != rows:
            raise ValueError("target length != rows")
        tptr = target.data_ptr()
    dev = z_student.device
    if grad_kd is None:
        grad_kd = torch.empty((rows, classes), dtype=torch.float32, device=dev)
    if ce_mode == CE_SPLIT and grad_ce is None:
        grad_ce = torch.empty((rows, classes), dtype=torch.float32, device=dev)
    if loss_out is None:
        loss_out = torch.empty(2, dtype=torch.float32, device=dev)
    ws = loss_workspace(dev)
    with torch.cuda.device(dev):
        rc = lib.mdb200_distill_loss(
            z_student.data_ptr(), z_teacher.data_ptr(), tptr, rows, classes,
            z_student.stride(0), z_teacher.stride(0), kind, alpha, beta, temperature,
            kd_scale, ce_mode, ce_scale, grad_kd.data_ptr(),
            grad_ce.data_ptr() if (grad_ce is not None and ce_mode == CE_SPLIT) else None,
            grad_kd.stride(0), loss_out.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    _lib.check(rc, "mdb200_distill_loss")
    launch_count += 1
    return loss_out, grad_kd, (grad_ce if ce_mode == CE_SPLIT else None)


def scale_by_device_scalar(grad: torch.Tensor, upstream: torch.Tensor,
                           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out = grad * upstream (0-dim device tensor); no host sync."""
    global launch_count
    lib = _lib.load()
    _need_cuda_f32(grad, "grad")
    _need_cuda_f32(upstream, "upstream")
    if not grad.is_contiguous():
        grad = grad.contiguous()
    if out is None:
        out = torch.empty_like(grad)
    with torch.cuda.device(grad.device):
        rc = lib.mdb200_scale_by_device_scalar(out.data_ptr(), grad.data_ptr(), upstream.data_ptr(),
                                               grad.numel(), _stream_ptr(grad.device))
    _lib.check(rc, "mdb200_scale_by_device_scalar")
    launch_count += 1
    return out


def _flat_ptr(t: Optional[torch.Tensor], name: str, n: int) -> Optional[int]:
    if t is None:
        return None
    _need_cuda_f32(t, name)
    if not t.is_contiguous() or t.numel() < n:
        raise ValueError(f"{name} must be contiguous with at least {n} elements")
    return t.data_ptr()


def dot_step(p: torch.Tensor, g_task: Optional[torch.Tensor], g_kd: Optional[torch.Tensor],
             v_task: Optional[torch.Tensor], v_kd: Optional[torch.Tensor], *, lr: float, momentum: float,
             momentum_kd: float, weight_decay: float = 0.0, dampening: float = 0.0, grad_scale: float = 1.0,
             first_task: bool = False, first_kd: bool = False, membership: int = DOT_BOTH,
             n: Optional[int] = None) -> None:
    """In-place DOT update of the flat range p[:n]."""
    global launch_count
    lib = _lib.load()
    n = p.numel() if n is None else n
    ptrs = (_flat_ptr(p, "p", n), _flat_ptr(g_task, "g_task", n), _flat_ptr(g_kd, "g_kd", n),
            _flat_ptr(v_task, "v_task", n), _flat_ptr(v_kd, "v_kd", n))
    with torch.cuda.device(p.device):
        rc = lib.mdb200_dot_step(*ptrs, n, lr, momentum, momentum_kd, weight_decay, dampening, grad_scale,
                                 int(first_task), int(first_kd), membership, _stream_ptr(p.device))
    _lib.check(rc, "mdb200_dot_step")
    launch_count += 1


def sgd_step(p: torch.Tensor, g: torch.Tensor, v: Optional[torch.Tensor], *, lr: float, momentum: float,
             weight_decay: float = 0.0, dampening: float = 0.0, grad_scale: float = 1.0,
             first_step: bool = False, n: Optional[int] = None,
             grad_scale_dev: Optional[torch.Tensor] = None) -> None:
    """In-place momentum-SGD update of the flat range p[:n]."""
    global launch_count
    lib = _lib.load()
    n = p.numel() if n is None else n
    ptrs = (_flat_ptr(p, "p", n), _flat_ptr(g, "g", n), _flat_ptr(v, "v", n))
    gs = _flat_ptr(grad_scale_dev, "grad_scale_dev", 1)
    with torch.cuda.device(p.device):
        rc = lib.mdb200_sgd_step(*ptrs, n, lr, momentum, weight_decay, dampening, grad_scale, gs, int(first_step),
                                 _stream_ptr(p.device))
    _lib.check(rc, "mdb200_sgd_step")
    launch_count += 1


def adam_step(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, *, lr: float, beta1: float,
              beta2: float, eps: float, weight_decay: float = 0.0, decoupled: bool = False, step: int = 1,
              grad_scale: float = 1.0, grad_scale_dev: Optional[torch.Tensor] = None,
              n: Optional[int] = None) -> None:
    """In-place Adam / AdamW update of the flat range p[:n] (step is 1-based)."""
    global launch_count
    lib = _lib.load()
    n = p.numel() if n is None else n
    ptrs = (_flat_ptr(p, "p", n), _flat_ptr(g, "g", n), _flat_ptr(m, "m", n), _flat_ptr(v, "v", n))
    gs = _flat_ptr(grad_scale_dev, "grad_scale_dev", 1)
    with torch.cuda.device(p.device):
        rc = lib.mdb200_adam_step(*ptrs, n, lr, beta1, beta2, eps, weight_decay, int(decoupled), int(step),
                                  grad_scale, gs, _stream_ptr(p.device))
    _lib.check(rc, "mdb200_adam_step")
    launch_count += 1


def sum_squares(x: torch.Tensor, scale: float = 1.0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out (1 x float64 on the device) = sum((scale * x)^2), deterministic."""
    global launch_count
    lib = _lib.load()
    _need_cuda_f32(x, "x")
    if not x.is_contiguous():
        raise ValueError("x must be contiguous")
    if out is None:
        out = torch.empty(1, dtype=torch.float64, device=x.device)
    ws = loss_workspace(x.device)
    with torch.cuda.device(x.device):
        rc = lib.mdb200_sum_squares(x.data_ptr(), x.numel(), scale, out.data_ptr(), ws.data_ptr(), ws.numel(),
                                    _stream_ptr(x.device))
    _lib.check(rc, "mdb200_sum_squares")
    launch_count += 1
    return out


def clip_coefficient(total_sumsq: torch.Tensor, max_norm: float, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out (1 x float32) = min(1, max_norm / (sqrt(total_sumsq) + 1e-6)) -- clip_grad_norm_'s factor."""
    global launch_count
    lib = _lib.load()
    if not total_sumsq.is_cuda or total_sumsq.dtype != torch.float64:
        raise TypeError("total_sumsq must be a CUDA float64 tensor")
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=total_sumsq.device)
    with torch.cuda.device(total_sumsq.device):
        rc = lib.mdb200_clip_coefficient(total_sumsq.data_ptr(), max_norm, out.data_ptr(),
                                         _stream_ptr(total_sumsq.device))
    _lib.check(rc, "mdb200_clip_coefficient")
    launch_count += 1
    return out


def amp_step_scale(total_sumsq: torch.Tensor, scale: torch.Tensor, growth_tracker: torch.Tensor, *, growth_factor: float,
                   backoff_factor: float, growth_interval: int, max_norm: float = 0.0,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """GradScaler's per-step decisions on the device (see mdb200_amp_step_scale): returns the device scalar to pass as
    ``grad_scale_dev`` (1/scale x clipping coefficient, negative = skip the update); ``scale`` and
    ``growth_tracker`` are updated in place."""
    global launch_count
    lib = _lib.load()
    if not (total_sumsq.is_cuda and total_sumsq.dtype == torch.float64 and scale.is_cuda and scale.dtype == torch.float32
            and growth_tracker.is_cuda and growth_tracker.dtype == torch.int32):
        raise TypeError("total_sumsq: CUDA float64, scale: CUDA float32, growth_tracker: CUDA int32")
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=scale.device)
    with torch.cuda.device(scale.device):
        rc = lib.mdb200_amp_step_scale(total_sumsq.data_ptr(), scale.data_ptr(), growth_tracker.data_ptr(),
                                       growth_factor, backoff_factor, int(growth_interval), float(max_norm or 0.0),
                                       out.data_ptr(), _stream_ptr(scale.device))
    _lib.check(rc, "mdb200_amp_step_scale")
    launch_count += 1
    return out


def metrics_record(logits: torch.Tensor, target: torch.Tensor, loss_ce: Optional[torch.Tensor],
                   loss_kd: Optional[torch.Tensor]) -> torch.Tensor:
    """5 floats on the device: {loss_ce, loss_kd, top-1 hits, top-5 hits, rows} -- one launch."""
    global launch_count
    lib = _lib.load()
    _need_cuda_f32(logits, "logits")
    if logits.dim() != 2 or logits.stride(1) != 1:
        logits = logits.reshape(logits.shape[0], -1).contiguous()
    if not target.is_cuda or target.dtype != torch.int64:
        raise TypeError("target must be a CUDA int64 tensor")
    target = target.reshape(-1).contiguous()
    rec = torch.zeros(5, dtype=torch.float32, device=logits.device)

    def scalar(t):
        if t is None:
            return None
        _need_cuda_f32(t, "loss")
        return t.data_ptr()
    with torch.cuda.device(logits.device):
        rc = lib.mdb200_metrics_record(logits.data_ptr(), target.data_ptr(), logits.shape[0], logits.shape[1],
                                       logits.stride(0), scalar(loss_ce), scalar(loss_kd), rec.data_ptr(),
                                       _stream_ptr(logits.device))
    _lib.check(rc, "mdb200_metrics_record")
    launch_count += 1
    return rec


def gather_flat(dst_flat: torch.Tensor, src_ptrs, chunk_table: torch.Tensor) -> None:
    """One launch: scatter-free copy of per-parameter gradient tensors (device addresses, 0 = none)
    into the flat bucket; see mdb200_gather_flat."""
    global launch_count
    import ctypes as C
    lib = _lib.load()
    _need_cuda_f32(dst_flat, "dst_flat")
    if not chunk_table.is_cuda or chunk_table.dtype != torch.int64 or not chunk_table.is_contiguous():
        raise TypeError("chunk_table must be a contiguous CUDA int64 tensor")
    n = len(src_ptrs)
    table = (C.c_void_p * n)(*src_ptrs)
    with torch.cuda.device(dst_flat.device):
        rc = lib.mdb200_gather_flat(dst_flat.data_ptr(), table, n, chunk_table.data_ptr(), chunk_table.shape[0],
                                    _stream_ptr(dst_flat.device))
    _lib.check(rc, "mdb200_gather_flat")
    launch_count += 1


def dot_step_fused(p_local: torch.Tensor, g_task_local: torch.Tensor, g_kd_local: torch.Tensor,
                   v_task: Optional[torch.Tensor], v_kd: torch.Tensor, *, multicast=None, peers=None,
                   kd_is_reduced: bool, shard_lo: int, shard_len: int, world: int, lr: float, momentum: float,
                   momentum_kd: float, weight_decay: float = 0.0, dampening: float = 0.0, grad_scale: float = 1.0,
                   first_task: bool = False, first_kd: bool = False, sync=None) -> None:
    """Fused reduce + DOT update + broadcast for this rank's shard (see mdb200_dot_step_fused).

    ``multicast`` = (p_mc, g_task_mc, g_kd_mc) integer multicast addresses or None;
    ``peers`` = (p_ptrs, g_task_ptrs, g_kd_ptrs), each a list of ``world`` integer base addresses;
    ``sync`` = (pad addresses of every rank, this rank, epoch) for the in-kernel handshake, or None when the
    caller brackets the launch with barriers itself.
    """
    global launch_count
    import ctypes as C
    lib = _lib.load()
    n = p_local.numel()
    ptrs = (_flat_ptr(p_local, "p_local", n), _flat_ptr(g_task_local, "g_task_local", n),
            _flat_ptr(g_kd_local, "g_kd_local", n), _flat_ptr(v_task, "v_task", shard_len),
            _flat_ptr(v_kd, "v_kd", shard_len))
    mc = multicast if multicast is not None else (None, None, None)

    def table(addrs):
        if addrs is None:
            return None
        return (C.c_void_p * world)(*[int(a) for a in addrs])

    pt = peers if peers is not None else (None, None, None)
    tabs = [table(t) for t in pt]
    with torch.cuda.device(p_local.device):
        rc = lib.mdb200_dot_step_fused(
            ptrs[0], mc[0] or None, tabs[0], ptrs[1], mc[1] or None, tabs[1], ptrs[2], mc[2] or None, tabs[2],
            int(kd_is_reduced), ptrs[3], ptrs[4], shard_lo, shard_len, world, lr, momentum, momentum_kd,
            weight_decay, dampening, grad_scale, int(first_task), int(first_kd),
            table(sync[0]) if sync is not None else None, int(sync[1]) if sync is not None else 0,
            int(sync[2]) if sync is not None else 0, _stream_ptr(p_local.device))
    _lib.check(rc, "mdb200_dot_step_fused")
    launch_count += 1


def sgd_step_fused(p_local: torch.Tensor, g_local: torch.Tensor, v: Optional[torch.Tensor], *, multicast=None, peers=None,
                   shard_lo: int, shard_len: int, world: int, lr: float, momentum: float, weight_decay: float = 0.0,
                   dampening: float = 0.0, grad_scale: float = 1.0, first_step: bool = False, sync=None) -> None:
    """Fused reduce + momentum SGD + broadcast for this rank's shard (see mdb200_sgd_step_fused).
    ``multicast`` = (p_mc, g_mc) or None; ``peers`` = (p_ptrs, g_ptrs); ``sync`` as in dot_step_fused."""
    global launch_count
    import ctypes as C
    lib = _lib.load()
    n = p_local.numel()
    ptrs = (_flat_ptr(p_local, "p_local", n), _flat_ptr(g_local, "g_local", n), _flat_ptr(v, "v", shard_len))
    mc = multicast if multicast is not None else (None, None)

    def table(addrs):
        if addrs is None:
            return None
        return (C.c_void_p * world)(*[int(a) for a in addrs])

    pt = peers if peers is not None else (None, None)
    tabs = [table(t) for t in pt]
    with torch.cuda.device(p_local.device):
        rc = lib.mdb200_sgd_step_fused(
            ptrs[0], mc[0] or None, tabs[0], ptrs[1], mc[1] or None, tabs[1], ptrs[2], shard_lo, shard_len, world,
            lr, momentum, weight_decay, dampening, grad_scale, int(first_step),
            table(sync[0]) if sync is not None else None, int(sync[1]) if sync is not None else 0,
            int(sync[2]) if sync is not None else 0, _stream_ptr(p_local.device))
    _lib.check(rc, "mdb200_sgd_step_fused")
    launch_count += 1
