"""Out-of-bounds write detection without compute-sanitizer (closed on this pool): every output lives
inside a larger buffer with sentinel words on both sides; the sentinels must survive the launch."""

import pytest
import torch

pytestmark = pytest.mark.gpu

SENTINEL = 12345.0
PAD = 64


def _guarded(n, fill=None):
    """A length-n fp32 view (16-byte aligned) with PAD sentinel words before and after it."""
    buf = torch.full((n + 2 * PAD,), SENTINEL, device="cuda")
    view = buf[PAD:PAD + n]
    if fill is not None:
        view.copy_(fill)
    return buf, view


def _intact(buf, n):
    return bool((buf[:PAD] == SENTINEL).all() and (buf[PAD + n:] == SENTINEL).all())


@pytest.mark.parametrize("B,C", [(1, 8), (5, 37), (67, 100), (9, 112), (33, 128), (7, 200), (6, 512), (5, 1000), (3, 2048), (2, 4099)])
def test_loss_kernel_writes_stay_inside_the_gradient_rows(B, C):
    from mdistiller_ddp_b200 import ops
    zs, zt = torch.randn(B, C, device="cuda") * 3, torch.randn(B, C, device="cuda") * 3
    y = torch.randint(0, C, (B,), device="cuda")
    for kind in (ops.LOSS_DKD, ops.LOSS_KD):
        for T, ce in ((1.0, ops.CE_SPLIT), (4.0, ops.CE_SPLIT), (4.0, ops.CE_COMBINED), (1.0, ops.CE_NONE)):
            bk, gk = _guarded(B * C)
            bc, gc = _guarded(B * C)
            bl, lo = _guarded(4)
            ops.distill_loss(zs, zt, y, kind=kind, alpha=1.0, beta=2.0, temperature=T, ce_mode=ce,
                             grad_kd=gk.view(B, C), grad_ce=gc.view(B, C), loss_out=lo[:2])
            torch.cuda.synchronize()
            assert _intact(bk, B * C) and _intact(bc, B * C) and _intact(bl, 2)
            assert torch.isfinite(gk).all() and (ce != ops.CE_SPLIT or torch.isfinite(gc).all())
            if ce != ops.CE_SPLIT:
                assert (gc == SENTINEL).all(), "grad_ce must not be touched unless CE_SPLIT"


@pytest.mark.parametrize("n", [1, 3, 4, 5, 1023, 4099, 100_003])
def test_update_kernels_write_only_their_range(n):
    from mdistiller_ddp_b200 import ops
    src = [torch.randn(n, device="cuda") for _ in range(5)]
    for mem in (ops.DOT_BOTH, ops.DOT_TASK_ONLY, ops.DOT_KD_ONLY):
        for first in (False, True):
            (bp, p), (bt, vt), (bk, vk) = _guarded(n, src[0]), _guarded(n, src[3]), _guarded(n, src[4])
            ops.dot_step(p, src[1], src[2], vt, vk, lr=0.1, momentum=0.8, momentum_kd=0.9, weight_decay=1e-4,
                         first_task=first, first_kd=first, membership=mem)
            torch.cuda.synchronize()
            assert _intact(bp, n) and _intact(bt, n) and _intact(bk, n)
            if mem == ops.DOT_TASK_ONLY:
                assert torch.equal(vk, src[4]), "KD momentum must be untouched for task-only ranges"
            if mem == ops.DOT_KD_ONLY:
                assert torch.equal(vt, src[3])
    (bp, p), (bt, m), (bk, v) = _guarded(n, src[0]), _guarded(n, src[3]), _guarded(n, src[4].abs())
    ops.adam_step(p, src[1], m, v, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True, step=2)
    ops.sgd_step(p, src[1], m, lr=0.1, momentum=0.9, weight_decay=1e-4)
    bo, out = _guarded(n)
    ops.scale_by_device_scalar(src[1], torch.tensor(0.5, device="cuda"), out=out)
    torch.cuda.synchronize()
    assert _intact(bp, n) and _intact(bt, n) and _intact(bk, n) and _intact(bo, n)
    assert torch.allclose(out, src[1] * 0.5)


def test_gather_writes_only_the_parameter_slices():
    from mdistiller_ddp_b200 import ops
    from mdistiller_ddp_b200.flat import FlatSpace
    params = [torch.nn.Parameter(torch.randn(*s, device="cuda")) for s in [(7, 3), (5,), (1,), (33001,), (4, 4, 3, 3)]]
    sp = FlatSpace(params)
    bb, bucket = _guarded(sp.total)
    bucket.fill_(-1.0)
    grads = [torch.randn_like(p) for p in params]
    grads[1] = None
    ops.gather_flat(bucket, [0 if g is None else g.data_ptr() for g in grads], sp.chunk_table(chunk_elems=1000))
    torch.cuda.synchronize()
    assert _intact(bb, sp.total)
    for o, e, p, g in zip(sp.offsets, sp.ends, params, grads):
        assert torch.equal(bucket[o:o + p.numel()], torch.zeros(p.numel(), device="cuda") if g is None else g.reshape(-1))
        assert (bucket[o + p.numel():e] == -1.0).all(), "alignment padding is never written"
