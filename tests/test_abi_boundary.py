"""The C-ABI library loads and exports every symbol the header declares (no GPU needed)."""

import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = []
    inc = os.path.join(ROOT, "include")
    for f in sorted(os.listdir(inc)):
        if f.endswith(".h"):
            text = open(os.path.join(inc, f)).read()
            names += re.findall(r"MDB200_API\s+[\w\s\*]+?\b(mdb200_\w+)\s*\(", text)
    return names


@pytest.fixture(scope="module")
def lib():
    import mdistiller_ddp_b200 as b200
    b200.build_library()
    from mdistiller_ddp_b200 import _lib
    return _lib.load()


def test_header_declares_the_expected_surface():
    assert set(_declared()) >= {"mdb200_abi_version", "mdb200_error_string", "mdb200_loss_workspace_bytes",
                                "mdb200_distill_loss", "mdb200_scale_by_device_scalar", "mdb200_dot_step",
                                "mdb200_sgd_step", "mdb200_dot_step_fused", "mdb200_sgd_step_fused", "mdb200_distill_loss_static_scratch",
                                "mdb200_adam_step", "mdb200_sum_squares", "mdb200_clip_coefficient", "mdb200_gather_flat", "mdb200_metrics_record"}


def test_every_declared_symbol_is_exported_and_bound(lib):
    from mdistiller_ddp_b200 import _lib
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in include/*.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"


def test_version_and_error_strings(lib):
    assert lib.mdb200_abi_version() == 3
    assert lib.mdb200_error_string(0) == b"ok"
    assert b"workspace" in lib.mdb200_error_string(-3)
    assert lib.mdb200_loss_workspace_bytes() >= 64 + 2 * 8 * 1184


def test_argument_errors_are_reported_before_any_cuda_work(lib):
    """Null pointers / bad modes are rejected with the documented codes (no device touched)."""
    rc = lib.mdb200_distill_loss(None, None, None, 4, 8, 8, 8, 0, 1.0, 1.0, 1.0, 1.0, 0, 1.0,
                                 None, None, 8, None, None, 0, None)
    assert rc == -1
    buf = (ctypes.c_float * 64)()
    p = ctypes.addressof(buf)
    rc = lib.mdb200_distill_loss(p, p, p, 4, 8, 4, 8, 0, 1.0, 1.0, 1.0, 1.0, 0, 1.0, p, None, 8, p, p, 1 << 20, None)
    assert rc == -2          # ld_s < classes
    rc = lib.mdb200_distill_loss(p, p, p, 4, 8, 8, 8, 0, 1.0, 1.0, 1.0, 1.0, 0, 1.0, p, None, 8, p, p, 16, None)
    assert rc == -3          # workspace too small
    rc = lib.mdb200_distill_loss(p, p, p, 4, 8, 8, 8, 7, 1.0, 1.0, 1.0, 1.0, 0, 1.0, p, None, 8, p, p, 1 << 20, None)
    assert rc == -1          # unknown loss kind
    assert lib.mdb200_dot_step(None, None, None, None, None, 16, .1, .9, .9, 0., 0., 1., 0, 0, 0, None) == -1
    assert lib.mdb200_dot_step(p, p, p, p, p, 16, .1, .9, .9, 0., 0., 1., 0, 0, 9, None) == -1
    assert lib.mdb200_dot_step(p, p, p, p, p, 0, .1, .9, .9, 0., 0., 1., 0, 0, 0, None) == 0   # empty range
    assert lib.mdb200_sgd_step(p, p, None, 16, .1, .9, 0., 0., 1., None, 0, None) == -1            # momentum needs v
    fused = lambda **kw: lib.mdb200_dot_step_fused(  # noqa: E731
        kw.get("p", p), None, None, p, None, None, p, None, None, 1, p, p, kw.get("lo", 0), kw.get("n", 16),
        kw.get("world", 2), .1, .9, .9, 0., 0., .5, 0, 0, kw.get("pads"), kw.get("rank", 0), 1, None)
    assert fused() == -1                       # neither multicast addresses nor peer tables
    assert fused(world=17) == -1               # more peers than the kernel addresses
    assert fused(lo=2) == -4                   # shard must start on a 16-byte slot
    assert fused(p=None) == -1
    sgd_fused = lambda **kw: lib.mdb200_sgd_step_fused(  # noqa: E731
        kw.get("p", p), kw.get("mc"), kw.get("peers"), p, kw.get("mc"), kw.get("peers"), kw.get("v", p), kw.get("lo", 0), 16,
        kw.get("world", 2), .1, kw.get("m", .9), 0., 0., .5, 0, kw.get("pads"), kw.get("rank", 0), 1, None)
    assert sgd_fused() == -1                   # neither multicast addresses nor peer tables
    assert sgd_fused(world=0) == -1 and sgd_fused(lo=6) == -4 and sgd_fused(p=None) == -1
    assert sgd_fused(v=None, mc=p) == -1       # momentum needs its buffer
    two = (ctypes.c_void_p * 2)(p, None)
    assert sgd_fused(peers=two) == -1          # a peer's base address is missing
    two = (ctypes.c_void_p * 2)(p, p)
    assert sgd_fused(peers=two, pads=two, rank=2) == -1      # rank outside the world
    assert lib.mdb200_adam_step(p, p, p, p, 16, 1e-3, .9, .999, 1e-8, 0., 0, 0, 1., None, None) == -1   # step is 1-based
    assert lib.mdb200_sum_squares(p, 16, 1., p, p, 16, None) == -3
    assert lib.mdb200_clip_coefficient(p, 0., p, None) == -1
    assert lib.mdb200_distill_loss_static_scratch(None, None, None, 4, 8, 8, 8, 0, 1., 1., 1., 1., 0, 1.,
                                                  None, None, 8, None, None) == -1


def test_ops_refuse_cpu_tensors():
    import torch
    from mdistiller_ddp_b200 import ops
    z = torch.zeros(4, 8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.distill_loss(z, z, torch.zeros(4, dtype=torch.long), kind=ops.LOSS_DKD)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.dot_step(torch.zeros(8), torch.zeros(8), torch.zeros(8), torch.zeros(8), torch.zeros(8),
                     lr=.1, momentum=.9, momentum_kd=.9)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from mdistiller_ddp_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.KernelLibraryError, match="no CPU/PyTorch fallback"):
        _lib.load()
