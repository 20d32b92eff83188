"""ctypes binding of libmdistiller_b200.so (declarations: include/mdistiller_b200.h).

There is no fallback: if the library is missing or a symbol is absent, importing the
kernels raises.  Build it with ``python __graft_entry__.py`` (or
``python mdistiller-ddp_b200/build.py``).
"""

from __future__ import annotations

import ctypes as C
import os
import threading

from .build import LIB_PATH

# constants mirrored from the header
ABI_VERSION = 3
LOSS_DKD, LOSS_KD = 0, 1
CE_NONE, CE_SPLIT, CE_COMBINED = 0, 1, 2
DOT_BOTH, DOT_TASK_ONLY, DOT_KD_ONLY = 0, 1, 2

_ptr = C.c_void_p
_i64 = C.c_int64
_f32 = C.c_float
_int = C.c_int

SIGNATURES = {
    "mdb200_abi_version": (_int, []),
    "mdb200_error_string": (C.c_char_p, [_int]),
    "mdb200_loss_workspace_bytes": (C.c_size_t, []),
    "mdb200_distill_loss": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _i64, _i64, _int, _f32, _f32, _f32,
                                   _f32, _int, _f32, _ptr, _ptr, _i64, _ptr, _ptr, C.c_size_t, _ptr]),
    "mdb200_distill_loss_static_scratch": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _i64, _i64, _int, _f32, _f32,
                                                  _f32, _f32, _int, _f32, _ptr, _ptr, _i64, _ptr, _ptr]),
    "mdb200_scale_by_device_scalar": (_int, [_ptr, _ptr, _ptr, _i64, _ptr]),
    "mdb200_dot_step": (_int, [_ptr, _ptr, _ptr, _ptr, _ptr, _i64, _f32, _f32, _f32, _f32, _f32, _f32,
                               _int, _int, _int, _ptr]),
    "mdb200_dot_step_fused": (_int, [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _int, _ptr, _ptr,
                                     _i64, _i64, _int, _f32, _f32, _f32, _f32, _f32, _f32, _int, _int,
                                     _ptr, _int, C.c_uint32, _ptr]),
    "mdb200_sgd_step_fused": (_int, [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i64, _int, _f32, _f32, _f32, _f32,
                                     _f32, _int, _ptr, _int, C.c_uint32, _ptr]),
    "mdb200_sgd_step": (_int, [_ptr, _ptr, _ptr, _i64, _f32, _f32, _f32, _f32, _f32, _ptr, _int, _ptr]),
    "mdb200_adam_step": (_int, [_ptr, _ptr, _ptr, _ptr, _i64, _f32, C.c_double, C.c_double, _f32, _f32, _int, _i64,
                                _f32, _ptr, _ptr]),
    "mdb200_gather_flat": (_int, [_ptr, _ptr, _int, _ptr, _i64, _ptr]),
    "mdb200_metrics_record": (_int, [_ptr, _ptr, _i64, _i64, _i64, _ptr, _ptr, _ptr, _ptr]),
    "mdb200_sum_squares": (_int, [_ptr, _i64, _f32, _ptr, _ptr, C.c_size_t, _ptr]),
    "mdb200_clip_coefficient": (_int, [_ptr, _f32, _ptr, _ptr]),
    "mdb200_amp_step_scale": (_int, [_ptr, _ptr, _ptr, _f32, _f32, _int, _f32, _ptr, _ptr]),
}

_lock = threading.Lock()
_lib = None


class KernelLibraryError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load (once) and type the shared library; raises KernelLibraryError if unusable."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = os.environ.get("MDB200_LIB", LIB_PATH)      # developer knob: A/B a differently built library
        if not os.path.exists(path):
            raise KernelLibraryError(
                f"{path} is missing: the sm_100a kernels are not built. "
                "Run `python __graft_entry__.py` (needs nvcc). There is no CPU/PyTorch fallback.")
        try:
            lib = C.CDLL(path)
        except OSError as e:  # pragma: no cover
            raise KernelLibraryError(f"cannot load {path}: {e}") from e
        for name, (res, args) in SIGNATURES.items():
            try:
                fn = getattr(lib, name)
            except AttributeError as e:
                raise KernelLibraryError(f"{path} does not export {name}; rebuild it") from e
            fn.restype, fn.argtypes = res, args
        got = lib.mdb200_abi_version()
        if got != ABI_VERSION:
            raise KernelLibraryError(f"ABI version {got} != expected {ABI_VERSION}; rebuild the library")
        _lib = lib
    return _lib


def check(code: int, what: str) -> None:
    if code != 0:
        msg = load().mdb200_error_string(code).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {code})")
