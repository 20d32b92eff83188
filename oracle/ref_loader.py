"""Import the *unmodified* reference: mounted tree here, staged archive on the GPU box.

TEST INFRASTRUCTURE.  ``/root/reference`` does not exist on the GPU box; what runs there
reads the archive ``oracle/make_ref.sh`` staged under the git-ignored ``oracle/_ref/``
(unpacked by ``mdistiller_ddp_b200.refhost``), never ``/root/reference``.  This module (a)
generates the committed fixtures under ``tests/golden/``, (b) lets tests cross-check the
oracle against the live reference, and (c) gives ``bench.py --impl reference`` the
reference's own functions to time.

Shims (no reference file is edited), following SURVEY.md section 0.1:
  D6  ``mdistiller/models/imagenet/vit.py`` needs timm and evaluates an
      undefined annotation -> pre-seed ``sys.modules`` with a stub.
  D5  CIFAR ResNet/WRN classes miss ``get_arch`` -> clear ``__abstractmethods__``.
  D8  ``mdistiller/engine/__init__.py`` pulls tensorboardX -> load
      ``engine/dot.py`` by file path instead.
"""

from __future__ import annotations

import importlib
import importlib.util
import os
import sys

import mdistiller_ddp_b200  # noqa: F401  (routes the hyphenated package directory)
from mdistiller_ddp_b200 import refhost


def available() -> bool:
    return refhost.available()


def _ensure_path() -> None:
    refhost.ensure_importable()


def __getattr__(name):          # REF_ROOT: resolved lazily (mounted tree, or the staged archive unpacked)
    if name == "REF_ROOT":
        return refhost.locate()
    raise AttributeError(name)


def loss_functions():
    """(dkd_loss, kd_loss) exactly as defined at DKD.py:8 and KD.py:8."""
    _ensure_path()
    from mdistiller.distillers.DKD import dkd_loss  # type: ignore
    from mdistiller.distillers.KD import kd_loss  # type: ignore

    return dkd_loss, kd_loss


def distiller_classes():
    """(DKD, KD) distiller classes (DKD.py:54, KD.py:16)."""
    _ensure_path()
    from mdistiller.distillers.DKD import DKD  # type: ignore
    from mdistiller.distillers.KD import KD  # type: ignore

    return DKD, KD


def dot_module():
    """``mdistiller/engine/dot.py`` loaded by path (avoids tensorboardX)."""
    _ensure_path()
    key = "_ref_engine_dot"
    if key in sys.modules:
        return sys.modules[key]
    path = os.path.join(refhost.locate(), "mdistiller", "engine", "dot.py")
    spec = importlib.util.spec_from_file_location(key, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[key] = mod
    spec.loader.exec_module(mod)
    return mod


def cifar_models():
    """cifar_model_dict with the D5 abstract-method shim applied."""
    return refhost.cifar_models()


def imagenet_models():
    return refhost.imagenet_models()
