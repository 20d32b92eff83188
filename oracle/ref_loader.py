"""Import the *unmodified* reference in place (build container only).

TEST INFRASTRUCTURE.  ``/root/reference`` does not exist on the GPU box, so
nothing that runs there (``-m gpu`` tests, ``smoke()``, ``bench.py``) may call
into this module; it exists to (a) generate the committed fixtures under
``tests/golden/`` and (b) let CPU-only tests cross-check the oracle against the
live reference when the tree happens to be mounted.

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
import types

REF_ROOT = os.environ.get("MDISTILLER_REF_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "mdistiller", "distillers", "DKD.py"))


def _ensure_path() -> None:
    if not available():
        raise FileNotFoundError(f"reference tree not mounted at {REF_ROOT}")
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    name = "mdistiller.models.imagenet.vit"
    if name not in sys.modules:
        stub = types.ModuleType(name)
        for n in ("tiny", "small", "base", "large"):
            setattr(stub, f"vit_{n}_patch16_224", None)
        sys.modules[name] = stub


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
    path = os.path.join(REF_ROOT, "mdistiller", "engine", "dot.py")
    spec = importlib.util.spec_from_file_location(key, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[key] = mod
    spec.loader.exec_module(mod)
    return mod


def cifar_models():
    """cifar_model_dict with the D5 abstract-method shim applied."""
    _ensure_path()
    for sub, cls in (("resnet", "ResNet"), ("wrn", "WideResNet"),
                     ("ShuffleNetv1", "ShuffleNet"), ("mobilenetv2", "MobileNetV2")):
        m = importlib.import_module(f"mdistiller.models.cifar.{sub}")
        getattr(m, cls).__abstractmethods__ = frozenset()
    from mdistiller.models.cifar import cifar_model_dict  # type: ignore

    return cifar_model_dict


def imagenet_models():
    _ensure_path()
    from mdistiller.models.imagenet import imagenet_model_dict  # type: ignore

    return imagenet_model_dict
