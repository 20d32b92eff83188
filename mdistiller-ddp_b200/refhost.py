"""Make an UNMODIFIED copy of the reference importable next to this package.

The tail drops into the reference (``install.install()`` patches its registries, then the
unchanged ``tools/train.py`` runs).  The reference as shipped does not import on this image
(SURVEY.md section 0.1): it needs ``yacs``, ``tensorboardX`` and ``timm`` (absent), evaluates an
undefined annotation in ``models/imagenet/vit.py`` (D6) and leaves ``get_arch`` abstract on its
CIFAR models (D5).  ``ensure_importable()`` supplies stand-ins for exactly those, without
editing a reference file; hosts that have the real packages get the real packages.

Where the reference is looked for, in order:
  1. ``$MDISTILLER_REF_ROOT``
  2. ``/root/reference``                       (the build container's read-only mount)
  3. ``oracle/_ref/mdistiller_ref.tar.gz``     (git-ignored archive written by ``oracle/make_ref.sh``;
     it travels to the GPU box with the built library and is unpacked into a temporary directory)
Nothing here is on the kernels' path; without a reference tree the package works the same,
only ``install()`` and the reference-hosted tests/benchmarks have nothing to attach to.
"""

from __future__ import annotations

import ast
import atexit
import copy
import importlib
import os
import shutil
import sys
import tarfile
import tempfile
import types
from typing import Optional

_REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ARCHIVE = os.environ.get("MDB200_REF_ARCHIVE") or os.path.join(_REPO_ROOT, "oracle", "_ref", "mdistiller_ref.tar.gz")
_MARKER = os.path.join("mdistiller", "distillers", "DKD.py")
_root: Optional[str] = None


def locate() -> Optional[str]:
    """Root directory of a reference tree (contains ``mdistiller/``, ``configs/``, ``tools/train.py``) or None."""
    global _root
    if _root is not None:
        return _root
    mounted = () if os.environ.get("MDISTILLER_REF_ARCHIVE_ONLY") == "1" else \
        (os.environ.get("MDISTILLER_REF_ROOT"), "/root/reference")
    for cand in mounted:
        if cand and os.path.isfile(os.path.join(cand, _MARKER)):
            _root = cand
            return _root
    if os.path.isfile(ARCHIVE):
        tmp = tempfile.mkdtemp(prefix="mdistiller_ref_")
        atexit.register(shutil.rmtree, tmp, ignore_errors=True)
        with tarfile.open(ARCHIVE) as tar:
            tar.extractall(tmp, filter="data")
        if os.path.isfile(os.path.join(tmp, _MARKER)):
            _root = tmp
            return _root
    return None


def available() -> bool:
    return locate() is not None


# ------------------------------------------------------------------ stand-ins for absent third-party modules

class CfgNode(dict):
    """The slice of ``yacs.config.CfgNode`` the reference touches (engine/cfg.py, tools/train.py)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def dump(self, **kw):
        import yaml

        def plain(x):
            return {k: plain(v) for k, v in x.items()} if isinstance(x, dict) else x
        return yaml.safe_dump(plain(self), **kw)

    def clone(self):
        return copy.deepcopy(self)

    def freeze(self):
        pass

    def defrost(self):
        pass

    def merge_from_list(self, opts):
        """KEY VALUE pairs from the command line; values are Python literals where they parse (yacs rule)."""
        for key, val in zip(opts[::2], opts[1::2]):
            node = self
            *path, leaf = key.split(".")
            for part in path:
                node = node[part]
            if isinstance(val, str):
                try:
                    val = ast.literal_eval(val)
                except (ValueError, SyntaxError):
                    pass
            old = node.get(leaf)
            if isinstance(old, float) and isinstance(val, int) and not isinstance(val, bool):
                val = float(val)
            node[leaf] = val

    def merge_from_file(self, path):
        import yaml

        def merge(dst, src):
            for k, v in src.items():
                if isinstance(v, dict):
                    merge(dst.setdefault(k, CfgNode()), v)
                else:
                    dst[k] = v
        with open(path) as f:
            merge(self, yaml.safe_load(f))


class SummaryWriter:
    """``tensorboardX.SummaryWriter`` stand-in: keeps the scalars in memory."""

    def __init__(self, *a, **k):
        self.scalars = []

    def add_scalar(self, *a):
        self.scalars.append(a)

    def flush(self):
        pass

    def close(self):
        pass


def install_import_shims() -> None:
    """yacs / tensorboardX stand-ins (only if the real modules are absent) and the timm-free ViT stub (D6, D8)."""
    if "yacs" not in sys.modules:
        try:
            import yacs  # noqa: F401
        except ImportError:
            y, yc = types.ModuleType("yacs"), types.ModuleType("yacs.config")
            yc.CfgNode = CfgNode
            y.config = yc
            sys.modules["yacs"], sys.modules["yacs.config"] = y, yc
    if "tensorboardX" not in sys.modules:
        try:
            import tensorboardX  # noqa: F401
        except ImportError:
            t = types.ModuleType("tensorboardX")
            t.SummaryWriter = SummaryWriter
            sys.modules["tensorboardX"] = t
    name = "mdistiller.models.imagenet.vit"
    if name not in sys.modules:
        stub = types.ModuleType(name)
        for n in ("tiny", "small", "base", "large"):
            setattr(stub, f"vit_{n}_patch16_224", None)
        sys.modules[name] = stub


def ensure_importable() -> str:
    """Put the reference on ``sys.path`` with the import shims in place; returns its root."""
    root = locate()
    if root is None:
        raise FileNotFoundError(
            "no reference tree: set MDISTILLER_REF_ROOT, mount /root/reference, or run oracle/make_ref.sh "
            f"(writes {ARCHIVE})")
    if root not in sys.path:
        sys.path.insert(0, root)
    install_import_shims()
    return root


def cifar_models():
    """The reference's ``cifar_model_dict`` with ``get_arch`` no longer abstract (D5)."""
    ensure_importable()
    for sub, cls in (("resnet", "ResNet"), ("wrn", "WideResNet"), ("ShuffleNetv1", "ShuffleNet"),
                     ("mobilenetv2", "MobileNetV2")):
        m = importlib.import_module(f"mdistiller.models.cifar.{sub}")
        getattr(m, cls).__abstractmethods__ = frozenset()
    from mdistiller.models.cifar import cifar_model_dict  # type: ignore
    return cifar_model_dict


def imagenet_models():
    ensure_importable()
    from mdistiller.models.imagenet import imagenet_model_dict  # type: ignore
    return imagenet_model_dict


def backbone(name: str, num_classes: int, dataset: str = "cifar100"):
    """Random-init reference backbone by its registry name (no checkpoint, no download):
    ``imagenet_model_dict`` (models/imagenet/__init__.py) or ``cifar_model_dict`` (models/cifar/__init__.py:32)."""
    if dataset == "imagenet":
        return imagenet_models()[name](pretrained=False, num_classes=num_classes)
    return cifar_models()[name][0](num_classes=num_classes)
