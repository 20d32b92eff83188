"""Stand-ins for the reference's missing third-party imports, test use only:
`yacs.config.CfgNode`, `tensorboardX.SummaryWriter` and the timm-dependent ViT module."""

import sys
import types

import yaml


class CfgNode(dict):
    """The slice of yacs' CfgNode the reference touches (engine/cfg.py, tools/train.py)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def dump(self, **kw):
        def plain(x):
            return {k: plain(v) for k, v in x.items()} if isinstance(x, dict) else x
        return yaml.safe_dump(plain(self), **kw)

    def clone(self):
        import copy
        return copy.deepcopy(self)

    def freeze(self):
        pass

    def merge_from_list(self, opts):
        for key, val in zip(opts[::2], opts[1::2]):
            node = self
            *path, leaf = key.split(".")
            for part in path:
                node = node[part]
            node[leaf] = type(node[leaf])(val) if leaf in node and not isinstance(node[leaf], (list, dict)) else val

    def merge_from_file(self, path):
        def merge(dst, src):
            for k, v in src.items():
                if isinstance(v, dict):
                    merge(dst.setdefault(k, CfgNode()), v)
                else:
                    dst[k] = v
        with open(path) as f:
            merge(self, yaml.safe_load(f))


class SummaryWriter:
    def __init__(self, *a, **k):
        self.scalars = []

    def add_scalar(self, *a):
        self.scalars.append(a)

    def flush(self):
        pass


def install_shims():
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
