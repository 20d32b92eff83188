"""Synthetic workloads for smoke / bench / tests: backbones with the reference's call
contract and random data of the benchmark shapes.

The distillation tail treats backbones as opaque ``nn.Module``s returning
``(logits, feature_dict)``; in a real run they are the reference's own modules
(out of scope here).  The GPU box has no copy of the reference, so
the benchmark instantiates architecture-equivalent stand-ins: torchvision's ResNet-18/
34/50 for the ImageNet configs (same parameter counts as the reference's
models/imagenet/resnet.py: 11 689 512 / 21 797 672 / 25 557 032) and small CIFAR-style
nets written here for the ``resnet8x4`` / ``resnet32x4`` (1 233 540 / 7 433 860 parameters),
``wrn_16_2`` / ``wrn_40_2`` (703 284 / 2 255 156) and ``vgg8`` / ``vgg13`` (3 965 028 / 9 462 180)
configs -- the reference's parameter counts.  Weights are random-init.
"""

from __future__ import annotations

from types import SimpleNamespace
from typing import Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


class _WithFeats(nn.Module):
    """Adapt a plain classifier to the ``(logits, feats)`` return convention."""

    def __init__(self, net: nn.Module):
        super().__init__()
        self.net = net

    def forward(self, x):
        return self.net(x), {}


def imagenet_resnet(name: str, num_classes: int = 1000) -> nn.Module:
    import torchvision.models as tvm

    ctor = {"ResNet18": tvm.resnet18, "ResNet34": tvm.resnet34, "ResNet50": tvm.resnet50}[name]
    return _WithFeats(ctor(weights=None, num_classes=num_classes))


class _Unit(nn.Module):
    """Two 3x3 conv-bn pairs with an identity or 1x1-projected skip."""

    def __init__(self, cin: int, cout: int, stride: int):
        super().__init__()
        self.c1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.n1 = nn.BatchNorm2d(cout)
        self.c2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.n2 = nn.BatchNorm2d(cout)
        self.skip = None
        if stride != 1 or cin != cout:
            self.skip = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = F.relu(self.n1(self.c1(x)))
        y = self.n2(self.c2(y))
        return F.relu(y + (x if self.skip is None else self.skip(x)))


class CifarResNet(nn.Module):
    """depth = 6n+2 CIFAR residual net; widths (w0, w1, w2, w3) = stem and three stages."""

    def __init__(self, depth: int, widths: Tuple[int, int, int, int], num_classes: int = 100):
        super().__init__()
        assert (depth - 2) % 6 == 0
        n = (depth - 2) // 6
        self.stem = nn.Sequential(nn.Conv2d(3, widths[0], 3, 1, 1, bias=False), nn.BatchNorm2d(widths[0]))
        units, cin = [], widths[0]
        for stage, cout in enumerate(widths[1:]):
            for u in range(n):
                units.append(_Unit(cin, cout, 2 if (u == 0 and stage > 0) else 1))
                cin = cout
        self.units = nn.Sequential(*units)
        self.head = nn.Linear(cin, num_classes)

    def forward(self, x):
        x = self.units(F.relu(self.stem(x)))
        pooled = torch.flatten(F.adaptive_avg_pool2d(x, 1), 1)
        return self.head(pooled), {"pooled_feat": pooled}


def cifar_resnet(name: str, num_classes: int = 100) -> nn.Module:
    depth = {"resnet8x4": 8, "resnet32x4": 32}[name]
    return CifarResNet(depth, (32, 64, 128, 256), num_classes)


class _PreActUnit(nn.Module):
    """Wide-ResNet unit: bn-relu-conv twice, 1x1 projection on the skip when the width changes."""

    def __init__(self, cin: int, cout: int, stride: int):
        super().__init__()
        self.n1, self.c1 = nn.BatchNorm2d(cin), nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.n2, self.c2 = nn.BatchNorm2d(cout), nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.proj = nn.Conv2d(cin, cout, 1, stride, 0, bias=False) if cin != cout else None

    def forward(self, x):
        a = F.relu(self.n1(x))
        y = self.c2(F.relu(self.n2(self.c1(a))))
        return y + (x if self.proj is None else self.proj(a))


class CifarWRN(nn.Module):
    """WRN-depth-k for 32x32 inputs (depth = 6n+4): wrn_16_2 has 703 284 parameters, wrn_40_2 2 255 156."""

    def __init__(self, depth: int, widen: int, num_classes: int = 100):
        super().__init__()
        assert (depth - 4) % 6 == 0
        n = (depth - 4) // 6
        widths = [16, 16 * widen, 32 * widen, 64 * widen]
        self.stem = nn.Conv2d(3, widths[0], 3, 1, 1, bias=False)
        units, cin = [], widths[0]
        for stage, cout in enumerate(widths[1:]):
            for u in range(n):
                units.append(_PreActUnit(cin, cout, 2 if (u == 0 and stage > 0) else 1))
                cin = cout
        self.units = nn.Sequential(*units)
        self.norm = nn.BatchNorm2d(cin)
        self.head = nn.Linear(cin, num_classes)

    def forward(self, x):
        x = F.relu(self.norm(self.units(self.stem(x))))
        pooled = torch.flatten(F.adaptive_avg_pool2d(x, 1), 1)
        return self.head(pooled), {"pooled_feat": pooled}


class CifarVGG(nn.Module):
    """VGG-bn for 32x32 inputs, one Linear head: vgg8 has 3 965 028 parameters, vgg13 9 462 180."""

    def __init__(self, convs_per_stage, num_classes: int = 100):
        super().__init__()
        layers, cin = [], 3
        for stage, (reps, cout) in enumerate(zip(convs_per_stage, (64, 128, 256, 512, 512))):
            for _ in range(reps):
                layers += [nn.Conv2d(cin, cout, 3, padding=1), nn.BatchNorm2d(cout), nn.ReLU(inplace=True)]
                cin = cout
            if stage < 4:
                layers.append(nn.MaxPool2d(2, 2))
        self.features = nn.Sequential(*layers)
        self.head = nn.Linear(512, num_classes)

    def forward(self, x):
        pooled = torch.flatten(F.adaptive_avg_pool2d(self.features(x), 1), 1)
        return self.head(pooled), {"pooled_feat": pooled}


def cifar_misc(name: str, num_classes: int = 100) -> nn.Module:
    if name.startswith("wrn"):
        depth, widen = {"wrn_16_2": (16, 2), "wrn_40_2": (40, 2)}[name]
        return CifarWRN(depth, widen, num_classes)
    return CifarVGG({"vgg8": (1, 1, 1, 1, 1), "vgg13": (2, 2, 2, 2, 2)}[name], num_classes)


class TinyNet(nn.Module):
    """A few thousand parameters; for smoke() and unit tests."""

    def __init__(self, num_classes: int = 100, width: int = 16):
        super().__init__()
        self.conv = nn.Conv2d(3, width, 3, 2, 1, bias=False)
        self.norm = nn.BatchNorm2d(width)
        self.head = nn.Linear(width, num_classes)

    def forward(self, x):
        f = torch.flatten(F.adaptive_avg_pool2d(F.relu(self.norm(self.conv(x))), 1), 1)
        return self.head(f), {"pooled_feat": f}


WORKLOADS = {
    # name: (teacher, student, builder, image hw, classes, distiller, trainer, per-GPU batch)
    "imagenet_r34_r18_dkd_dot": ("ResNet34", "ResNet18", imagenet_resnet, 224, 1000, "DKD", "dot", 256),
    "imagenet_r34_r18_kd_dot": ("ResNet34", "ResNet18", imagenet_resnet, 224, 1000, "KD", "dot", 256),
    "cifar_res32x4_res8x4_kd_dot": ("resnet32x4", "resnet8x4", cifar_resnet, 32, 100, "KD", "dot", 64),
    "cifar_res32x4_res8x4_dkd_base": ("resnet32x4", "resnet8x4", cifar_resnet, 32, 100, "DKD", "base", 64),
    # BASELINE.json configs[2]: DKD, base trainer, 128 images per GPU
    "cifar_wrn40_2_wrn16_2_dkd_base": ("wrn_40_2", "wrn_16_2", cifar_misc, 32, 100, "DKD", "base", 128),
    "cifar_vgg13_vgg8_dkd_base": ("vgg13", "vgg8", cifar_misc, 32, 100, "DKD", "base", 128),
}


def default_cfg(workload: str = "imagenet_r34_r18_dkd_dot") -> SimpleNamespace:
    """The hyper-parameters of the reference yaml each workload stands for
    (configs/imagenet/r34_r18/{dkd,dot}.yaml, configs/cifar100/{dkd,dot}/res32x4_res8x4.yaml,
    defaults from mdistiller/engine/cfg.py:50-59,82-86,184-194)."""
    imagenet = workload.startswith("imagenet")
    ns = SimpleNamespace
    return ns(
        EXPERIMENT=ns(AMP=False),
        DISTILLER=ns(TYPE=WORKLOADS[workload][5]),
        DKD=ns(CE_WEIGHT=1.0, ALPHA=1.0,
               BETA=0.5 if imagenet else (6.0 if ("wrn" in workload or "vgg" in workload) else 8.0),   # yaml values
               T=1.0 if imagenet else 4.0,
               WARMUP=1 if imagenet else 20),
        KD=ns(TEMPERATURE=1 if imagenet else 4,
              LOSS=ns(CE_WEIGHT=0.5 if imagenet else 0.1, KD_WEIGHT=0.5 if imagenet else 0.9)),
        SOLVER=ns(TRAINER=WORKLOADS[workload][6], TYPE="SGD", LR=0.2 if imagenet else 0.05,
                  WEIGHT_DECAY=1e-4 if imagenet else 5e-4, GRAD_CLIP=0.0,
                  SGD=ns(MOMENTUM=0.9), DOT=ns(DELTA=0.09 if imagenet else 0.075)),
    )


# registry names of the same architectures in the reference (models/cifar/__init__.py:32, models/imagenet/__init__.py)
_REFERENCE_NAMES = {"vgg8": "vgg8", "vgg13": "vgg13"}

backbone_source = "stand-in"      # what the last build_pair() built: "reference" | "stand-in"


def build_pair(workload: str, seed: int = 0, reference: str = "auto"):
    """(teacher, student, image size, classes), random-init.  ``reference="auto"``: the reference's OWN backbone
    modules when a reference tree is available (``refhost.locate()``: mounted, or staged by oracle/make_ref.sh),
    else the architecture-equivalent stand-ins above; "never" / "require" force one or the other."""
    global backbone_source
    teacher_name, student_name, builder, hw, classes = WORKLOADS[workload][:5]
    from . import refhost
    use_ref = reference == "require" or (reference == "auto" and refhost.available())
    torch.manual_seed(seed)
    if use_ref:
        dataset = "imagenet" if workload.startswith("imagenet") else "cifar100"
        teacher = refhost.backbone(_REFERENCE_NAMES.get(teacher_name, teacher_name), classes, dataset)
        student = refhost.backbone(_REFERENCE_NAMES.get(student_name, student_name), classes, dataset)
        backbone_source = "reference"
    else:
        teacher = builder(teacher_name, classes)
        student = builder(student_name, classes)
        backbone_source = "stand-in"
    return teacher, student, hw, classes


def synthetic_batch(batch: int, hw: int, classes: int, seed: int, device="cpu", pin: bool = False):
    """(image, target, index) like the reference's train batches (dataset/cifar100.py:14-19)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    image = torch.randn(batch, 3, hw, hw, generator=g)
    target = torch.randint(0, classes, (batch,), generator=g)
    index = torch.arange(batch)
    if pin:
        image, target, index = image.pin_memory(), target.pin_memory(), index.pin_memory()
    if str(device) != "cpu":
        image, target, index = image.to(device), target.to(device), index.to(device)
    return image, target, index
