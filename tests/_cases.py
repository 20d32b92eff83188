"""Seeded input recipes shared by the golden generator and the parity tests.

Everything is produced by the torch *CPU* generator so that the same tensors
come out in the build container (where the fixtures are made from the reference)
and on the GPU box (where they are replayed through the CUDA path).  Each
fixture stores a checksum of its inputs; a mismatch means the RNG stream
changed and the fixture must be regenerated, not that the kernels are wrong.
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, FrozenSet, List, Optional, Tuple

import numpy as np
import torch

FULL_TENSOR_LIMIT = 8192     # fixtures keep whole tensors up to this many elements
SAMPLE_POINTS = 512          # ... and this many sampled elements beyond it


@dataclass(frozen=True)
class LossCase:
    name: str
    seed: int
    B: int
    C: int
    sigma: float = 3.0
    alpha: float = 1.0
    beta: float = 8.0
    T: float = 4.0
    boost: float = 0.0        # added to the label logit of both nets ("confident" rows)


# First four rows are SURVEY.md Appendix C.1's known-answer vectors.
LOSS_CASES: List[LossCase] = [
    LossCase("c1_cifar_64x100", 0, 64, 100, 3.0, 1.0, 8.0, 4.0),
    LossCase("c1_cifar_128x100", 1, 128, 100, 3.0, 1.0, 6.0, 4.0),
    LossCase("c1_imnet_256x1000", 2, 256, 1000, 3.0, 1.0, 0.5, 1.0),
    LossCase("c1_imnet_1024x1000", 3, 1024, 1000, 3.0, 1.0, 0.5, 1.0),
    # edge shapes: single row, two classes, C not a multiple of 4, C just past a vector
    LossCase("edge_1x2", 10, 1, 2, 1.0, 1.0, 8.0, 4.0),
    LossCase("edge_5x3", 11, 5, 3, 2.0, 1.0, 8.0, 4.0),
    LossCase("edge_33x37", 12, 33, 37, 3.0, 1.0, 6.0, 4.0),
    LossCase("edge_7x101", 13, 7, 101, 3.0, 1.0, 8.0, 2.0),
    LossCase("edge_3x1001", 14, 3, 1001, 3.0, 1.0, 0.5, 1.0),
    LossCase("edge_2x4099", 15, 2, 4099, 3.0, 2.0, 3.0, 1.5),
    # confident rows (label logit boosted in both nets), wide logits
    LossCase("conf15_64x100", 20, 64, 100, 3.0, 1.0, 8.0, 4.0, 15.0),
    LossCase("conf8_32x1000_T1", 21, 32, 1000, 3.0, 1.0, 0.5, 1.0, 8.0),
    LossCase("wide_48x100_s10", 22, 48, 100, 10.0, 1.0, 8.0, 4.0),
    LossCase("big_4096x100", 23, 4096, 100, 3.0, 1.0, 8.0, 4.0),
    # 113..128 classes: the 8-lanes-per-row register layout (launch_shape's <8,4> branch)
    LossCase("mid_96x116", 24, 96, 116, 3.0, 1.0, 8.0, 4.0),
    LossCase("mid_64x120_T1", 25, 64, 120, 3.0, 1.0, 0.5, 1.0),
    LossCase("mid_200x128", 26, 200, 128, 3.0, 1.0, 6.0, 2.0),
]


def make_logits(case: LossCase) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    g = torch.Generator(device="cpu")
    g.manual_seed(case.seed)
    zs = torch.randn(case.B, case.C, generator=g) * case.sigma
    zt = torch.randn(case.B, case.C, generator=g) * case.sigma
    y = torch.randint(0, case.C, (case.B,), generator=g)
    if case.boost:
        r = torch.arange(case.B)
        zs[r, y] += case.boost
        zt[r, y] += case.boost
    return zs, zt, y


def input_checksum(zs, zt, y) -> np.ndarray:
    return np.array([zs.double().sum().item(), zt.double().sum().item(), float(y.sum().item())])


def sample_index(n: int) -> np.ndarray:
    rs = np.random.RandomState(12345)
    return np.sort(rs.choice(n, size=min(SAMPLE_POINTS, n), replace=False))


# --------------------------------------------------------------------------

def _cifar_like_shapes() -> List[Tuple[int, ...]]:
    """A res8x4-flavoured list (conv/bn/fc shapes, 1.2 M elements, sizes multiple of 4)."""
    shapes: List[Tuple[int, ...]] = [(32, 3, 3, 3), (32,), (32,)]
    cin = 32
    for cout in (64, 128, 256):
        shapes += [(cout, cin, 3, 3), (cout,), (cout,), (cout, cout, 3, 3), (cout,), (cout,),
                   (cout, cin, 1, 1), (cout,), (cout,)]
        cin = cout
    shapes += [(100, 256), (100,)]
    return shapes


SMALL_SHAPES: List[Tuple[int, ...]] = [(16, 3, 3, 3), (16,), (16,), (24, 16, 3, 3), (24,), (10, 24), (10,), (7,)]


@dataclass(frozen=True)
class DotCase:
    name: str
    seed: int
    kind: str = "dot"                       # "dot" | "sgd" | "adam" | "adamw"
    clip: float = 0.0                       # SOLVER.GRAD_CLIP (global-norm clipping before the step)
    shapes: Tuple[Tuple[int, ...], ...] = tuple(SMALL_SHAPES)
    steps: int = 100
    lr: float = 0.05
    momentum: float = 0.9
    delta: float = 0.075
    wd: float = 5e-4
    dampening: float = 0.0
    lr_drop_at: int = 40
    kd_missing: FrozenSet[int] = frozenset()     # params that never get a KD gradient
    task_missing: FrozenSet[int] = frozenset()   # params that never get a task gradient
    kd_from_step: Tuple[Tuple[int, int], ...] = ()   # (param index, first step with a KD grad)


DOT_CASES: List[DotCase] = [
    DotCase("dot_cifar_small", 100),
    DotCase("dot_imnet_small", 101, lr=0.2, delta=0.09, wd=1e-4),
    DotCase("dot_membership", 102, kd_missing=frozenset({1}), task_missing=frozenset({4}),
            kd_from_step=((6, 3),)),
    DotCase("dot_dampening", 103, dampening=0.1),
    DotCase("dot_no_task_momentum", 104, momentum=0.075, delta=0.075),   # m_task == 0 bypass
    DotCase("dot_no_decay", 105, wd=0.0),
    DotCase("dot_cifar_medium", 106, shapes=tuple(_cifar_like_shapes())),
    DotCase("sgd_cifar_small", 110, kind="sgd"),
    DotCase("sgd_medium", 111, kind="sgd", shapes=tuple(_cifar_like_shapes()), wd=1e-4, lr=0.2),
    # optimizer-path variants of the base trainer (trainer.py:92-107, 316-320)
    DotCase("sgd_clip_small", 112, kind="sgd", clip=0.5),
    DotCase("adam_small", 113, kind="adam", lr=1e-3, wd=1e-4),
    DotCase("adamw_small", 114, kind="adamw", lr=1e-3, wd=5e-2),
    DotCase("adamw_clip_partial", 115, kind="adamw", lr=1e-3, wd=5e-2, clip=0.5, task_missing=frozenset({4})),
    DotCase("adam_medium", 116, kind="adam", shapes=tuple(_cifar_like_shapes()), lr=1e-3, wd=0.0, steps=50),
]

ADAM_BETAS, ADAM_EPS = (0.9, 0.999), 1e-8


def make_params(case: DotCase) -> List[torch.Tensor]:
    g = torch.Generator(device="cpu")
    g.manual_seed(case.seed)
    return [torch.randn(*s, generator=g) * 0.1 for s in case.shapes]


def make_grads(case: DotCase, step: int, shapes) -> Tuple[List[Optional[torch.Tensor]], List[Optional[torch.Tensor]]]:
    """(kd grads, task grads) for one step; ``None`` where the parameter has no gradient."""
    g = torch.Generator(device="cpu")
    g.manual_seed(case.seed * 100003 + step)
    kd = [torch.randn(*s, generator=g) * 0.1 for s in shapes]
    task = [torch.randn(*s, generator=g) * 0.1 for s in shapes]
    late: Dict[int, int] = dict(case.kd_from_step)
    kd_out: List[Optional[torch.Tensor]] = []
    for i, t in enumerate(kd):
        drop = i in case.kd_missing or step < late.get(i, 0)
        kd_out.append(None if drop else t)
    task_out = [None if i in case.task_missing else t for i, t in enumerate(task)]
    return kd_out, task_out


def dot_input_checksum(case: DotCase) -> np.ndarray:
    p = make_params(case)
    kd, task = make_grads(case, case.steps - 1, [t.shape for t in p])
    s = lambda ts: float(sum(t.double().sum().item() for t in ts if t is not None))
    return np.array([s(p), s(kd), s(task)])
