"""B200-native distillation tail for mdistiller-ddp.

Fused DKD/KD(+CE) logit loss+gradient kernel, DOT dual-momentum / SGD flat update kernel
and the reduce-scatter -> update -> all-gather data-parallel step, behind the reference's
``Distiller.forward_train`` and ``DistillationOrientedTrainer`` interfaces.  Kernels are
hand-written CUDA for sm_100a reached through a C-ABI library
(``include/mdistiller_b200.h``); there is no CPU or PyTorch fallback.
"""

from .build import LIB_PATH, build as build_library  # noqa: F401
from . import ops  # noqa: F401
from .losses import dkd_loss, kd_loss, distill_tail, TailResult  # noqa: F401
from .distillers import Distiller, DKD, KD  # noqa: F401
from .optim import DistillationOrientedTrainer, FlatAdam, FlatSGD  # noqa: F401
from .amp import DeviceLossScaler  # noqa: F401
from .comm import ShardComm  # noqa: F401
from .flat import FlatSpace  # noqa: F401
from .trainer import TailStepMixin, StandaloneTrainer, new_train_meters, silence_ddp_reducer  # noqa: F401
from .install import install, uninstall  # noqa: F401

distiller_dict = {"DKD": DKD, "KD": KD}

__all__ = [
    "dkd_loss", "kd_loss", "distill_tail", "TailResult", "Distiller", "DKD", "KD",
    "DistillationOrientedTrainer", "FlatSGD", "FlatAdam", "ShardComm", "FlatSpace", "TailStepMixin",
    "StandaloneTrainer", "new_train_meters", "silence_ddp_reducer", "install", "uninstall", "distiller_dict",
    "build_library", "LIB_PATH", "ops", "DeviceLossScaler",
]
