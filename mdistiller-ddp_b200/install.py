"""Graft the B200 tail into an importable copy of the reference, leaving its files unchanged.

    import mdistiller_ddp_b200 as b200
    b200.install()                       # before tools/train.py builds distiller / trainer
    runpy.run_path("tools/train.py", run_name="__main__")

What gets replaced (and only this):
  mdistiller.distillers.distiller_dict["DKD"|"KD"]       -> fused-tail distillers
  mdistiller.distillers.DKD.dkd_loss / KD.kd_loss          -> kernel-backed functions (same signatures)
  mdistiller.engine.dot.DistillationOrientedTrainer        -> flat / sharded optimizer
  mdistiller.engine.trainer_dict["dot"|"base"]             -> subclasses of the reference's BaseTrainer
                                                             with init_optimizer / train_iter from here
"""

from __future__ import annotations

import importlib
import sys

_saved = {}


def install(replace_base: bool = True):
    """Returns the dict of patched objects; raises ImportError if `mdistiller` is not importable."""
    from . import distillers, losses, optim
    from .trainer import TailStepMixin

    ref_distillers = importlib.import_module("mdistiller.distillers")
    ref_engine = importlib.import_module("mdistiller.engine")
    ref_trainer = importlib.import_module("mdistiller.engine.trainer")
    ref_dot = importlib.import_module("mdistiller.engine.dot")

    if not _saved:
        _saved.update(
            DKD=ref_distillers.distiller_dict["DKD"], KD=ref_distillers.distiller_dict["KD"],
            dkd_loss=sys.modules["mdistiller.distillers.DKD"].dkd_loss,
            kd_loss=sys.modules["mdistiller.distillers.KD"].kd_loss,
            opt_dot=ref_dot.DistillationOrientedTrainer, opt_trainer=ref_trainer.DistillationOrientedTrainer,
            dot=ref_engine.trainer_dict["dot"], base=ref_engine.trainer_dict["base"])
    ref_distillers.distiller_dict["DKD"] = distillers.DKD
    ref_distillers.distiller_dict["KD"] = distillers.KD
    # `mdistiller.distillers.DKD` is the class (the package __init__ shadows the module name),
    # so reach the modules through sys.modules to patch the free functions.
    sys.modules["mdistiller.distillers.DKD"].dkd_loss = losses.dkd_loss
    sys.modules["mdistiller.distillers.KD"].kd_loss = losses.kd_loss
    ref_dot.DistillationOrientedTrainer = optim.DistillationOrientedTrainer
    ref_trainer.DistillationOrientedTrainer = optim.DistillationOrientedTrainer

    base = ref_trainer.BaseTrainer

    class _RefHosted(TailStepMixin, base):
        def __init__(self, experiment_name, distiller, train_loader, val_loader, cfg):
            self.grad_clip = cfg.SOLVER.GRAD_CLIP       # needed by init_optimizer's callers early (D1)
            self.use_amp = cfg.EXPERIMENT.AMP
            base.__init__(self, experiment_name, distiller, train_loader, val_loader, cfg)
            # lets train_iter know when an epoch ends: lagged metrics are settled and the sharded optimizer
            # state is snapshotted on every rank, so train_epoch (trainer.py:188-255 of the reference, unchanged)
            # can read exact averages and call optimizer.state_dict() on the master rank alone
            self.iters_per_epoch = len(train_loader)

    class DOT(_RefHosted):
        trainer_kind = "dot"

    class Base(_RefHosted):
        trainer_kind = "base"

    ref_engine.trainer_dict["dot"] = DOT
    patched = {"DKD": distillers.DKD, "KD": distillers.KD, "dot": DOT}
    if replace_base:
        ref_engine.trainer_dict["base"] = Base
        patched["base"] = Base
    return patched


def uninstall() -> None:
    """Put the reference's own objects back (used by tests that share a process with the oracle)."""
    if not _saved:
        return
    ref_distillers = importlib.import_module("mdistiller.distillers")
    ref_engine = importlib.import_module("mdistiller.engine")
    ref_trainer = importlib.import_module("mdistiller.engine.trainer")
    ref_dot = importlib.import_module("mdistiller.engine.dot")
    ref_distillers.distiller_dict["DKD"], ref_distillers.distiller_dict["KD"] = _saved["DKD"], _saved["KD"]
    sys.modules["mdistiller.distillers.DKD"].dkd_loss = _saved["dkd_loss"]
    sys.modules["mdistiller.distillers.KD"].kd_loss = _saved["kd_loss"]
    ref_dot.DistillationOrientedTrainer = _saved["opt_dot"]
    ref_trainer.DistillationOrientedTrainer = _saved["opt_trainer"]
    ref_engine.trainer_dict["dot"], ref_engine.trainer_dict["base"] = _saved["dot"], _saved["base"]
    _saved.clear()
