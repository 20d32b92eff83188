"""Run the reference's UNCHANGED entry script, ``tools/train.py``, on synthetic data.

TEST INFRASTRUCTURE (used by ``tests/`` and by ``bench.py``'s comparator legs only).  One call =
one ``runpy.run_path("tools/train.py", run_name="__main__")`` in this process, from the reference
root as working directory (D8), with

  * ``get_dataset`` replaced by in-memory synthetic loaders of the configured shape (there are no
    datasets on the box): ``train`` yields ``(image, target, index)``, ``val`` yields ``(image, target)``;
  * teacher checkpoints replaced by a seeded random initialisation (``load_checkpoint`` returns the
    state of a freshly constructed teacher; there are no checkpoints on the box);
  * ``patched=True``: ``mdistiller_ddp_b200.install()`` first -- the B200 tail behind the reference's
    registries;  ``patched=False``: the reference's own losses, optimizer and trainer, with only the two
    constructor defects repaired that stop its DOT trainer from being built at all (D1: ``use_amp`` read
    before assignment, D2: ``cfg.SOLVER.MOMENTUM`` does not exist).

Both variants see identical seeds, initial weights and batches, so their results can be compared.
Returns what the run left on disk (the checkpoints the reference's epoch loop writes).
"""

from __future__ import annotations

import copy
import os
import runpy
import sys
from typing import Dict, List, Optional, Sequence

import torch

import mdistiller_ddp_b200 as b200
from mdistiller_ddp_b200 import refhost


class SyntheticLoader:
    """A list of pre-generated CPU batches with the two things the reference's loops need: len() and iteration."""

    def __init__(self, batches: List[tuple]):
        self.batches = batches

    def __len__(self):
        return len(self.batches)

    def __iter__(self):
        return iter(self.batches)


def synthetic_loaders(batch: int, hw: int, classes: int, iters: int, val_iters: int, seed: int):
    g = torch.Generator().manual_seed(seed)
    train = [(torch.randn(batch, 3, hw, hw, generator=g), torch.randint(0, classes, (batch,), generator=g),
              torch.arange(batch)) for _ in range(iters)]
    val = [(torch.randn(batch, 3, hw, hw, generator=g), torch.randint(0, classes, (batch,), generator=g))
           for _ in range(val_iters)]
    return SyntheticLoader(train), SyntheticLoader(val)


def _repair_reference_dot_constructor(ref_trainer):
    """D1 + D2 only: let the reference's own DOT trainer be constructed; its train_iter, dot() and losses stay."""
    DOT = ref_trainer.DOT
    if getattr(DOT, "_mdb200_repaired", False):
        return

    def init_optimizer(self, cfg):
        if cfg.EXPERIMENT.AMP:
            raise NotImplementedError("AMP is not supported for DOT.")
        m, d = cfg.SOLVER.SGD.MOMENTUM, cfg.SOLVER.DOT.DELTA
        return ref_trainer.DistillationOrientedTrainer(
            self.distiller.module.get_learnable_parameters(), lr=cfg.SOLVER.LR, momentum=m - d, momentum_kd=m + d,
            weight_decay=cfg.SOLVER.WEIGHT_DECAY)

    DOT.init_optimizer = init_optimizer
    DOT._mdb200_repaired = True


def run_train(config: str, *, patched: bool, out_dir: str, iters: int = 4, val_iters: int = 1, epochs: int = 1,
              batch: Optional[int] = None, data_seed: int = 1234, init_seed: int = 0,
              opts: Sequence[str] = (), allow_tf32: bool = False) -> Dict:
    """Run ``tools/train.py --cfg <config> ...`` (config relative to the reference root).  Needs CUDA and the
    env of a torchrun worker (LOCAL_RANK, WORLD_SIZE, MASTER_ADDR, MASTER_PORT; defaults = one process)."""
    root = refhost.ensure_importable()
    refhost.cifar_models()
    os.environ.setdefault("LOCAL_RANK", "0")
    os.environ.setdefault("WORLD_SIZE", "1")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29431")
    rank = int(os.environ["LOCAL_RANK"])

    import mdistiller.dataset as ref_dataset
    import mdistiller.engine.cfg as ref_cfg
    import mdistiller.engine.utils as ref_utils
    import mdistiller.engine.trainer as ref_trainer

    b200.uninstall()
    if patched:
        b200.install()
    else:
        _repair_reference_dot_constructor(ref_trainer)

    saved_cfg = copy.deepcopy(dict(ref_cfg.CFG))
    saved = (ref_dataset.get_dataset, ref_utils.load_checkpoint, os.getcwd(), list(sys.argv),
             torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)

    def get_dataset(cfg):
        hw, classes = (224, 1000) if cfg.DATASET.TYPE == "imagenet" else (32, 100)
        tr, va = synthetic_loaders(cfg.SOLVER.BATCH_SIZE, hw, classes, iters, val_iters, data_seed + 1000 * rank)
        return tr, va, iters * cfg.SOLVER.BATCH_SIZE, classes

    def load_checkpoint(path):
        if os.path.exists(path):
            return saved[1](path)
        name = os.path.basename(os.path.dirname(path)).replace("_vanilla", "")     # .../<teacher>_vanilla/ckpt_epoch_240.pth
        gen_state = torch.random.get_rng_state()
        torch.manual_seed(init_seed + 7)
        model = refhost.cifar_models()[name][0](num_classes=100)
        torch.random.set_rng_state(gen_state)
        return {"model": model.state_dict()}

    argv = ["tools/train.py", "--cfg=" + config,          # "=": --cfg is nargs="*" and would swallow the options
            "SOLVER.EPOCHS", str(epochs), "LOG.PREFIX", out_dir, "LOG.WANDB", "False"]
    if batch is not None:
        argv += ["SOLVER.BATCH_SIZE", str(batch * int(os.environ["WORLD_SIZE"]))]      # train.py divides by the world size
    argv += list(opts)
    launches0 = b200.ops.launch_count
    try:
        ref_dataset.get_dataset, ref_utils.load_checkpoint = get_dataset, load_checkpoint
        os.chdir(root)
        sys.argv = argv
        torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = allow_tf32
        torch.manual_seed(init_seed)
        torch.cuda.manual_seed_all(init_seed)
        runpy.run_path(os.path.join(root, "tools", "train.py"), run_name="__main__")
        torch.cuda.synchronize()
    finally:
        ref_dataset.get_dataset, ref_utils.load_checkpoint = saved[0], saved[1]
        os.chdir(saved[2])
        sys.argv = saved[3]
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved[4], saved[5]
        ref_cfg.CFG.clear()
        ref_cfg.CFG.update(copy.deepcopy(saved_cfg))
        if patched:
            b200.uninstall()
    result = {"launches": b200.ops.launch_count - launches0, "log_dirs": []}
    for dirpath, _, files in os.walk(out_dir):
        if "latest" in files:
            result["log_dirs"].append(dirpath)
            result["latest"] = torch.load(os.path.join(dirpath, "latest"), map_location="cpu", weights_only=False)
            result["student"] = torch.load(os.path.join(dirpath, "student_latest"), map_location="cpu",
                                           weights_only=False)["model"]
            with open(os.path.join(dirpath, "worklog.txt")) as f:
                result["worklog"] = f.read()
    return result
