"""`install()` against the real, unmodified reference (build container only; CPU).

Proves the drop-in boundary of SURVEY.md section 8b: the registries the unchanged
`tools/train.py` reads are patched, the reference's own `BaseTrainer.__init__` constructs our
optimizer through `init_optimizer` (reference defects D1/D2 handled), and the reference's
`adjust_learning_rate` drives it through `param_groups`.  Kernels are not called (no GPU here).
"""

import os
import sys

import pytest
import torch

from oracle import ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not mounted (GPU box)")


@pytest.fixture()
def patched(tmp_path, monkeypatch):
    import _ref_shims
    _ref_shims.install_shims()
    ref_loader._ensure_path()
    ref_loader.cifar_models()
    import mdistiller_ddp_b200 as b200
    patched = b200.install()
    monkeypatch.chdir(ref_loader.REF_ROOT)          # BaseTrainer copies ./mdistiller/distillers/<type>.py (D8)
    monkeypatch.setenv("IS_MASTER_NODE", "1")
    yield b200, patched, tmp_path
    b200.uninstall()                                 # the oracle tests in this process use the reference's own functions


def test_registries_are_patched_and_reference_files_untouched(patched):
    b200, p, _ = patched
    import mdistiller.distillers as D
    import mdistiller.engine as E
    from mdistiller.engine import trainer as T
    assert D.distiller_dict["DKD"] is b200.DKD and D.distiller_dict["KD"] is b200.KD
    assert sys.modules["mdistiller.distillers.DKD"].dkd_loss is b200.dkd_loss
    assert sys.modules["mdistiller.distillers.KD"].kd_loss is b200.kd_loss
    assert T.DistillationOrientedTrainer is b200.DistillationOrientedTrainer
    for key in ("dot", "base"):
        cls = E.trainer_dict[key]
        assert issubclass(cls, T.BaseTrainer) and issubclass(cls, b200.TailStepMixin)
        assert cls.train_iter is b200.TailStepMixin.train_iter and cls.train_epoch is T.BaseTrainer.train_epoch
    assert set(D.distiller_dict) >= {"NONE", "AT", "CRD", "REVIEWKD"}       # everything else left alone


def test_reference_base_trainer_constructs_our_dot_optimizer(patched):
    b200, p, tmp = patched
    import mdistiller.engine as E
    from mdistiller.engine.cfg import CFG
    from mdistiller.engine.utils import adjust_learning_rate
    from mdistiller_ddp_b200 import workloads as W
    cfg = CFG.clone()
    cfg.DISTILLER.TYPE, cfg.SOLVER.TRAINER = "DKD", "dot"
    cfg.LOG.PREFIX = str(tmp)
    cfg.SOLVER.SCHEDULE.MULTISTEP.STAGES = [2, 4]
    import mdistiller.distillers as D
    distiller = D.distiller_dict["DKD"](W.TinyNet(100, 16), W.TinyNet(100, 24), cfg)
    trainer = E.trainer_dict["dot"]("exp", distiller, [], [], cfg)          # D1 would raise AttributeError here
    opt = trainer.optimizer
    assert isinstance(opt, b200.DistillationOrientedTrainer)
    g = opt.param_groups[0]
    assert abs(g["momentum"] - (0.9 - 0.075)) < 1e-12 and abs(g["momentum_kd"] - (0.9 + 0.075)) < 1e-12   # D2
    assert g["lr"] == cfg.SOLVER.LR and g["weight_decay"] == cfg.SOLVER.WEIGHT_DECAY
    assert os.path.isfile(os.path.join(str(tmp), "exp", "code", "distiller.py"))
    # the reference's scheduler mutates param_groups[*]["lr"] (engine/utils.py:83-97)
    assert adjust_learning_rate(3, 0, cfg, opt) == pytest.approx(cfg.SOLVER.LR * 0.1)
    assert opt.param_groups[0]["lr"] == pytest.approx(cfg.SOLVER.LR * 0.1)
    # student parameters now live in one flat buffer, the teacher's do not
    flat = opt.flat_params
    lo, hi = flat.data_ptr(), flat.data_ptr() + flat.numel() * 4
    assert all(lo <= q.data_ptr() < hi for q in distiller.student.parameters())
    assert not any(lo <= q.data_ptr() < hi for q in distiller.teacher.parameters())
    # state_dict is what BaseTrainer.train_epoch saves into the checkpoint (trainer.py:231-236)
    sd = opt.state_dict()
    assert sd["param_groups"][0]["momentum_kd"] == g["momentum_kd"] and sd["state"] == {}


def test_reference_base_trainer_gets_flat_sgd(patched):
    b200, p, tmp = patched
    import mdistiller.engine as E
    import mdistiller.distillers as D
    from mdistiller.engine.cfg import CFG
    from mdistiller_ddp_b200 import workloads as W
    cfg = CFG.clone()
    cfg.DISTILLER.TYPE, cfg.SOLVER.TRAINER = "KD", "base"
    cfg.LOG.PREFIX = str(tmp)
    distiller = D.distiller_dict["KD"](W.TinyNet(100, 16), W.TinyNet(100, 24), cfg)
    trainer = E.trainer_dict["base"]("exp2", distiller, [], [], cfg)
    assert isinstance(trainer.optimizer, b200.FlatSGD) and distiller.combined_tail
    assert trainer.optimizer.param_groups[0]["momentum"] == cfg.SOLVER.SGD.MOMENTUM
    cfg.SOLVER.TYPE, cfg.SOLVER.GRAD_CLIP = "AdamW", 1.0          # configs/imagenet/optim/adamw.yaml
    tr = E.trainer_dict["base"]("exp3", distiller, [], [], cfg)
    assert isinstance(tr.optimizer, b200.FlatAdam) and tr.optimizer.decoupled and tr.optimizer.max_grad_norm == 1.0
    assert tr.optimizer.param_groups[0]["betas"] == tuple(cfg.SOLVER.ADAM.BETAS)
    cfg.EXPERIMENT.AMP, cfg.SOLVER.TRAINER = True, "dot"          # DOT refuses AMP, as the reference (trainer.py:377-378)
    with pytest.raises(NotImplementedError):
        E.trainer_dict["dot"]("exp4", distiller, [], [], cfg)


def test_cfgnode_stand_in_parses_command_line_options_like_yacs():
    """tools/train.py feeds KEY VALUE strings through cfg.merge_from_list: literals must come out typed
    ("False" is not a true string), floats stay floats, paths stay strings."""
    from mdistiller_ddp_b200.refhost import CfgNode
    cfg = CfgNode(LOG=CfgNode(WANDB=True, PREFIX="./output"), SOLVER=CfgNode(LR=0.05, EPOCHS=240, BATCH_SIZE=64,
                                                                            SCHEDULE=CfgNode(STAGES=[150, 180])))
    cfg.merge_from_list(["LOG.WANDB", "False", "LOG.PREFIX", "/tmp/x,y", "SOLVER.LR", "1", "SOLVER.EPOCHS", "2",
                         "SOLVER.SCHEDULE.STAGES", "[1, 2]"])
    assert cfg.LOG.WANDB is False and cfg.LOG.PREFIX == "/tmp/x,y"
    assert cfg.SOLVER.LR == 1.0 and isinstance(cfg.SOLVER.LR, float) and cfg.SOLVER.EPOCHS == 2
    assert cfg.SOLVER.SCHEDULE.STAGES == [1, 2] and cfg.clone() is not cfg and cfg.clone() == cfg
