"""The drop-in, for real: the reference's UNCHANGED ``tools/train.py`` on a GPU, with and without this package.

``north_star``: "drops into tools/train.py unchanged ... backbones stay as the reference's PyTorch modules".
Each test runs the reference's entry script twice from the staged reference tree (``oracle/make_ref.sh`` ->
``oracle/_ref``, unpacked by ``refhost``; never ``/root/reference`` on the box) through ``oracle/ref_harness``:
once as the reference itself (its own ``dkd_loss``/``kd_loss``, ``DistillationOrientedTrainer`` / ``optim.SGD``,
DDP reducer) and once after ``mdistiller_ddp_b200.install()`` -- same yaml, same seeds, same synthetic batches,
the reference's own ResNet32x4 -> ResNet8x4 backbones, its epoch loop, validation, logging and checkpointing.
What the two runs leave in ``student_latest`` / ``latest`` must agree.

Tolerance.  Two runs of the REFERENCE ITSELF do not agree bit for bit on a GPU: cuDNN's backward kernels sum in
a run-dependent order, and a few tensors amplify that (BatchNorm biases start at zero and their gradients are
small differences of large sums: measured on B200 after 5 iterations, reference vs reference differs by up to
4e-2 there, 2e-3 on the stem convolution, 3e-4 on the median weight tensor; profiles/r2_dropin_noise.txt).  So
every variant is run TWICE and each tensor is held to
    |ours - ref| <= max(1e-4, 8 x the larger of the two run-to-run distances of that tensor)       (rel = / max|ref|)
i.e. the patched run must be indistinguishable from the reference's own repeatability, tensor by tensor; with one
iteration, where that noise is ~1e-6, this is a tight comparison, with five it follows the drift.  The file sorts
near the end of the GPU tests so that a `-x` run has been through the kernel parity tests first.
"""

import os

import pytest
import torch
import torch.multiprocessing as mp

import oracle

pytestmark = pytest.mark.gpu

TOL = 1e-4
NOISE_FACTOR = 8.0
DOT_YAML = "configs/cifar100/dot/res32x4_res8x4.yaml"      # KD + DOT trainer   (BASELINE.json configs[1])
DKD_YAML = "configs/cifar100/dkd/res32x4_res8x4.yaml"      # DKD + base trainer (BASELINE.json configs[0])


def _staged():
    from mdistiller_ddp_b200 import refhost
    if not refhost.available():
        pytest.skip("no reference tree staged (run oracle/make_ref.sh in the build container)")


def _tensors(run, momenta):
    """Everything the run left behind as {name: tensor}: student weights + BatchNorm statistics + optimizer momenta."""
    out = dict(run["student"])
    for i, st in run["latest"]["optimizer"]["state"].items():
        assert set(st) >= set(momenta), (i, set(st))
        for key in momenta:
            out[f"optimizer.state[{i}].{key}"] = st[key]
    return out


def _compare(ours_pair, ref_pair, momenta):
    """ours vs reference, tensor by tensor, against the run-to-run repeatability of both (see the module docstring)."""
    o1, o2, r1, r2 = (_tensors(x, momenta) for x in (*ours_pair, *ref_pair))
    assert set(o1) == set(r1)
    assert ours_pair[0]["latest"]["optimizer"]["param_groups"][0]["lr"] == ref_pair[0]["latest"]["optimizer"]["param_groups"][0]["lr"]
    report = []
    for k, v in r1.items():
        if not v.dtype.is_floating_point:
            assert torch.equal(o1[k], v), k                        # num_batches_tracked
            continue
        noise = max(oracle.max_rel_err(r2[k], v), oracle.max_rel_err(o2[k], o1[k]))
        err = oracle.max_rel_err(o1[k], v)
        report.append((err / max(TOL, NOISE_FACTOR * noise), err, noise, k))
    report.sort(reverse=True)
    assert report[0][0] <= 1.0, "ours vs reference beyond the reference's own repeatability: " + "; ".join(
        f"{k}: err {e:.2e}, run-to-run {n:.2e}" for _, e, n, k in report[:4])
    return report


@pytest.mark.parametrize("yaml_path,momenta", [(DOT_YAML, ("momentum_buffer", "momentum_kd_buffer")),
                                               (DKD_YAML, ("momentum_buffer",))], ids=["kd_dot", "dkd_base"])
def test_unchanged_train_py_with_install_matches_the_reference_run(yaml_path, momenta, tmp_path):
    _staged()
    from oracle import ref_harness
    for iters in (1, 5):
        run = lambda tag, patched: ref_harness.run_train(yaml_path, patched=patched, out_dir=str(tmp_path / f"{tag}_{iters}"),
                                                         iters=iters, batch=64)
        ref, ref_again = run("ref_a", False), run("ref_b", False)
        ours, ours_again = run("ours_a", True), run("ours_b", True)
        assert ref["launches"] == 0 and ours["launches"] >= iters * 3, "the patched run must go through the sm_100a kernels"
        report = _compare((ours, ours_again), (ref, ref_again), momenta)
        if iters == 1:      # one step in: tensors whose repeatability is tight are matched tightly
            tight = [e for _, e, n, _ in report if n <= 1e-6]
            assert not tight or max(tight) <= TOL
    # the progress of both runs is the same story: same losses in the epoch log (worklog.txt, 4 decimals)
    pick = lambda w: [l.strip() for l in w.splitlines() if l.strip().startswith(("ce:", "kd:", "train_acc"))]
    for a, b in zip(pick(ours["worklog"]), pick(ref["worklog"])):
        ka, va = a.split(":")
        kb, vb = b.split(":")
        assert ka == kb and abs(float(va) - float(vb)) <= 2e-3 * max(1.0, abs(float(vb))), (a, b)


def _two_gpu_worker(rank, port, out):
    os.environ.update(LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    from oracle import ref_harness
    res = {}
    next_port = iter(range(port, port + 5))
    use_port = lambda: os.environ.update(MASTER_PORT=str(next(next_port)))     # a fresh rendezvous per run
    # base trainer: the reference's DDP averages the one gradient set correctly, so its 2-GPU run IS the oracle
    for tag, patched in (("ref", False), ("ref_again", False), ("ours", True), ("ours_again", True)):
        use_port()
        res[tag] = ref_harness.run_train(DKD_YAML, patched=patched, out_dir=os.path.join(out, f"{tag}{rank}"), iters=4, batch=32)
    # DOT trainer: the reference's own 2-GPU run leaves the task gradient rank-local, so only ours
    # runs -- through the epoch end, where the reference's loop asks rank 0 ALONE for optimizer.state_dict()
    use_port()
    res["dot"] = ref_harness.run_train(DOT_YAML, patched=True, out_dir=os.path.join(out, f"dot{rank}"), iters=4, batch=32,
                                       epochs=2)
    if rank == 0:
        torch.save(res, os.path.join(out, "rank0.pt"))


def test_unchanged_train_py_on_two_gpus_matches_the_reference_ddp_run(tmp_path):
    _staged()
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    mp.spawn(_two_gpu_worker, args=(29750 + 5 * (os.getpid() % 40), str(tmp_path)), nprocs=2, join=True)
    res = torch.load(str(tmp_path / "rank0.pt"), weights_only=False)
    _compare((res["ours"], res["ours_again"]), (res["ref"], res["ref_again"]), ("momentum_buffer",))
    dot = res["dot"]
    assert dot["latest"]["epoch"] == 2 and dot["launches"] > 0
    st = dot["latest"]["optimizer"]["state"]
    shapes = [v.shape for k, v in dot["student"].items() if k.endswith(("weight", "bias"))]
    assert len(st) == len(shapes) and all(set(s) == {"momentum_buffer", "momentum_kd_buffer"} for s in st.values())
    assert all(torch.isfinite(v).all() for v in dot["student"].values() if v.dtype.is_floating_point)
