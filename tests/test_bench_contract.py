"""bench.py's CPU (reference) arm prints one JSON line with the contract's keys -- runs without a GPU."""

import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*extra, env=None):
    e = dict(os.environ, **(env or {}))
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "1", "--cpu-batch", "2", *extra], capture_output=True, text=True, env=e,
                         timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    return lines


def test_reference_arm_line_has_the_contract_keys():
    lines = _run("--workload", "cifar_res32x4_res8x4_kd_dot")
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "train_images_per_sec" and d["unit"] == "images/s"
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["value"] > 0 and d["steps"] == 1 and d["warmup"] == 1 and d["data"] == "synthetic"
    assert d["config"]["workload"].startswith("cifar_res32x4_res8x4_kd_dot")
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["config"]["per_gpu_batch"] == 2
    assert d["e2e"] == {"value": d["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_defaults_to_the_workloads_own_batch_and_honours_warmup():
    """Same config as our arm: the per-step batch is the workload's per-GPU batch unless overridden."""
    e = dict(os.environ)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "0",
                          "--workload", "cifar_res32x4_res8x4_dkd_base"], capture_output=True, text=True, env=e,
                         timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][0])
    assert d["config"]["per_gpu_batch"] == 64 and d["steps"] == 2 and d["warmup"] == 0


def test_reference_arm_falls_back_to_the_oracle_port_without_a_reference_tree(tmp_path):
    lines = _run("--workload", "cifar_res32x4_res8x4_kd_dot",
                 env={"MDISTILLER_REF_ARCHIVE_ONLY": "1", "MDB200_REF_ARCHIVE": str(tmp_path / "none.tar.gz")})
    assert json.loads(lines[0])["cpu_baseline"]["kind"] == "port"


def test_staged_archive_is_the_mounted_reference_byte_for_byte(tmp_path):
    """oracle/make_ref.sh packs mdistiller/, configs/ and tools/train.py unmodified: every file the archive
    unpacks hashes to the manifest, and (where the tree is mounted) to the mounted file."""
    import hashlib
    import tarfile
    sys.path.insert(0, ROOT)
    from mdistiller_ddp_b200 import refhost
    if not os.path.isfile(refhost.ARCHIVE):
        import pytest
        pytest.skip("no staged archive (run oracle/make_ref.sh)")
    with tarfile.open(refhost.ARCHIVE) as tar:
        tar.extractall(tmp_path, filter="data")
    manifest = dict(reversed(l.split(maxsplit=1)) for l in
                    open(os.path.join(os.path.dirname(refhost.ARCHIVE), "MANIFEST.sha256")).read().splitlines())
    assert "tools/train.py" in manifest and "mdistiller/engine/dot.py" in manifest and len(manifest) > 50
    for rel, digest in manifest.items():
        got = hashlib.sha256(open(os.path.join(tmp_path, rel), "rb").read()).hexdigest()
        assert got == digest, rel
        mounted = os.path.join("/root/reference", rel)
        if os.path.isfile(mounted):
            assert hashlib.sha256(open(mounted, "rb").read()).hexdigest() == digest, rel


def test_reference_arm_other_ranks_print_nothing():
    assert _run("--workload", "cifar_res32x4_res8x4_kd_dot", env={"RANK": "1", "WORLD_SIZE": "2"}) == []


def test_stand_in_backbones_have_the_reference_parameter_counts():
    """SURVEY.md Appendix B (counted on the reference's own constructors)."""
    import sys
    sys.path.insert(0, ROOT)
    from mdistiller_ddp_b200 import workloads as W
    want = {"resnet8x4": 1_233_540, "resnet32x4": 7_433_860, "wrn_16_2": 703_284, "wrn_40_2": 2_255_156,
            "vgg8": 3_965_028, "vgg13": 9_462_180}
    for name, n in want.items():
        net = W.cifar_resnet(name) if name.startswith("resnet") else W.cifar_misc(name)
        assert sum(p.numel() for p in net.parameters()) == n, name
    for name, n in (("ResNet18", 11_689_512), ("ResNet34", 21_797_672)):
        assert sum(p.numel() for p in W.imagenet_resnet(name).parameters()) == n
    assert set(W.WORKLOADS) >= {"imagenet_r34_r18_dkd_dot", "cifar_res32x4_res8x4_kd_dot", "cifar_res32x4_res8x4_dkd_base",
                                "cifar_wrn40_2_wrn16_2_dkd_base", "cifar_vgg13_vgg8_dkd_base"}
