#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2_pytest_gpu_3.log 2>&1; tail -6 gpurun_out/r2_pytest_gpu_3.log | cut -c1-300
grep -E "^(FAILED|ERROR)" gpurun_out/r2_pytest_gpu_3.log | head
for wl in cifar_res32x4_res8x4_kd_dot cifar_res32x4_res8x4_dkd_base cifar_vgg13_vgg8_dkd_base cifar_wrn40_2_wrn16_2_dkd_base; do
  python bench.py --steps 40 --warmup 10 --no-sweep --workload $wl > gpurun_out/r2_n1_$wl.json 2> gpurun_out/r2_n1_$wl.err
  echo "$wl rc=$?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r2_n1_$wl.json") if l.startswith("{")][-1])
    t=d.get("torch_on_b200_baseline",{}).get("torch_train_step",{})
    print("  ", round(d["value"]), "img/s e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],3), "| torch-on-b200", t.get("images/s"), "| cpu", d["cpu_baseline"]["value"] if d.get("cpu_baseline") else None, d["config"]["backbones"][:20])
except Exception as e:
    print("   no line:", e); print(open("gpurun_out/r2_n1_$wl.err").read()[-1200:])
PY
done
