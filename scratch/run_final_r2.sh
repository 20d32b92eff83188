#!/bin/bash
# final single-GPU verification of round 2: the driver's own test command, smoke, graph-replayed CIFAR lines
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest_gpu_final.log 2>&1; tail -3 gpurun_out/r2_pytest_gpu_final.log | cut -c1-200
python __graft_entry__.py --smoke 2>&1 | tail -1 | cut -c1-250
for wl in cifar_res32x4_res8x4_kd_dot cifar_res32x4_res8x4_dkd_base cifar_vgg13_vgg8_dkd_base cifar_wrn40_2_wrn16_2_dkd_base; do
  python bench.py --steps 40 --warmup 10 --no-sweep --no-cpu --workload $wl 2>/dev/null | grep "^{" > gpurun_out/r2_n1_graph_$wl.json
  python - <<PY
import json
d = json.load(open("gpurun_out/r2_n1_graph_$wl.json"))
print("$wl", round(d["value"]), round(d["e2e"]["value"]), round(d["ms_per_step"], 3), d["cuda_graph"]["replays"], round(d["roofline"]["us_per_launch"], 1), d["gpu_launches"])
PY
done
