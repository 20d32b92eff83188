#!/bin/bash
# 2-GPU lease: multi-GPU parity tests, the real drop-in under torchrun-like spawn, and comm-mode comparison on the short-step configs
mkdir -p gpurun_out
if [ "$1" = "tests" ]; then
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_zdropin_reference.py -m gpu -q > gpurun_out/r2_pytest_2gpu.log 2>&1
tail -12 gpurun_out/r2_pytest_2gpu.log | cut -c1-300
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
i=0
for wl in cifar_res32x4_res8x4_kd_dot cifar_vgg13_vgg8_dkd_base cifar_wrn40_2_wrn16_2_dkd_base; do
  for mode in auto barrier; do
    i=$((i+1))
    if [ $mode = barrier ]; then export MDB200_FUSED_SYNC=barrier; c=auto; else unset MDB200_FUSED_SYNC; c=$mode; fi
    timeout 300 $TR --master-port $((29500+i)) bench.py --gpus 2 --steps 30 --warmup 10 --no-sweep --workload $wl --comm $c \
      > gpurun_out/r2_n2_${wl}_${mode}.json 2> gpurun_out/r2_n2_${wl}_${mode}.err
    echo "$wl $mode rc=$?"
    python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r2_n2_${wl}_${mode}.json") if l.startswith("{")][-1])
    print("  ", round(d["value"]), "img/s e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],3), "opt_step_us", round(d["optimizer_step"]["us"]), d["optimizer_step"]["comm"], d["optimizer_step"]["fused"], "parity", d["dp_parity"]["max_rel_err"], d["dp_parity"]["ranks_bit_equal"])
except Exception as e:
    print("   no line:", e); print(open("gpurun_out/r2_n2_${wl}_${mode}.err").read()[-1500:])
PY
  done
done
