#!/bin/bash
# 8-GPU lease (round 2): full multi-GPU parity test log, then comm-mode comparison on the headline and short-step configs
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -q -rs > gpurun_out/r2_pytest_multi.log 2>&1
tail -6 gpurun_out/r2_pytest_multi.log | cut -c1-300
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
i=0
run() {   # workload n mode steps warmup
  i=$((i+1))
  timeout 400 $TR --nproc-per-node $2 --master-port $((29500+i)) bench.py --gpus $2 --steps $4 --warmup $5 --no-sweep --workload $1 --comm $3 \
      > gpurun_out/r2_n$2_$1_$3.json 2> gpurun_out/r2_n$2_$1_$3.err
  echo "$1 N=$2 $3 rc=$?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r2_n$2_$1_$3.json") if l.startswith("{")][-1])
    print("  ", round(d["value"]), "img/s e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],3), "opt_step_us", round(d["optimizer_step"]["us"]), d["optimizer_step"]["comm"], d["optimizer_step"]["fused"], "roofline", round(d["roofline"]["us_per_launch"],1), "us", round(d["roofline"]["frac"],3), "parity", d["dp_parity"]["max_rel_err"], d["dp_parity"]["ranks_bit_equal"])
except Exception as e:
    print("   no line:", e); print(open("gpurun_out/r2_n$2_$1_$3.err").read()[-1200:])
PY
}
run imagenet_r34_r18_dkd_dot 8 auto 10 3
run imagenet_r34_r18_dkd_dot 8 nccl 10 3
run cifar_res32x4_res8x4_kd_dot 8 auto 40 10
run cifar_res32x4_res8x4_kd_dot 8 nccl 40 10
run cifar_res32x4_res8x4_kd_dot 4 auto 40 10
run cifar_vgg13_vgg8_dkd_base 8 auto 40 10
run cifar_vgg13_vgg8_dkd_base 8 nccl 40 10
run cifar_wrn40_2_wrn16_2_dkd_base 8 auto 40 10
run cifar_wrn40_2_wrn16_2_dkd_base 8 nccl 40 10
