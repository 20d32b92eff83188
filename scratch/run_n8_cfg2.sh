#!/bin/bash
mkdir -p gpurun_out
for w in cifar_wrn40_2_wrn16_2_dkd_base cifar_vgg13_vgg8_dkd_base; do
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --workload $w --steps 30 --warmup 5 --no-sweep > gpurun_out/c2_n8_$w.json 2> gpurun_out/c2_n8.err || tail -5 gpurun_out/c2_n8.err
  python scratch/show_n.py gpurun_out/c2_n8_$w.json $w
done
