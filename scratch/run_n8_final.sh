#!/bin/bash
mkdir -p gpurun_out
run() { # n comm
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 10 --warmup 3 --no-sweep --comm $2 > gpurun_out/final_n$1_$2.json 2> gpurun_out/final_n$1_$2.err || tail -5 gpurun_out/final_n$1_$2.err
  python scratch/show_n.py gpurun_out/final_n$1_$2.json $2
}
run 8 auto
run 8 nccl
run 4 auto
run 2 auto
