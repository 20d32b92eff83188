#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "8-" 2>&1 | tail -4
run() { # n comm
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 10 --warmup 3 --no-sweep --comm $2 > gpurun_out/bench_n$1_$2.json 2> gpurun_out/bench_n$1_$2.err || tail -5 gpurun_out/bench_n$1_$2.err
  python scratch/show_n.py gpurun_out/bench_n$1_$2.json $2
}
run 8 nccl
run 8 fused
run 8 fused_p2p
run 8 fused_kd
run 4 nccl
run 4 fused_p2p
