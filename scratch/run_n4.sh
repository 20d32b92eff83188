#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "4-" 2>&1 | tail -3
run() { # n comm
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 10 --warmup 3 --no-sweep --comm $2 > gpurun_out/bench_n$1_$2.json 2> gpurun_out/bench_n$1_$2.err || tail -5 gpurun_out/bench_n$1_$2.err
  python scratch/show_n.py gpurun_out/bench_n$1_$2.json $2
}
run 4 auto
run 2 auto
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 4 --steps 2 --warmup 1 | cut -c1-200
