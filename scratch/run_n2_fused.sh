#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "fused" 2>&1 | tail -2
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 10 --warmup 3 --no-sweep --comm $2 > gpurun_out/f_n$1_$2.json 2> gpurun_out/f_n$1_$2.err || tail -5 gpurun_out/f_n$1_$2.err; python scratch/show_n.py gpurun_out/f_n$1_$2.json $2; }
run 2 fused_p2p
run 2 fused
run 2 fused_kd
