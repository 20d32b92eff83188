#!/bin/bash
mkdir -p gpurun_out
run() { MDB200_FUSED_UNROLL=$3 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 10 --warmup 3 --no-sweep --comm $2 > gpurun_out/u_n$1_$2_$3.json 2> gpurun_out/u.err || tail -5 gpurun_out/u.err; echo -n "unroll=$3 "; python scratch/show_n.py gpurun_out/u_n$1_$2_$3.json $2; }
run 2 fused_p2p 1
run 2 fused_p2p 2
run 2 fused_p2p 4
run 2 fused 1
run 2 fused 2
