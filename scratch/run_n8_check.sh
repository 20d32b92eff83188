#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "8-6-fused_kd]" 2>&1 | tail -2
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 --no-sweep > gpurun_out/final2_n8_auto.json 2> gpurun_out/final2_n8.err || tail -5 gpurun_out/final2_n8.err
python scratch/show_n.py gpurun_out/final2_n8_auto.json auto
