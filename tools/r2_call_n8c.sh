#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR --nproc-per-node 8 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu > gpurun_out/n8c_bench_cfg3.json 2> gpurun_out/n8c_bench_cfg3.err
