#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR --nproc-per-node 8 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu > gpurun_out/n8b_bench_cfg3.json 2> gpurun_out/n8b_bench_cfg3.err
timeout 600 $TR --nproc-per-node 4 bench.py --gpus 4 --steps 4 --warmup 3 --no-cpu > gpurun_out/n8b_bench_cfg3_n4.json 2> gpurun_out/n8b_bench_cfg3_n4.err
timeout 600 $TR --nproc-per-node 2 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu > gpurun_out/n8b_bench_cfg3_n2.json 2> gpurun_out/n8b_bench_cfg3_n2.err
timeout 900 $TR --nproc-per-node 8 bench.py --gpus 8 --config cfg4 --steps 3 --warmup 3 --no-cpu > gpurun_out/n8b_bench_cfg4.json 2> gpurun_out/n8b_bench_cfg4.err
