#!/bin/bash
# the other named shapes on 2 GPUs (BASELINE names them on 1 or 8: --force)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR --nproc-per-node 2 bench.py --gpus 2 --config cfg1 --force --steps 5 --warmup 3 --no-cpu > gpurun_out/n2b_bench_cfg1.json 2> gpurun_out/n2b_bench_cfg1.err
timeout 300 $TR --nproc-per-node 2 bench.py --gpus 2 --config cfg2 --force --steps 5 --warmup 3 --no-cpu > gpurun_out/n2b_bench_cfg2.json 2> gpurun_out/n2b_bench_cfg2.err
timeout 600 $TR --nproc-per-node 2 bench.py --gpus 2 --config cfg5 --force --steps 2 --warmup 3 --no-cpu > gpurun_out/n2b_bench_cfg5.json 2> gpurun_out/n2b_bench_cfg5.err
timeout 900 $TR --nproc-per-node 2 bench.py --gpus 2 --config cfg4 --force --steps 2 --warmup 3 --no-cpu > gpurun_out/n2b_bench_cfg4.json 2> gpurun_out/n2b_bench_cfg4.err
