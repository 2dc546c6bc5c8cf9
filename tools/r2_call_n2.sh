#!/bin/bash
# round 2, 2 GPUs: the multi-GPU call and the exchange, correctness first; then the N=2 bench line and the CLI on 2 GPUs
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
nvidia-smi -L > gpurun_out/n2_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_exchange.py tests/test_host_cli.py -q -m gpu -s -k "multi or exchange or two_gpus" > gpurun_out/n2_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/n2_tests.log
timeout 600 $TR --nproc-per-node 2 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu > gpurun_out/n2_bench_cfg3.json 2> gpurun_out/n2_bench_cfg3.err
timeout 900 bash tools/cli_cfg3.sh 2 > gpurun_out/n2_cli_cfg3.log 2>&1
