#!/bin/bash
# round 2, 8 GPUs: headline line, partition re-calibration, the exchange diagnosis, config 4, the CLI on 8 GPUs
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
nvidia-smi -L > gpurun_out/n8_gpus.txt
timeout 600 $TR --nproc-per-node 8 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu > gpurun_out/n8_bench_cfg3.json 2> gpurun_out/n8_bench_cfg3.err
GCTA_B200_E2E_PREFETCH=1 timeout 600 $TR --nproc-per-node 8 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu --no-parity > gpurun_out/n8_bench_cfg3_prefetch.json 2> gpurun_out/n8_bench_cfg3_prefetch.err
GCTA_B200_PART_ALPHA=1.33 GCTA_B200_PART_BETA=0.83 timeout 600 $TR --nproc-per-node 8 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu --no-e2e --no-parity > gpurun_out/n8_bench_cfg3_partold.json 2> gpurun_out/n8_bench_cfg3_partold.err
timeout 900 $TR --nproc-per-node 8 bench.py --gpus 8 --config cfg4 --steps 3 --warmup 3 --no-cpu > gpurun_out/n8_bench_cfg4.json 2> gpurun_out/n8_bench_cfg4.err
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu -s > gpurun_out/n8_tests.log 2>&1
timeout 900 bash tools/cli_cfg3.sh 8 > gpurun_out/n8_cli_cfg3.log 2>&1
