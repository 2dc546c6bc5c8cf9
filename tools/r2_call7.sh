#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/c7_all_tests.log 2>&1
echo "all tests exit $?" >> gpurun_out/c7_all_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/c7_bench_cfg3.json 2> gpurun_out/c7_bench_cfg3.err
AB="--steps 3 --warmup 3 --no-e2e --no-parity --no-cpu --no-cold"
GCTA_B200_CHAIN_SNPS=16384 timeout 300 python bench.py $AB > gpurun_out/c7_ab_chain16k.json 2>&1
GCTA_B200_CHAIN_SNPS=0 timeout 300 python bench.py $AB > gpurun_out/c7_ab_chain0.json 2>&1
timeout 600 python bench.py --config cfg2 --steps 5 --warmup 3 --no-cpu > gpurun_out/c7_bench_cfg2.json 2> gpurun_out/c7_bench_cfg2.err
