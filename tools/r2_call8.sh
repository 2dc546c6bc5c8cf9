#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c8_all_tests.log 2>&1
echo "all tests exit $?" >> gpurun_out/c8_all_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/c8_bench_cfg3.json 2> gpurun_out/c8_bench_cfg3.err
