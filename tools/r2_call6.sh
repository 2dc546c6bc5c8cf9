#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/decode_alone.py > gpurun_out/c6_decode_alone.json 2> gpurun_out/c6_decode_alone.err
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-parity --no-cold > gpurun_out/c6_bench_cfg3.json 2> gpurun_out/c6_bench_cfg3.err
timeout 900 python -m pytest tests/test_gpu_decode2.py -q -m gpu > gpurun_out/c6_tests.log 2>&1
