#!/bin/bash
# round 2, GPU call 1: new decode kernels + relatedness parity + first bench lines and A/B switches (one GPU)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/c1_gpu.txt 2>&1
timeout 1500 python -m pytest tests/test_gpu_decode2.py tests/test_gpu_related.py tests/test_zz_no_alloc.py -x -q -s -m gpu > gpurun_out/c1_new_tests.log 2>&1
echo "new tests exit $?" >> gpurun_out/c1_new_tests.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/c1_bench_cfg3.json 2> gpurun_out/c1_bench_cfg3.err
AB="--steps 2 --warmup 3 --no-e2e --no-parity --no-cpu --no-cold"
GCTA_B200_DECODE=legacy timeout 300 python bench.py $AB > gpurun_out/c1_ab_legacy_decode.json 2>&1
GCTA_B200_DECODE_OCC=6 timeout 300 python bench.py $AB > gpurun_out/c1_ab_occ6.json 2>&1
GCTA_B200_EPI_RED=0 timeout 300 python bench.py $AB > gpurun_out/c1_ab_rmw_epilogue.json 2>&1
GCTA_B200_CHAIN_SNPS=0 timeout 300 python bench.py $AB > gpurun_out/c1_ab_chain_whole.json 2>&1
GCTA_B200_CHAIN_SNPS=16384 timeout 300 python bench.py $AB > gpurun_out/c1_ab_chain_16k.json 2>&1
GCTA_B200_DECODE_ROWS=96 timeout 300 python bench.py $AB > gpurun_out/c1_ab_rows96.json 2>&1
GCTA_B200_DECODE_ROWS=1536 timeout 300 python bench.py $AB > gpurun_out/c1_ab_rows1536.json 2>&1
timeout 1500 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_related.py > gpurun_out/c1_all_tests.log 2>&1
echo "all tests exit $?" >> gpurun_out/c1_all_tests.log
