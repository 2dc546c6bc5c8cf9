#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/c9_all_tests.log 2>&1
echo "all tests exit $?" >> gpurun_out/c9_all_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-cold > gpurun_out/c9_bench_cfg3.json 2> gpurun_out/c9_bench_cfg3.err
GCTA_B200_CHAIN_LAUNCHES=1 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-cold --no-e2e --no-parity > gpurun_out/c9_bench_cfg3_chainlaunches.json 2> gpurun_out/c9_bench_cfg3_chainlaunches.err
PROF_M=171136 PROF_CHUNK=171136 PROF_RUNS=1 timeout 300 python tools/profile_kernels.py > gpurun_out/c9_prof_plain.log 2>&1
PROF_M=171136 PROF_CHUNK=171136 PROF_RUNS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:syrk_tiles -c 1 -f -o gpurun_out/r02_syrk_final_n50000_m171136 python tools/profile_kernels.py > gpurun_out/c9_ncu_syrk.log 2>&1
