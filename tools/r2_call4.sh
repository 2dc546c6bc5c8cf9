#!/bin/bash
# round 2, GPU call 4: inline counting, drainer, decode capture at the bench's chunk length, CLI wall at config 3
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_decode2.py tests/test_gpu_path.py tests/test_gpu_stream_bed.py tests/test_gpu_related.py tests/test_zz_no_alloc.py -q -m gpu -s > gpurun_out/c4_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/c4_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/c4_bench_cfg3.json 2> gpurun_out/c4_bench_cfg3.err
timeout 600 python bench.py --config cfg2 --steps 5 --warmup 3 --no-cpu > gpurun_out/c4_bench_cfg2.json 2> gpurun_out/c4_bench_cfg2.err
GCTA_B200_COUNT_INLINE=0 timeout 600 python bench.py --config cfg2 --steps 5 --warmup 3 --no-cpu --no-e2e --no-parity --no-cold > gpurun_out/c4_bench_cfg2_separate.json 2>&1
PROF_M=171136 PROF_CHUNK=171136 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"decode_emit_diag|snp_stats" -c 4 -f -o gpurun_out/r02_decode_final_n50000_m171136 python tools/profile_kernels.py > gpurun_out/c4_ncu_decode.log 2>&1
timeout 1500 bash tools/cli_cfg3.sh 1 > gpurun_out/c4_cli_cfg3.log 2>&1
