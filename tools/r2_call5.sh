#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode2.py tests/test_gpu_path.py tests/test_gpu_stream_bed.py tests/test_zz_no_alloc.py -q -m gpu -s > gpurun_out/c5_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/c5_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/c5_bench_cfg3.json 2> gpurun_out/c5_bench_cfg3.err
timeout 600 python bench.py --config cfg2 --steps 5 --warmup 3 --no-cpu > gpurun_out/c5_bench_cfg2.json 2> gpurun_out/c5_bench_cfg2.err
PROF_M=171136 PROF_CHUNK=171136 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"decode_emit_diag|snp_stats" -c 4 -f -o gpurun_out/r02_decode_final2_n50000_m171136 python tools/profile_kernels.py > gpurun_out/c5_ncu_decode.log 2>&1
