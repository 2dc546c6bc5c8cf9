#!/bin/bash
# round 2, GPU call 3: decode with prefetch + wave-aligned grid; new tests
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode2.py tests/test_gpu_multi.py tests/test_zz_no_alloc.py tests/test_gpu_path.py tests/test_gpu_stream_bed.py tests/test_gpu_related.py tests/test_host_cli.py -q -m gpu -s > gpurun_out/c3_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/c3_tests.log
AB="--steps 2 --warmup 3 --no-e2e --no-parity --no-cpu --no-cold"
for v in 0 1; do for o in 4 6; do
GCTA_B200_DECODE_VAR=$v GCTA_B200_DECODE_OCC=$o timeout 300 python bench.py $AB > gpurun_out/c3_ab_var${v}_occ${o}.json 2>&1
done; done
GCTA_B200_DECODE_ROWS=768 timeout 300 python bench.py $AB > gpurun_out/c3_ab_rows768.json 2>&1
GCTA_B200_DECODE_ROWS=192 timeout 300 python bench.py $AB > gpurun_out/c3_ab_rows192.json 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"decode_emit_diag|snp_stats" -c 4 -f -o gpurun_out/r02_decode_v2_n50000_m32768 python tools/profile_kernels.py > gpurun_out/c3_ncu_decode.log 2>&1
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/c3_bench_cfg3.json 2> gpurun_out/c3_bench_cfg3.err
GCTA_B200_STAGE=0 timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/c3_bench_cfg3_nostage.json 2> gpurun_out/c3_bench_cfg3_nostage.err
