#!/bin/bash
# round 2, GPU call 2: whole GPU suite, ncu captures of the decode kernels and of K2, launch list, A/B of the decode variants, other configs
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -q -m gpu -s > gpurun_out/c2_all_tests.log 2>&1
echo "all tests exit $?" >> gpurun_out/c2_all_tests.log
AB="--steps 2 --warmup 3 --no-e2e --no-parity --no-cpu --no-cold"
for v in 0 1; do for o in 4 6; do
GCTA_B200_DECODE_VAR=$v GCTA_B200_DECODE_OCC=$o timeout 300 python bench.py $AB > gpurun_out/c2_ab_var${v}_occ${o}.json 2>&1
done; done
timeout 300 python tools/profile_kernels.py > gpurun_out/c2_prof_plain.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"decode_emit_diag|snp_stats" -c 4 -f -o gpurun_out/r02_decode_n50000_m32768 python tools/profile_kernels.py > gpurun_out/c2_ncu_decode.log 2>&1
PROF_RUNS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:syrk_tiles -c 2 -f -o gpurun_out/r02_syrk_n50000_m32768 python tools/profile_kernels.py > gpurun_out/c2_ncu_syrk.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02_launches_cfg3_n50000_m1000000.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-parity --no-cpu --no-cold > gpurun_out/c2_ncu_launches.log 2>&1
timeout 600 python bench.py --config cfg1 --steps 5 --warmup 3 > gpurun_out/c2_bench_cfg1.json 2> gpurun_out/c2_bench_cfg1.err
timeout 600 python bench.py --config cfg2 --steps 5 --warmup 3 > gpurun_out/c2_bench_cfg2.json 2> gpurun_out/c2_bench_cfg2.err
timeout 900 python bench.py --config cfg5 --steps 2 --warmup 3 > gpurun_out/c2_bench_cfg5.json 2> gpurun_out/c2_bench_cfg5.err
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/c2_bench_cfg3.json 2> gpurun_out/c2_bench_cfg3.err
ls -la gpurun_out > gpurun_out/c2_ls.txt
