#!/bin/bash
# round 2, final one-GPU capture: the bench lines of every single-GPU configuration, the reference arm, the launch list and the
# full-set captures of the final kernels (each ncu command only after the same command has exited 0 without ncu)
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 --cpu-full > gpurun_out/f_bench_cfg3.json 2> gpurun_out/f_bench_cfg3.err
timeout 300 python bench.py --config cfg1 --steps 5 --warmup 3 --no-cpu > gpurun_out/f_bench_cfg1.json 2> gpurun_out/f_bench_cfg1.err
timeout 300 python bench.py --config cfg2 --steps 5 --warmup 3 --no-cpu > gpurun_out/f_bench_cfg2.json 2> gpurun_out/f_bench_cfg2.err
timeout 600 python bench.py --config cfg5 --steps 2 --warmup 3 --no-cpu > gpurun_out/f_bench_cfg5.json 2> gpurun_out/f_bench_cfg5.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/f_bench_reference.json 2> gpurun_out/f_bench_reference.err
timeout 300 python tools/decode_alone.py > gpurun_out/f_decode_alone.json 2> gpurun_out/f_decode_alone.err
AB="--steps 1 --warmup 3 --no-e2e --no-parity --no-cpu --no-cold"
timeout 300 python bench.py $AB > gpurun_out/f_bench_cfg3_plain.json 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_launches_cfg3_n50000_m1000000.csv python bench.py $AB > gpurun_out/f_ncu_launches.log 2>&1
PROF_M=171136 PROF_CHUNK=171136 timeout 300 python tools/profile_kernels.py > gpurun_out/f_prof_plain.log 2>&1
PROF_M=171136 PROF_CHUNK=171136 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"decode_emit_diag|snp_stats" -c 4 -f -o gpurun_out/r02_decode_final_n50000_m171136 python tools/profile_kernels.py > gpurun_out/f_ncu_decode.log 2>&1
PROF_RUNS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:syrk_tiles -c 2 -f -o gpurun_out/r02_syrk_final_n50000_m32768 python tools/profile_kernels.py > gpurun_out/f_ncu_syrk.log 2>&1
