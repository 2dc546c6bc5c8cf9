#!/bin/bash
# what the driver runs at round end, on one GPU: the GPU suite, smoke(), the default bench line, the reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/s_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/s_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/s_smoke.log
timeout 900 python bench.py > gpurun_out/s_bench.json 2> gpurun_out/s_bench.err; echo "bench exit $?" >> gpurun_out/s_bench.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s_bench_ref.json 2> gpurun_out/s_bench_ref.err
