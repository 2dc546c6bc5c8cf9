#!/bin/bash
# Round-2 closing check on a GPU box: the CLI and drop-in GPU tests after the writer changes, then the writer alone at
# config 3's size (fabricated GRM, no GPU work): text with N constant, text with N varying, binary.
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_host_cli.py tests/test_dropin.py -q -m gpu > gpurun_out/s4_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/s4_tests.log
nproc > gpurun_out/s4_writer.log
W=/tmp/wsel; mkdir -p $W
GCTA_B200_SELFTEST_CONST_N=1 timeout 200 ./gcta_b200/gcta64_b200 --selftest-writer 50000 --out $W/c 2>&1 | grep selftest >> gpurun_out/s4_writer.log; ls -la $W/c.grm.gz >> gpurun_out/s4_writer.log; rm -f $W/c.grm.gz
timeout 200 ./gcta_b200/gcta64_b200 --selftest-writer 50000 --out $W/b --make-grm-bin 2>&1 | grep selftest >> gpurun_out/s4_writer.log; rm -f $W/b.grm.bin
