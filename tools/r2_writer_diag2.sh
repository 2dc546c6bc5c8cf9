#!/bin/bash
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; L=gpurun_out/s6_writer_diag.log; : > $L
W=/tmp/wsel; mkdir -p $W
export GCTA_B200_SELFTEST_CONST_N=1
( time timeout 120 ./gcta_b200/gcta64_b200 --selftest-writer 50000 --out $W/a ) 2>&1 | grep -E "selftest|real|user|sys" >> $L; ls -la $W/a.grm.gz >> $L
