#!/bin/bash
# Where the text writer's time goes on a GPU box's host (no GPU work): default, without the file writes, onto tmpfs.
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; L=gpurun_out/s5_writer_diag.log; : > $L
(lscpu | grep -E "Model name|^CPU\(s\)|Thread|Socket"; df -T /tmp /dev/shm | tail -2; free -g | head -2) >> $L 2>&1
W=/tmp/wsel; mkdir -p $W
export GCTA_B200_SELFTEST_CONST_N=1
run() { echo "== $1" >> $L; ( time timeout 120 env $2 ./gcta_b200/gcta64_b200 --selftest-writer 50000 --out $3 ) 2>&1 | grep -E "selftest|real|user|sys" >> $L; rm -f $3.grm.gz; }
run "default, /tmp" "X=1" $W/a
run "dry run (no pwrite)" "GCTA_B200_GZ_DRY_RUN=1" $W/b
run "tmpfs" "X=1" /dev/shm/c
