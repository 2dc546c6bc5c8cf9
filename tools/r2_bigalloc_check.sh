#!/bin/bash
# The CLI with its large host buffers in huge-page mappings against the same run through operator new (GPU box).
ROOT=$(cd "$(dirname "$0")/.." && pwd); BIN=$ROOT/gcta_b200/gcta64_b200; L=$ROOT/gpurun_out/s7_bigalloc.log; mkdir -p $ROOT/gpurun_out; : > $L
W=/tmp/ba; mkdir -p $W; cd $W
n=8000; m=40000
( printf '\x6c\x1b\x01'; head -c $((2000*m)) /dev/urandom ) > r.bed
seq 1 $n | awk '{print "F"$1" I"$1" 0 0 1 -9"}' > r.fam
seq 1 $m | awk '{print 1+($1%22)"\trs"$1"\t0\t"$1"\tA\tG"}' > r.bim
export GCTA_B200_TIMING=1
timeout 20 $BIN --bfile r --make-grm-bin --out a > a.out 2> a.err; tail -3 a.out >> $L; cat a.err >> $L
GCTA_B200_BIG_ALLOC=0 timeout 20 $BIN --bfile r --make-grm-bin --out b > b.out 2> b.err; cat b.err >> $L
timeout 20 $BIN --bfile r --make-grm --out c > c.out 2> c.err; cat c.err >> $L
python - >> $L 2>&1 <<'PY'
import numpy as np, gzip
a = np.fromfile('/tmp/ba/a.grm.bin'); b = np.fromfile('/tmp/ba/b.grm.bin')
n = 8000
print('bin: shapes', a.shape, b.shape, 'max|a-b|', float(np.abs(a - b).max()), 'diag mean', float(a.reshape(n, n).diagonal().mean()), 'symmetric', bool((a.reshape(n, n) == a.reshape(n, n).T).all()))
rows = [ln.split() for ln in gzip.open('/tmp/ba/c.grm.gz', 'rt').readlines()[:50000]]
A = a.reshape(n, n)
d = max(abs(float(r[3]) - A[int(r[0]) - 1, int(r[1]) - 1]) for r in rows)
print('text vs bin on the first 50000 lines: max diff', d, 'last', rows[-1])
PY
