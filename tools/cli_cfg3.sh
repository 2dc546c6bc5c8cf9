#!/bin/bash
# Wall clock of the real program, gcta64_b200, on BASELINE config 3 (n = 50,000 x M = 1,000,000; a 12.5 GB .bed):
# read, device compute, .grm.bin / .grm.gz write, with the phase split the program prints under GCTA_B200_TIMING=1.
# Usage (GPU box): bash tools/cli_cfg3.sh [gpus]      -> JSON lines in gpurun_out/r02_cli_cfg3_gpus<g>.jsonl
ROOT=$(cd "$(dirname "$0")/.." && pwd)
G=${1:-1}
N=${CLI_N:-50000}; M=${CLI_M:-1000000}
W=${CLI_DIR:-/tmp/clicfg3}; rm -rf $W; mkdir -p $W; cd $W
OUT=$ROOT/gpurun_out/r02_cli_cfg3_gpus$G.jsonl; : > $OUT
df -h $W | tail -1
python - <<PY
import sys, time, numpy as np; sys.path.insert(0, '$ROOT')
from gcta_b200 import abi, synth
n, m = $N, $M
t = time.time()
stride = (n + 3) // 4
bed = np.empty((m, stride), dtype=np.uint8)
with abi.Plan(n, m) as p:
    p.synth(1236, synth.snp_thresholds(1236, m))
    p.download_bed(bed)
with open('big.bed', 'wb') as f:
    f.write(synth.BED_MAGIC); f.write(bed.data)
with open('big.fam', 'w') as f:
    f.write(''.join(f'F{i} I{i} 0 0 1 -9\n' for i in range(n)))
with open('big.bim', 'w') as f:
    f.write(''.join(f'{1 + k % 22}\trs{k}\t0\t{k + 1}\tA\tG\n' for k in range(m)))
print('{"generate_and_write_inputs_s": %.1f}' % (time.time() - t))
PY
ls -la big.bed
export GCTA_B200_TIMING=1
run() {   # name, flags...
  local name=$1; shift
  local t0=$(date +%s.%N)
  $ROOT/gcta_b200/gcta64_b200 --bfile big "$@" --gpus $G --out $name > $name.out 2> $name.err
  local t1=$(date +%s.%N)
  python - <<PY >> $OUT
import json
phases = {}
for ln in open('$W/$name.err'):
    ln = ln.strip()
    if ln.startswith('{'):
        try: d = json.loads(ln)
        except Exception: continue
        if 'gpu' in d: phases.setdefault('per_gpu', []).append(d)
        else: phases.update(d)
    elif ln: phases.setdefault('stderr', []).append(ln[:200])
print(json.dumps({"run": "$name", "flags": "$*", "gpus": $G, "n": $N, "m": $M, "wall_s": round($t1 - $t0, 2), **phases}))
PY
  tail -2 $name.out
}
run o_bin --make-grm-bin
ls -la o_bin.grm.bin o_bin.grm.id
python - <<PY >> $OUT
import numpy as np, json
n = $N
a = np.memmap('$W/o_bin.grm.bin', dtype=np.float64, mode='r', shape=(n, n))
d = np.array([a[i, i] for i in range(0, n, 97)])
print(json.dumps({"check": "o_bin.grm.bin", "diag_mean_sample": float(d.mean()), "symmetric_sample": bool(a[123, 45678 % n] == a[45678 % n, 123])}))
PY
rm -f o_bin.grm.bin
FREE=$(df --output=avail -BG $W | tail -1 | tr -dc 0-9)
if [ "${FREE:-0}" -gt 40 ]; then
  run o_txt --make-grm
  ls -la o_txt.grm.gz
  echo "{\"check\": \"o_txt.grm.gz\", \"first_lines\": \"$(zcat o_txt.grm.gz | head -3 | tr '\t\n' ' ;')\", \"bytes\": $(stat -c %s o_txt.grm.gz)}" >> $OUT
else
  echo "{\"skipped\": \"--make-grm text output: only ${FREE}G free\"}" >> $OUT
fi
cat $OUT
rm -rf $W
