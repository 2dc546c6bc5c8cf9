# End-to-end timing of the gcta64_b200 CLI at n=20000, M=20000 (1 % missing): read, device compute, .grm.gz / .grm.bin write.
# Usage (GPU box): bash tools/cli_scale.sh
ROOT=$(cd "$(dirname "$0")/.." && pwd)
W=/tmp/cliscale; rm -rf $W; mkdir -p $W; cd $W
python - <<PY
import sys, time; sys.path.insert(0, '$ROOT')
from gcta_b200 import synth
n, m = 20000, 20000
t=time.time(); bed = synth.make_bed_payload(5, n, m, miss_rate=0.01); print('generate', round(time.time()-t, 1), 's')
synth.write_plink('big', bed, n)
PY
ls -la big.bed
export GCTA_B200_TIMING=1
echo "== --make-grm (text)"; SECONDS=0
$ROOT/gcta_b200/gcta64_b200 --bfile big --make-grm --out o1 > out1.txt 2> err1.txt; echo "wall ${SECONDS}s"
tail -4 out1.txt; cat err1.txt
echo "== --make-grm-bin"; SECONDS=0
$ROOT/gcta_b200/gcta64_b200 --bfile big --make-grm-bin --out o2 > out2.txt 2> err2.txt; echo "wall ${SECONDS}s"
tail -3 out2.txt; cat err2.txt
ls -la o1.* o2.*
zcat o1.grm.gz | head -2; echo "lines: $(zcat o1.grm.gz | wc -l)"
