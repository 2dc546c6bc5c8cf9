#!/bin/bash
# Experiments queued for the next round ("Next round").  Each block is one gpurun call; run from the repo root.
# Nothing here is part of the product or of the tests.
set -u
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
pick='import sys,json; d=json.loads(sys.stdin.read()); e=d["e2e"]; print(d["n_gpus"], round(d["tflops"]), d["ms_per_step"], e["ms_per_step"], json.dumps(e.get("last_step_ms_per_rank")))'

# 1. Why is K2 slower while the peer exchange runs?  Same e2e harness, three ways, 4 GPUs (cheap) then 8:
#    a) default: pulls overlap the compute; b) PREFETCH: the whole exchange first, then the compute; c) NCCL all-gather.
#    If (b)'s "device syrk" equals the device-resident ms_syrk, the concurrent copies are the cause; if it is as slow as
#    (a), look at the pipelined last chunk / host side instead.
exp1() {
  local n=$1
  gpurun --gpus $n --timeout 600 -- "for mode in '' 'GCTA_B200_E2E_PREFETCH=1' 'GCTA_B200_E2E_EXCHANGE=nccl'; do env \$mode timeout 180 $TR --nproc-per-node $n bench.py --gpus $n --steps 3 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c '$pick'; done"
}

# 2. The wave-counted partition: 8-GPU step against the model's 236 ms; per-rank waves in per_rank.
exp2() {
  gpurun --gpus 8 --timeout 420 -- "timeout 300 $TR --nproc-per-node 8 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu 2>&1 | tail -1 > gpurun_out/n8_wave_partition.json; cut -c1-3000 gpurun_out/n8_wave_partition.json"
}

# 3. CLI wall at scale with the new writer / reader (n = 20k x M = 20k and n = 40k x M = 100k), one GPU.
exp3() {
  gpurun --timeout 900 -- "bash tools/cli_scale.sh 2>&1 | tail -20"
}

case "${1:-}" in
  1) exp1 "${2:-4}" ;;
  2) exp2 ;;
  3) exp3 ;;
  *) echo "usage: $0 1 [ngpus] | 2 | 3" ;;
esac
