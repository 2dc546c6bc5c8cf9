"""One small run of the device path for ncu (tools/r2_call*.sh): n = 50,000 individuals x a few chain-lengths of SNPs,
genotypes generated on the device.  The kernels captured here have the launch shapes of the headline workload
(same n, same SNPs per launch); only the number of launches is smaller."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gcta_b200 import abi, synth  # noqa: E402

n = int(os.environ.get("PROF_N", "50000"))
m = int(os.environ.get("PROF_M", "65536"))
chunk = int(os.environ.get("PROF_CHUNK", "32768"))
miss = float(os.environ.get("PROF_MISS", "0"))
keep = None
if os.environ.get("PROF_KEEP"):   # a keep list that drops every 10th individual (--remove of 10 % of the sample)
    import numpy as np
    keep = np.array([i for i in range(n) if i % 10 != 9], dtype=np.int32)
with abi.Plan(n, m, chunk_snps=chunk, keep=keep) as p:
    p.synth(1236, synth.snp_thresholds(1236, m, miss_rate=miss))
    for _ in range(int(os.environ.get("PROF_RUNS", "2"))):
        o = p.run()
    print(f"n={n} m={m} chunk={chunk}: {o.ms_total:.2f} ms (decode {o.ms_decode:.2f}, syrk {o.ms_syrk:.2f}), launches {o.launches}")
