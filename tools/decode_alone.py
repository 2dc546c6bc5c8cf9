"""The standardisation stage timed ALONE (no SYRK between the launches, so the board is far below its power cap and the
SM clock sits at its maximum): K1s + K1f on one 171,136-SNP chunk of n = 50,000 individuals, CUDA events on the
launching stream, 3 warm-up passes, then 20 timed.  bench.py times the same kernels inside the step, where they follow
a power-capped SYRK launch; the difference between the two rates is the clock, not the kernel.
Prints one JSON line."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gcta_b200 import abi, synth  # noqa: E402

n = int(os.environ.get("PROF_N", "50000"))
kc = int(os.environ.get("PROF_M", "171136"))
P, I64, I32 = C.c_void_p, C.c_int64, C.c_int32
with abi.Plan(n, kc, chunk_snps=kc) as p:
    p.synth(1236, synth.snp_thresholds(1236, kc))
    pitch = p.use_device_bed(None)
    d_bed = C.c_void_p(abi.lib().gcta_b200_plan_own_bed(p._h)) if False else None
    abi.lib().gcta_b200_plan_own_bed.restype = C.c_void_p
    bed_ptr = abi.lib().gcta_b200_plan_own_bed(p._h)
    ldz = (n + 255) // 256 * 256
    stride = (n + 3) // 4
    km = np.zeros(pitch // 4, dtype=np.uint32)
    full = np.uint32(0x55555555)
    km[: n // 16] = full
    for f in range(n // 16 * 16, n):
        km[f >> 4] |= np.uint32(1 << ((f & 15) * 2))
    d_km = torch.from_numpy(km).cuda()
    Z = torch.empty(((kc + 63) // 64 * 64, ldz), dtype=torch.int16, device="cuda")
    mu = torch.zeros(kc, dtype=torch.float64, device="cuda")
    cnt = torch.zeros(kc, dtype=torch.int32, device="cuda")
    sm = torch.zeros(kc, dtype=torch.int32, device="cuda")
    zlut = torch.zeros(kc * 8, dtype=torch.float32, device="cuda")
    tot = torch.zeros(1, dtype=torch.int64, device="cuda")
    diag = torch.zeros(n, dtype=torch.float64, device="cuda")
    nii = torch.zeros(n, dtype=torch.int32, device="cuda")
    st = P(torch.cuda.current_stream().cuda_stream)

    def once(with_diag=True):
        abi.call("gcta_b200_snp_stats", P(bed_ptr), I64(pitch), I64(n), P(d_km.data_ptr()), I64(n), P(None), I64(0), I64(kc), P(None), I32(0),
                 P(mu.data_ptr()), P(cnt.data_ptr()), P(sm.data_ptr()), P(zlut.data_ptr()), P(tot.data_ptr()), st)
        abi.call("gcta_b200_decode_emit_diag", P(bed_ptr), I64(pitch), I64(n), P(None), I64(0), I64(kc), P(zlut.data_ptr()), I32(abi.FP16),
                 P(Z.data_ptr()), I64(ldz), P(diag.data_ptr() if with_diag else None), P(nii.data_ptr() if with_diag else None), I64(0), I64(n), P(None), st)

    out = {"n": n, "snps": kc}
    for name, wd in (("with_diag", True), ("no_diag", False)):
        for _ in range(3):
            once(wd)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        reps = 20
        for _ in range(reps):
            once(wd)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        alg = kc * (stride + 2 * ldz)
        out[name] = {"ms_per_chunk": ms, "algorithmic_GBps": alg / ms / 1e6, "ms_per_1M_snps": ms * 1e6 / kc}
    print(json.dumps(out))
