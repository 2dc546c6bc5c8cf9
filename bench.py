#!/usr/bin/env python
"""bench.py - GRM pair.SNP/s of the --make-grm hot path on B200 ("Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n N --m M --miss R]

One step = one pass of the whole device path (decode + SYRK + [pair counts] + finalize) over
the synthetic workload.  Default workload = BASELINE.json configs[2], the configuration the
metric is quoted on: n=50,000 x M=1,000,000, MAF U(0.01,0.5), no missing genotypes.  With N>1
(launched by torchrun, one rank per GPU) the SAME problem is split into N row bands of the
lower triangle (strong scaling, no data-path collective); value = total pair.SNP / max-over-ranks
device time.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "grm_pair_snp_per_s"
UNIT = "pair*SNP/s"
# BASELINE.json configs[0..4]: name -> (n, M, missing rate, seed = 1234 + config index, GPUs it is named on)
CONFIGS = {
    "cfg1": (2000, 50000, 0.0, 1234, (1,)),
    "cfg2": (10000, 500000, 0.02, 1235, (1,)),
    "cfg3": (50000, 1000000, 0.0, 1236, (1, 2, 4, 8)),
    "cfg4": (100000, 600000, 0.0, 1237, (8,)),
    "cfg5": (20000, 5000000, 0.0, 1238, (1,)),
}
SEED = 1236


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1384.3), d.get("bf16_tflops", 1650.6), d.get("hbm_gbs", 6546.2), "measured"
    return 1400.0, 1590.0, 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def run_reference_sample(n_s, m_s, miss, seed, workdir):
    """Time the compiled, unmodified reference (oracle/_ref/gcta_ref, single thread) on an
    n_s x m_s sample of the workload.  -> (pair.SNP/s, seconds, cores)."""
    from gcta_b200 import synth
    from oracle import oracle
    if not oracle.have_ref():
        raise RuntimeError("oracle/_ref/gcta_ref is not built")
    bed = synth.make_bed_payload(seed, n_s, m_s, miss_rate=miss)
    synth.write_plink(os.path.join(workdir, "s"), bed, n_s)
    t0 = time.perf_counter()
    r = oracle.run_ref(["--bfile", "s", "--make-grm-bin", "--out", "o"], cwd=workdir)
    dt = time.perf_counter() - t0
    if "Analysis finished" not in r.stdout or not os.path.exists(os.path.join(workdir, "o.grm.bin")):
        raise RuntimeError("reference run failed: " + r.stderr[-300:])
    return n_s * (n_s + 1) / 2 * m_s / dt, dt, 1


def run_reference_jobs(n_s, m_s, procs, workdir):
    """`procs` concurrent runs of the unmodified reference CLI on the sample already written to workdir/s.*, each
    with its own output prefix - the way GCTA is spread over cores in practice: one --make-grm job per SNP set
    (chromosome), merged afterwards with --mgrm.  -> seconds until the last job has finished."""
    from oracle import oracle
    t0 = time.perf_counter()
    jobs = [subprocess.Popen([oracle.REF_BIN, "--bfile", "s", "--make-grm-bin", "--out", f"o{k}"], cwd=workdir,
                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for k in range(procs)]
    outs = [j.communicate()[0] for j in jobs]
    dt = time.perf_counter() - t0
    for k, out in enumerate(outs):
        if "Analysis finished" not in out or not os.path.exists(os.path.join(workdir, f"o{k}.grm.bin")):
            raise RuntimeError("reference run failed")
        os.remove(os.path.join(workdir, f"o{k}.grm.bin"))
    return dt


def main_reference(a, rank, world):
    if rank != 0:
        return
    from gcta_b200 import synth
    from oracle import oracle
    if not oracle.have_ref():
        raise RuntimeError("oracle/_ref/gcta_ref is not built")
    n_s, m_s = a.ref_n, a.ref_m
    if a.warmup + a.steps > 20:   # keep the whole run within a few minutes: ~8 s per step at the default sample
        m_s = max(1000, int(m_s * 20 / (a.warmup + a.steps)))
    # the reference has no threading of its own; all host cores are used by running one job per core
    procs = a.ref_procs if a.ref_procs > 0 else len(os.sched_getaffinity(0))
    secs = []
    with tempfile.TemporaryDirectory() as d:
        synth.write_plink(os.path.join(d, "s"), synth.make_bed_payload(SEED, n_s, m_s, miss_rate=a.miss), n_s)
        for i in range(a.warmup + a.steps):
            dt = run_reference_jobs(n_s, m_s, procs, d)
            if i >= a.warmup:
                secs.append(dt)
    value = procs * n_s * (n_s + 1) / 2 * m_s * len(secs) / sum(secs)
    cores = procs
    sample = (f"{procs} concurrent runs (one per host core; the reference is single-threaded, so cores are used the way GCTA jobs are "
              f"split per SNP set and merged with --mgrm) of an n={n_s} x M={m_s} sub-sample of the workload (same generator, seed {SEED}); "
              "the rate of the O(n^2 M) loop is scale-free")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": 1e3 * sum(secs) / len(secs), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"synthetic .bed n={a.n} x M={a.m}, miss={a.miss}, --make-grm (BASELINE {a.config_name})", "timed_sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def main_ours(a, rank, world, local_rank):
    import torch
    from gcta_b200 import abi, synth
    from gcta_b200 import dist as gdist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - gcta_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    ranks = gdist.Ranks(backend="nccl", device=torch.device("cuda", local_rank))   # same class the gloo CPU test drives

    def barrier():
        torch.cuda.synchronize()
        ranks.barrier()
        torch.cuda.synchronize()

    max_over_ranks, sum_over_ranks = ranks.max, ranks.sum

    n, m = a.n, a.m
    pairs_total = n * (n + 1) / 2
    thr = synth.snp_thresholds(SEED, m, miss_rate=a.miss)
    plan = abi.Plan(n, m, device=local_rank, part_rank=rank, part_world=world, operand=abi.BF16 if a.operand == "bf16" else abi.FP16)
    plan.synth(SEED, thr)  # synthetic genotypes generated in HBM (untimed set-up)

    # ---------------- device-resident throughput ("value") ----------------
    for _ in range(a.warmup):
        plan.run()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    dev_ms = syrk_ms = dec_ms = pc_ms = fin_ms = 0.0
    launches = 0
    for _ in range(a.steps):
        o = plan.run()
        dev_ms += o.ms_total; syrk_ms += o.ms_syrk; dec_ms += o.ms_decode; pc_ms += o.ms_paircount; fin_ms += o.ms_finalize
        launches += o.launches
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    row_begin, row_end = o.row_begin, o.row_end
    band_pairs = row_end * (row_end + 1) / 2 - row_begin * (row_begin + 1) / 2
    dev_ms_max = max_over_ranks(dev_ms)
    wall_ms_max = max_over_ranks(wall_ms)
    launches_total = int(sum_over_ranks(launches))
    per_rank = ranks.gather([row_begin, row_end, dev_ms / a.steps, syrk_ms / a.steps, dec_ms / a.steps, clocks.get("sm_mhz") or 0.0])
    value = pairs_total * m * a.steps / (dev_ms_max * 1e-3)

    # roofline of the dominant kernel (K2, tensor bound): algorithmic flops of this rank's band / K2 device time
    sust, burst, hbm, which = peaks()
    syrk_tflops = 2.0 * band_pairs * m * a.steps / (syrk_ms * 1e-3) / 1e12 if syrk_ms > 0 else 0.0
    syrk_tflops_min = -max_over_ranks(-syrk_tflops)
    chain = int(os.environ.get("GCTA_B200_CHAIN_SNPS", "32768")) or (1 << 40)
    chain = max(128, chain // 128 * 128)
    k2_chains = sum(-(-((e - b + 63) // 64 * 64) // chain) for b, e in plan.chunk_rows())   # TMEM chains per step on this rank
    k2_launches = len(plan.chunk_rows())   # all chains of a Z chunk run in one launch
    # Decode stage (K1s + K1f, HBM bound), SURVEY.md 8d: algorithmic bytes of THIS rank = M x (packed row read once +
    # 16-bit Z row of the band's ldz columns written once); the second read of the packed rows by K1f is not algorithmic
    ldz_rank = (max(row_end, 1) + 255) // 256 * 256
    k1_bytes = m * ((n + 3) // 4 + 2 * ldz_rank)
    k1_gbps = k1_bytes * a.steps / (dec_ms * 1e-3) / 1e9 if dec_ms > 0 else 0.0
    k1_gbps_min = -max_over_ranks(-k1_gbps)
    # the same two kernels timed alone (outside every timed region)
    alone = None
    if not a.no_decode_alone and a.miss == 0:
        time.sleep(0.5)   # let the board fall below its power cap
        alone = decode_alone(plan, n, m, row_begin, row_end)
    alone_gbps_min = -max_over_ranks(-(alone[1] if alone else 0.0))

    # ---------------- end to end through the C ABI with host buffers ("e2e") ----------------
    e2e = None
    stride = (n + 3) // 4
    base = row_begin * (row_begin + 1) // 2
    cnt = int(band_pairs)
    if not (a.no_e2e and a.no_parity):
        h_bed = torch.empty(stride * m, dtype=torch.uint8, pin_memory=True)
        plan.download_bed(h_bed.data_ptr())
        h_A = torch.empty(max(cnt, 1), dtype=torch.float32, pin_memory=True)
        h_N = np.empty(max(cnt, 1), dtype=np.int32)
        A_ptr = h_A.data_ptr() - 4 * base  # the ABI indexes the whole packed triangle; only the band is touched
        N_ptr = h_N.ctypes.data - 4 * base
    if not a.no_e2e:
        e2e_steps = max(1, min(a.steps, a.e2e_steps))
        ex = None
        if world == 1:
            def e2e_step():
                # H2D of chunk c+1 overlaps the compute of chunk c; finalize + D2H overlap the last chunk's SYRK
                plan.run_to_host(h_bed.data_ptr(), A_ptr, N_ptr)
            api = "gcta_b200_plan_run_to_host (what gcta_b200_make_grm runs: upload, compute and download pipelined), pinned host buffers"
            h2d = stride * m
        else:
            import torch.distributed as tdist
            host = tdist.new_group(backend="gloo")   # host barriers only; keeps NCCL kernels out of the timed region
            peer_err = ""
            if os.environ.get("GCTA_B200_E2E_EXCHANGE", "peer") != "nccl":
                # set the peer exchange up; if CUDA IPC is not available on this box (any rank), every rank falls back to
                # the NCCL all-gather below instead of losing the run
                try:
                    ex = abi.Exchange(plan, rank, world)      # the plan now reads the exchange's (zeroed) gather buffer
                except Exception as e:   # noqa: BLE001
                    peer_err = str(e)
                blobs = [None] * world
                tdist.all_gather_object(blobs, ex.blob if ex is not None else b"", group=host)
                if ex is not None and all(len(b) == len(ex.blob) for b in blobs):
                    try:
                        ex.connect(b"".join(blobs))
                    except Exception as e:   # noqa: BLE001
                        peer_err = str(e)
                elif ex is not None:
                    peer_err = "a peer could not create its exchange"
                if ranks.min(0.0 if peer_err else 1.0) < 1.0:
                    peer_err = peer_err or "a peer could not connect"
                    if ex is not None:
                        ex.disconnect()
                    barrier()
                    if ex is not None:
                        ex.close()
                    ex = None
            if ex is None:
                # every rank pulls 1/N of the SNP rows over its own PCIe link, one NCCL all-gather of the packed rows over
                # NVLink into the buffer the plan reads, then each computes its row band (nothing overlaps the exchange)
                rows_per = -(-m // world)
                pitch = plan.use_device_bed(None)
                d_all = torch.zeros(rows_per * world * pitch, dtype=torch.uint8, device="cuda")
                plan.use_device_bed(d_all.data_ptr())
                r0, r1 = min(m, rank * rows_per), min(m, (rank + 1) * rows_per)
                mine = d_all[rank * rows_per * pitch:(rank + 1) * rows_per * pitch]

                def e2e_step():
                    plan.upload_rows(h_bed.data_ptr(), r0, r1)
                    tdist.all_gather_into_tensor(d_all, mine)
                    torch.cuda.synchronize()
                    plan.run_to_host(None, A_ptr, N_ptr)   # genotypes resident after the all-gather; finalize + D2H pipelined
                api = ("gcta_b200_plan_upload_rows (1/N of the rows per rank) + NCCL all-gather of the packed rows over NVLink + "
                       "plan_run_to_host (finalize + download pipelined behind the last SYRK), pinned host buffers")
                h2d = stride * (r1 - r0)
            else:
                # gcta_b200_exchange_*: the file rows travel in 32 segments; of each, a rank uploads 1/N over its own PCIe
                # link and pulls the rest out of its peers' HBM with the copy engines over NVLink (CUDA IPC, no SM used), and
                # the plan starts chunk c as soon as the segments it reads are in: the exchange hides behind the compute
                e2e_phases = {}

                def e2e_step():
                    t = [time.perf_counter()]
                    tdist.barrier(group=host)   # no peer still reads my buffer from the step before
                    ex.post(h_bed.data_ptr())
                    tdist.barrier(group=host)   # every rank has recorded this step's segment events
                    ex.pull()
                    if os.environ.get("GCTA_B200_E2E_PREFETCH") == "1":
                        ex.finish()   # diagnosis only: the whole exchange first, then the compute (nothing concurrent)
                    t.append(time.perf_counter())
                    oo = plan.run_to_host(None, A_ptr, N_ptr)   # waits per chunk; finalize + D2H pipelined behind the last SYRK
                    t.append(time.perf_counter())
                    ex.finish()
                    e2e_phases.update({"barriers+post+pull (host)": 1e3 * (t[1] - t[0]), "plan_run_to_host (host wall)": 1e3 * (t[2] - t[1]),
                                       "device, first kernel to last": oo.ms_total, "device syrk": oo.ms_syrk, "device decode+diag": oo.ms_decode})
                api = ("gcta_b200_exchange_post/pull (per rank 1/N of the rows over PCIe, the rest peer-to-peer over NVLink by the copy "
                       "engines, 32 segments, overlapped with the compute) + plan_run_to_host (finalize + download pipelined), pinned host buffers")
                h2d = stride * sum(e - b for b, e in abi.exchange_slices(m, world, rank))
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        if ex is not None:
            ex.disconnect()
            barrier()     # no gather buffer is freed while a peer still maps it
            ex.close()    # the plan goes back to its own buffer
        d2h = sum_over_ranks(cnt * 4 * (2 if o.total_missing > 0 else 1))
        e2e = {"value": pairs_total * m * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(sum_over_ranks(h2d)),
               "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps, "api": api}
        if world > 1 and peer_err:
            e2e["peer_exchange_unavailable"] = peer_err[:300]
        if ex is not None:
            ph = ranks.gather([e2e_phases[k] for k in sorted(e2e_phases)])
            e2e["last_step_ms_per_rank"] = {k: [round(float(r[i]), 1) for r in ph] for i, k in enumerate(sorted(e2e_phases))}

    # ---------------- parity of THIS run's output (SURVEY.md 8c sub-block identity) ----------------
    # a seeded set of individuals that straddles every band boundary; the oracle (checker only, rank 0, float64 numpy)
    # builds their GRM block from the host copy of the device-generated .bed with the device's 2p; every rank compares
    # the rows of its own band.  N must be equal, A within the north-star's 1e-3, the diagonal mean within 1e-5.
    parity = None
    if not a.no_parity:
        from oracle import oracle
        mu_dev = np.empty(m, dtype=np.float64)
        if a.no_e2e:
            plan.run()
        plan.download(A_ptr, N_ptr, mu_dev.ctypes.data)   # the band as the last step left it on the device (+ 2p)
        sub, bounds = gdist.parity_subset(n, world, a.parity_sub, SEED)
        ns = len(sub)
        if rank == 0:
            t_or = time.perf_counter()
            Aref, Nref, _ = oracle.make_grm_numpy_blocked(h_bed.numpy().reshape(m, stride), n, keep=sub, mu=mu_dev)
            t_or = time.perf_counter() - t_or
            ref = np.concatenate([Aref.ravel(), Nref.astype(np.float64).ravel()])
        else:
            ref = np.zeros(2 * ns * ns)
        ref = ranks.broadcast(ref, 0)
        Aref, Nref = ref[:ns * ns].reshape(ns, ns), ref[ns * ns:].reshape(ns, ns)
        dA_max, n_bad, dsum, dn = gdist.band_deltas(h_A.numpy(), h_N, base, row_begin, row_end, sub, Aref, Nref)
        dA_max = max_over_ranks(dA_max)
        n_bad = int(sum_over_ranks(n_bad))
        dmean = sum_over_ranks(dsum) / max(1.0, sum_over_ranks(dn))
        parity = {"max_abs_dA": dA_max, "diag_mean_delta": abs(dmean), "N_equal": n_bad == 0, "sub": ns,
                  "pairs_checked": ns * (ns + 1) // 2, "band_boundaries_straddled": bounds,
                  "ok": bool(dA_max <= 1e-3 and abs(dmean) <= 1e-5 and n_bad == 0),
                  "oracle": "oracle.make_grm_numpy_blocked (float64) on the sub-block with the device's 2p; rank 0, not timed"}
        if rank == 0:
            parity["oracle_seconds"] = round(t_or, 1)

    # ---------------- the one-shot plugin call on pageable buffers, plan creation included ("e2e_cold") ----------------
    e2e_cold = None
    if world == 1 and not a.no_cold and not a.no_parity:
        bed_pg = np.array(h_bed.numpy())            # pageable copies: what a host program that malloc'ed its buffers passes
        A_pg = np.empty(max(cnt, 1), dtype=np.float32)
        N_pg = np.empty(max(cnt, 1), dtype=np.int32)
        args, ka = abi.make_args(bed_pg, n, m, operand=abi.BF16 if a.operand == "bf16" else abi.FP16, device=local_rank)
        import ctypes as C
        oc = abi.Out()
        oc.A_lower, oc.N_lower, oc.mu_out = A_pg.ctypes.data, N_pg.ctypes.data, None
        err = C.create_string_buffer(abi.ERRCAP)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = abi.lib().gcta_b200_make_grm(C.byref(args), C.byref(oc), err, abi.ERRCAP)
        cold_s = time.perf_counter() - t0
        if rc:
            e2e_cold = {"failed": err.value.decode(errors="replace")[:200]}
        else:
            same = float(np.abs(A_pg - h_A.numpy()[:len(A_pg)]).max())   # against the band downloaded for the parity check
            e2e_cold = {"value": pairs_total * m / cold_s, "unit": UNIT, "ms_per_step": 1e3 * cold_s, "steps": 1,
                        "api": "gcta_b200_make_grm, one call: plan creation (all cudaMalloc), upload from PAGEABLE host memory, compute, "
                               "download into pageable host memory, plan destruction",
                        "device_ms": oc.ms_total, "max_abs_diff_vs_plan_output": same}
        del bed_pg, A_pg, N_pg

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dev_ms_max / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": a.operand + " operands, f32 accumulate", "data": "synthetic",
            "config": {"workload": f"synthetic .bed n={n} x M={m}, MAF U(0.01,0.5), miss={a.miss}, --make-grm (BASELINE {a.config_name})",
                       "partition": f"{world} row band(s) of the lower triangle, packed genotypes replicated per GPU",
                       "l2": "inputs (packed bed + Z chunks) far larger than the 126 MB L2; no flush needed",
                       "timing": "CUDA events on the plan's stream, max over ranks"},
            "tflops": 2 * value / 1e12, "wall_ms_per_step": wall_ms_max / a.steps,
            "phases_ms_per_step_rank0": {"decode+diag": dec_ms / a.steps, "syrk": syrk_ms / a.steps, "paircount": pc_ms / a.steps,
                                         "finalize": fin_ms / a.steps},
            # frac is quoted against the burst peak: the kernel beats the "sustained" figure (a long cuBLAS loop under the
            # power cap), so that one is no ceiling; both are given
            "roofline": {"kernel": "syrk_tiles_kernel (K2)", "bound": "tensor", "achieved": syrk_tflops_min, "peak": burst, "unit": "TFLOP/s",
                         "frac": syrk_tflops_min / burst, "frac_of_sustained_peak": syrk_tflops_min / sust, "sustained_peak": sust,
                         "peak_source": which + " (burst bf16; min over ranks of algorithmic flops / K2 event time)",
                         "traffic": k2_traffic(n, world, chain, m), "launches_per_step": k2_launches, "tmem_chains_per_step": k2_chains,
                         "traffic_note": "dram bytes of the captured full-length launch (171,136 SNPs = 6 chains of 32,768) when this run launches that shape"},
            "roofline_decode": {"kernel": "snp_stats_kernel + decode_emit_diag_kernel (K1s + K1f)", "bound": "hbm", "achieved": k1_gbps_min,
                                "peak": hbm, "unit": "GB/s", "frac": k1_gbps_min / hbm, "peak_source": which + " (copy bandwidth)",
                                "bytes_per_step_rank0": k1_bytes, "traffic": k1_traffic(n, world),
                                "achieved_alone": alone_gbps_min or None, "frac_alone": (alone_gbps_min / hbm) if alone_gbps_min else None,
                                "alone": (f"the same kernels on one {alone[2]}-SNP chunk with no SYRK launch between them, so not under the step's power cap "
                                          f"(SM at its maximum clock, the condition the copy peak was measured under): rank 0 {alone[0]:.3f} ms per chunk; "
                                          "CUDA events, 3 warm-up + 10 timed passes, outside every timed region") if alone else None,
                                "note": "min over ranks; per-rank algorithmic bytes M x (ceil(n/4) + 2 ldz_rank) / that rank's decode event time"},
            "clocks": clocks, "gpu_launches": launches_total,
            "per_rank": [{"rows": [int(r[0]), int(r[1])], "ms_total": round(r[2], 2), "ms_syrk": round(r[3], 2), "ms_decode": round(r[4], 2),
                          "sm_mhz": r[5]} for r in per_rank],
        }
        if e2e:
            line["e2e"] = e2e
        if parity:
            line["parity"] = parity
        if e2e_cold:
            line["e2e_cold"] = e2e_cold
    # CPU baseline: the compiled reference on a bounded sample, rank 0 at N=1 only
    if rank == 0 and world == 1 and not a.no_cpu:
        try:
            with tempfile.TemporaryDirectory() as d:
                v, dt, cores = run_reference_sample(a.ref_n, a.ref_m, a.miss, SEED, d)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "seconds": dt,
                                    "sample": f"oracle/_ref/gcta_ref --make-grm-bin on n={a.ref_n} x M={a.ref_m} of the same generator (single thread: the reference has no threading)"}
        except Exception as ex:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 1, "kind": "reference", "sample": f"failed: {ex}"}
    # the reference on BASELINE config 1 IN FULL (n = 2,000 x M = 50,000, ~7 minutes of one core): once per round, on request
    if rank == 0 and world == 1 and a.cpu_full:
        try:
            with tempfile.TemporaryDirectory() as d:
                v, dt, cores = run_reference_sample(2000, 50000, 0.0, 1234, d)
            line["cpu_baseline_cfg1_full"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "seconds": dt,
                                              "sample": "oracle/_ref/gcta_ref --make-grm-bin on BASELINE config 1 in full (n=2000 x M=50000), single thread"}
        except Exception as ex:  # noqa: BLE001
            line["cpu_baseline_cfg1_full"] = {"value": None, "sample": f"failed: {ex}"}
    if rank == 0:
        print(json.dumps(line))
    plan.close()
    ranks.close()


def decode_alone(plan, n, m, row_begin, row_end):
    """K1s + K1f of this rank on one Z chunk, timed ALONE (CUDA events on torch's stream, 3 warm-up passes, 10 timed): no SYRK
    between the launches, so the board is below its power cap and the SM clock at its maximum - the condition the HBM
    copy peak of MEASURED_PEAKS.json was measured under.  -> (ms per chunk, algorithmic GB/s, SNPs in the chunk)"""
    import ctypes as C
    import torch
    from gcta_b200 import abi
    P, I64, I32 = C.c_void_p, C.c_int64, C.c_int32
    L = abi.lib()
    L.gcta_b200_plan_own_bed.restype = C.c_void_p
    bed_ptr = L.gcta_b200_plan_own_bed(plan._h)
    if not bed_ptr:
        return None
    pitch = plan.use_device_bed(None)
    kc = min(m, 171136)
    ldz = (max(row_end, 1) + 255) // 256 * 256
    km = np.zeros(pitch // 4, dtype=np.uint32)
    km[: n // 16] = np.uint32(0x55555555)
    for f in range(n // 16 * 16, n):
        km[f >> 4] |= np.uint32(1 << ((f & 15) * 2))
    d_km = torch.from_numpy(km).cuda()
    try:
        Z = torch.empty(((kc + 63) // 64 * 64, ldz), dtype=torch.int16, device="cuda")
    except RuntimeError:
        return None   # not enough free HBM for a second Z chunk next to the plan's
    mu = torch.zeros(kc, dtype=torch.float64, device="cuda")
    cnt = torch.zeros(kc, dtype=torch.int32, device="cuda")
    sm = torch.zeros(kc, dtype=torch.int32, device="cuda")
    zlut = torch.zeros(kc * 8, dtype=torch.float32, device="cuda")
    tot = torch.zeros(1, dtype=torch.int64, device="cuda")
    diag = torch.zeros(n, dtype=torch.float64, device="cuda")
    nii = torch.zeros(n, dtype=torch.int32, device="cuda")
    st = P(torch.cuda.current_stream().cuda_stream)

    def once():
        abi.call("gcta_b200_snp_stats", P(bed_ptr), I64(pitch), I64(n), P(d_km.data_ptr()), I64(n), P(None), I64(0), I64(kc), P(None), I32(0),
                 P(mu.data_ptr()), P(cnt.data_ptr()), P(sm.data_ptr()), P(zlut.data_ptr()), P(tot.data_ptr()), st)
        abi.call("gcta_b200_decode_emit_diag", P(bed_ptr), I64(pitch), I64(n), P(None), I64(0), I64(kc), P(zlut.data_ptr()), I32(abi.FP16),
                 P(Z.data_ptr()), I64(ldz), P(diag.data_ptr()), P(nii.data_ptr()), I64(row_begin), I64(row_end), P(None), st)

    for _ in range(3):
        once()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        once()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    del Z
    torch.cuda.empty_cache()
    return ms, kc * ((n + 3) // 4 + 2 * ldz) / ms / 1e6, kc


def k1_traffic(n, world):
    """dram bytes (read + write) of K1s + K1f per full-length Z chunk from the committed ncu capture, when this run has
    the captured shape (n individuals, whole triangle on one GPU); else None."""
    p = os.path.join(ROOT, "profiles", "k1_traffic.json")
    if world != 1 or not os.path.exists(p):
        return None
    d = json.load(open(p))
    return d["dram_bytes_total"] if d.get("n") == n else None   # per launch pair of 171,136 SNPs; algorithmic: d["algorithmic_bytes"]


def k2_traffic(n, world, chain, m):
    """dram__bytes_read.sum + dram__bytes_write.sum of one K2 launch from the committed ncu capture (profiles/),
    only when this run launches the captured shape (whole triangle of n individuals on one GPU, same SNPs per launch);
    else None."""
    p = os.path.join(ROOT, "profiles", "k2_traffic.json")
    if world != 1 or not os.path.exists(p):
        return None
    d = json.load(open(p))
    if d.get("n") != n or d.get("chain_snps") != chain or d.get("snps_per_launch") > m:
        return None
    return d["dram_bytes_read"] + d["dram_bytes_write"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--individuals", dest="n", type=int, default=None)   # --individuals: unambiguous under torchrun
    ap.add_argument("--m", "--snps", dest="m", type=int, default=None)
    ap.add_argument("--miss", type=float, default=None)
    ap.add_argument("--operand", default="fp16", choices=["fp16", "bf16"])
    ap.add_argument("--ref-n", type=int, default=600, help="individuals in the CPU reference sample")
    ap.add_argument("--ref-m", type=int, default=20000, help="SNPs in the CPU reference sample")
    ap.add_argument("--ref-procs", type=int, default=0, help="--impl reference: concurrent reference jobs, 0 = one per host core")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-full", action="store_true", help="also time the reference on BASELINE config 1 in full (about 7 minutes)")
    ap.add_argument("--config", default="cfg3", choices=sorted(CONFIGS), help="BASELINE.json configuration (cfg3 = the one the metric is quoted on)")
    ap.add_argument("--force", action="store_true", help="run a configuration on a GPU count BASELINE.json does not name it on")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-cold", action="store_true")
    ap.add_argument("--no-decode-alone", action="store_true")
    ap.add_argument("--parity-sub", type=int, default=256, help="individuals in the sub-block parity check")
    a = ap.parse_args()
    global SEED
    cn, cm, cmiss, SEED, cgpus = CONFIGS[a.config]
    a.n = cn if a.n is None else a.n
    a.m = cm if a.m is None else a.m
    a.miss = cmiss if a.miss is None else a.miss
    a.config_name = a.config if (a.n, a.m, a.miss) == (cn, cm, cmiss) else "custom"
    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    if a.impl == "ours" and a.config_name != "custom" and world_env not in cgpus and not a.force:
        raise SystemExit(f"bench.py: {a.config} is named on {cgpus} GPU(s) in BASELINE.json, this run has {world_env}; pass --force to run it anyway")
    a.warmup = max(a.warmup, 3) if a.impl == "ours" else a.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if a.impl == "reference":
        main_reference(a, rank, world)
    else:
        main_ours(a, rank, world, local_rank)


if __name__ == "__main__":
    main()
