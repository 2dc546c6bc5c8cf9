"""Packed genotypes larger than their HBM allowance (GPU): the plan keeps no resident copy and streams the rows of one
SNP chunk at a time from host memory through two ring buffers (the north-star's "chunk-streamed from pinned host memory
when they exceed 180 GB"; the reference holds n x M floats in host RAM instead, data.cpp:989-991).

The allowance is forced down with GCTA_B200_BED_CAP_GIB; results must equal the resident run's: same chunks, same
kernels, same order of additions - N, 2p and the off-diagonal A bit for bit (the diagonal is a sum of float64 atomics
whose order is not fixed: 1e-7).
"""
import numpy as np
import pytest

from gcta_b200 import abi, synth
from oracle import oracle
from tests.helpers import check_grm, pack_lower

pytestmark = pytest.mark.gpu


def _diag_index(n):
    i = np.arange(n, dtype=np.int64)
    return i * (i + 1) // 2 + i


def _same(A0, N0, mu0, A1, N1, mu1, n):
    assert np.array_equal(N0, N1) and np.array_equal(mu0, mu1)
    off = np.ones(len(A0), dtype=bool)
    off[_diag_index(n)] = False
    assert np.array_equal(A0[off], A1[off])
    assert np.abs(A0[~off].astype(np.float64) - A1[~off]).max() <= 1e-7


@pytest.mark.parametrize("kw", [
    dict(n=700, m=9000, miss=0.02, chunk=1024),                  # fused decode, the indicator pass re-stages every chunk
    dict(n=700, m=9000, miss=0.0, chunk=1024),                   # nothing missing: pipelined last chunk
    dict(n=300, m=3000, miss=0.05, chunk=512),                   # fewer than 8192 SNPs: split-FP16 planes, row kernels
    dict(n=333, m=9000, miss=0.03, chunk=2048, keep=True, include=True),
    dict(n=260, m=2500, miss=0.04, chunk=640, alg=1),            # --grm-alg 1: a third pass over the packed rows
])
def test_streamed_genotypes_equal_resident(monkeypatch, kw):
    n, m = kw["n"], kw["m"]
    bed = synth.make_bed_payload(91 + n, n, m, miss_rate=kw["miss"])
    rng = np.random.default_rng(n)
    keep = np.sort(rng.choice(n, n * 2 // 3, replace=False)).astype(np.int32) if kw.get("keep") else None
    include = np.sort(rng.choice(m, m * 3 // 5, replace=False)).astype(np.int32) if kw.get("include") else None
    args = dict(keep=keep, include=include, alg=kw.get("alg", 0), chunk_snps=kw["chunk"])
    monkeypatch.delenv("GCTA_B200_BED_CAP_GIB", raising=False)
    A0, N0, mu0, o0 = abi.make_grm(bed, n, m, **args)
    monkeypatch.setenv("GCTA_B200_BED_CAP_GIB", "0.00001")       # ~10 KB: nothing fits, every chunk goes through the ring
    A1, N1, mu1, o1 = abi.make_grm(bed, n, m, **args)
    nn = n if keep is None else len(keep)
    _same(A0, N0, mu0, A1, N1, mu1, nn)
    assert o1.total_missing == o0.total_missing
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n, keep=keep, include=include, alg=kw.get("alg", 0))
    check_grm(A1, N1, pack_lower(Aref), pack_lower(Nref))


def test_streamed_plan_refuses_calls_that_need_a_resident_copy(monkeypatch):
    monkeypatch.setenv("GCTA_B200_BED_CAP_GIB", "0.00001")
    n, m = 100, 9000
    bed = synth.make_bed_payload(3, n, m)
    with abi.Plan(n, m) as p:
        for call in (lambda: p.upload(bed), lambda: p.run(), lambda: p.synth(3, synth.snp_thresholds(3, m))):
            with pytest.raises(abi.GctaB200Error) as e:
                call()
            assert e.value.code == 1 and "streamed" in str(e.value)
        A = np.zeros(n * (n + 1) // 2, dtype=np.float32)
        p.run_to_host(bed, A)       # the streaming entry points work
        Aref, _, _ = oracle.make_grm_numpy(bed, n)
        assert np.abs(A - pack_lower(Aref)).max() <= 1e-3


def test_config5_shape_streamed_through_a_4_gib_allowance(monkeypatch):
    """BASELINE config 5's shape (n = 20,000 x M = 5,000,000, a 25 GB .bed) with the allowance forced to 4 GiB: equal to
    the resident run."""
    import torch
    n, m, seed = 20_000, 5_000_000, 1238
    stride = (n + 3) // 4
    thr = synth.snp_thresholds(seed, m)
    h_bed = torch.empty(stride * m, dtype=torch.uint8, pin_memory=True)
    tri = n * (n + 1) // 2
    A0 = np.zeros(tri, dtype=np.float32)
    mu0 = np.zeros(m)
    monkeypatch.delenv("GCTA_B200_BED_CAP_GIB", raising=False)
    with abi.Plan(n, m, chunk_snps=262144) as p:
        p.synth(seed, thr)
        p.download_bed(h_bed.data_ptr())
        p.run()
        p.download(A0, None, mu0)
    monkeypatch.setenv("GCTA_B200_BED_CAP_GIB", "4")
    A1 = np.zeros(tri, dtype=np.float32)
    mu1 = np.zeros(m)
    with abi.Plan(n, m, chunk_snps=262144) as p:
        o = p.run_to_host(h_bed.data_ptr(), A1, None, mu1)
        print(f"config 5 shape streamed: device {o.ms_total:.0f} ms (decode {o.ms_decode:.0f}, syrk {o.ms_syrk:.0f}) for 25 GB through two ring buffers")
    assert np.array_equal(mu0, mu1)
    off = np.ones(tri, dtype=bool)
    off[_diag_index(n)] = False
    assert np.array_equal(A0[off], A1[off])
    assert np.abs(A0[~off].astype(np.float64) - A1[~off]).max() <= 1e-7
    assert abs(A1[~off].astype(np.float64).mean() - 1.0) < 1e-3


@pytest.mark.parametrize("miss", [0.01, 0.0])
def test_pageable_host_buffers_go_through_pinned_staging(monkeypatch, miss):
    """pageable buffers of 70 .. 150 MB (numpy memory): the one-shot call moves the genotypes in and the results out through
    pinned slots with helper threads (GCTA_B200_STAGE=0: the driver's own staging); same result either way.
    miss = 0 takes the pipelined last chunk (segments drained behind the SYRK), miss > 0 the plain download of A and N."""
    n, m, seed = 6000, 100_000, 321
    bed = np.empty((m, (n + 3) // 4), dtype=np.uint8)
    with abi.Plan(n, m) as p:
        p.synth(seed, synth.snp_thresholds(seed, m, miss_rate=miss))
        p.download_bed(bed)
    inc = np.sort(np.random.default_rng(1).choice(m, 60_000, replace=False)).astype(np.int32)
    monkeypatch.setenv("GCTA_B200_STAGE_MIN_MB", "32")   # the helpers normally only take transfers of GBs
    for include in (None, inc):
        monkeypatch.setenv("GCTA_B200_STAGE", "0")
        A0, N0, mu0, _ = abi.make_grm(bed, n, m, include=include)
        monkeypatch.delenv("GCTA_B200_STAGE")
        A1, N1, mu1, _ = abi.make_grm(bed, n, m, include=include)
        _same(A0, N0, mu0, A1, N1, mu1, n)
    # and against the oracle on a sub-block (the last run had the include list)
    sub = np.arange(0, n, 40).astype(np.int64)
    Aref, Nref, _ = oracle.make_grm_numpy_blocked(bed[inc], n, keep=sub, mu=mu1)
    ii, jj = np.tril_indices(len(sub))
    idx = sub[ii] * (sub[ii] + 1) // 2 + sub[jj]
    assert np.array_equal(N1[idx], Nref[ii, jj]) and np.abs(A1[idx].astype(np.float64) - Aref[ii, jj]).max() <= 1e-3
