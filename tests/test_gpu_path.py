"""Whole-path parity (GPU) through gcta_b200_make_grm / the plan API against the reference.

Tolerances are the north-star's: N_ij bit-exact, max|dA| <= 1e-3, |mean diag delta| <= 1e-5.
Golden fixtures are outputs of the compiled, unmodified reference (tests/golden/).
"""
import os

import numpy as np
import pytest

from gcta_b200 import abi, grmio, synth
from oracle import oracle
from tests.helpers import GOLDEN, check_grm, load_plink, pack_lower

pytestmark = pytest.mark.gpu


def _names(path):
    return [ln.split()[0] for ln in open(path) if ln.strip()]


@pytest.mark.parametrize("case,txt,binf,kw", [
    ("g1_basic", "o", "ob", {}),
    ("g2_edge", "o", "ob", {}),
    ("g2_edge", "o_alg1", "ob_alg1", {"alg": 1}),
    ("g4_xchr", "o", "ob", {"xchr": 1}),
    ("g5_subset", "o", "ob", {"subset": True}),
    ("g7_nomiss", "o", "ob", {}),
    ("g10_related", "o", "ob", {}),   # duplicates, parent-offspring and sib pairs: off-diagonals of 0.4 .. 0.9
])
@pytest.mark.parametrize("operand", [abi.FP16, abi.BF16])
def test_golden(case, txt, binf, kw, operand):
    d = os.path.join(GOLDEN, case)
    bed, n, m, chrom, sex, snps, ids = load_plink(d + "/t")
    keep = include = None
    if kw.get("subset"):
        kn = set(_names(d + "/keep.txt"))
        keep = np.array([i for i, (f, _) in enumerate(ids) if f in kn], dtype=np.int32)
        en = set(_names(d + "/extract.txt"))
        include = np.array([k for k, s in enumerate(snps) if s in en], dtype=np.int32)
    A, N, mu, out = abi.make_grm(bed, n, m, keep=keep, include=include, sex=sex, xchr=kw.get("xchr", 0), alg=kw.get("alg", 0), operand=operand)
    Aref = pack_lower(grmio.read_grm_bin(f"{d}/{binf}.grm.bin"))
    _, Nref, _ = grmio.read_grm_gz(f"{d}/{txt}.grm.gz")
    # fewer than 8192 SNPs: either 16-bit operand type runs in the split-FP16 mode, so both meet the north-star's bar
    check_grm(A, N, Aref, Nref, atol=1e-3)
    assert out.launches > 0


def test_update_freq_sub_block():
    d = os.path.join(GOLDEN, "g5_subset")
    bed, n, m, chrom, sex, snps, ids = load_plink(d + "/t")
    kn = set(_names(d + "/keep.txt"))
    keep = np.array([i for i, (f, _) in enumerate(ids) if f in kn], dtype=np.int32)
    freq = np.array([float(ln.split()[2]) for ln in open(d + "/full.freq")])
    A, N, mu, _ = abi.make_grm(bed, n, m, keep=keep, mu=2.0 * freq)
    Aref = pack_lower(grmio.read_grm_bin(d + "/ob_uf.grm.bin"))
    assert np.abs(A - Aref).max() <= 1e-3
    assert np.array_equal(mu, 2.0 * freq)


def test_chunked_equals_single_pass_and_bands_tile_the_triangle():
    n, m = 700, 1500
    bed = synth.make_bed_payload(42, n, m, miss_rate=0.02)
    A0, N0, mu0, _ = abi.make_grm(bed, n, m)
    A1, N1, mu1, _ = abi.make_grm(bed, n, m, chunk_snps=256)  # 6 chunks accumulate into S
    assert np.array_equal(N0, N1) and np.array_equal(mu0, mu1)
    assert np.abs(A0 - A1).max() < 1e-5
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n)
    check_grm(A0, N0, pack_lower(Aref), pack_lower(Nref))
    # 3 row bands (the multi-GPU partition) written into one buffer reproduce the whole triangle
    A = np.full(n * (n + 1) // 2, np.nan, dtype=np.float32)
    N = np.full(n * (n + 1) // 2, -1, dtype=np.int32)
    covered = 0
    for r in range(3):
        with abi.Plan(n, m, part_rank=r, part_world=3) as p:
            p.upload(bed)
            o = p.run()
            p.download(A, N)
            covered += o.row_end - o.row_begin
    assert covered == n
    assert np.array_equal(A, A0) and np.array_equal(N, N0)


def test_bf16_operands_meet_the_bar_on_a_long_contraction():
    """BF16 proper (no upgrade at 20,000 SNPs): rounding errors of the 8-bit significands average out over M"""
    n, m = 1500, 20000
    bed = synth.make_bed_payload(55, n, m, miss_rate=0.01)
    A, N, mu, out = abi.make_grm(bed, n, m, operand=abi.BF16)
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n)
    dA, dd = check_grm(A, N, pack_lower(Aref), pack_lower(Nref))
    print(f"bf16, M=20000: max|dA|={dA:.3e} diag-mean delta={dd:.3e}")


def test_device_generator_matches_numpy_generator():
    n, m = 515, 333
    thr = synth.snp_thresholds(9, m, miss_rate=0.03)
    bed = synth.make_bed_payload(9, n, m, miss_rate=0.03)
    A0, N0, mu0, _ = abi.make_grm(bed, n, m)
    A = np.zeros_like(A0)
    N = np.zeros_like(N0)
    mu = np.zeros_like(mu0)
    with abi.Plan(n, m) as p:
        p.synth(9, thr)
        p.run()
        p.download(A, N, mu)
    assert np.array_equal(mu, mu0) and np.array_equal(N, N0) and np.array_equal(A, A0)


def test_config1_shape_no_missing():
    """BASELINE config 1 (n=2000, M=50000, no missing): N == M everywhere, A within tolerance."""
    n, m = 2000, 50000
    bed = synth.make_bed_payload(1235, n, m)
    A, N, mu, out = abi.make_grm(bed, n, m)
    assert out.total_missing == 0 and (N == m).all()
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n)
    dA, dd = check_grm(A, N, pack_lower(Aref), pack_lower(Nref))
    print(f"config1: max|dA|={dA:.3e} diag-mean delta={dd:.3e} syrk {out.ms_syrk:.2f} ms decode {out.ms_decode:.2f} ms")


def test_larger_with_missing_sub_block_check():
    """n=6000 x M=20000, 2 % missing, multi-tile persistent schedule (1128 tiles > 148 SMs).
    Oracle on a sub-block of individuals with the full-sample mu (SURVEY.md 8c identity)."""
    n, m = 6000, 20000
    bed = synth.make_bed_payload(77, n, m, miss_rate=0.02)
    A, N, mu, out = abi.make_grm(bed, n, m)
    assert out.total_missing > 0
    rng = np.random.default_rng(1)
    sub = np.sort(rng.choice(n, 500, replace=False))
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n, keep=sub, mu=mu)
    ii, jj = np.tril_indices(len(sub))
    gi, gj = sub[ii], sub[jj]
    idx = gi.astype(np.int64) * (gi + 1) // 2 + gj
    check_grm(A[idx], N[idx], Aref[ii, jj], Nref[ii, jj])
    mu_ref, _, _ = oracle.snp_mu(bed[:200], n)
    assert np.array_equal(mu[:200], mu_ref)


def test_tensor_core_counts_equal_popc_counts(monkeypatch):
    """N via the indicator SYRK (default) and via AND+popc over bit planes (GCTA_B200_N_MODE=popc) are the same
    integers, with a keep list (gather path), several chunks and heavy missingness."""
    n_file, m = 900, 2100
    bed = synth.make_bed_payload(63, n_file, m, miss_rate=0.2)
    keep = np.sort(np.random.default_rng(2).choice(n_file, 611, replace=False)).astype(np.int32)
    A0, N0, _, o0 = abi.make_grm(bed, n_file, m, keep=keep, chunk_snps=512)
    monkeypatch.setenv("GCTA_B200_N_MODE", "popc")
    A1, N1, _, o1 = abi.make_grm(bed, n_file, m, keep=keep, chunk_snps=512)
    assert np.array_equal(N0, N1) and np.array_equal(A0, A1)
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n_file, keep=keep)
    check_grm(A0, N0, pack_lower(Aref), pack_lower(Nref))


@pytest.mark.parametrize("with_keep", [False, True])
def test_row_bands_on_the_wide_row_decode_path(with_keep):
    """n >= 16384 takes K1's one-CTA-per-SNP path (byte-table emit loop); row bands then standardise only the
    first row_end individuals (narrower Z) and sum the diagonal only for their own slice.  Four bands written into
    one buffer must equal the single-GPU result bit for bit, with and without a keep list (gather path)."""
    n_file, m = 17000, 320
    bed = synth.make_bed_payload(23, n_file, m, miss_rate=0.01)
    keep = None
    if with_keep:
        keep = np.sort(np.random.default_rng(5).choice(n_file, 16500, replace=False)).astype(np.int32)
    n = n_file if keep is None else len(keep)
    A0, N0, mu0, _ = abi.make_grm(bed, n_file, m, keep=keep)
    A = np.full(n * (n + 1) // 2, np.nan, dtype=np.float32)
    N = np.full(n * (n + 1) // 2, -1, dtype=np.int32)
    for r in range(4):
        with abi.Plan(n_file, m, keep=keep, part_rank=r, part_world=4) as p:
            p.upload(bed)
            p.run()
            p.download(A, N)
    assert np.array_equal(N, N0) and np.array_equal(A, A0)
    # oracle on a sub-block that straddles band boundaries
    sub = np.sort(np.random.default_rng(6).choice(n, 300, replace=False))
    ksub = sub if keep is None else keep[sub]
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n_file, keep=ksub, mu=mu0)
    ii, jj = np.tril_indices(len(sub))
    gi, gj = sub[ii].astype(np.int64), sub[jj].astype(np.int64)
    idx = gi * (gi + 1) // 2 + gj
    check_grm(A[idx], N[idx], Aref[ii, jj], Nref[ii, jj])


@pytest.mark.parametrize("seed", range(12))
def test_randomised_shapes_against_the_oracle(seed):
    """Seeded random sweep over the corners the fixed cases do not pin: n around multiples of 256 / 512, M not a
    multiple of the 64 / 128-SNP K blocks, random keep / include lists, chunk sizes that split M unevenly, missing
    rates from 0 to 30 %, alg 0 / 1, X chromosome, 1-3 row bands.  N exact, A within the north-star tolerance."""
    rng = np.random.default_rng(1000 + seed)
    n_file = int(rng.choice([255, 256, 257, 300, 511, 512, 513, 640, 769, 1030]))
    m_file = int(rng.integers(70, 900))
    miss = float(rng.choice([0.0, 0.0, 0.01, 0.1, 0.3]))
    bed = synth.make_bed_payload(500 + seed, n_file, m_file, miss_rate=miss)
    keep = include = None
    if rng.random() < 0.5:
        keep = np.sort(rng.choice(n_file, int(rng.integers(n_file // 2, n_file)), replace=False)).astype(np.int32)
    if rng.random() < 0.5:
        include = np.sort(rng.choice(m_file, int(rng.integers(m_file // 2, m_file)), replace=False)).astype(np.int32)
    xchr = int(rng.random() < 0.25)
    alg = int(rng.random() < 0.3)
    sex = rng.integers(1, 3, size=n_file).astype(np.int32)
    chunk = int(rng.choice([0, 64, 128, 192, 320]))
    world = int(rng.choice([1, 1, 2, 3]))
    n = n_file if keep is None else len(keep)
    A = np.full(n * (n + 1) // 2, np.nan, dtype=np.float32)
    N = np.full(n * (n + 1) // 2, -1, dtype=np.int32)
    for r in range(world):
        with abi.Plan(n_file, m_file, keep=keep, include=include, sex=sex, xchr=xchr, alg=alg, chunk_snps=chunk,
                      part_rank=r, part_world=world) as p:
            p.upload(bed)
            p.run()
            p.download(A, N)
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n_file, keep=keep, include=include, sex=sex, xchr=xchr, alg=alg)
    check_grm(A, N, pack_lower(Aref), pack_lower(Nref))


@pytest.mark.parametrize("n,world", [(7000, 1), (9000, 2), (9216, 1), (14000, 3)])
def test_pipelined_download_equals_plain_download(n, world):
    """plan_run_to_host cuts the last chunk into segments (3-4 raster groups here) and streams each segment's rows to
    the host behind the remaining SYRK work; the result must be the very same bytes as run + download."""
    m = 9000   # above the 8192-SNP threshold, so the single-plane FP16 path (the pipelined one) is taken
    bed = synth.make_bed_payload(33, n, m)
    tri = n * (n + 1) // 2
    A0 = np.full(tri, np.nan, dtype=np.float32); N0 = np.full(tri, -1, dtype=np.int32)
    A1 = np.full(tri, np.nan, dtype=np.float32); N1 = np.full(tri, -1, dtype=np.int32)
    for r in range(world):
        with abi.Plan(n, m, chunk_snps=4096, part_rank=r, part_world=world) as p:
            p.upload(bed)
            p.run()
            p.download(A0, N0)
        with abi.Plan(n, m, chunk_snps=4096, part_rank=r, part_world=world) as p:
            o = p.run_to_host(bed, A1, N1)
            assert o.total_missing == 0
    assert np.array_equal(A0, A1) and np.array_equal(N0, N1) and (N1 == m).all()
    sub = np.sort(np.random.default_rng(8).choice(n, 200, replace=False))
    Aref, Nref, _ = oracle.make_grm_numpy(bed[:, :], n, keep=sub, mu=oracle.snp_mu(bed, n)[0])
    ii, jj = np.tril_indices(len(sub))
    idx = sub[ii].astype(np.int64) * (sub[ii] + 1) // 2 + sub[jj]
    check_grm(A1[idx], N1[idx], Aref[ii, jj], Nref[ii, jj])


def test_chunk_events_gate_each_chunk_on_its_rows():
    """plan_set_chunk_events: the caller fills the genotype buffer chunk by chunk on its own stream (late on purpose -
    a long spin kernel goes first) and records one event per chunk; the plan must wait on event c before reading
    chunk c, so the result equals the plain upload + run + download bytes."""
    import torch
    n, m = 1500, 13000
    bed = synth.make_bed_payload(5, n, m, miss_rate=0.02)
    tri = n * (n + 1) // 2
    A0 = np.empty(tri, dtype=np.float32); N0 = np.empty(tri, dtype=np.int32)
    A1 = np.full(tri, np.nan, dtype=np.float32); N1 = np.full(tri, -1, dtype=np.int32)
    with abi.Plan(n, m, chunk_snps=4096) as p:
        p.upload(bed); p.run(); p.download(A0, N0)
    with abi.Plan(n, m, chunk_snps=4096) as p:
        chunks = p.chunk_rows()
        assert len(chunks) == 4 and chunks[0][0] == 0 and chunks[-1][1] == m
        assert all(a[1] == b[0] for a, b in zip(chunks, chunks[1:]))
        pitch = p.use_device_bed(None)
        d_bed = torch.zeros(m * pitch, dtype=torch.uint8, device="cuda")
        p.use_device_bed(d_bed.data_ptr())
        h = torch.from_numpy(bed).pin_memory()
        side = torch.cuda.Stream()
        events = [torch.cuda.Event() for _ in chunks]
        torch.cuda.synchronize()
        with torch.cuda.stream(side):
            torch.cuda._sleep(400_000_000)   # ~0.2 s: the plan is launched long before the rows arrive
            for ev, (c0, c1) in zip(events, chunks):
                p.upload_rows_async(h.data_ptr(), c0, c1, side.cuda_stream)
                ev.record(side)
        p.set_chunk_events([ev.cuda_event for ev in events])
        p.run_to_host(None, A1, N1)
        p.set_chunk_events([])
        side.synchronize()
    assert np.array_equal(N0, N1) and np.array_equal(A0, A1)
    with abi.Plan(n, m, chunk_snps=4096) as p:
        with pytest.raises(abi.GctaB200Error):
            p.set_chunk_events([0, 0])   # count must be 0 or the number of chunks


def test_split_fp16_operand_mode_is_tighter():
    """GCTA_B200_FP16X2 (hi + lo planes, three tensor passes): agreement with the reference at FP32-accumulation
    level, an order of magnitude tighter than plain FP16 operands; several chunks and missing data."""
    n, m = 900, 12000
    bed = synth.make_bed_payload(17, n, m, miss_rate=0.01)
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n)
    ref = pack_lower(Aref)
    A1, N1, _, _ = abi.make_grm(bed, n, m, operand=abi.FP16, chunk_snps=4096)
    A2, N2, _, o2 = abi.make_grm(bed, n, m, operand=abi.FP16X2, chunk_snps=4096)
    assert np.array_equal(N1, pack_lower(Nref)) and np.array_equal(N2, N1)
    e1, e2 = np.abs(A1 - ref).max(), np.abs(A2 - ref).max()
    print(f"max|dA| fp16 {e1:.2e}  fp16x2 {e2:.2e}")
    assert e2 <= 5e-7 and e2 < e1 / 5
    check_grm(A2, N2, ref, pack_lower(Nref), atol=1e-6)


def test_errors_are_reported_not_thrown():
    bed = synth.make_bed_payload(1, 40, 50)
    with pytest.raises(abi.GctaB200Error) as e:
        abi.make_grm(bed, 40, 50, keep=np.array([], dtype=np.int32))
    assert "No individual is retained" in str(e.value)
    with pytest.raises(abi.GctaB200Error):
        abi.make_grm(bed, 40, 50, xchr=1, sex=np.zeros(40, dtype=np.int32))


@pytest.mark.parametrize("miss,xchr", [(0.0, 0), (0.03, 0), (0.03, 1)])
def test_grm_alg1(miss, xchr):
    """--grm-alg 1 (grm.cpp:230-237, 257-264, 278-285): centred genotypes, denominators sum_k ok ok d_k.
    Without missing data the denominator is the constant sum_k d_k; with missing data it comes from a second
    SYRK over ok*sqrt(d_k); chunked so that S, N and D all accumulate over several launches."""
    n, m = 420, 1300
    bed = synth.make_bed_payload(91, n, m, miss_rate=miss)
    sex = np.random.default_rng(3).integers(1, 3, size=n).astype(np.int32)
    A, N, mu, out = abi.make_grm(bed, n, m, alg=1, xchr=xchr, sex=sex, chunk_snps=512)
    Aref, Nref, _ = oracle.make_grm(bed, n, sex=sex, xchr=xchr, alg=1)
    check_grm(A, N, Aref, Nref)
    # and it is a different matrix from the default algorithm
    A0, _, _, _ = abi.make_grm(bed, n, m, xchr=xchr, sex=sex)
    assert np.abs(A0 - A).max() > 1e-3


def test_counts_chunk_by_chunk_equal_the_separate_pass(monkeypatch):
    """With missing genotypes in the first SNP chunk the indicators are written with Z and counted chunk by chunk; when
    they only turn up later, a separate indicator pass counts after the loop.  Same N (and A) either way."""
    n, m = 1100, 9000
    bed = synth.make_bed_payload(71, n, m, miss_rate=0.03)
    A0, N0, mu0, o0 = abi.make_grm(bed, n, m, chunk_snps=1024)                 # chunk by chunk
    monkeypatch.setenv("GCTA_B200_COUNT_INLINE", "0")
    A1, N1, mu1, o1 = abi.make_grm(bed, n, m, chunk_snps=1024)                 # separate pass
    monkeypatch.delenv("GCTA_B200_COUNT_INLINE")
    assert np.array_equal(N0, N1) and np.array_equal(A0, A1) and o0.launches < o1.launches
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n)
    check_grm(A0, N0, pack_lower(Aref), pack_lower(Nref))
    # nothing missing in the first two chunks, 5 % afterwards
    late = np.concatenate([synth.make_bed_payload(72, n, 2048, miss_rate=0.0), synth.make_bed_payload(73, n, m - 2048, miss_rate=0.05)])
    A2, N2, _, o2 = abi.make_grm(late, n, m, chunk_snps=1024)
    assert o2.total_missing > 0
    Aref, Nref, _ = oracle.make_grm_numpy(late, n)
    check_grm(A2, N2, pack_lower(Aref), pack_lower(Nref))


def test_short_tmem_chains_and_the_cta_pair_kernel(monkeypatch):
    """GCTA_B200_CHAIN_SNPS cuts the contraction into TMEM chains (all in one launch by default, one launch per chain with
    GCTA_B200_CHAIN_LAUNCHES=1 or with the CTA-pair kernel, GCTA_B200_2CTA=1): the same GRM up to FP32 summation order"""
    n, m = 9000, 9000      # 630 tiles: enough waves for the one-launch form (smaller tile lists always launch per chain)
    bed = synth.make_bed_payload(81, n, m, miss_rate=0.02)
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n)
    res = {}
    for name, env in [("whole", {"GCTA_B200_CHAIN_SNPS": "0"}), ("c256", {"GCTA_B200_CHAIN_SNPS": "256"}),
                      ("c256_launches", {"GCTA_B200_CHAIN_SNPS": "256", "GCTA_B200_CHAIN_LAUNCHES": "1"}),
                      ("pairs_c512", {"GCTA_B200_CHAIN_SNPS": "512", "GCTA_B200_2CTA": "1"})]:
        for k in ("GCTA_B200_CHAIN_SNPS", "GCTA_B200_CHAIN_LAUNCHES", "GCTA_B200_2CTA"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        A, N, _, _ = abi.make_grm(bed, n, m, chunk_snps=4096)
        check_grm(A, N, pack_lower(Aref), pack_lower(Nref))
        res[name] = A
    assert np.array_equal(res["c256"], res["c256_launches"])          # same chains, same order of additions
    assert np.abs(res["c256"].astype(np.float64) - res["whole"]).max() < 2e-6
    assert np.abs(res["pairs_c512"].astype(np.float64) - res["whole"]).max() < 2e-6


def test_keep_list_goes_through_compaction(monkeypatch):
    """--keep / --remove on a long contraction: each chunk's rows are compacted to the kept individuals and then take the
    plain fused path; same N and off-diagonal A as the per-individual gathers of the row kernels (GCTA_B200_COMPACT=0)"""
    n_file, m = 2500, 9000
    bed = synth.make_bed_payload(91, n_file, m, miss_rate=0.02)
    rng = np.random.default_rng(9)
    keep = np.sort(rng.choice(n_file, 1801, replace=False)).astype(np.int32)
    inc = np.sort(rng.choice(m, 8500, replace=False)).astype(np.int32)
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n_file, keep=keep, include=inc)
    A0, N0, mu0, o0 = abi.make_grm(bed, n_file, m, keep=keep, include=inc, chunk_snps=2048)
    check_grm(A0, N0, pack_lower(Aref), pack_lower(Nref))
    monkeypatch.setenv("GCTA_B200_COMPACT", "0")
    A1, N1, mu1, o1 = abi.make_grm(bed, n_file, m, keep=keep, include=inc, chunk_snps=2048)
    monkeypatch.delenv("GCTA_B200_COMPACT")
    assert np.array_equal(N0, N1) and np.array_equal(mu0, mu1)
    n = len(keep)
    off = np.ones(len(A0), dtype=bool)
    i = np.arange(n, dtype=np.int64)
    off[i * (i + 1) // 2 + i] = False
    assert np.array_equal(A0[off], A1[off]) and np.abs(A0[~off].astype(np.float64) - A1[~off]).max() <= 1e-6
    # two row bands with a keep list (the multi-GPU partition)
    A = np.full_like(A0, np.nan)
    N = np.full_like(N0, -1)
    for r in range(2):
        with abi.Plan(n_file, m, keep=keep, include=inc, chunk_snps=2048, part_rank=r, part_world=2) as p:
            p.upload(bed)
            p.run()
            p.download(A, N)
    assert np.array_equal(N, N0) and np.array_equal(A[off], A0[off])
