"""Kernel-level parity (GPU): each CUDA stage through the C ABI against the CPU oracle.

Integer / bit work (allele counts, N, planes, 16-bit Z codes) must be bit-exact; the SYRK is
checked exactly on integer-valued operands (every partial sum is representable in FP32, so any
descriptor / swizzle / pipeline error shows as a wrong integer) and to rounding on random data.
"""
import ctypes as C

import numpy as np
import pytest

from gcta_b200 import abi, synth
from oracle import oracle

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

P, I64, I32 = C.c_void_p, C.c_int64, C.c_int32


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def stream():
    return P(torch.cuda.current_stream().cuda_stream)


def pitched(bed, n_file):
    """host rows re-pitched to a multiple of 16 bytes + keep mask words"""
    m, stride = bed.shape
    pitch = (stride + 15) // 16 * 16
    out = np.zeros((m, pitch), dtype=np.uint8)
    out[:, :stride] = bed
    return out, pitch


def mask_words(idx, pitch):
    km = np.zeros(pitch // 4, dtype=np.uint32)
    for f in idx:
        km[f >> 4] |= np.uint32(1 << ((f & 15) * 2))
    return km


def run_decode(bed, n_file, keep=None, include=None, mu_in=None, sex=None, xchr=0, operand=abi.FP16, k0=0, kc=None):
    bedp, pitch = pitched(bed, n_file)
    n = n_file if keep is None else len(keep)
    m = bed.shape[0] if include is None else len(include)
    kc = m - k0 if kc is None else kc
    kc_pad = (kc + 63) // 64 * 64
    ldz = (n + 255) // 256 * 256
    kidx = np.arange(n_file) if keep is None else np.asarray(keep)
    d_bed = dev(bedp)
    d_km = dev(mask_words(kidx, pitch))
    d_keep = None if keep is None else dev(np.asarray(keep, dtype=np.int32))
    d_male = None
    if xchr:
        d_male = dev(mask_words([f for f in kidx if sex[f] == 1], pitch))
    d_inc = None if include is None else dev(np.asarray(include, dtype=np.int32))
    d_mu_in = None if mu_in is None else dev(np.asarray(mu_in, dtype=np.float64))
    Z = torch.full((kc_pad, ldz), 7, dtype=torch.int16, device="cuda")  # poison: padding must be overwritten
    mu = torch.zeros(m, dtype=torch.float64, device="cuda")
    cnt = torch.zeros(m, dtype=torch.int32, device="cuda")
    sm = torch.zeros(m, dtype=torch.int32, device="cuda")
    zlut = torch.zeros(m * 8, dtype=torch.float32, device="cuda")
    tot = torch.zeros(1, dtype=torch.int64, device="cuda")
    ptr = lambda t: P(None if t is None else t.data_ptr())
    abi.call("gcta_b200_bed_decode", ptr(d_bed), I64(pitch), I64(n_file), ptr(d_km), ptr(d_keep), I64(n), ptr(d_male),
             ptr(d_inc), I64(k0), I64(kc), ptr(d_mu_in), I32(xchr), I32(0), I32(operand), ptr(Z), I64(ldz), ptr(mu),
             ptr(cnt), ptr(sm), ptr(zlut), ptr(tot), stream())
    torch.cuda.synchronize()
    return dict(Z=Z, mu=mu.cpu().numpy(), cnt=cnt.cpu().numpy(), sum=sm.cpu().numpy(), zlut=zlut.cpu().numpy().reshape(m, 8),
                total_missing=int(tot.item()), d_bed=d_bed, pitch=pitch, ldz=ldz, kc_pad=kc_pad, d_inc=d_inc, d_keep=d_keep,
                d_male=d_male, d_zlut=zlut)


def z16_expected(z32, operand):
    t = torch.from_numpy(z32)
    t = t.to(torch.float16 if operand == abi.FP16 else torch.bfloat16)
    return t.view(torch.int16).numpy()


@pytest.mark.parametrize("n,m,miss,operand", [(203, 300, 0.03, abi.FP16), (1000, 130, 0.0, abi.BF16),
                                              (16500, 70, 0.02, abi.FP16), (37, 64, 0.3, abi.FP16)])
def test_bed_decode_bit_exact(n, m, miss, operand):
    bed = synth.make_bed_payload(11 + n, n, m, miss_rate=miss)
    r = run_decode(bed, n, operand=operand)
    mu, cnt, s = oracle.snp_mu(bed, n)
    assert np.array_equal(r["cnt"], cnt) and np.array_equal(r["sum"], s)
    assert np.array_equal(r["mu"], mu)  # float64, exact
    assert r["total_missing"] == int(n * m - cnt.sum())
    x = oracle.decode_x(bed, n)
    z, ok, _ = oracle.standardise(x)
    for k in range(0, m, max(1, m // 17)):
        zl = oracle.z_lut(mu[k])
        assert np.array_equal(r["zlut"][k, [3, 2, 0]].view(np.uint32), zl.view(np.uint32))  # x=0,1,2 <- codes 3,2,0
        assert r["zlut"][k, 1] == 0.0
    Z = r["Z"].cpu().numpy()
    assert np.array_equal(Z[:m, :n], z16_expected(z, operand))
    assert not Z[m:, :].any() and not Z[:, n:].any()  # padding rows / columns are zero


def test_bed_decode_edge_golden(golden_dir):
    from tests.helpers import load_plink
    bed, n, m, *_ = load_plink(golden_dir + "/g2_edge/t")  # monomorphic, all-missing SNP / individual, pad garbage
    r = run_decode(bed, n)
    mu, cnt, s = oracle.snp_mu(bed, n)
    assert np.array_equal(r["cnt"], cnt) and np.array_equal(r["sum"], s) and np.array_equal(r["mu"], mu)
    z, ok, _ = oracle.standardise(oracle.decode_x(bed, n))
    assert np.array_equal(r["Z"].cpu().numpy()[:m, :n], z16_expected(z, abi.FP16))


def test_bed_decode_keep_include_mu_override_xchr():
    n_file, m_file = 333, 257
    bed = synth.make_bed_payload(5, n_file, m_file, miss_rate=0.05)
    rng = np.random.default_rng(0)
    keep = np.sort(rng.choice(n_file, 190, replace=False)).astype(np.int32)
    inc = np.sort(rng.choice(m_file, 150, replace=False)).astype(np.int32)
    sex = rng.integers(1, 3, size=n_file).astype(np.int32)
    r = run_decode(bed, n_file, keep=keep, include=inc)
    mu, cnt, s = oracle.snp_mu(bed, n_file, keep=keep, include=inc)
    assert np.array_equal(r["cnt"], cnt) and np.array_equal(r["sum"], s) and np.array_equal(r["mu"], mu)
    z, _, _ = oracle.standardise(oracle.decode_x(bed, n_file, keep, inc))
    assert np.array_equal(r["Z"].cpu().numpy()[:150, :190], z16_expected(z, abi.FP16))
    # mu override (what --update-freq leaves in _mu)
    mu_o = np.clip(mu * 0.9 + 0.05, 0.02, 1.9)
    r = run_decode(bed, n_file, keep=keep, include=inc, mu_in=mu_o)
    z, _, _ = oracle.standardise(oracle.decode_x(bed, n_file, keep, inc), mu=mu_o)
    assert np.array_equal(r["Z"].cpu().numpy()[:150, :190], z16_expected(z, abi.FP16))
    # X chromosome: sex-weighted mu, male rows * sqrt(0.5); identity keep and gather keep
    for kp in (None, keep):
        r = run_decode(bed, n_file, keep=kp, sex=sex, xchr=1)
        mu_x, _, _ = oracle.snp_mu(bed, n_file, keep=kp, sex=sex, xchr=1)
        assert np.array_equal(r["mu"], mu_x)
        sx = sex if kp is None else sex[kp]
        z, _, _ = oracle.standardise(oracle.decode_x(bed, n_file, kp), sex=sx, xchr=1)
        nn = n_file if kp is None else len(kp)
        assert np.array_equal(r["Z"].cpu().numpy()[:m_file, :nn], z16_expected(z, abi.FP16))


def s_tiles_to_dense(S, tiles, n_pad):
    """[ntiles][64][256][4] tile storage -> dense [n_pad, n_pad] (only listed tiles filled)"""
    out = np.zeros((n_pad, n_pad), dtype=S.dtype)
    T = S.reshape(len(tiles), 64, 256, 4)
    for t, (bi, bj) in enumerate(tiles):
        blk = T[t].transpose(1, 0, 2).reshape(256, 256)  # [r][c4][4] -> [r][c]
        out[bi * 256:(bi + 1) * 256, bj * 256:(bj + 1) * 256] = blk
    return out


def run_syrk(Zt, n, operand, S=None):
    kc_pad, ldz = Zt.shape
    tiles = abi.tile_list(n)
    d_tiles = dev(tiles)
    if S is None:
        S = torch.zeros(len(tiles) * 256 * 256, dtype=torch.float32, device="cuda")
    abi.call("gcta_b200_syrk_tiles", P(Zt.data_ptr()), I64(ldz), I64(kc_pad), I32(operand), P(d_tiles.data_ptr()),
             I64(len(tiles)), P(S.data_ptr()), stream())
    torch.cuda.synchronize()
    return S, tiles


@pytest.mark.parametrize("n,kc,operand", [(256, 64, abi.FP16), (300, 128, abi.FP16), (1000, 512, abi.BF16),
                                          (4096, 256, abi.FP16), (2500, 4096, abi.FP16)])
def test_syrk_exact_on_integer_operands(n, kc, operand):
    ldz = (n + 255) // 256 * 256
    g = torch.Generator(device="cuda").manual_seed(n + kc)
    Zi = torch.randint(-3, 4, (kc, ldz), generator=g, device="cuda", dtype=torch.int32)
    Zi[:, n:] = 0
    Zt = Zi.to(torch.float16 if operand == abi.FP16 else torch.bfloat16).contiguous()
    S, tiles = run_syrk(Zt, n, operand)
    ref = (Zi.double().T @ Zi.double()).cpu().numpy()  # exact integers, |S| <= 9*kc < 2^24
    got = s_tiles_to_dense(S.cpu().numpy(), tiles, ldz)
    for (bi, bj) in tiles:
        r = slice(bi * 256, (bi + 1) * 256)
        c = slice(bj * 256, (bj + 1) * 256)
        assert np.array_equal(got[r, c], ref[r, c]), f"tile {(bi, bj)} wrong, max diff {np.abs(got[r, c] - ref[r, c]).max()}"
    # every element of the lower triangle is covered by some tile
    cov = np.zeros((ldz, ldz), dtype=bool)
    for (bi, bj) in tiles:
        cov[bi * 256:(bi + 1) * 256, bj * 256:(bj + 1) * 256] = True
    ii, jj = np.tril_indices(n)
    assert cov[ii, jj].all()


def test_syrk_accumulates_over_chunks_and_random_data():
    n, kc = 777, 320
    ldz = 1024
    g = torch.Generator(device="cuda").manual_seed(3)
    Z1 = torch.randn(kc, ldz, generator=g, device="cuda")
    Z2 = torch.randn(kc, ldz, generator=g, device="cuda")
    Z1[:, n:] = 0
    Z2[:, n:] = 0
    h1, h2 = Z1.half().contiguous(), Z2.half().contiguous()
    S, tiles = run_syrk(h1, n, abi.FP16)
    S, tiles = run_syrk(h2, n, abi.FP16, S)
    ref = (h1.double().T @ h1.double() + h2.double().T @ h2.double()).cpu().numpy()
    got = s_tiles_to_dense(S.cpu().numpy(), tiles, ldz)
    ii, jj = np.tril_indices(n)
    assert np.abs(got[ii, jj] - ref[ii, jj]).max() < 2e-3  # fp32 accumulation of 640 terms of O(1)


@pytest.mark.parametrize("n,kc,operand", [(300, 128, abi.FP16), (1000, 512, abi.BF16), (6000, 256, abi.FP16), (2500, 4096, abi.FP16)])
def test_syrk_cta_pairs_exact_on_integer_operands(n, kc, operand):
    """cta_group::2 variant (csrc/syrk2.cu): same tile storage, exact on integer operands; n=6000 gives 300 tiles =
    more pair items than SM pairs, so the persistent loop, the wave barrier and odd row-block counts are exercised;
    a second launch in counting mode (int32 epilogue) accumulates on top of the first."""
    ldz = (n + 255) // 256 * 256
    g = torch.Generator(device="cuda").manual_seed(n + kc + 1)
    Zi = torch.randint(-3, 4, (kc, ldz), generator=g, device="cuda", dtype=torch.int32)
    Zi[:, n:] = 0
    Zt = Zi.to(torch.float16 if operand == abi.FP16 else torch.bfloat16).contiguous()
    tiles = abi.tile_list(n)
    L = abi.lib()
    L.gcta_b200_pair_list.restype = C.c_int64
    npairs = L.gcta_b200_pair_list(P(tiles.ctypes.data), I64(len(tiles)), I32(12), P(None))
    pairs = np.zeros((npairs, 4), dtype=np.int32)
    L.gcta_b200_pair_list(P(tiles.ctypes.data), I64(len(tiles)), I32(12), P(pairs.ctypes.data))
    # every tile appears exactly once as an upper or lower member of a pair
    seen = sorted([t for t in pairs[:, 2:].ravel() if t >= 0])
    assert seen == list(range(len(tiles)))
    d_pairs = dev(pairs)
    S = torch.zeros(len(tiles) * 256 * 256, dtype=torch.float32, device="cuda")
    abi.call("gcta_b200_syrk_pairs", P(Zt.data_ptr()), P(Zt.data_ptr()), I64(ldz), I64(kc), I32(operand), P(d_pairs.data_ptr()),
             I64(npairs), P(S.data_ptr()), I32(0), stream())
    torch.cuda.synchronize()
    ref = (Zi.double().T @ Zi.double()).cpu().numpy()
    got = s_tiles_to_dense(S.cpu().numpy(), tiles, ldz)
    for (bi, bj) in tiles:
        r = slice(bi * 256, (bi + 1) * 256)
        c = slice(bj * 256, (bj + 1) * 256)
        assert np.array_equal(got[r, c], ref[r, c]), f"tile {(bi, bj)} wrong, max diff {np.abs(got[r, c] - ref[r, c]).max()}"
    # counting mode on |Z| in {0,1}
    Ok = (Zi != 0).to(torch.float16 if operand == abi.FP16 else torch.bfloat16).contiguous()
    Nt = torch.zeros(len(tiles) * 256 * 256, dtype=torch.int32, device="cuda")
    for _ in range(2):
        abi.call("gcta_b200_syrk_pairs", P(Ok.data_ptr()), P(Ok.data_ptr()), I64(ldz), I64(kc), I32(operand), P(d_pairs.data_ptr()),
                 I64(npairs), P(Nt.data_ptr()), I32(1), stream())
    torch.cuda.synchronize()
    okd = (Zi != 0).double()
    refn = (2 * (okd.T @ okd)).cpu().numpy()
    gotn = s_tiles_to_dense(Nt.cpu().numpy(), tiles, ldz)
    for (bi, bj) in tiles:
        r = slice(bi * 256, (bi + 1) * 256)
        c = slice(bj * 256, (bj + 1) * 256)
        assert np.array_equal(gotn[r, c], refn[r, c])


def test_diag_planes_paircount_finalize():
    n, m = 600, 1000
    bed = synth.make_bed_payload(21, n, m, miss_rate=0.04)
    r = run_decode(bed, n)
    ptr = lambda t: P(None if t is None else t.data_ptr())
    # K1d
    diag = torch.zeros(n, dtype=torch.float64, device="cuda")
    nii = torch.zeros(n, dtype=torch.int32, device="cuda")
    abi.call("gcta_b200_diag_accum", ptr(r["d_bed"]), I64(r["pitch"]), I64(n), P(None), P(None), I64(0), I64(m), ptr(r["d_zlut"]),
             ptr(diag), ptr(nii), stream())
    x = oracle.decode_x(bed, n)
    z, ok, mu = oracle.standardise(x)
    assert np.array_equal(nii.cpu().numpy(), ok.sum(0).astype(np.int32))
    dref = (z.astype(np.float64) ** 2).sum(0)
    # squares are rounded once to f32 (6e-8, unbiased) and summed in f32 per 192-SNP tile, f64 across tiles
    err = diag.cpu().numpy() - dref
    assert np.abs(err).max() <= 2e-6 * dref.max() and abs(err.mean()) <= 2e-7 * dref.mean()
    # K1m
    ldm = r["ldz"]
    nwords = (m + 31) // 32
    planes = torch.full((nwords, ldm), -1, dtype=torch.int32, device="cuda")
    abi.call("gcta_b200_nonmiss_planes", ptr(r["d_bed"]), I64(r["pitch"]), I64(n), P(None), I64(n), P(None), I64(0), I64(m),
             ptr(planes), I64(ldm), stream())
    okp = np.zeros((nwords * 32, ldm), dtype=np.uint32)
    okp[:m, :n] = ok
    exp = np.zeros((nwords, ldm), dtype=np.uint32)
    for b in range(32):
        exp |= okp[b::32] << np.uint32(b)
    assert np.array_equal(planes.cpu().numpy().view(np.uint32), exp)
    # K3
    tiles = abi.tile_list(n)
    d_tiles = dev(tiles)
    Nt = torch.zeros(len(tiles) * 256 * 256, dtype=torch.int32, device="cuda")
    abi.call("gcta_b200_paircount_tiles", ptr(planes), I64(ldm), I64(nwords), ptr(d_tiles), I64(len(tiles)), ptr(Nt), stream())
    Nref = (ok.astype(np.int64).T @ ok.astype(np.int64))
    Nd = s_tiles_to_dense(Nt.cpu().numpy(), tiles, ldm)
    ii, jj = np.tril_indices(n)
    assert np.array_equal(Nd[ii, jj], Nref[ii, jj])
    # K1n + K2 in counting mode: the tensor-core route must give the very same integers as AND+popc
    for operand, dt in ((abi.FP16, torch.float16), (abi.BF16, torch.bfloat16)):
        OK = torch.full((r["kc_pad"], ldm), 7, dtype=torch.int16, device="cuda")
        abi.call("gcta_b200_nonmiss_matrix", ptr(r["d_bed"]), I64(r["pitch"]), I64(n), P(None), I64(n), P(None), I64(0), I64(m),
                 I32(operand), ptr(OK), I64(ldm), stream())
        okm = OK.view(dt).float().cpu().numpy()
        assert np.array_equal(okm[:m, :n], ok.astype(np.float32)) and not okm[m:].any() and not okm[:, n:].any()
        Nt2 = torch.zeros_like(Nt)
        for _ in range(2):   # two "chunks" of the same data: counts accumulate in int32
            abi.call("gcta_b200_count_tiles", ptr(OK), I64(ldm), I64(r["kc_pad"]), I32(operand), ptr(d_tiles), I64(len(tiles)), ptr(Nt2), stream())
        Nd2 = s_tiles_to_dense(Nt2.cpu().numpy(), tiles, ldm)
        assert np.array_equal(Nd2[ii, jj], 2 * Nref[ii, jj])
    # the same counts from 8-bit E4M3 indicators (kind::f8f6f4, K block 128)
    E4M3 = 3
    kp8 = (m + 127) // 128 * 128
    OK8 = torch.full((kp8, ldm), 7, dtype=torch.uint8, device="cuda")
    abi.call("gcta_b200_nonmiss_matrix", ptr(r["d_bed"]), I64(r["pitch"]), I64(n), P(None), I64(n), P(None), I64(0), I64(m),
             I32(E4M3), ptr(OK8), I64(ldm), stream())
    ok8 = OK8.cpu().numpy()
    assert np.array_equal(ok8[:m, :n], ok.astype(np.uint8) * 0x38) and not ok8[m:].any() and not ok8[:, n:].any()
    Nt8 = torch.zeros_like(Nt)
    for _ in range(2):
        abi.call("gcta_b200_count_tiles", ptr(OK8), I64(ldm), I64(kp8), I32(E4M3), ptr(d_tiles), I64(len(tiles)), ptr(Nt8), stream())
    assert np.array_equal(s_tiles_to_dense(Nt8.cpu().numpy(), tiles, ldm)[ii, jj], 2 * Nref[ii, jj])
    # K2 + K4
    S, _ = run_syrk(r["Z"].view(torch.float16), n, abi.FP16)
    A = torch.zeros(n * (n + 1) // 2, dtype=torch.float32, device="cuda")
    Nl = torch.zeros(n * (n + 1) // 2, dtype=torch.int32, device="cuda")
    abi.call("gcta_b200_finalize_tiles", ptr(S), ptr(Nt), I32(m), ptr(d_tiles), I64(len(tiles)), I64(n), P(None), ptr(diag),
             ptr(nii), I64(0), I64(n), ptr(A), ptr(Nl), stream())
    torch.cuda.synchronize()
    Aref, Nref2, _ = oracle.make_grm_numpy(bed, n)
    from tests.helpers import check_grm
    check_grm(A.cpu().numpy(), Nl.cpu().numpy(), Aref[ii, jj], Nref2[ii, jj])
