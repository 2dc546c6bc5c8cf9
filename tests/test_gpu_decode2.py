"""K1s + K1f (csrc/decode2.cu) through the C ABI against the CPU oracle (GPU).

Bit-exact: allele counts, 2p, the per-SNP standardised values, every 16-bit code of Z, the zero padding,
the per-individual non-missing counts.  The diagonal sums are float32 partial sums folded into float64:
compared with the oracle's float64 sum of the same float32 squares to 1e-6 relative.
"""
import ctypes as C

import numpy as np
import pytest

from gcta_b200 import abi, synth
from oracle import oracle
from tests.test_gpu_kernels import dev, mask_words, pitched, run_decode, stream, z16_expected

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

P, I64, I32 = C.c_void_p, C.c_int64, C.c_int32


def run_fused(bed, n, include=None, mu_in=None, operand=abi.FP16, k0=0, kc=None, ldz=None, diag_range=None, want_diag=True, want_ok=False):
    bedp, pitch = pitched(bed, n)
    m = bed.shape[0] if include is None else len(include)
    kc = m - k0 if kc is None else kc
    kc_pad = (kc + 63) // 64 * 64
    ldz = (n + 255) // 256 * 256 if ldz is None else ldz
    d_bed = dev(bedp)
    d_km = dev(mask_words(np.arange(n), pitch))
    d_inc = None if include is None else dev(np.asarray(include, dtype=np.int32))
    d_mu_in = None if mu_in is None else dev(np.asarray(mu_in, dtype=np.float64))
    Z = torch.full((kc_pad, ldz), 7, dtype=torch.int16, device="cuda")   # poison: every element must be overwritten
    mu = torch.zeros(m, dtype=torch.float64, device="cuda")
    cnt = torch.zeros(m, dtype=torch.int32, device="cuda")
    sm = torch.zeros(m, dtype=torch.int32, device="cuda")
    zlut = torch.zeros(m * 8, dtype=torch.float32, device="cuda")
    tot = torch.zeros(1, dtype=torch.int64, device="cuda")
    diag = torch.zeros(n, dtype=torch.float64, device="cuda")
    nii = torch.zeros(n, dtype=torch.int32, device="cuda")
    OK = torch.full(((kc + 127) // 128 * 128, ldz), 9, dtype=torch.uint8, device="cuda") if want_ok else None   # poisoned too
    ptr = lambda t: P(None if t is None else t.data_ptr())
    abi.call("gcta_b200_snp_stats", ptr(d_bed), I64(pitch), I64(n), ptr(d_km), I64(n), ptr(d_inc), I64(k0), I64(kc), ptr(d_mu_in),
             I32(0), ptr(mu), ptr(cnt), ptr(sm), ptr(zlut), ptr(tot), stream())
    f0, f1 = (0, n) if diag_range is None else diag_range
    abi.call("gcta_b200_decode_emit_diag", ptr(d_bed), I64(pitch), I64(n), ptr(d_inc), I64(k0), I64(kc), ptr(zlut), I32(operand),
             ptr(Z), I64(ldz), ptr(diag if want_diag else None), ptr(nii if want_diag else None), I64(f0), I64(f1), ptr(OK), stream())
    torch.cuda.synchronize()
    return dict(OK=None if OK is None else OK.cpu().numpy(), Z=Z.cpu().numpy(), mu=mu.cpu().numpy(), cnt=cnt.cpu().numpy(), sum=sm.cpu().numpy(),
                zlut=zlut.cpu().numpy().reshape(m, 8), total_missing=int(tot.item()), diag=diag.cpu().numpy(), nii=nii.cpu().numpy())


def oracle_diag(z, ok):
    z64 = z.astype(np.float64)
    return (z64 * z64).sum(0), ok.sum(0).astype(np.int32)


@pytest.mark.parametrize("n,m,miss,operand", [
    (203, 300, 0.03, abi.FP16),        # one partly filled slab, n not a multiple of 16
    (1000, 130, 0.0, abi.BF16),
    (2048, 97, 0.1, abi.FP16),         # exactly one slab
    (5003, 700, 0.02, abi.FP16),       # three slabs, several table blocks, row groups
    (16500, 70, 0.02, abi.FP16),
    (37, 64, 0.3, abi.FP16),
    (4100, 49, 0.5, abi.FP16),         # 49 rows = one table block + one row
])
def test_fused_decode_bit_exact(n, m, miss, operand):
    bed = synth.make_bed_payload(11 + n, n, m, miss_rate=miss)
    r = run_fused(bed, n, operand=operand)
    mu, cnt, s = oracle.snp_mu(bed, n)
    assert np.array_equal(r["cnt"], cnt) and np.array_equal(r["sum"], s)
    assert np.array_equal(r["mu"], mu)   # float64, exact
    assert r["total_missing"] == int(n * m - cnt.sum())
    z, ok, _ = oracle.standardise(oracle.decode_x(bed, n))
    for k in range(0, m, max(1, m // 17)):
        zl = oracle.z_lut(mu[k])
        assert np.array_equal(r["zlut"][k, [3, 2, 0]].view(np.uint32), zl.view(np.uint32))   # x = 0, 1, 2 <- codes 3, 2, 0
        assert r["zlut"][k, 1] == 0.0
    assert np.array_equal(r["Z"][:m, :n], z16_expected(z, operand))
    assert not r["Z"][m:, :].any() and not r["Z"][:, n:].any()   # padding rows / columns are zero
    d_ref, n_ref = oracle_diag(z, ok)
    assert np.array_equal(r["nii"], n_ref)
    assert np.allclose(r["diag"], d_ref, rtol=1e-6, atol=1e-9)


def test_fused_decode_equals_the_row_kernels():
    """same Z codes, per-SNP outputs and diagonal as gcta_b200_bed_decode + gcta_b200_diag_accum"""
    n, m = 3001, 513
    bed = synth.make_bed_payload(21, n, m, miss_rate=0.04)
    a = run_decode(bed, n)
    b = run_fused(bed, n)
    assert np.array_equal(a["Z"].cpu().numpy(), b["Z"])
    assert np.array_equal(a["mu"], b["mu"]) and np.array_equal(a["cnt"], b["cnt"]) and np.array_equal(a["sum"], b["sum"])
    assert np.array_equal(a["zlut"].view(np.uint32), b["zlut"].view(np.uint32))
    assert a["total_missing"] == b["total_missing"]


def test_fused_decode_include_mu_override_and_offset_chunk():
    n, m_file = 777, 400
    bed = synth.make_bed_payload(5, n, m_file, miss_rate=0.05)
    rng = np.random.default_rng(0)
    inc = np.sort(rng.choice(m_file, 250, replace=False)).astype(np.int32)
    mu, cnt, s = oracle.snp_mu(bed, n, include=inc)
    mu_o = np.clip(mu * 0.9 + 0.05, 0.02, 1.9)
    # a chunk in the middle of the included SNPs, with the 2p override (what --update-freq leaves in _mu)
    k0, kc = 64, 130
    r = run_fused(bed, n, include=inc, mu_in=mu_o, k0=k0, kc=kc)
    z, ok, _ = oracle.standardise(oracle.decode_x(bed, n, None, inc), mu=mu_o)
    assert np.array_equal(r["Z"][:kc, :n], z16_expected(z[k0:k0 + kc], abi.FP16))
    assert not r["Z"][kc:, :].any()
    assert np.array_equal(r["mu"][k0:k0 + kc], mu_o[k0:k0 + kc]) and np.array_equal(r["cnt"][k0:k0 + kc], cnt[k0:k0 + kc])
    d_ref, n_ref = oracle_diag(z[k0:k0 + kc], ok[k0:k0 + kc])
    assert np.array_equal(r["nii"], n_ref) and np.allclose(r["diag"], d_ref, rtol=1e-6, atol=1e-9)


def test_fused_decode_row_band():
    """a row band: Z narrower than n (only the first ldz individuals), diagonal of the band's own individuals only"""
    n, m = 6000, 333
    bed = synth.make_bed_payload(8, n, m, miss_rate=0.02)
    ldz, f0, f1 = 2560, 2304, 2560
    r = run_fused(bed, n, ldz=ldz, diag_range=(f0, f1))
    z, ok, _ = oracle.standardise(oracle.decode_x(bed, n))   # statistics over all n individuals
    assert np.array_equal(r["Z"][:m, :], z16_expected(z[:, :ldz], abi.FP16))
    assert not r["Z"][m:, :].any()
    d_ref, n_ref = oracle_diag(z, ok)
    assert np.array_equal(r["nii"][f0:f1], n_ref[f0:f1]) and np.allclose(r["diag"][f0:f1], d_ref[f0:f1], rtol=1e-6)
    assert not r["diag"][:f0].any() and not r["diag"][f1:].any() and not r["nii"][:f0].any() and not r["nii"][f1:].any()
    # no diagonal at all
    r2 = run_fused(bed, n, ldz=ldz, want_diag=False)
    assert np.array_equal(r2["Z"], r["Z"])


def test_fused_decode_long_run_flushes_partial_sums(monkeypatch):
    """more rows per CTA than one flush interval (64 table blocks of 48 rows) and 16-bit missing counters in use"""
    n, m = 130, 7000
    monkeypatch.setenv("GCTA_B200_DECODE_ROWS", "7008")   # one CTA walks all rows
    bed = synth.make_bed_payload(31, n, m, miss_rate=0.2)
    r = run_fused(bed, n)
    z, ok, _ = oracle.standardise(oracle.decode_x(bed, n))
    assert np.array_equal(r["Z"][:m, :n], z16_expected(z, abi.FP16))
    d_ref, n_ref = oracle_diag(z, ok)
    assert np.array_equal(r["nii"], n_ref) and np.allclose(r["diag"], d_ref, rtol=1e-6)


@pytest.mark.parametrize("n,m,miss,want_diag", [(203, 300, 0.1, True), (5003, 259, 0.3, True), (4100, 130, 0.02, False)])
def test_fused_decode_indicator_plane(n, m, miss, want_diag):
    """the E4M3 non-missing indicators written in the same pass: 0x38 (1.0) / 0 per genotype, zero padding to a multiple
    of 128 rows and behind n; Z and the diagonal unchanged by asking for them"""
    bed = synth.make_bed_payload(41 + n, n, m, miss_rate=miss)
    r = run_fused(bed, n, want_ok=True, want_diag=want_diag)
    r0 = run_fused(bed, n, want_ok=False, want_diag=want_diag)
    z, ok, _ = oracle.standardise(oracle.decode_x(bed, n))
    assert np.array_equal(r["OK"][:m, :n], np.where(ok, 0x38, 0).astype(np.uint8))
    assert not r["OK"][m:, :].any() and not r["OK"][:, n:].any()
    assert np.array_equal(r["Z"], r0["Z"]) and np.array_equal(r["nii"], r0["nii"])
    assert np.allclose(r["diag"], r0["diag"], rtol=1e-12)


@pytest.mark.parametrize("n_keep", [611, 985, 1003])
def test_compact_rows_kernel(n_keep):
    """K1c: the packed rows restricted to a keep list equal numpy's re-pack of the kept columns (with an include list,
    a chunk in the middle, n not a multiple of 16).  611 of 1003 kept: many short runs of file neighbours per thread (the
    per-individual walk); 985: a few long runs (the funnel-shift path); 1003: everyone, one run"""
    n_file, m_file = 1003, 300
    bed = synth.make_bed_payload(77, n_file, m_file, miss_rate=0.1)
    rng = np.random.default_rng(3)
    keep = np.sort(rng.choice(n_file, n_keep, replace=False)).astype(np.int32)
    inc = np.sort(rng.choice(m_file, 200, replace=False)).astype(np.int32)
    bedp, pitch = pitched(bed, n_file)
    n = len(keep)
    pitch_out = ((n + 3) // 4 + 15) // 16 * 16
    k0, kc = 32, 150
    out = torch.full((kc, pitch_out), 0xAB, dtype=torch.uint8, device="cuda")
    d_bed, d_keep, d_inc = dev(bedp), dev(keep), dev(inc)
    abi.call("gcta_b200_compact_rows", P(d_bed.data_ptr()), I64(pitch), P(d_inc.data_ptr()), I64(k0), I64(kc), P(d_keep.data_ptr()), I64(n),
             P(out.data_ptr()), I64(pitch_out), stream())
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    x = oracle.decode_x(bed, n_file, keep, inc)[k0:k0 + kc]                 # copies of allele 1, -1 = missing
    codes = np.array([3, 2, 0, 1], dtype=np.uint8)[x.astype(np.int64)]      # 0 -> 11, 1 -> 10b, 2 -> 00, missing -> 01b
    pad = np.zeros((kc, pitch_out * 4), dtype=np.uint8)
    pad[:, :n] = codes
    pad = pad.reshape(kc, pitch_out, 4)
    want = pad[:, :, 0] | (pad[:, :, 1] << 2) | (pad[:, :, 2] << 4) | (pad[:, :, 3] << 6)
    nw = (n + 15) // 16 * 4          # bytes of the words that hold individuals; behind them the row is not written
    assert np.array_equal(got[:, :nw], want[:, :nw])
