"""gcta_b200_diag_accum_static: the variant of the per-individual diagonal step for hosts that
cannot allocate device memory (all of HBM already belongs to their own pool, or their loop may
never enter the CUDA allocator because cudaMalloc/cudaFree stall the device).  Its per-SNP table
sits in a fixed 16 MiB array inside the library instead of a cudaMalloc'ed buffer.

Checked against the existing pair gcta_b200_bed_decode -> gcta_b200_diag_accum on two calls in a
row with different inputs (the array keeps the previous call's contents, which must not leak),
with the tolerance of tests/test_gpu_kernels.py: counts exact, float64 sums to 1e-9 relative.
Kept in its own file, collected last, so an -x run has exercised everything else first.
"""
import ctypes as C

import numpy as np
import pytest

from gcta_b200 import abi, synth
from oracle import oracle

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

P, I64, I32 = C.c_void_p, C.c_int64, C.c_int32


def _existing(bed_d, pitch, n, m, km_d, k0, kc):
    """diag / nii through the existing entry points (table in torch-allocated memory)."""
    ldz = (n + 255) // 256 * 256
    kc_pad = (kc + 63) // 64 * 64
    Z = torch.empty((kc_pad, ldz), dtype=torch.int16, device="cuda")
    mu = torch.zeros(m, dtype=torch.float64, device="cuda")
    cnt = torch.zeros(m, dtype=torch.int32, device="cuda")
    sm = torch.zeros(m, dtype=torch.int32, device="cuda")
    zlut = torch.zeros(m * 8, dtype=torch.float32, device="cuda")
    tot = torch.zeros(1, dtype=torch.int64, device="cuda")
    st = P(torch.cuda.current_stream().cuda_stream)
    abi.call("gcta_b200_bed_decode", P(bed_d.data_ptr()), I64(pitch), I64(n), P(km_d.data_ptr()), P(None), I64(n), P(None),
             P(None), I64(k0), I64(kc), P(None), I32(0), I32(0), I32(abi.FP16), P(Z.data_ptr()), I64(ldz), P(mu.data_ptr()),
             P(cnt.data_ptr()), P(sm.data_ptr()), P(zlut.data_ptr()), P(tot.data_ptr()), st)
    diag = torch.zeros(n, dtype=torch.float64, device="cuda")
    nii = torch.zeros(n, dtype=torch.int32, device="cuda")
    abi.call("gcta_b200_diag_accum", P(bed_d.data_ptr()), I64(pitch), I64(n), P(None), P(None), I64(k0), I64(kc),
             P(zlut.data_ptr()), P(diag.data_ptr()), P(nii.data_ptr()), st)
    torch.cuda.synchronize()
    return diag.cpu().numpy(), nii.cpu().numpy()


def _static(bed_d, pitch, n, km_d, k0, kc, inc_d=None):
    diag = torch.zeros(n, dtype=torch.float64, device="cuda")
    nii = torch.zeros(n, dtype=torch.int32, device="cuda")
    abi.call("gcta_b200_diag_accum_static", P(bed_d.data_ptr()), I64(pitch), I64(n), P(km_d.data_ptr()),
             P(None if inc_d is None else inc_d.data_ptr()), I64(k0),
             I64(kc), P(diag.data_ptr()), P(nii.data_ptr()), P(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return diag.cpu().numpy(), nii.cpu().numpy()


def _device_inputs(seed, n, m, miss, kept=None, want_host=False):
    bed = synth.make_bed_payload(seed, n, m, miss_rate=miss)
    stride = bed.shape[1]
    pitch = (stride + 15) // 16 * 16
    bp = np.zeros((m, pitch), dtype=np.uint8)
    bp[:, :stride] = bed
    km = np.zeros(pitch // 4, dtype=np.uint32)
    for f in (range(n) if kept is None else kept):
        km[f >> 4] |= np.uint32(1 << ((f & 15) * 2))
    out = (torch.from_numpy(bp).cuda(), pitch, torch.from_numpy(km).cuda())
    return out + (bed,) if want_host else out


def _oracle_diag(bed, n, k0, kc, kept=None, include=None):
    """float64 sum of the float32 squares and the non-missing counts, per FILE individual, for included SNPs
    [k0, k0+kc); 2p from the kept individuals only (what the keep mask selects)"""
    inc = None if include is None else np.asarray(include)[k0:k0 + kc]
    rows = bed[k0:k0 + kc] if include is None else bed[inc]
    mu, _, _ = oracle.snp_mu(rows, n, keep=None if kept is None else np.asarray(kept, dtype=np.int32))
    z, ok, _ = oracle.standardise(oracle.decode_x(rows, n), mu=mu)
    z64 = z.astype(np.float64)
    return (z64 * z64).sum(0), ok.sum(0).astype(np.int32)


def test_static_table_variant_matches_existing_entry_point_twice():
    # first call: larger range; second call: different data, shorter range starting at k0 > 0, so any
    # stale table entry from call 1 that leaked into call 2 would change the sums
    for seed, n, m, miss, k0, kc in [(301, 777, 900, 0.05, 0, 900), (302, 500, 640, 0.01, 128, 300)]:
        bed_d, pitch, km_d, bed = _device_inputs(seed, n, m, miss, want_host=True)
        d_ref, n_ref = _existing(bed_d, pitch, n, m, km_d, k0, kc)
        d_new, n_new = _static(bed_d, pitch, n, km_d, k0, kc)
        assert np.array_equal(n_new, n_ref)
        assert np.abs(d_new - d_ref).max() <= 1e-9 * max(1.0, d_ref.max())
        assert d_ref.max() > 0
        # and against the oracle directly (float32 partial sums over up to 768 SNPs: a few 1e-6 relative)
        d_or, n_or = _oracle_diag(bed, n, k0, kc)
        assert np.array_equal(n_new, n_or) and np.allclose(d_new, d_or, rtol=1e-5, atol=1e-9)


def test_static_table_variant_with_a_keep_mask_subset_and_an_include_list():
    """2p comes from the kept individuals only; the SNP rows are reached through d_include"""
    n, m_file = 333, 500
    rng = np.random.default_rng(7)
    kept = np.sort(rng.choice(n, 200, replace=False))
    inc = np.sort(rng.choice(m_file, 320, replace=False)).astype(np.int32)
    bed_d, pitch, km_d, bed = _device_inputs(303, n, m_file, 0.04, kept=kept, want_host=True)
    inc_d = torch.from_numpy(inc).cuda()
    for k0, kc in [(0, 320), (64, 100)]:
        d_new, n_new = _static(bed_d, pitch, n, km_d, k0, kc, inc_d)
        d_or, n_or = _oracle_diag(bed, n, k0, kc, kept=kept, include=inc)
        assert np.array_equal(n_new, n_or) and np.allclose(d_new, d_or, rtol=1e-5, atol=1e-9)


def test_static_table_variant_rejects_ranges_beyond_its_capacity():
    bed_d, pitch, km_d = _device_inputs(5, 64, 8, 0.0)
    with pytest.raises(abi.GctaB200Error) as e:
        _static(bed_d, pitch, 64, km_d, 0, 524288 + 1)
    assert e.value.code == 1 and "524288" in str(e.value)
