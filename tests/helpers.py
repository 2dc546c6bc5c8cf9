"""Shared helpers for the parity tests (harness only)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_plink(prefix):
    """-> (bed payload uint8 [m, ceil(n/4)], n, m, chrom int[m], sex int[n], snp names, ids)."""
    fam = [ln.split() for ln in open(prefix + ".fam") if ln.strip()]
    bim = [ln.split() for ln in open(prefix + ".bim") if ln.strip()]
    n, m = len(fam), len(bim)
    raw = np.fromfile(prefix + ".bed", dtype=np.uint8)
    bed = raw[3:].reshape(m, (n + 3) // 4)
    chrom = np.array([int(b[0]) for b in bim], dtype=np.int32)
    sex = np.array([int(f[4]) for f in fam], dtype=np.int32)
    return bed, n, m, chrom, sex, [b[1] for b in bim], [(f[0], f[1]) for f in fam]


def pack_lower(M):
    i, j = np.tril_indices(M.shape[0])
    return M[i, j]


def check_grm(A, N, A_ref, N_ref, atol=1e-3, diag_tol=1e-5):
    """North-star tolerances: N exact; max|dA| <= 1e-3; |mean diag delta| <= 1e-5."""
    n = int((np.sqrt(8 * len(A_ref) + 1) - 1) / 2 + 0.5)
    assert np.array_equal(np.asarray(N), np.asarray(N_ref)), "N differs from the reference"
    dA = np.abs(np.asarray(A, dtype=np.float64) - A_ref)
    assert dA.max() <= atol, f"max|dA| = {dA.max()}"
    di = np.arange(n) * (np.arange(n) + 1) // 2 + np.arange(n)
    dd = abs(np.asarray(A, dtype=np.float64)[di].mean() - A_ref[di].mean())
    assert dd <= diag_tol, f"diag mean delta = {dd}"
    return dA.max(), dd
