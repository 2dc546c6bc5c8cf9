"""ctypes front for oracle/grm_oracle.c plus a numpy float64 restatement for larger cases.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product path.  Parity status: PINNED
against the compiled, unmodified reference (oracle/_ref/gcta_ref) - see
tests/test_oracle_pin.py and tests/golden/.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgrm_oracle.so")
REF_BIN = os.path.join(HERE, "_ref", "gcta_ref")
_lib = None


def build(ref: bool = True) -> None:
    """Compile the C restatement, and the reference binary when /root/reference is present."""
    subprocess.run(["make", "-s", "-C", HERE, "port"], check=True)
    if ref and os.path.isdir(os.environ.get("GCTA_REFERENCE_DIR", "/root/reference")):
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(os.path.join(HERE, "grm_oracle.c")):
            build(ref=False)
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.grm_oracle_make_grm.restype = ctypes.c_int
        _lib.grm_oracle_mu.restype = ctypes.c_int
    return _lib


def _p(a, ct):
    return None if a is None else a.ctypes.data_as(ctypes.POINTER(ct))


def make_grm(bed: np.ndarray, n_file: int, keep=None, include=None, sex=None, mu=None,
             xchr: int = 0, alg: int = 0):
    """bed: uint8 [m_file, ceil(n_file/4)].  Returns (A_lower f64, N_lower i32, mu f64).

    Follows grm.cpp:219-301 exactly (O(n^2 M), scalar) - keep n*n*M below ~1e9."""
    bed = np.ascontiguousarray(bed, dtype=np.uint8)
    keep = None if keep is None else np.ascontiguousarray(keep, dtype=np.int32)
    include = None if include is None else np.ascontiguousarray(include, dtype=np.int32)
    sex = None if sex is None else np.ascontiguousarray(sex, dtype=np.int32)
    mu = None if mu is None else np.ascontiguousarray(mu, dtype=np.float64)
    n = n_file if keep is None else len(keep)
    m = bed.shape[0] if include is None else len(include)
    A = np.empty(n * (n + 1) // 2, dtype=np.float64)
    N = np.empty(n * (n + 1) // 2, dtype=np.int32)
    mu_out = np.empty(m, dtype=np.float64)
    rc = lib().grm_oracle_make_grm(
        _p(bed, ctypes.c_uint8), ctypes.c_int64(n_file), _p(keep, ctypes.c_int32), ctypes.c_int64(n),
        _p(include, ctypes.c_int32), ctypes.c_int64(m), _p(sex, ctypes.c_int32), _p(mu, ctypes.c_double),
        ctypes.c_int(xchr), ctypes.c_int(alg), _p(A, ctypes.c_double), _p(N, ctypes.c_int32),
        _p(mu_out, ctypes.c_double))
    if rc != 0:
        raise MemoryError("grm_oracle_make_grm failed")
    return A, N, mu_out


def snp_mu(bed: np.ndarray, n_file: int, keep=None, include=None, sex=None, xchr: int = 0):
    """(mu f64, cnt i32, sum i32) per included SNP - calcu_mu, data.cpp:759-803."""
    bed = np.ascontiguousarray(bed, dtype=np.uint8)
    keep = None if keep is None else np.ascontiguousarray(keep, dtype=np.int32)
    include = None if include is None else np.ascontiguousarray(include, dtype=np.int32)
    sex = None if sex is None else np.ascontiguousarray(sex, dtype=np.int32)
    n = n_file if keep is None else len(keep)
    m = bed.shape[0] if include is None else len(include)
    mu = np.empty(m, dtype=np.float64)
    cnt = np.empty(m, dtype=np.int32)
    s = np.empty(m, dtype=np.int32)
    lib().grm_oracle_mu(_p(bed, ctypes.c_uint8), ctypes.c_int64(n_file), _p(keep, ctypes.c_int32),
                        ctypes.c_int64(n), _p(include, ctypes.c_int32), ctypes.c_int64(m),
                        _p(sex, ctypes.c_int32), ctypes.c_int(xchr), _p(mu, ctypes.c_double),
                        _p(cnt, ctypes.c_int32), _p(s, ctypes.c_int32))
    return mu, cnt, s


def z_lut(mu: float, divide: bool = True) -> np.ndarray:
    z = (ctypes.c_float * 3)()
    lib().grm_oracle_z_lut(ctypes.c_double(mu), ctypes.c_int(int(divide)), z)
    return np.array(list(z), dtype=np.float32)


# ---------------------------------------------------------------------------------------------
# numpy restatement (float64 BLAS contraction) for sizes the scalar loop cannot finish.
# Same arithmetic up to summation order (|dA| ~ 1e-13); N exact.
# ---------------------------------------------------------------------------------------------
def decode_x(bed: np.ndarray, n_file: int, keep=None, include=None) -> np.ndarray:
    """int8 [m, n]: copies of allele 1 (0/1/2), -1 = missing.  data.cpp:176-181, 1001-1006."""
    rows = bed if include is None else bed[np.asarray(include)]
    if keep is not None and len(keep) * 4 < n_file:
        # few individuals of many: fetch only their bytes (sub-block checks at the BASELINE shapes)
        kp = np.asarray(keep, dtype=np.int64)
        codes = (rows[:, kp >> 2] >> ((kp & 3) * 2).astype(np.uint8)[None, :]) & 3
    else:
        codes = np.stack([(rows >> s) & 3 for s in (0, 2, 4, 6)], axis=-1).reshape(rows.shape[0], -1)[:, :n_file]
        if keep is not None:
            codes = codes[:, np.asarray(keep)]
    lut = np.array([2, -1, 1, 0], dtype=np.int8)  # code 00->2, 01b->missing, 10b->1, 11->0
    return lut[codes]


def standardise(x: np.ndarray, mu=None, divide: bool = True, sex=None, xchr: int = 0):
    """x int8 [m, n] -> (z f32 [m, n] with 0 at missing, ok bool [m, n], mu f64 [m]).
    data.cpp:759-803 (mu) and 1016-1056 (two float roundings, Appendix A step 3)."""
    ok = x >= 0
    xv = np.where(ok, x, 0).astype(np.float64)
    if mu is None:
        if xchr:
            fac = np.where(np.asarray(sex) == 1, 0.5, np.where(np.asarray(sex) == 2, 1.0, 0.0))[None, :]
        else:
            fac = np.ones((1, x.shape[1]))
        num = (xv * fac * ok).sum(1)
        den = (fac * ok).sum(1)
        mu = np.where(den > 0, num / np.where(den > 0, den, 1), 0.0)
    v = mu * (1.0 - 0.5 * mu)
    with np.errstate(divide="ignore"):
        isd = np.where(np.abs(v) < 1e-50, 0.0, np.sqrt(1.0 / np.where(np.abs(v) < 1e-50, 1.0, v)))
    z = (x.astype(np.float32).astype(np.float64) - mu[:, None]).astype(np.float32)
    if divide:
        z = (z.astype(np.float64) * isd[:, None]).astype(np.float32)
    if xchr:
        male = (np.asarray(sex) == 1)[None, :]
        z = np.where(male, (z.astype(np.float64) * np.sqrt(0.5)).astype(np.float32), z)
    z = np.where(ok, z, np.float32(0)).astype(np.float32)
    return z, ok, mu


def make_grm_numpy(bed: np.ndarray, n_file: int, keep=None, include=None, mu=None, sex=None,
                   xchr: int = 0, alg: int = 0):
    """Full n x n (A f64, N i32, mu) via float64 GEMMs; grm.cpp:249-298 up to summation order."""
    x = decode_x(bed, n_file, keep, include)
    if sex is not None and keep is not None:
        sex = np.asarray(sex)[np.asarray(keep)]
    z, ok, mu = standardise(x, mu, divide=(alg == 0), sex=sex, xchr=xchr)
    z64 = z.astype(np.float64)
    okf = ok.astype(np.float64)
    S = z64.T @ z64
    N = np.rint(okf.T @ okf).astype(np.int32)
    if alg == 1:
        cnt = ok.sum(1)
        d = np.where(cnt > 0, (z64 * z64).sum(1) / np.maximum(cnt, 1), 0.0)
        D = (okf * d[:, None]).T @ okf
    else:
        D = N.astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        A = np.where(N > 0, S / np.where(N > 0, D, 1.0), 0.0)
    dn = np.diag(N) == 0
    A[np.diag_indices_from(A)] = np.where(dn, 1.0, np.diag(A))
    return A, N, mu


def make_grm_numpy_blocked(bed: np.ndarray, n_file: int, keep=None, mu=None, block: int = 8192):
    """make_grm_numpy (alg 0, autosomes, all SNPs) accumulated over SNP blocks, for sub-blocks of individuals at SNP
    counts whose standardised matrix would not fit in memory at once.  With `keep` and the FULL-sample `mu` this is the
    sub-block identity of SURVEY.md 8c.  grm.cpp:249-298 up to summation order."""
    n = n_file if keep is None else len(keep)
    m = bed.shape[0]
    S = np.zeros((n, n))
    Nf = np.zeros((n, n))
    mu_all = np.empty(m)
    for k0 in range(0, m, block):
        k1 = min(m, k0 + block)
        x = decode_x(bed[k0:k1], n_file, keep)
        z, ok, mu_b = standardise(x, None if mu is None else np.asarray(mu)[k0:k1])
        mu_all[k0:k1] = mu_b
        z64 = z.astype(np.float64)
        S += z64.T @ z64
        okf = ok.astype(np.float32)
        Nf += okf.T @ okf          # 0/1 products, block sums below 2^24: exact in float32
    N = np.rint(Nf).astype(np.int32)
    with np.errstate(divide="ignore", invalid="ignore"):
        A = np.where(N > 0, S / np.where(N > 0, N, 1), 0.0)
    dn = np.diag(N) == 0
    A[np.diag_indices_from(A)] = np.where(dn, 1.0, np.diag(A))
    return A, N, mu_all


def pack_lower(Mx: np.ndarray) -> np.ndarray:
    i, j = np.tril_indices(Mx.shape[0])
    return Mx[i, j]


# ---------------------------------------------------------------------------------------------
# running the compiled reference (this container, or anywhere oracle/_ref/gcta_ref travelled)
# ---------------------------------------------------------------------------------------------
def have_ref() -> bool:
    return os.access(REF_BIN, os.X_OK)


def run_ref(args, cwd=None, timeout=None) -> subprocess.CompletedProcess:
    return subprocess.run([REF_BIN] + [str(a) for a in args], cwd=cwd, timeout=timeout,
                          stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
