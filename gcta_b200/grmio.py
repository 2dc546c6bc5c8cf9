"""Readers for the reference's GRM output formats (harness only).

.grm.gz  : gzip text ``i<TAB>j<TAB>N<TAB>A`` 1-based, j <= i, row-major lower triangle
           (grm.cpp:321-331)
.grm.bin : n*n float64 row-major symmetric, no header, no N (grm.cpp:309-317)
.grm.id  : ``FID<TAB>IID`` per kept individual (grm.cpp:335-339)
"""
from __future__ import annotations

import gzip

import numpy as np


def read_grm_gz(path: str):
    """-> (A_lower f64, N_lower i32, raw A strings) in file order; verifies the index columns."""
    with gzip.open(path, "rt") as f:
        rows = [ln.split("\t") for ln in f.read().splitlines()]
    n_lines = len(rows)
    n = int((np.sqrt(8 * n_lines + 1) - 1) / 2 + 0.5)
    assert n * (n + 1) // 2 == n_lines, f"{path}: {n_lines} lines is not a triangle number"
    ii, jj = np.tril_indices(n)
    fi = np.array([int(r[0]) for r in rows])
    fj = np.array([int(r[1]) for r in rows])
    assert (fi == ii + 1).all() and (fj == jj + 1).all(), "index columns out of order"
    N = np.array([int(r[2]) for r in rows], dtype=np.int32)
    A = np.array([float(r[3]) for r in rows], dtype=np.float64)
    return A, N, [r[3] for r in rows]


def read_grm_bin(path: str) -> np.ndarray:
    a = np.fromfile(path, dtype=np.float64)
    n = int(np.sqrt(a.size) + 0.5)
    assert n * n == a.size
    return a.reshape(n, n)


def read_grm_id(path: str):
    with open(path) as f:
        return [tuple(ln.rstrip("\n").split("\t")) for ln in f if ln.strip()]
