"""Seeded synthetic PLINK .bed/.bim/.fam generator (harness only, no GPU work).

The genotype of individual ``i`` at SNP ``k`` is a pure function of ``(seed, k, i)`` through a
counter-based 64-bit mix, so the same bytes can be produced in SNP blocks on the host (numpy,
here) or on the device (``gcta_b200_synth_bed`` in csrc/synth.cu) without ever holding an
M x n matrix.  Distribution: per SNP ``p ~ U(maf_lo, maf_hi)``, genotype =
Binomial(2, p) copies of the .bim allele 1 under HWE, independent SNPs and individuals,
missing i.i.d. Bernoulli(miss_rate).

File format pinned by the reference: magic ``6C 1B 01`` then M rows of ceil(n/4) bytes,
2 bits per individual LSB first (raw_geno.cpp:142-170, data.cpp:164-185); codes
``00`` = 2 copies of A1, ``10b`` = 1 copy, ``11`` = 0 copies, ``01b`` = missing
(data.cpp:176-181 with the missing test at data.cpp:792).
"""
from __future__ import annotations

import os

import numpy as np

_C1 = np.uint64(0x9E3779B97F4A7C15)
_C2 = np.uint64(0xC2B2AE3D27D4EB4F)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
BED_MAGIC = bytes([0x6C, 0x1B, 0x01])


def mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on uint64 arrays (wrapping arithmetic)."""
    with np.errstate(over="ignore"):
        x = x ^ (x >> np.uint64(30))
        x = x * _M1
        x = x ^ (x >> np.uint64(27))
        x = x * _M2
        x = x ^ (x >> np.uint64(31))
    return x


def snp_thresholds(seed: int, m: int, maf_lo: float = 0.01, maf_hi: float = 0.5,
                   miss_rate: float = 0.0, k0: int = 0) -> np.ndarray:
    """uint32 [m, 4] = (T_hom2, T_het_cum, T_miss, 0) for SNPs k0 .. k0+m-1.

    genotype draw g (low 32 hash bits): g < T_hom2 -> 2 copies; g < T_het_cum -> 1 copy; else 0.
    missing draw (high 32 hash bits) < T_miss -> missing.
    """
    k = np.arange(k0, k0 + m, dtype=np.uint64)
    with np.errstate(over="ignore"):
        h = mix64(np.uint64(seed) ^ (k * _C1 + np.uint64(0x5851F42D4C957F2D)))
    u = (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
    p = maf_lo + (maf_hi - maf_lo) * u
    two32 = 4294967296.0
    t2 = np.minimum(np.floor(p * p * two32), two32 - 1)
    t1 = np.minimum(np.floor((p * p + 2 * p * (1 - p)) * two32), two32 - 1)
    tm = np.full(m, min(np.floor(miss_rate * two32), two32 - 1))
    out = np.zeros((m, 4), dtype=np.uint32)
    out[:, 0] = t2.astype(np.uint32)
    out[:, 1] = t1.astype(np.uint32)
    out[:, 2] = tm.astype(np.uint32)
    return out


def bed_rows(seed: int, n: int, k0: int, m: int, thr: np.ndarray) -> np.ndarray:
    """Packed rows uint8 [m, ceil(n/4)] for SNPs k0..k0+m-1 (pad bits 0)."""
    stride = (n + 3) // 4
    k = np.arange(k0, k0 + m, dtype=np.uint64)[:, None]
    i = np.arange(stride * 4, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        h = mix64(np.uint64(seed) + k * _C1 + i * _C2)
    g = (h & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    r = (h >> np.uint64(32)).astype(np.uint32)
    code = np.full(h.shape, 3, dtype=np.uint8)            # 0 copies -> 11
    code[g < thr[:, 1:2]] = 2                             # 1 copy   -> 10b
    code[g < thr[:, 0:1]] = 0                             # 2 copies -> 00
    code[r < thr[:, 2:3]] = 1                             # missing  -> 01b
    code[:, n:] = 0                                       # pad bits
    code = code.reshape(m, stride, 4)
    return (code[:, :, 0] | (code[:, :, 1] << 1 << 1) | (code[:, :, 2] << 4) | (code[:, :, 3] << 6)).astype(np.uint8)


def make_bed_payload(seed: int, n: int, m: int, miss_rate: float = 0.0, maf_lo: float = 0.01,
                     maf_hi: float = 0.5, block: int = 4096) -> np.ndarray:
    """Whole payload (without the 3 magic bytes) as uint8 [m, ceil(n/4)]."""
    stride = (n + 3) // 4
    out = np.empty((m, stride), dtype=np.uint8)
    for k0 in range(0, m, block):
        mm = min(block, m - k0)
        thr = snp_thresholds(seed, mm, maf_lo, maf_hi, miss_rate, k0)
        out[k0:k0 + mm] = bed_rows(seed, n, k0, mm, thr)
    return out


def write_plink(prefix: str, payload: np.ndarray, n: int, chrom=None, sex=None, fid=None, iid=None) -> None:
    """Write prefix.bed/.bim/.fam.  .fam: ``F<i> I<i> 0 0 <sex> -9``; .bim: chr 1+(k mod 22),
    ``rs<k>``, 0 cM, bp k+1, alleles A G (single-character alleles, data.cpp:104-118)."""
    m = payload.shape[0]
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    with open(prefix + ".bed", "wb") as f:
        f.write(BED_MAGIC)
        f.write(np.ascontiguousarray(payload).tobytes())
    with open(prefix + ".fam", "w") as f:
        for i in range(n):
            s = 1 if sex is None else int(sex[i])
            f.write(f"{fid[i] if fid is not None else 'F%d' % i} {iid[i] if iid is not None else 'I%d' % i} 0 0 {s} -9\n")
    with open(prefix + ".bim", "w") as f:
        for k in range(m):
            c = 1 + (k % 22) if chrom is None else int(chrom[k])
            f.write(f"{c}\trs{k}\t0\t{k + 1}\tA\tG\n")


def _unpack_copies(rows: np.ndarray, n: int) -> np.ndarray:
    """packed rows -> int8 [m, n] copies of allele 1 (0/1/2), -1 = missing"""
    codes = np.stack([(rows >> s) & 3 for s in (0, 2, 4, 6)], axis=-1).reshape(rows.shape[0], -1)[:, :n]
    return np.array([2, -1, 1, 0], dtype=np.int8)[codes]


def _pack_copies(x: np.ndarray) -> np.ndarray:
    """int8 [m, n] copies (-1 = missing) -> packed rows uint8 [m, ceil(n/4)] (pad bits 0)"""
    m, n = x.shape
    stride = (n + 3) // 4
    code = np.zeros((m, stride * 4), dtype=np.uint8)
    code[:, :n] = np.array([3, 2, 0, 1], dtype=np.uint8)[x.astype(np.int64)]   # 0 -> 11, 1 -> 10b, 2 -> 00, -1 -> 01b
    code = code.reshape(m, stride, 4)
    return (code[:, :, 0] | (code[:, :, 1] << 2) | (code[:, :, 2] << 4) | (code[:, :, 3] << 6)).astype(np.uint8)


def make_related_payload(seed: int, n_founders: int, m: int, n_dup: int = 0, n_po: int = 0, n_sib: int = 0,
                         miss_rate: float = 0.0, block: int = 4096):
    """Founders as in make_bed_payload plus relatives appended behind them (a GRM's large off-diagonals):

      n_dup  exact duplicates of founders 0 .. n_dup-1                       (A_ij ~ 1)
      n_po   children of founder pairs (2i, 2i+1), Mendelian sampling        (parent-offspring, A_ij ~ 0.5)
      n_sib  sib pairs: two children each of founder pairs (2i, 2i+1) taken from the top of the founder list
             (full sibs, A_ij ~ 0.5)

    Duplicates copy the founder's missing calls; the children's are drawn independently (seeded).  Returns (payload uint8 [m, ceil(n/4)], n,
    pairs) with pairs = list of (i, j, kind), i > j, the related pairs in individual indices."""
    n = n_founders + n_dup + n_po + 2 * n_sib
    assert n_dup <= n_founders and 2 * n_po <= n_founders and 2 * n_sib <= n_founders
    stride = (n + 3) // 4
    out = np.empty((m, stride), dtype=np.uint8)
    rng = np.random.default_rng(seed ^ 0x5EED)
    for k0 in range(0, m, block):
        mm = min(block, m - k0)
        thr = snp_thresholds(seed, mm, miss_rate=miss_rate, k0=k0)
        f = _unpack_copies(bed_rows(seed, n_founders, k0, mm, thr), n_founders)   # founders, -1 = missing
        g = np.where(f < 0, 0, f)   # a missing parental genotype transmits like a 0 (only matters when miss_rate > 0)

        def child(a, b):
            ta = (rng.random((mm, len(a))) * 2 < g[:, a]).astype(np.int8)   # allele 1 transmitted w.p. copies / 2
            tb = (rng.random((mm, len(b))) * 2 < g[:, b]).astype(np.int8)
            c = ta + tb
            if miss_rate > 0:
                c[rng.random(c.shape) < miss_rate] = -1
            return c

        cols = [f]
        if n_dup:
            cols.append(f[:, :n_dup])
        if n_po:
            cols.append(child(np.arange(0, 2 * n_po, 2), np.arange(1, 2 * n_po, 2)))
        if n_sib:
            pa = n_founders - 2 * n_sib + np.arange(0, 2 * n_sib, 2)
            cols.append(child(pa, pa + 1))
            cols.append(child(pa, pa + 1))
        out[k0:k0 + mm] = _pack_copies(np.concatenate(cols, axis=1))
    pairs = []
    base = n_founders
    pairs += [(base + i, i, "dup") for i in range(n_dup)]
    base += n_dup
    for i in range(n_po):
        pairs += [(base + i, 2 * i, "po"), (base + i, 2 * i + 1, "po")]
    base += n_po
    pairs += [(base + n_sib + i, base + i, "sib") for i in range(n_sib)]
    return out, n, pairs
