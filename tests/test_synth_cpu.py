"""Harness pieces that the GPU parity tests lean on, checked on the CPU: the related-individuals generator and the
SNP-blocked sub-block oracle."""
import numpy as np

from gcta_b200 import synth
from oracle import oracle


def test_related_generator_is_mendelian_and_reproducible():
    n_f, m = 120, 1500
    bed, n, pairs = synth.make_related_payload(4, n_f, m, n_dup=6, n_po=5, n_sib=4, miss_rate=0.0)
    bed2, _, _ = synth.make_related_payload(4, n_f, m, n_dup=6, n_po=5, n_sib=4, miss_rate=0.0)
    assert np.array_equal(bed, bed2) and n == n_f + 6 + 5 + 8
    x = oracle.decode_x(bed, n)                                  # copies of allele 1 per SNP x individual
    assert np.array_equal(oracle.decode_x(synth.make_bed_payload(4, n_f, m), n_f), x[:, :n_f])   # founders = the plain generator
    kinds = {}
    for i, j, k in pairs:
        kinds.setdefault(k, []).append((i, j))
    for i, j in kinds["dup"]:
        assert np.array_equal(x[:, i], x[:, j])
    # a child carries one allele of each parent: impossible trios are those where both parents are the same homozygote
    # and the child differs, or the child is a homozygote for an allele one parent lacks
    po = {}
    for i, j in kinds["po"]:
        po.setdefault(i, []).append(j)
    for c, (pa, pb) in po.items():
        lo = (x[:, pa] == 2).astype(int) + (x[:, pb] == 2).astype(int)      # alleles that must be transmitted
        hi = (x[:, pa] >= 1).astype(int) + (x[:, pb] >= 1).astype(int)      # alleles that can be transmitted
        assert ((x[:, c] >= lo) & (x[:, c] <= hi)).all()
    assert len(kinds["sib"]) == 4 and all(i > j for i, j, _ in pairs)
    # the kinship shows in the oracle's GRM: duplicates ~1, parent-offspring and sibs ~0.5
    A, N, _ = oracle.make_grm_numpy(bed, n)
    assert min(A[i, j] for i, j in kinds["dup"]) > 0.8
    assert 0.3 < np.mean([A[i, j] for i, j in kinds["po"]]) < 0.7
    assert 0.2 < np.mean([A[i, j] for i, j in kinds["sib"]]) < 0.8


def test_related_generator_with_missing_genotypes():
    bed, n, pairs = synth.make_related_payload(9, 60, 800, n_dup=3, n_po=3, n_sib=3, miss_rate=0.05)
    x = oracle.decode_x(bed, n)
    frac = (x < 0).mean()
    assert 0.03 < frac < 0.07
    assert not (x[:, 60:63] != x[:, :3]).any()          # duplicates copy the founder's missing calls too


def test_blocked_sub_block_oracle_equals_the_whole_matrix():
    n, m = 240, 5000
    bed = synth.make_bed_payload(3, n, m, miss_rate=0.03)
    A, N, mu = oracle.make_grm_numpy(bed, n)
    sub = np.array(sorted(np.random.default_rng(1).choice(n, 37, replace=False)))
    for block in (777, 5000, 8192):
        A2, N2, mu2 = oracle.make_grm_numpy_blocked(bed, n, keep=sub, mu=mu, block=block)
        assert np.array_equal(N2, N[np.ix_(sub, sub)]) and np.array_equal(mu2, mu)
        assert np.abs(A2 - A[np.ix_(sub, sub)]).max() < 1e-12
    # without the full-sample 2p the sub-block is standardised with its own frequencies: a different (valid) GRM
    A3, N3, mu3 = oracle.make_grm_numpy_blocked(bed, n, keep=sub)
    Asub, Nsub, musub = oracle.make_grm_numpy(bed, n, keep=sub)
    assert np.array_equal(N3, Nsub) and np.abs(A3 - Asub).max() < 1e-12 and np.abs(A3 - A2).max() > 1e-3


def test_sparse_keep_decode_equals_the_dense_path():
    n, m = 401, 300
    bed = synth.make_bed_payload(8, n, m, miss_rate=0.1)
    for keep in (np.array([0, 7, 8, 9, 255, 256, 400]), np.arange(0, n, 5), np.arange(n)):
        assert np.array_equal(oracle.decode_x(bed, n, keep), oracle.decode_x(bed, n)[:, keep])
