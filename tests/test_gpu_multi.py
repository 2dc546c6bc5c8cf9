"""gcta_b200_make_grm_multi (GPU): the one call per process on several devices - row bands, 1/G uploads, peer pushes.
On a box with one GPU the call degrades to gcta_b200_make_grm (still checked); the band logic proper needs two."""
import numpy as np
import pytest

from gcta_b200 import abi, synth
from oracle import oracle
from tests.helpers import check_grm, pack_lower

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,m,miss,chunk", [(900, 9000, 0.02, 1024), (2100, 20000, 0.0, 0), (700, 3000, 0.05, 512)])
def test_multi_gpu_call_equals_single_gpu_call(n, m, miss, chunk):
    bed = synth.make_bed_payload(17 + n, n, m, miss_rate=miss)
    ndev = abi.lib().gcta_b200_device_count()
    G = min(ndev, 4)
    A0, N0, mu0, o0 = abi.make_grm(bed, n, m, chunk_snps=chunk)
    A1, N1, mu1, o1, per = abi.make_grm_multi(bed, n, m, max(G, 2), chunk_snps=chunk)   # asks for 2 even on a 1-GPU box
    assert np.array_equal(N0, N1) and np.array_equal(mu0, mu1)
    assert np.abs(A0.astype(np.float64) - A1).max() <= 1e-5
    Aref, Nref, _ = oracle.make_grm_numpy(bed, n)
    check_grm(A1, N1, pack_lower(Aref), pack_lower(Nref))
    assert o1.row_begin == 0 and o1.row_end == n and o1.launches > 0
    if G >= 2:
        bands = [(p.row_begin, p.row_end) for p in per[:G]]
        assert bands[0][0] == 0 and bands[-1][1] == n and all(bands[i][1] == bands[i + 1][0] for i in range(G - 1))
        assert all(p.launches > 0 for p in per[:G] if p.row_end > p.row_begin)


def test_multi_gpu_call_with_keep_include_and_xchr():
    n, m = 640, 2600
    bed = synth.make_bed_payload(5, n, m, miss_rate=0.03)
    rng = np.random.default_rng(1)
    keep = np.sort(rng.choice(n, 500, replace=False)).astype(np.int32)
    inc = np.sort(rng.choice(m, 1800, replace=False)).astype(np.int32)
    sex = rng.integers(1, 3, size=n).astype(np.int32)
    for kw in (dict(keep=keep, include=inc), dict(sex=sex, xchr=1), dict(alg=1)):
        A1, N1, mu1, o1, per = abi.make_grm_multi(bed, n, m, 2, **kw)
        Aref, Nref, _ = oracle.make_grm_numpy(bed, n, **kw)
        check_grm(A1, N1, pack_lower(Aref), pack_lower(Nref))
