"""CPU-side checks of the C-ABI library: it loads, exports every symbol the header declares,
its host-only helpers (row-band partition, tile list) are right, and compute entry points fail
loudly - with the reference's error-string convention - when there is no CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from gcta_b200 import abi, synth
from tests.helpers import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "gcta_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gcta_b200_[a-z0-9_]+)\s*\(", src)) - {"gcta_b200_s_tile_index"})


def test_library_exports_every_declared_symbol():
    L = abi.lib()
    names = _declared()
    assert len(names) >= 18
    for nme in names:
        assert hasattr(L, nme), f"{nme} declared in include/gcta_b200.h but not exported"
    assert sorted(abi.EXPORTS) == [x for x in names if x in abi.EXPORTS]
    assert b"sm_100a" in L.gcta_b200_version()


@pytest.mark.parametrize("n", [1, 127, 128, 129, 256, 257, 1000, 5000])
def test_tile_list_covers_lower_triangle_once(n):
    t = abi.tile_list(n)
    assert len({tuple(x) for x in t}) == len(t)
    n_pad = (n + 255) // 256 * 256
    cov = np.zeros((n_pad, n_pad), dtype=np.int32)
    for bi, bj in t:
        assert bi * 256 < n and bj * 256 < n
        cov[bi * 256:(bi + 1) * 256, bj * 256:(bj + 1) * 256] += 1
    ii, jj = np.tril_indices(n)
    assert (cov[ii, jj] == 1).all()
    # no tile lies entirely above the diagonal
    for bi, bj in t:
        assert min(bi * 256 + 255, n - 1) >= bj * 256


@pytest.mark.parametrize("n,world", [(50000, 8), (10000, 4), (2000, 2), (300, 8), (100, 3)])
def test_partition_bands_are_contiguous_and_balanced(n, world):
    bands = [abi.partition(n, r, world) for r in range(world)]
    assert bands[0][0] == 0 and bands[-1][1] == n
    for (b0, e0), (b1, e1) in zip(bands, bands[1:]):
        assert e0 == b1 and b0 <= e0
    counts = [len(abi.tile_list(n, b, e)) for b, e in bands if e > b]
    assert sum(counts) == len(abi.tile_list(n))
    if n >= 10000:
        assert max(counts) <= 1.05 * (sum(counts) / world) + 64
    # all boundaries except the last are multiples of the tile height
    assert all(b % 256 == 0 for b, e in bands if e > b)


def test_partition_counts_whole_waves_on_the_headline_workload():
    """K2 takes whole waves of 148 tiles per launch: with 8 GPUs on n = 50,000 no band may need more than 17 waves
    (19,306 tiles cannot be done in fewer: 8 x 16 waves hold only 18,944)."""
    bands = [abi.partition(50000, r, 8) for r in range(8)]
    waves = [-(-len(abi.tile_list(50000, b, e)) // 148) for b, e in bands]
    assert max(waves) == 17 and sum(w * 148 for w in waves) >= 19306
    # later bands standardise more columns, so they never get more tensor work than earlier ones
    assert waves == sorted(waves, reverse=True)


@pytest.mark.parametrize("m,world", [(1000000, 8), (13000, 2), (31, 4), (1, 3), (64, 64), (100003, 7)])
def test_exchange_slices_cover_every_file_row_once(m, world):
    """gcta_b200_exchange_slice: per segment the ranks' slices are disjoint, ascending and together the whole segment;
    the segments follow each other and end at the last row - what exchange_post uploads and exchange_pull fetches."""
    per_rank = [abi.exchange_slices(m, world, r) for r in range(world)]
    nseg = len(per_rank[0])
    assert 1 <= nseg <= 32 and all(len(p) == nseg for p in per_rank)
    pos = 0
    for s in range(nseg):
        for r in range(world):
            b, e = per_rank[r][s]
            assert b == pos and e >= b
            pos = e
    assert pos == m
    sizes = [sum(e - b for b, e in p) for p in per_rank]
    assert sum(sizes) == m and (m < 32 * world or max(sizes) - min(sizes) <= 32 + m // (world * 16))


def test_header_is_plain_c_and_links_from_c(tmp_path):
    """The boundary is a C ABI: the header must compile as strict C99, and a C program must link against the library
    and call it (device-free entry points only: there is no GPU here)."""
    import subprocess
    hdr = os.path.join(ROOT, "include", "gcta_b200.h")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c", hdr], check=True)
    src = tmp_path / "t.c"
    src.write_text('#include <stdio.h>\n#include "gcta_b200.h"\n'
                   "int main(void) {\n"
                   "    int64_t b = -1, e = -1, r0 = -1, r1 = -1;\n"
                   "    if (gcta_b200_partition(50000, 1, 2, &b, &e)) return 1;\n"
                   "    int64_t nseg = gcta_b200_exchange_slice(1000000, 8, 3, 5, &r0, &r1);\n"
                   '    printf("%s|%lld|%lld|%lld|%lld|%lld|%lld\\n", gcta_b200_version(), (long long)b, (long long)e, (long long)nseg,\n'
                   "           (long long)r0, (long long)r1, (long long)gcta_b200_s_tile_index(5, 7));\n"
                   "    return 0;\n}\n")
    exe = tmp_path / "t"
    libdir = os.path.join(ROOT, "gcta_b200")
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe), "-L", libdir, "-lgcta_b200",
                    "-Wl,-rpath," + libdir], check=True)
    out = subprocess.run([str(exe)], check=True, stdout=subprocess.PIPE, text=True).stdout.strip().split("|")
    assert out[0].startswith("gcta_b200 ")
    assert (int(out[1]), int(out[2])) == abi.partition(50000, 1, 2)
    assert int(out[3]) == 32 and (int(out[4]), int(out[5])) == abi.exchange_slices(1000000, 8, 3)[5]
    assert int(out[6]) == ((7 >> 2) * 256 + 5) * 4 + (7 & 3)


def test_no_cpu_fallback():
    if abi.lib().gcta_b200_device_count() > 0:
        pytest.skip("a CUDA device is present")
    bed = synth.make_bed_payload(1, 40, 50)
    with pytest.raises(abi.GctaB200Error) as e:
        abi.make_grm(bed, 40, 50)
    assert e.value.code == 2 and "no CUDA device" in str(e.value)


def test_argument_errors_use_reference_messages():
    bed = synth.make_bed_payload(1, 40, 50)
    with pytest.raises(abi.GctaB200Error) as e:
        abi.make_grm(bed, 40, 50, include=np.array([], dtype=np.int32))
    assert str(e.value) == "Error: No SNP is retained for analysis."  # data.cpp:149
