#!/usr/bin/env python
"""Generate the golden fixtures in tests/golden/ by running the UNMODIFIED reference.

Run in the build container (needs oracle/_ref/gcta_ref, built by `make -C oracle ref` from
/root/reference).  The reference ships no tests or fixtures of its own, so these
outputs of the compiled reference are the pin for the oracle and for the CUDA path.  Inputs
are produced by gcta_b200.synth (seeded) plus the hand-made edge cases below; the small
.bed/.bim/.fam inputs are committed next to the reference outputs so the tests do not depend
on the generator staying unchanged.

    python tests/golden/make_golden.py
"""
import os
import shutil
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gcta_b200 import synth  # noqa: E402
from oracle import oracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def run(case, args):
    d = os.path.join(OUT, case)
    r = oracle.run_ref(args, cwd=d)
    with open(os.path.join(d, "ref_stdout_%s.txt" % args[-1]), "w") as f:
        # drop the progress lines ("\r"-separated) and the time stamps
        keep = [ln for ln in r.stdout.replace("\r", "\n").split("\n")
                if "individuals." not in ln or "Genotype" in ln or "GRM of" in ln]
        keep = [ln for ln in keep if not ln.startswith(("Analysis started", "Analysis finished", "Computational time"))]
        f.write("\n".join(keep))
        if r.stderr:
            f.write("\n[stderr]\n" + r.stderr)
    return r


def fresh(case):
    d = os.path.join(OUT, case)
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    return d


def edge_payload(n=37, m=300, seed=77):
    bed = synth.make_bed_payload(seed, n, m, miss_rate=0.05)
    rng = np.random.default_rng(seed)

    def set_code(k, i, c):
        b, s = i // 4, (i % 4) * 2
        bed[k, b] = (int(bed[k, b]) & ~(3 << s) & 0xFF) | (c << s)

    for i in range(n):
        set_code(5, i, 0)       # monomorphic: everyone 2 copies
        set_code(6, i, 3)       # monomorphic: everyone 0 copies
        set_code(7, i, 1)       # SNP missing in everyone
    for i in range(n):          # monomorphic among the non-missing only
        set_code(8, i, 3 if i % 3 else 1)
    for k in range(m):
        set_code(k, 11, 1)      # individual 11 missing everywhere
    # garbage in the pad bits of the last byte (n=37 -> 3 unused slots)
    pad = rng.integers(0, 4, size=(m, 3))
    for k in range(m):
        for t in range(3):
            set_code(k, n + t, int(pad[k, t]))
    return bed


def main():
    assert oracle.have_ref(), "build oracle/_ref/gcta_ref first (make -C oracle ref)"

    # g1: plain case with 2 % missing, both output formats
    d = fresh("g1_basic")
    n, m = 203, 1000
    synth.write_plink(d + "/t", synth.make_bed_payload(1234, n, m, miss_rate=0.02), n)
    run("g1_basic", ["--bfile", "t", "--make-grm", "--out", "o"])
    run("g1_basic", ["--bfile", "t", "--make-grm-bin", "--out", "ob"])
    run("g1_basic", ["--bfile", "t", "--freq", "--out", "of"])

    # g2: edge cases: monomorphic SNPs, all-missing SNP / individual, pad garbage
    d = fresh("g2_edge")
    synth.write_plink(d + "/t", edge_payload(), 37)
    run("g2_edge", ["--bfile", "t", "--make-grm", "--out", "o"])
    run("g2_edge", ["--bfile", "t", "--make-grm-bin", "--out", "ob"])
    # g3: same input, --grm-alg 1
    run("g2_edge", ["--bfile", "t", "--make-grm", "--grm-alg", "1", "--out", "o_alg1"])
    run("g2_edge", ["--bfile", "t", "--make-grm-bin", "--grm-alg", "1", "--out", "ob_alg1"])

    # g4: X chromosome
    d = fresh("g4_xchr")
    n, m = 50, 200
    rng = np.random.default_rng(4)
    sex = rng.integers(1, 3, size=n)
    synth.write_plink(d + "/t", synth.make_bed_payload(4, n, m, miss_rate=0.03), n, chrom=[23] * m, sex=sex)
    run("g4_xchr", ["--bfile", "t", "--make-grm-xchr", "--out", "o"])
    run("g4_xchr", ["--bfile", "t", "--make-grm-xchr-bin", "--out", "ob"])

    # g5: --keep / --extract subsets and the sub-block identity with --update-freq
    d = fresh("g5_subset")
    n, m = 120, 500
    synth.write_plink(d + "/t", synth.make_bed_payload(5, n, m, miss_rate=0.02), n)
    rng = np.random.default_rng(5)
    keep = np.sort(rng.choice(n, 70, replace=False))
    ext = np.sort(rng.choice(m, 300, replace=False))
    with open(d + "/keep.txt", "w") as f:
        for i in keep:
            f.write(f"F{i} I{i}\n")
    with open(d + "/extract.txt", "w") as f:
        for k in ext:
            f.write(f"rs{k}\n")
    run("g5_subset", ["--bfile", "t", "--keep", "keep.txt", "--extract", "extract.txt", "--make-grm", "--out", "o"])
    run("g5_subset", ["--bfile", "t", "--keep", "keep.txt", "--extract", "extract.txt", "--make-grm-bin", "--out", "ob"])
    run("g5_subset", ["--bfile", "t", "--freq", "--out", "full"])
    run("g5_subset", ["--bfile", "t", "--keep", "keep.txt", "--update-freq", "full.freq", "--make-grm-bin", "--out", "ob_uf"])
    run("g5_subset", ["--bfile", "t", "--make-grm-bin", "--out", "ob_full"])
    run("g5_subset", ["--bfile", "t", "--maf", "0.1", "--make-grm", "--out", "o_maf"])

    # g7: no missing data (N == M everywhere), tile-aligned sizes
    d = fresh("g7_nomiss")
    n, m = 64, 512
    synth.write_plink(d + "/t", synth.make_bed_payload(7, n, m, miss_rate=0.0), n)
    run("g7_nomiss", ["--bfile", "t", "--make-grm", "--out", "o"])
    run("g7_nomiss", ["--bfile", "t", "--make-grm-bin", "--out", "ob"])

    # error conventions: chr > autosome_num without --autosome (grm.cpp:15-20)
    d = fresh("g8_errors")
    synth.write_plink(d + "/t", synth.make_bed_payload(8, 12, 30), 12, chrom=[1] * 29 + [23])
    run("g8_errors", ["--bfile", "t", "--make-grm", "--out", "o"])
    run("g8_errors", ["--bfile", "missing_prefix", "--make-grm", "--out", "o2"])
    run("g8_errors", ["--bfile", "t", "--make-grm", "--bogus-flag", "--out", "o3"])

    for root, _, files in os.walk(OUT):
        for fn in files:
            if fn.endswith(".log"):
                os.remove(os.path.join(root, fn))
    print("golden fixtures written under", OUT)


def round2():
    """Fixtures added in round 2 (the earlier ones are left untouched):
    g9_refallele - --update-ref-allele flips the reference allele of every other SNP, then --maf / --update-freq leave
                   2p of the REFERENCE allele in _mu (data.cpp:794, 915-916): what a caller must convert before it
                   hands 2p of allele 1 to the library;
    g10_related  - duplicates, parent-offspring and sib pairs next to unrelated founders (large off-diagonals)."""
    d = fresh("g9_refallele")
    n, m = 60, 240
    synth.write_plink(d + "/t", synth.make_bed_payload(9, n, m, miss_rate=0.02), n)
    with open(d + "/refA.txt", "w") as f:
        for k in range(m):
            f.write(f"rs{k} {'G' if k % 2 else 'A'}\n")
    run("g9_refallele", ["--bfile", "t", "--update-ref-allele", "refA.txt", "--maf", "0.05", "--make-grm", "--out", "o_maf"])
    run("g9_refallele", ["--bfile", "t", "--update-ref-allele", "refA.txt", "--freq", "--out", "full"])
    run("g9_refallele", ["--bfile", "t", "--update-ref-allele", "refA.txt", "--update-freq", "full.freq", "--make-grm-bin", "--out", "ob_uf"])
    run("g9_refallele", ["--bfile", "t", "--update-ref-allele", "refA.txt", "--update-freq", "full.freq", "--make-grm", "--out", "o_uf"])

    d = fresh("g10_related")
    bed, n, pairs = synth.make_related_payload(10, 40, 600, n_dup=4, n_po=4, n_sib=4, miss_rate=0.01)
    synth.write_plink(d + "/t", bed, n)
    with open(d + "/pairs.txt", "w") as f:
        for i, j, kind in pairs:
            f.write(f"{i} {j} {kind}\n")
    run("g10_related", ["--bfile", "t", "--make-grm", "--out", "o"])
    run("g10_related", ["--bfile", "t", "--make-grm-bin", "--out", "ob"])
    run("g10_related", ["--bfile", "t", "--grm-cutoff", "0.3", "--make-grm", "--out", "o_cut"])   # --grm-cutoff only acts on --grm input: ignored here
    for dd in ("g9_refallele", "g10_related"):
        for fn in os.listdir(os.path.join(OUT, dd)):
            if fn.endswith(".log"):
                os.remove(os.path.join(OUT, dd, fn))


if __name__ == "__main__":
    if "--round2" in sys.argv:
        round2()
    else:
        main()
