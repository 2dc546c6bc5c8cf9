"""The drop-in for real: integration/_dropin/gcta64_dropin is the reference program - its own option
parser, readers, filters, writers, compiled from /root/reference - with the body of gcta::make_grm replaced by one call
of gcta_b200_make_grm (integration/make_grm_dropin.cpp).  Built by __graft_entry__.build() where /root/reference exists;
the binary travels to the GPU box.  On the GPU its output is compared with the unmodified reference's (the goldens); on
the CPU box the call must come back as the library's no-device error through the reference's own error channel."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from gcta_b200 import abi, grmio
from tests.helpers import GOLDEN, ROOT, check_grm

DROPIN = os.path.join(ROOT, "integration", "_dropin", "gcta64_dropin")
needs_dropin = pytest.mark.skipif(not os.access(DROPIN, os.X_OK), reason="integration/_dropin/gcta64_dropin not built")


def _run(args, cwd):
    return subprocess.run([DROPIN] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)


def _ref_log(path):
    out, _, err = open(path).read().partition("\n[stderr]\n")
    lines = [ln for ln in out.split("\n") if ln.strip()]
    return lines[lines.index("Options:"):], err.strip()


def _log(stdout):
    lines = [ln for ln in stdout.split("\n") if ln.strip()]
    return [ln for ln in lines[lines.index("Options:"):] if not ln.startswith(("Analysis started", "Analysis finished", "Computational time"))]


@needs_dropin
def test_dropin_reports_the_missing_device_like_any_gcta_error(tmp_path):
    if abi.lib().gcta_b200_device_count() > 0:
        pytest.skip("a CUDA device is present")
    d = tmp_path / "g1_basic"
    shutil.copytree(os.path.join(GOLDEN, "g1_basic"), d)
    r = _run(["--bfile", "t", "--make-grm", "--out", "d"], d)
    assert r.returncode == 0                                   # option.cpp:42-50: message on stderr, exit 0
    assert r.stderr.strip() == "Error: no CUDA device is available (gcta_b200 has no CPU fallback)."
    assert "Analysis finished" in r.stdout and not os.path.exists(d / "d.grm.gz")
    ref, _ = _ref_log(os.path.join(GOLDEN, "g1_basic", "ref_stdout_o.txt"))
    ours = [ln.replace("[d.", "[o.").replace("--out d", "--out o") for ln in _log(r.stdout)]
    stop = next(i for i, ln in enumerate(ref) if ln.startswith("Calculating the genetic relationship matrix"))
    assert ours[:stop + 1] == ref[:stop + 1]                   # same program up to the call


CASES = [
    ("g1_basic", ["--bfile", "t", "--make-grm"], "o", False),
    ("g1_basic", ["--bfile", "t", "--make-grm-bin"], "ob", True),
    ("g2_edge", ["--bfile", "t", "--make-grm"], "o", False),
    ("g2_edge", ["--bfile", "t", "--make-grm", "--grm-alg", "1"], "o_alg1", False),
    ("g4_xchr", ["--bfile", "t", "--make-grm-xchr"], "o", False),
    ("g5_subset", ["--bfile", "t", "--keep", "keep.txt", "--extract", "extract.txt", "--make-grm"], "o", False),
    ("g5_subset", ["--bfile", "t", "--keep", "keep.txt", "--update-freq", "full.freq", "--make-grm-bin"], "ob_uf", True),   # 2p handed over
    ("g5_subset", ["--bfile", "t", "--maf", "0.1", "--make-grm"], "o_maf", False),   # the reference's own calcu_mu + filter first
    # reference allele flipped for every other SNP: _mu then holds 2p of the flipped allele and must be turned back
    ("g9_refallele", ["--bfile", "t", "--update-ref-allele", "refA.txt", "--maf", "0.05", "--make-grm"], "o_maf", False),
    ("g9_refallele", ["--bfile", "t", "--update-ref-allele", "refA.txt", "--update-freq", "full.freq", "--make-grm-bin"], "ob_uf", True),
    ("g10_related", ["--bfile", "t", "--make-grm"], "o", False),   # duplicates, parent-offspring, sibs
]


@pytest.mark.gpu
@needs_dropin
@pytest.mark.parametrize("case,args,ref_out,is_bin", CASES, ids=[c[0] + ":" + c[2] for c in CASES])
def test_dropin_matches_the_unmodified_reference(tmp_path, case, args, ref_out, is_bin):
    d = tmp_path / case
    shutil.copytree(os.path.join(GOLDEN, case), d)
    r = _run(args + ["--out", "d"], d)
    assert r.stderr.strip() == "", r.stderr
    assert open(d / "d.grm.id").read() == open(os.path.join(GOLDEN, case, ref_out + ".grm.id")).read()
    if is_bin:
        A = grmio.read_grm_bin(str(d / "d.grm.bin"))
        Aref = grmio.read_grm_bin(os.path.join(GOLDEN, case, ref_out + ".grm.bin"))
        assert A.shape == Aref.shape and np.abs(A - Aref).max() <= 1e-3 and abs(np.diag(A).mean() - np.diag(Aref).mean()) <= 1e-5
    else:
        A, N, _ = grmio.read_grm_gz(str(d / "d.grm.gz"))
        Aref, Nref, _ = grmio.read_grm_gz(os.path.join(GOLDEN, case, ref_out + ".grm.gz"))
        check_grm(A, N, Aref, Nref)
