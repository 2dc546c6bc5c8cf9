"""gcta64_b200 (C++ front end) against the reference's own CLI behaviour.

CPU part: option echo, reader / filter log lines and error strings are compared with the logs the
compiled reference printed for the golden cases (tests/golden/*/ref_stdout_*.txt).  GPU part:
complete runs - .grm.gz (N exact, A within tolerance), .grm.bin, .grm.id, .freq byte-identical -
and the reference's own GRM reader consuming our multi-member .grm.gz.
"""
import os
import shutil
import subprocess

import numpy as np
import pytest

from gcta_b200 import grmio
from oracle import oracle
from tests.helpers import GOLDEN, ROOT, check_grm, pack_lower

BIN = os.path.join(ROOT, "gcta_b200", "gcta64_b200")


def run_cli(args, cwd):
    return subprocess.run([BIN] + [str(a) for a in args], cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)


def body(stdout):
    """log lines after the 'Options:' header, without time stamps and progress lines"""
    lines = stdout.replace("\r", "\n").split("\n")
    i = lines.index("Options:")
    keep = [ln for ln in lines[i:] if not ln.startswith(("Analysis started", "Analysis finished", "Computational time"))]
    return [ln for ln in keep if ln.strip() and ("of" not in ln or "individuals." not in ln or "Genotype" in ln or "GRM of" in ln)]


def ref_body(path):
    txt = open(path).read()
    out, _, err = txt.partition("\n[stderr]\n")
    lines = out.split("\n")
    i = lines.index("Options:")
    return [ln for ln in lines[i:] if ln.strip()], err.strip()


def test_binary_is_built():
    assert os.access(BIN, os.X_OK), "python -m gcta_b200.build builds gcta_b200/gcta64_b200"


def test_fast_number_formatter_equals_printf():
    """The .grm.gz writer formats n(n+1)/2 lines with its own %.6e routine; the binary checks it against snprintf
    on 4.5 million values (float-valued like A, 7-digit decimals, specials)."""
    r = subprocess.run([BIN, "--selftest-format"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert r.returncode == 0 and "0 mismatches" in r.stdout, r.stdout + r.stderr


def test_structured_gz_encoder_on_awkward_inputs():
    """--selftest-gz: single lines, constant fields (runs merged across lines), NaN/inf/denormal A, extreme N, row
    ranges from the middle of the triangle, an empty range - each member inflated by zlib and compared with the text."""
    r = subprocess.run([BIN, "--selftest-gz"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert r.returncode == 0 and " 0 failures" in r.stdout, r.stdout + r.stderr


def _selftest_writer(n, prefix, cwd, env=None, extra=()):
    e = dict(os.environ)
    e.pop("GCTA_B200_GZ_LEVEL", None)
    e.update(env or {})
    r = subprocess.run([BIN, "--selftest-writer", str(n), "--out", prefix, *extra], cwd=cwd, env=e, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=300)
    assert r.returncode == 0 and "writer selftest" in r.stdout, r.stdout + r.stderr
    return r


def test_structured_gz_encoder_writes_the_same_text_as_zlib(tmp_path):
    """The default .grm.gz codec (host/grm_text_gz.cpp: shared runs with the previous line as deflate matches, one
    dynamic Huffman code per member) must decode - with a stock gzip reader, CRC and length checked - to exactly the
    text the formatter + zlib route gives; several members (n = 1600: 1.28 M lines), first rows included."""
    import gzip
    _selftest_writer(1600, "a", tmp_path)
    _selftest_writer(1600, "b", tmp_path, env={"GCTA_B200_GZ_LEVEL": "6"})
    raw = open(tmp_path / "a.grm.gz", "rb").read()
    assert raw[:4] == b"\x1f\x8b\x08\x04" and raw.count(b"\x1f\x8b\x08\x04") >= 2   # FEXTRA: member length
    ta, tb = gzip.decompress(raw), gzip.decompress(open(tmp_path / "b.grm.gz", "rb").read())
    assert ta == tb
    lines = ta.split(b"\n")
    assert len(lines) == 1600 * 1601 // 2 + 1 and lines[-1] == b""
    assert lines[0].split(b"\t")[:2] == [b"1", b"1"] and lines[-2].split(b"\t")[:2] == [b"1600", b"1600"]
    assert len(raw) < 0.45 * len(ta)   # random 7-digit mantissas cap any deflate stream near 2.6-2.9x
    assert open(tmp_path / "a.grm.id").read().splitlines()[:2] == ["F0\tI0", "F1\tI1"]
    # both routes announce each member's length in its header, so the --grm reader takes the parallel path for
    # either; the re-saved matrices must agree
    rawb = open(tmp_path / "b.grm.gz", "rb").read()
    assert rawb[:4] == b"\x1f\x8b\x08\x04" and rawb[12:14] == b"GM" and int.from_bytes(rawb[16:20], "little") > 20
    for p in ("a", "b"):
        r = run_cli(["--grm", p, "--make-grm", "--out", "re_" + p], tmp_path)
        assert r.stderr.strip() == "", r.stderr
    assert gzip.decompress(open(tmp_path / "re_a.grm.gz", "rb").read()) == gzip.decompress(open(tmp_path / "re_b.grm.gz", "rb").read())
    # without missing genotypes N is the same on every line: the encoder then takes "\t N \t" as one run without looking
    const = {"GCTA_B200_SELFTEST_CONST_N": "1"}
    _selftest_writer(1100, "c", tmp_path, env=const)
    _selftest_writer(1100, "d", tmp_path, env={**const, "GCTA_B200_GZ_LEVEL": "1"})
    tc = gzip.decompress(open(tmp_path / "c.grm.gz", "rb").read())
    assert tc == gzip.decompress(open(tmp_path / "d.grm.gz", "rb").read())
    assert tc.startswith(b"1\t1\t1000000\t") and tc.count(b"\n") == 1100 * 1101 // 2


def test_single_stream_gz_is_read_in_blocks_by_several_parsers(tmp_path):
    """A .grm.gz that is one plain gzip stream (what the reference writes) is inflated by one thread and parsed in 8 MB
    blocks by several: same matrix as the member-indexed route, N == 0 warnings in file order, and of two bad lines in
    different blocks the first one is reported (n = 1300: 25 MB of text, four blocks)."""
    import gzip
    n = 1300
    _selftest_writer(n, "a", tmp_path)
    text = gzip.decompress(open(tmp_path / "a.grm.gz", "rb").read())
    assert len(text) > 2.9 * (8 << 20)
    ids = open(tmp_path / "a.grm.id").read()

    def put(name, t):
        open(tmp_path / (name + ".grm.gz"), "wb").write(gzip.compress(t, 1))
        open(tmp_path / (name + ".grm.id"), "w").write(ids)

    put("plain", text)
    ra = run_cli(["--grm", "a", "--make-grm", "--out", "ra"], tmp_path)
    rp = run_cli(["--grm", "plain", "--make-grm", "--out", "rp"], tmp_path)
    assert ra.stderr.strip() == "" and rp.stderr.strip() == ""
    assert gzip.decompress(open(tmp_path / "ra.grm.gz", "rb").read()) == gzip.decompress(open(tmp_path / "rp.grm.gz", "rb").read())
    lines = text.split(b"\n")
    zero = [5, 300000, 610000, 840000]
    for k in zero:
        f = lines[k].split(b"\t")
        lines[k] = f[0] + b"\t" + f[1] + b"\t0\t0.000000e+00"
    put("warn", b"\n".join(lines))
    rw = run_cli(["--grm", "warn", "--make-grm", "--out", "rw"], tmp_path)
    warned = [ln for ln in rw.stdout.split("\n") if ln.startswith("Warning: ")]
    assert rw.stderr.strip() == "" and warned == ["Warning: " + lines[k].decode() for k in zero]
    lines[700000] = b"oops"
    lines[420000] = b"12\t34\t56"
    put("bad", b"\n".join(lines))
    rb = run_cli(["--grm", "bad", "--make-grm", "--out", "rb"], tmp_path)
    assert rb.stderr.strip().endswith("Error occurs in line:\n12\t34\t56")
    warned = [ln for ln in rb.stdout.split("\n") if ln.startswith("Warning: ")]
    assert warned == ["Warning: " + lines[k].decode() for k in zero[:2]]      # only what precedes the bad line
    assert not os.path.exists(tmp_path / "rb.grm.gz")


def test_parallel_bin_writer_lays_out_the_full_square(tmp_path):
    """.grm.bin is written by a pool of threads in 16-row bands with pwrite (host/grm_bin_writer.h): n x n float64,
    symmetric, and element for element the matrix the text writer prints (n = 333: 21 bands, the last one ragged)."""
    n = 333
    _selftest_writer(n, "t", tmp_path)
    _selftest_writer(n, "b", tmp_path, extra=["--make-grm-bin"])
    A, _, _ = grmio.read_grm_gz(str(tmp_path / "t.grm.gz"))
    B = grmio.read_grm_bin(str(tmp_path / "b.grm.bin"))
    assert B.shape == (n, n) and np.array_equal(B, B.T)
    low = B[np.tril_indices(n)]
    assert np.abs(low - A).max() <= 5.1e-7 * np.abs(low).max()
    assert np.array_equal(low.astype(np.float32).astype(np.float64), low)   # float32 values widened, nothing else


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/gcta_ref not built")
def test_reference_reader_accepts_the_structured_gz(tmp_path):
    """SURVEY.md 8f-1 without a GPU: the reference's read_grm_gz (grm.cpp:381-448, gzifstream of zfstream.cpp) reads the
    encoder's members and re-saves the same matrix."""
    _selftest_writer(1500, "mine", tmp_path)
    rr = oracle.run_ref(["--grm", "mine", "--make-grm", "--out", "re"], cwd=tmp_path)
    assert "Analysis finished" in rr.stdout and rr.stderr.strip() == "", rr.stderr
    A, N, _ = grmio.read_grm_gz(str(tmp_path / "mine.grm.gz"))
    A2, N2, _ = grmio.read_grm_gz(str(tmp_path / "re.grm.gz"))
    assert np.array_equal(N, N2) and np.abs(A - A2).max() <= 1e-6 * max(1.0, np.abs(A).max())


@pytest.mark.parametrize("case,args,log", [
    ("g8_errors", ["--bfile", "t", "--make-grm", "--out", "o"], "ref_stdout_o.txt"),                # chr 23 without --autosome
    ("g8_errors", ["--bfile", "missing_prefix", "--make-grm", "--out", "o2"], "ref_stdout_o2.txt"),  # cannot open .fam
    ("g8_errors", ["--bfile", "t", "--make-grm", "--bogus-flag", "--out", "o3"], "ref_stdout_o3.txt"),
])
def test_error_convention_matches_reference(case, args, log, tmp_path):
    d = tmp_path / case
    shutil.copytree(os.path.join(GOLDEN, case), d)
    r = run_cli(args, d)
    assert r.returncode == 0                       # the reference exits 0 after an error (option.cpp:50)
    ref_lines, ref_err = ref_body(os.path.join(GOLDEN, case, log))
    assert r.stderr.strip() == ref_err
    assert "Analysis finished" in r.stdout
    ours = body(r.stdout)
    # up to the point of failure the logs are identical, except that we print make_XMat's line before the guard fails
    assert ours[:len(ref_lines)] == ref_lines or ours == ref_lines


def test_range_errors_and_echo(tmp_path):
    r = run_cli(["--bfile", "x", "--maf", "0.7"], tmp_path)
    assert "--maf 0.7" in r.stdout and "Error: --maf should be within the range from 0 to 0.5." in r.stderr
    r = run_cli(["--bfile", "x", "--grm-alg", "3"], tmp_path)
    assert "Error: --grm-alg should be 0 or 1." in r.stderr
    r = run_cli(["--out", "zzz"], tmp_path)
    assert "Error: no analysis has been launched by the option(s)." in r.stderr


def test_reader_and_filter_log_lines_match_reference(tmp_path):
    """keep / extract / bed reading lines are produced before any device work, so they can be
    compared on the CPU box: everything up to the GRM banner must equal the reference's log."""
    case = "g5_subset"
    d = tmp_path / case
    shutil.copytree(os.path.join(GOLDEN, case), d)
    r = run_cli(["--bfile", "t", "--keep", "keep.txt", "--extract", "extract.txt", "--make-grm", "--out", "o"], d)
    ref_lines, _ = ref_body(os.path.join(GOLDEN, case, "ref_stdout_o.txt"))
    ours = body(r.stdout)
    stop = next(i for i, ln in enumerate(ref_lines) if ln.startswith("Calculating the genetic relationship matrix"))
    assert ours[:stop + 1] == ref_lines[:stop + 1]


# ------------------------------------------------------------------------------------------------
# GPU: complete runs
# ------------------------------------------------------------------------------------------------
def _full_case(tmp_path, case, args_txt, ref_txt, args_bin=None, ref_bin=None, same_log=True):
    d = tmp_path / case
    shutil.copytree(os.path.join(GOLDEN, case), d, dirs_exist_ok=True)
    r = run_cli(args_txt + ["--out", "mine"], d)
    assert r.stderr.strip() == "", r.stderr
    A, N, _ = grmio.read_grm_gz(str(d / "mine.grm.gz"))
    Aref, Nref, _ = grmio.read_grm_gz(os.path.join(GOLDEN, case, ref_txt + ".grm.gz"))
    check_grm(A, N, Aref, Nref)
    assert open(d / "mine.grm.id").read() == open(os.path.join(GOLDEN, case, ref_txt + ".grm.id")).read()
    if same_log:
        ref_lines, _ = ref_body(os.path.join(GOLDEN, case, f"ref_stdout_{ref_txt}.txt"))
        assert body(r.stdout) == [ln.replace(f"[{ref_txt}.", "[mine.") for ln in ref_lines if not ln.startswith("--out")] or \
            [ln for ln in body(r.stdout) if not ln.startswith("--out")] == [ln.replace(f"[{ref_txt}.", "[mine.") for ln in ref_lines if not ln.startswith("--out")]
    if args_bin:
        r = run_cli(args_bin + ["--out", "mineb"], d)
        assert r.stderr.strip() == "", r.stderr
        Ab = grmio.read_grm_bin(str(d / "mineb.grm.bin"))
        Arefb = grmio.read_grm_bin(os.path.join(GOLDEN, case, ref_bin + ".grm.bin"))
        assert Ab.shape == Arefb.shape and np.abs(Ab - Arefb).max() <= 1e-3
        assert np.array_equal(Ab, Ab.T)
    return d


@pytest.mark.gpu
def test_cli_basic_text_and_binary(tmp_path):
    _full_case(tmp_path, "g1_basic", ["--bfile", "t", "--make-grm"], "o", ["--bfile", "t", "--make-grm-bin"], "ob")


@pytest.mark.gpu
def test_cli_edge_cases(tmp_path):
    _full_case(tmp_path, "g2_edge", ["--bfile", "t", "--make-grm"], "o", ["--bfile", "t", "--make-grm-bin"], "ob")


@pytest.mark.gpu
def test_cli_keep_extract_maf_update_freq(tmp_path):
    d = _full_case(tmp_path, "g5_subset", ["--bfile", "t", "--keep", "keep.txt", "--extract", "extract.txt", "--make-grm"], "o",
                   ["--bfile", "t", "--keep", "keep.txt", "--extract", "extract.txt", "--make-grm-bin"], "ob")
    # --maf runs calcu_mu on the device, filters, then make_grm reuses it
    r = run_cli(["--bfile", "t", "--maf", "0.1", "--make-grm", "--out", "mine_maf"], d)
    assert r.stderr.strip() == ""
    A, N, _ = grmio.read_grm_gz(str(d / "mine_maf.grm.gz"))
    Aref, Nref, _ = grmio.read_grm_gz(os.path.join(GOLDEN, "g5_subset", "o_maf.grm.gz"))
    check_grm(A, N, Aref, Nref)
    ref_lines, _ = ref_body(os.path.join(GOLDEN, "g5_subset", "ref_stdout_o_maf.txt"))
    assert [ln for ln in body(r.stdout) if "MAF" in ln] == [ln for ln in ref_lines if "MAF" in ln]
    # --freq output is byte-identical; --update-freq sub-block identity
    r = run_cli(["--bfile", "t", "--freq", "--out", "mine_full"], d)
    assert open(d / "mine_full.freq").read() == open(os.path.join(GOLDEN, "g5_subset", "full.freq")).read()
    r = run_cli(["--bfile", "t", "--keep", "keep.txt", "--update-freq", "mine_full.freq", "--make-grm-bin", "--out", "mine_uf"], d)
    assert r.stderr.strip() == ""
    Ab = grmio.read_grm_bin(str(d / "mine_uf.grm.bin"))
    assert np.abs(Ab - grmio.read_grm_bin(os.path.join(GOLDEN, "g5_subset", "ob_uf.grm.bin"))).max() <= 1e-3


@pytest.mark.gpu
def test_cli_update_ref_allele_and_related(tmp_path):
    _full_case(tmp_path, "g9_refallele", ["--bfile", "t", "--update-ref-allele", "refA.txt", "--update-freq", "full.freq", "--make-grm"], "o_uf",
               ["--bfile", "t", "--update-ref-allele", "refA.txt", "--update-freq", "full.freq", "--make-grm-bin"], "ob_uf")
    _full_case(tmp_path, "g9_refallele", ["--bfile", "t", "--update-ref-allele", "refA.txt", "--maf", "0.05", "--make-grm"], "o_maf")
    _full_case(tmp_path, "g10_related", ["--bfile", "t", "--make-grm"], "o", ["--bfile", "t", "--make-grm-bin"], "ob")


@pytest.mark.gpu
def test_cli_xchr(tmp_path):
    _full_case(tmp_path, "g4_xchr", ["--bfile", "t", "--make-grm-xchr"], "o", ["--bfile", "t", "--make-grm-xchr-bin"], "ob")


@pytest.mark.gpu
def test_cli_freq_basic(tmp_path):
    d = tmp_path / "g1_basic"
    shutil.copytree(os.path.join(GOLDEN, "g1_basic"), d)
    r = run_cli(["--bfile", "t", "--freq", "--out", "minef"], d)
    assert r.stderr.strip() == ""
    assert open(d / "minef.freq").read() == open(os.path.join(GOLDEN, "g1_basic", "of.freq")).read()


@pytest.mark.gpu
@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/gcta_ref not built")
def test_reference_reader_accepts_our_multi_member_gz(tmp_path):
    """the reference's read_grm_gz (grm.cpp:381-448) must consume our .grm.gz.
    n = 1500 gives 1,125,750 lines = two gzip members."""
    from gcta_b200 import synth
    n, m = 1500, 3000
    synth.write_plink(str(tmp_path / "t"), synth.make_bed_payload(31, n, m, miss_rate=0.01), n)
    r = run_cli(["--bfile", "t", "--make-grm", "--out", "mine"], tmp_path)
    assert r.stderr.strip() == ""
    raw = open(tmp_path / "mine.grm.gz", "rb").read()
    assert raw.count(b"\x1f\x8b\x08") >= 2
    rr = oracle.run_ref(["--grm", "mine", "--make-grm", "--out", "re"], cwd=tmp_path)   # save_grm -> read_grm_gz -> rewrite
    assert "Analysis finished" in rr.stdout and rr.stderr.strip() == "", rr.stderr
    A, N, _ = grmio.read_grm_gz(str(tmp_path / "mine.grm.gz"))
    A2, N2, _ = grmio.read_grm_gz(str(tmp_path / "re.grm.gz"))
    assert np.array_equal(N, N2) and np.abs(A - A2).max() <= 1e-6 * max(1.0, np.abs(A).max())


def _flag_dataset(d):
    """n=90, M=400 with chromosomes 1..23 mixed, both sexes, 3 % missing, plus the list files the flags need."""
    from gcta_b200 import synth
    n, m = 90, 400
    rng = np.random.default_rng(77)
    chrom = rng.integers(1, 24, size=m)
    sex = rng.integers(1, 3, size=n)
    synth.write_plink(str(d / "t"), synth.make_bed_payload(77, n, m, miss_rate=0.03), n, chrom=chrom, sex=sex)
    open(d / "remove.txt", "w").write("".join(f"F{i} I{i}\n" for i in range(0, n, 7)))
    open(d / "keep.txt", "w").write("".join(f"F{i} I{i} extra columns ignored\n" for i in range(0, n, 2)))
    open(d / "exclude.txt", "w").write("".join(f"rs{k}\n" for k in range(0, m, 5)))
    open(d / "sex.txt", "w").write("".join(f"F{i} I{i} {'M' if i % 3 else 'F'}\n" for i in range(n)))
    open(d / "refA.txt", "w").write("".join(f"rs{k} {'G' if k % 2 else 'A'}\n" for k in range(m)))
    return n, m


FLAG_SETS = [
    ["--autosome", "--remove", "remove.txt", "--exclude", "exclude.txt", "--make-grm"],
    ["--chr", "3", "--make-grm"],
    ["--autosome", "--keep", "keep.txt", "--remove", "remove.txt", "--make-grm"],
    ["--autosome", "--exclude-snp", "rs7", "--max-maf", "0.3", "--make-grm"],
    ["--extract-snp", "rs5", "--autosome-num", "23", "--make-grm"],
    ["--update-sex", "sex.txt", "--make-grm-xchr"],
    ["--autosome", "--update-ref-allele", "refA.txt", "--make-grm"],
    ["--autosome", "--update-ref-allele", "refA.txt", "--freq"],
    ["--autosome-num", "10", "--make-grm"],                     # chromosomes above 10 -> "autosomal SNPs only" error
    ["--autosome", "--maf", "0.05", "--grm-alg", "1", "--make-grm"],
    ["--make-grm-xchr"],                                        # autosomal data is filtered to chr 23 first (option.cpp:571)
]


HOST_ONLY_FLAG_SETS = FLAG_SETS + [
    ["--keep", "nosuchfile.txt", "--make-grm"],                        # cannot open the list
    ["--autosome", "--extract-snp", "rs_unknown", "--make-grm"],       # SNP not found
    ["--autosome", "--exclude-snp", "rs_unknown", "--make-grm"],
    ["--chr", "24", "--make-grm"],                                     # no SNP left
    ["--update-sex", "remove.txt", "--make-grm-xchr"],                 # malformed sex file
    ["--autosome", "--rm-high-ld", "1.5", "--make-grm"],               # range error while parsing
    ["--autosome", "--save-ram", "--make-grm-bin"],
]


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/gcta_ref not built")
@pytest.mark.parametrize("flags", HOST_ONLY_FLAG_SETS, ids=[" ".join(f) for f in HOST_ONLY_FLAG_SETS])
def test_host_side_of_the_flags_matches_the_reference_without_a_gpu(flags, tmp_path):
    """Everything the front end does before it needs the device - option echo, .fam/.bim/.bed reading, the selection
    lists, the guards and their error strings - must produce the reference's log line for line; where the reference
    fails before any arithmetic, stderr must be the same too.  Runs on the CPU box."""
    _flag_dataset(tmp_path)
    rr = oracle.run_ref(["--bfile", "t"] + flags + ["--out", "ref"], cwd=tmp_path)
    rm = run_cli(["--bfile", "t"] + flags + ["--out", "mine"], tmp_path)
    ours = [ln.replace("[mine.", "[X.") for ln in body(rm.stdout) if not ln.startswith("--out")]
    theirs = [ln.replace("[ref.", "[X.") for ln in body(rr.stdout) if not ln.startswith("--out")]
    device_lines = ("Calculating allele frequencies", "Recoding genotypes", "Pruning SNPs", "Writing allele frequencies",
                    "Calculating the genetic relationship")
    stop = next((i for i, ln in enumerate(theirs) if ln.startswith(device_lines)), len(theirs))
    assert stop > 0 and ours[:stop] == theirs[:stop]
    if stop == len(theirs):     # the reference never reached the arithmetic: same error, same place
        assert rm.stderr.strip() == rr.stderr.strip() and rr.stderr.strip() != ""
        assert ours == theirs


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/gcta_ref not built")
@pytest.mark.parametrize("cut", [1, 23 * 17, 23 * 399 + 5])
def test_truncated_bed_is_reported_like_the_reference(tmp_path, cut):
    """The .bed reader uses pread from several threads; a file that ends early (mid-row, or rows missing) must give the
    reference's message (data.cpp:173) - also when only some rows are wanted (--exclude)."""
    n, m = _flag_dataset(tmp_path)
    raw = open(tmp_path / "t.bed", "rb").read()
    assert len(raw) == 3 + m * 23
    open(tmp_path / "t.bed", "wb").write(raw[:len(raw) - cut])
    for extra in ([], ["--exclude", "exclude.txt"]):
        rr = oracle.run_ref(["--bfile", "t", "--autosome"] + extra + ["--make-grm", "--out", "ref"], cwd=tmp_path)
        rm = run_cli(["--bfile", "t", "--autosome"] + extra + ["--make-grm", "--out", "mine"], tmp_path)
        assert rr.stderr.strip() == "Error: problem with the BED file...has the FAM/BIM file been changed?"
        assert rm.stderr.strip() == rr.stderr.strip()


@pytest.mark.gpu
@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/gcta_ref not built")
@pytest.mark.parametrize("flags", FLAG_SETS, ids=[" ".join(f) for f in FLAG_SETS])
def test_cli_flags_live_against_the_reference(flags, tmp_path):
    """Same command line through the compiled reference and through gcta64_b200: stderr, log lines, .grm.id, .freq
    identical; N identical; A within tolerance."""
    _flag_dataset(tmp_path)
    rr = oracle.run_ref(["--bfile", "t"] + flags + ["--out", "ref"], cwd=tmp_path)
    rm = run_cli(["--bfile", "t"] + flags + ["--out", "mine"], tmp_path)
    assert rm.stderr.strip() == rr.stderr.strip()
    ours = [ln.replace("[mine.", "[X.") for ln in body(rm.stdout) if not ln.startswith("--out")]
    theirs = [ln.replace("[ref.", "[X.") for ln in body(rr.stdout) if not ln.startswith("--out")]
    if not rr.stderr.strip():
        assert ours == theirs
    if os.path.exists(tmp_path / "ref.grm.gz"):
        A, N, _ = grmio.read_grm_gz(str(tmp_path / "mine.grm.gz"))
        Aref, Nref, _ = grmio.read_grm_gz(str(tmp_path / "ref.grm.gz"))
        check_grm(A, N, Aref, Nref)   # also with M = 1 or 17 SNPs: short contractions run in the split-FP16 mode
        assert open(tmp_path / "mine.grm.id").read() == open(tmp_path / "ref.grm.id").read()
    else:
        assert not os.path.exists(tmp_path / "mine.grm.gz")
    if os.path.exists(tmp_path / "ref.freq"):
        assert open(tmp_path / "mine.freq").read() == open(tmp_path / "ref.freq").read()


@pytest.mark.gpu
def test_cli_two_gpus_when_present(tmp_path):
    """--gpus 2: one host thread per GPU, each computing its row band into the shared host buffers."""
    from gcta_b200 import abi
    if abi.lib().gcta_b200_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    d = tmp_path / "g1_basic"
    shutil.copytree(os.path.join(GOLDEN, "g1_basic"), d)
    r = run_cli(["--bfile", "t", "--make-grm", "--gpus", "2", "--out", "mine2"], d)
    assert r.stderr.strip() == "", r.stderr
    A, N, _ = grmio.read_grm_gz(str(d / "mine2.grm.gz"))
    Aref, Nref, _ = grmio.read_grm_gz(os.path.join(GOLDEN, "g1_basic", "o.grm.gz"))
    check_grm(A, N, Aref, Nref)


@pytest.mark.gpu
def test_cli_grm_alg1(tmp_path):
    _full_case(tmp_path, "g2_edge", ["--bfile", "t", "--make-grm", "--grm-alg", "1"], "o_alg1",
               ["--bfile", "t", "--make-grm-bin", "--grm-alg", "1"], "ob_alg1", same_log=False)
