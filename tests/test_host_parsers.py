"""Malformed and unusual .fam / .bim / list files: the front end must read them the way the reference's iostream loops do
(data.cpp:39-137, 520-533, 657-673) - same SNP and individual counts, same log lines, same errors - because pipelines
that feed GCTA such files see those messages.  Compared live with the compiled reference, up to the first line that
needs the device.  No GPU needed."""
import os
import subprocess

import pytest

from oracle import oracle
from tests.helpers import ROOT

BIN = os.path.join(ROOT, "gcta_b200", "gcta64_b200")
pytestmark = pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/gcta_ref not built")
N, M = 6, 5
BED = bytes([0x6c, 0x1b, 1]) + bytes([0b10011100, 0b00001110] * M)


def _fam():
    return ["F%d I%d 0 0 %d -9" % (i, i, 1 + i % 2) for i in range(N)]


def _bim():
    return ["%d\trs%d\t0\t%d\tA\tG" % (1 + k, k, k + 1) for k in range(M)]


def _cases():
    c = {}
    f = _fam(); f[2] = "F2 I2 0 0 1 -9 extra tokens"; c["fam_extra_tokens"] = (f, _bim(), [])
    f = _fam(); f[2] = "F2\tI2\t0\t0\tX\tNA"; c["fam_tabs_and_letters"] = (f, _bim(), [])
    f = _fam(); f[5] = "F5 I5 0 0 1"; c["fam_last_record_short"] = (f, _bim(), [])
    f = _fam(); f.insert(3, ""); c["fam_blank_line"] = (f, _bim(), [])
    c["fam_crlf"] = ([ln + "\r" for ln in _fam()], _bim(), [])
    f = _fam(); f[1] = f[0]; c["fam_duplicate_id"] = (f, _bim(), [])
    c["fam_no_final_newline"] = (_fam(), _bim(), ["fam_nonl"])
    b = _bim(); b[1] = "X\trs1\t0\t2\tA\tG"; c["bim_chr_letter"] = (_fam(), b, [])
    b = _bim(); b[1] = "2\trs1\t0\t2\tAC\tG"; c["bim_two_letter_allele"] = (_fam(), b, [])
    b = _bim(); b[1] = "2 rs1 0.5 2 A G extra"; c["bim_extra_token"] = (_fam(), b, [])
    b = _bim(); b[3] = "4\trs2\t0\t4\tA\tG"; c["bim_duplicate_name"] = (_fam(), b, [])
    b = _bim(); b.insert(2, ""); c["bim_blank_line"] = (_fam(), b, [])
    c["bim_crlf"] = (_fam(), [ln + "\r" for ln in _bim()], [])
    b = _bim(); b[4] = "5\trs4\t0\t5\tA"; c["bim_last_record_short"] = (_fam(), b, [])
    b = _bim(); b[0] = "1\trs0\t0\t1e3\tA\tG"; c["bim_fractional_position"] = (_fam(), b, [])
    b = _bim(); b[2] = "30\trs2\t0\t3\tA\tG"; c["bim_chr_30"] = (_fam(), b, [])
    c["bim_no_final_newline"] = (_fam(), _bim(), ["bim_nonl"])
    c["bim_empty"] = (_fam(), [], [])
    c["fam_empty"] = ([], _bim(), [])
    return c


CASES = _cases()


def _log(stdout, tag):
    lines = [ln for ln in stdout.replace("\r", "\n").split("\n") if ln.strip()]
    lines = lines[lines.index("Options:"):] if "Options:" in lines else lines
    lines = [ln.replace(tag, "X") for ln in lines if not ln.startswith(("Analysis started", "Analysis finished", "Computational time", "--out"))]
    stop = [k for k, ln in enumerate(lines) if ln.startswith(("Calculating allele", "Recoding genotypes", "Calculating the genetic"))]
    return lines[:stop[0]] if stop else lines


LIST_CASES = {
    "keep_no_final_newline": ("--keep", "F1 I1\nF2 I2"), "keep_one_column_last": ("--keep", "F1 I1\nF2"),
    "keep_one_column_last_nl": ("--keep", "F1 I1\nF2\n"), "keep_one_column_middle": ("--keep", "F1\nF2 I2\nF3 I3\n"),
    "keep_blank_lines": ("--keep", "F1 I1\n\n\nF2 I2\n"), "keep_crlf": ("--keep", "F1 I1\r\nF2 I2\r\n"), "keep_empty": ("--keep", ""),
    "keep_unknown": ("--keep", "Q Q\n"), "remove_everybody": ("--remove", "".join("F%d I%d\n" % (i, i) for i in range(N))),
    "remove_no_final_newline": ("--remove", "F0 I0"),
    "extract_no_final_newline": ("--extract", "rs1\nrs2"), "extract_crlf": ("--extract", "rs1\r\nrs2\r\n"),
    "extract_blank_and_extra": ("--extract", "rs1\n\nrs3 extra words\n"), "extract_empty": ("--extract", ""), "extract_unknown": ("--extract", "zz\n"),
    "exclude_everything": ("--exclude", "".join("rs%d\n" % k for k in range(M))),
    "sex_no_final_newline": ("--update-sex", "".join("F%d I%d %d\n" % (i, i, 1 + i % 2) for i in range(N)).rstrip("\n")),
    "sex_bad_code": ("--update-sex", "F0 I0 3\n"), "sex_partial": ("--update-sex", "F0 I0 1\n"),
    "sex_letters": ("--update-sex", "".join("F%d I%d %s\n" % (i, i, "MF"[i % 2]) for i in range(N))),
    "sex_last_record_short": ("--update-sex", "".join("F%d I%d %d\n" % (i, i, 1 + i % 2) for i in range(N - 1)) + "F5 I5"),
    "refA_no_final_newline": ("--update-ref-allele", "".join("rs%d G\n" % k for k in range(M)).rstrip("\n")),
    "refA_invalid": ("--update-ref-allele", "rs0 T\n"), "refA_partial": ("--update-ref-allele", "rs0 A\n"),
    "refA_two_letters": ("--update-ref-allele", "".join("rs%d GA\n" % k for k in range(M))),
    "freq_invalid_allele": ("--update-freq", "rs0 T 0.1\n"), "freq_out_of_range": ("--update-freq", "rs0 A 1.5\n"),
    "freq_partial": ("--update-freq", "rs0 A 0.3\n"),
    "freq_no_final_newline": ("--update-freq", "".join("rs%d A 0.%d\n" % (k, k + 1) for k in range(M)).rstrip("\n")),
    "freq_text_value": ("--update-freq", "rs0 A abc\n"),
}


@pytest.mark.parametrize("name", sorted(LIST_CASES), ids=sorted(LIST_CASES))
def test_unusual_list_files(tmp_path, name):
    flag, content = LIST_CASES[name]
    (tmp_path / "t.fam").write_text("\n".join(_fam()) + "\n")
    (tmp_path / "t.bim").write_text("\n".join(_bim()) + "\n")
    (tmp_path / "t.bed").write_bytes(BED)
    (tmp_path / "l.txt").write_bytes(content.encode())
    args = ["--bfile", "t", flag, "l.txt", "--make-grm", "--out"]
    r = oracle.run_ref(args + ["zzr"], cwd=tmp_path, timeout=60)
    m = subprocess.run([BIN] + args + ["zzm"], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60)
    assert _log(m.stdout, "zzm") == _log(r.stdout, "zzr")
    er, em = r.stderr.strip(), m.stderr.strip()
    assert er == em or (er == "" and "no CUDA device" in em)


OPTION_CASES = [
    ["--bfile", "t", "--make-grm", "--make-grm"], ["--bfile", "t", "--make-grm", "--out"], ["--bfile", "t", "--make-grm", "--maf"],
    ["--bfile", "t", "--maf", "--make-grm"], ["--bfile", "t", "--chr", "abc", "--make-grm"], ["--bfile", "t", "--chr", "0", "--make-grm"],
    ["--bfile", "t", "--chr", "101", "--make-grm"], ["--bfile", "t", "--autosome-num", "0", "--make-grm"],
    ["--bfile", "t", "--autosome-num", "101", "--make-grm"], ["--bfile", "t", "--make-grm", "--keep"], ["--bfile"], ["--make-grm"], [],
    ["--bfile", "t"], ["--bfile", "t", "--autosome"], ["--bfile", "t", "--maf", "-0.1", "--make-grm"], ["--bfile", "t", "--max-maf", "0.6", "--make-grm"],
    ["--bfile", "t", "--grm-alg", "2", "--make-grm"], ["--bfile", "t", "--grm-alg", "x", "--make-grm"], ["--bfile", "t", "--make-grm-bin", "--make-grm"],
    ["--bfile", "t", "--make-grm", "--make-grm-xchr"], ["--bfile", "t", "--rm-high-ld", "0", "--make-grm"], ["--bfile", "t", "--rm-high-ld", "1", "--make-grm"],
    ["--bfile", "t", "--freq", "--make-grm"], ["--bfile", "t", "--extract-snp", "rs1", "--exclude-snp", "rs1", "--make-grm"], ["--bfile", "t", "-make-grm"],
    ["--bfile", "t", "make-grm"], ["--BFILE", "t"], ["--bfile", "t", "--autosome", "--chr", "2", "--make-grm"],
    ["--bfile", "t", "--chr", "2", "--autosome", "--make-grm"], ["--bfile", "t", "--save-ram", "--autosome", "--make-grm"], ["--grm", "x"],
    ["--grm", "x", "--make-grm", "--grm-cutoff", "2"], ["--grm", "x", "--pca", "0"], ["--mgrm", "x", "--grm-adj", "-1", "--make-grm"],
    ["--grm", "x", "--dc", "2", "--make-grm"], ["--bfile", "t", "--out", "a", "--out", "b", "--make-grm"], ["--bfile", "t", "--update-sex", "nofile", "--make-grm"],
    ["--bfile", "t", "--update-freq", "nofile", "--make-grm"], ["--bfile", "t", "--update-ref-allele", "nofile", "--make-grm"],
]


@pytest.mark.parametrize("args", OPTION_CASES, ids=[" ".join(a) or "(none)" for a in OPTION_CASES])
def test_option_handling(tmp_path, args):
    """option.cpp:17-51, 111-135, 170-327, 547-571: echo lines, range errors, flags without their value (the next
    token or the sentinel is taken), repeated and unknown flags, nothing to do."""
    (tmp_path / "t.fam").write_text("\n".join(_fam()) + "\n")
    (tmp_path / "t.bim").write_text("\n".join(_bim()) + "\n")
    (tmp_path / "t.bed").write_bytes(BED)
    out = [] if "--out" in args else ["--out"]
    r = oracle.run_ref(args + (out and out + ["zzr"]), cwd=tmp_path, timeout=60)
    m = subprocess.run([BIN] + args + (out and out + ["zzm"]), cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60)

    def log(stdout, tag):
        lines = [ln for ln in stdout.replace("\r", "\n").split("\n") if ln.strip()]
        lines = lines[lines.index("Options:"):] if "Options:" in lines else lines
        lines = [ln.replace(tag, "X") for ln in lines if not ln.startswith(("Analysis started", "Analysis finished", "Computational time"))]
        stop = [k for k, ln in enumerate(lines) if ln.startswith(("Calculating allele", "Recoding genotypes", "Calculating the genetic"))]
        return lines[:stop[0]] if stop else lines

    assert log(m.stdout, "zzm") == log(r.stdout, "zzr")
    er, em = r.stderr.strip().replace("zzr", "X"), m.stderr.strip().replace("zzm", "X")
    assert er == em or (er == "" and "no CUDA device" in em)


@pytest.mark.parametrize("name", sorted(CASES), ids=sorted(CASES))
def test_unusual_fam_and_bim_files(tmp_path, name):
    fam, bim, flags = CASES[name]
    (tmp_path / "t.fam").write_text("\n".join(fam) + ("" if "fam_nonl" in flags or not fam else "\n"))
    (tmp_path / "t.bim").write_text("\n".join(bim) + ("" if "bim_nonl" in flags or not bim else "\n"))
    (tmp_path / "t.bed").write_bytes(BED)
    r = oracle.run_ref(["--bfile", "t", "--make-grm", "--out", "zzr"], cwd=tmp_path, timeout=60)
    m = subprocess.run([BIN, "--bfile", "t", "--make-grm", "--out", "zzm"], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                       text=True, timeout=60)
    assert _log(m.stdout, "zzm") == _log(r.stdout, "zzr")
    er, em = r.stderr.strip(), m.stderr.strip()
    assert er == em or (er == "" and "no CUDA device" in em) or (er == "" and em == "")
