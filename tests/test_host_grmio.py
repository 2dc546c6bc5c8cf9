"""Consumer side of the .grm.gz (grm.cpp:343-654): `--grm` / `--mgrm` with `--make-grm[-bin]`, and
--keep / --remove / --grm-cutoff / --grm-adj / --dc on the way, run through gcta64_b200 and through the compiled
reference on the same files.  Everything is host code in both, so the bar is byte identity: the decompressed text,
the .grm.bin, the .grm.id and the log lines.  No GPU needed.
"""
import gzip
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle
from tests.helpers import ROOT

BIN = os.path.join(ROOT, "gcta_b200", "gcta64_b200")
pytestmark = pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/gcta_ref not built")


def _write_grm(d, prefix, ids, seed, nbase, zero_n_at=None):
    r = np.random.default_rng(seed)
    n = len(ids)
    with gzip.open(d / (prefix + ".grm.gz"), "wt") as f:
        for i in range(n):
            for j in range(i + 1):
                a = 1.0 + 0.05 * r.standard_normal() if i == j else 0.03 * r.standard_normal()
                N = nbase - int(r.integers(0, 50))
                if zero_n_at == (i, j):
                    N, a = 0, 0.0
                f.write("%d\t%d\t%d\t%.6e\n" % (i + 1, j + 1, N, np.float32(a)))
    with open(d / (prefix + ".grm.id"), "w") as f:
        f.write("".join("%s\t%s\n" % t for t in ids))


@pytest.fixture(scope="module")
def data(tmp_path_factory):
    d = tmp_path_factory.mktemp("grmio")
    ids = [("F%d" % i, "I%d" % i) for i in range(60)]
    _write_grm(d, "a", ids, 1, 5000)
    _write_grm(d, "b", ids, 2, 3000, zero_n_at=(7, 3))   # one pair never observed: "Warning:" line, N = 0 in the merge
    perm = [ids[i] for i in np.random.default_rng(5).permutation(60)][:45]
    _write_grm(d, "c", perm, 3, 2000)                      # other order, fewer individuals
    (d / "list.txt").write_text("a\nb\n")
    (d / "list3.txt").write_text("a\nb\n\nc\nnot a single token\n")
    (d / "keep.txt").write_text("".join("%s %s\n" % t for t in ids[5:50:2]))
    (d / "sex.txt").write_text("".join("%s %s %d\n" % (t[0], t[1], 1 + (k % 2)) for k, t in enumerate(ids)))
    (d / "bad.grm.gz").write_bytes(gzip.compress(b"1\t1\t100\t1.0\n2\t1\t100\tabc\n2\t2\t100\t1.0\n"))
    (d / "bad.grm.id").write_text("F0\tI0\nF1\tI1\n")
    (d / "short.grm.gz").write_bytes(gzip.compress(b"1\t1\t100\t1.0\n2\t1\t100\t0.1\n"))
    (d / "short.grm.id").write_text("F0\tI0\nF1\tI1\n")
    # files in gcta64_b200's own layout (every gzip member announces its length, read by all cores): N == 0 warnings
    # spread over members, a bad line in a later member, members that cut a line in two (sequential fallback)
    text = gzip.decompress((d / "b.grm.gz").read_bytes())
    lines = text.split(b"\n")[:-1]
    lines[40] = lines[40].split(b"\t")[0] + b"\t" + lines[40].split(b"\t")[1] + b"\t0\t0.000000e+00"
    lines[1500] = lines[1500].split(b"\t")[0] + b"\t" + lines[1500].split(b"\t")[1] + b"\t0\t0.000000e+00"
    cuts = [0, 300, 900, 1400, len(lines)]
    chunks = [b"".join(ln + b"\n" for ln in lines[a:b]) for a, b in zip(cuts, cuts[1:])]
    (d / "idxwarn.grm.gz").write_bytes(b"".join(_member(c) for c in chunks))
    bad = list(lines)
    bad[1000] = b"47\t12\tx\t1.0"
    bad[1600] = b"nonsense"
    (d / "idxbad.grm.gz").write_bytes(b"".join(_member(b"".join(ln + b"\n" for ln in bad[a:b])) for a, b in zip(cuts, cuts[1:])))
    whole = b"".join(chunks)
    assert whole[7000:7001] != b"\n" and whole[20010:20011] != b"\n"
    (d / "idxsplit.grm.gz").write_bytes(_member(whole[:7001]) + _member(whole[7001:20011]) + _member(whole[20011:]))
    for name in ("idxwarn", "idxbad", "idxsplit"):
        (d / (name + ".grm.id")).write_text((d / "b.grm.id").read_text())
    return d


def _member(text: bytes) -> bytes:
    """one gzip member with the FEXTRA 'GM' subfield = its own length (host/grm_text_gz.h)"""
    import struct
    import zlib
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    body = c.compress(text) + c.flush()
    size = 20 + len(body) + 8
    head = bytes([0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 3, 8, 0]) + b"GM" + bytes([4, 0]) + struct.pack("<I", size)
    return head + body + struct.pack("<II", zlib.crc32(text), len(text) & 0xFFFFFFFF)


def _body(txt, prefix):
    lines = [ln for ln in txt.split("\n") if ln.strip()]
    i = lines.index("Options:")
    return [ln.replace(prefix, "X") for ln in lines[i:] if not ln.startswith(("Analysis started", "Analysis finished", "Computational time"))]


CASES = [
    ("resave", ["--grm", "a", "--make-grm"]),
    ("bin", ["--grm", "a", "--make-grm-bin"]),
    ("warn", ["--grm", "b", "--make-grm"]),
    ("cut", ["--grm", "a", "--grm-cutoff", "0.05", "--make-grm"]),
    ("keep", ["--grm", "a", "--keep", "keep.txt", "--make-grm"]),
    ("rm", ["--grm", "a", "--remove", "keep.txt", "--make-grm"]),
    ("adj", ["--grm", "a", "--grm-adj", "0", "--make-grm"]),
    ("dc1", ["--grm", "a", "--dc", "1", "--update-sex", "sex.txt", "--make-grm"]),
    ("dc0", ["--grm", "a", "--dc", "0", "--update-sex", "sex.txt", "--make-grm"]),
    ("mg", ["--mgrm", "list.txt", "--make-grm"]),
    ("mg3", ["--mgrm", "list3.txt", "--make-grm"]),
    ("mgbin", ["--mgrm", "list.txt", "--make-grm-bin"]),
    ("mgsup", ["--grm", "a", "--mgrm", "list.txt", "--grm-cutoff", "0.05", "--make-grm"]),   # both suppressed with warnings
    ("mgsup2", ["--mgrm", "list.txt", "--grm", "a", "--make-grm"]),   # the later name wins and is then read as a list: error
    ("all", ["--grm", "a", "--keep", "keep.txt", "--grm-cutoff", "0.04", "--grm-adj", "0.1", "--make-grm"]),
    ("dcnosex", ["--grm", "a", "--dc", "1", "--make-grm"]),            # error: --update-sex required
    ("adjrange", ["--grm", "a", "--grm-adj", "1.5", "--make-grm"]),    # error: range
    ("missing", ["--grm", "nothere", "--make-grm"]),                   # error: cannot open .grm.id
    ("badline", ["--grm", "bad", "--make-grm"]),                       # error: format, line echoed
    ("short", ["--grm", "short", "--make-grm"]),                       # error: incorrect number of lines
    ("idxwarn", ["--grm", "idxwarn", "--make-grm"]),                   # member-indexed file: warnings stay in file order
    ("idxbad", ["--grm", "idxbad", "--make-grm"]),                     # ... the first bad line in file order is reported
    ("idxsplit", ["--grm", "idxsplit", "--make-grm"]),                 # ... members that cut lines: sequential fallback
]


ODD_LINES = [
    b"1\t1\t5\t1.0", b"1 1 5 1.0", b"  1   1  5   1.0  ", b"1\t1\t5\t1.0\r", b"1\t1\t5\t.5", b"1\t1\t5\t5.", b"1\t1\t5\t1e", b"1\t1\t5\t1e+",
    b"1\t1\t5\t1.0e-3x", b"+1\t1\t5\t1.0", b"1\t+1\t5\t1.0", b"1\t1\t+5\t+1.0", b"1\t1\t5\t-0.0", b"1\t1\t5\tnan", b"1\t1\t5\tinf", b"1\t1\t5\t0x10",
    b"1\t1\t5.7\t1.0", b"1\t1\t5e2\t1.0", b"1.0\t1\t5\t1.0", b"1\t1\t5\t1.0\t", b"1\t1\t5\t1.0 extra", b"1\t1\t5", b"", b"1\t1\t-3\t1.0", b"1\t1\t0\t1.0",
    b"01\t01\t005\t001.500", b"1\t1\t5\t1.23456789012345678901234567890", b"1\t1\t5\t1e-400", b"1\t1\t5\t1e400",
    b"1\t1\t5\t123456789012345678901234567890", b"1\t1\t5\t0.000000000000000000000000000001", b"1\t1\t99999999999\t1.0", b"1\t1\t5\t1,0",
    b"2\t1\t5\t1.0", b"1\t2\t5\t1.0", b"1\t1\t5\t1.0E+00", b"1\t1\t5\t1D0", b"1\t1\t5\t 1.0\n", b"1\t1\t5\t1.0\n\n", b"\n1\t1\t5\t1.0",
    b"99999999999\t1\t5\t1.0",
]


def test_odd_lines_are_accepted_or_rejected_like_the_reference(tmp_path):
    """The tokenizer against `stringstream >> int >> int >> double >> double` (grm.cpp:428-441) on 41 one-line files:
    separators, signs, bare dots, dangling exponents, hex, nan/inf, overflow and underflow, too few / too many tokens,
    CR LF, empty lines, out-of-range indices.  Same verdict, same message, same re-saved value.  (Not compared: indices
    below 1, which the reference does not check and writes out of bounds with.)"""
    (tmp_path / "x.grm.id").write_text("F0\tI0\n")
    for k, c in enumerate(ODD_LINES):
        (tmp_path / "x.grm.gz").write_bytes(gzip.compress(c if c.endswith(b"\n") or c == b"" else c + b"\n"))
        res = []
        for tag, run in (("zzr", lambda a: oracle.run_ref(a, cwd=tmp_path, timeout=60)),
                         ("zzm", lambda a: subprocess.run([BIN] + a, cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60))):
            for ext in (".grm.gz", ".grm.id"):
                if (tmp_path / (tag + ext)).exists():
                    os.remove(tmp_path / (tag + ext))
            p = run(["--grm", "x", "--make-grm", "--out", tag])
            out = gzip.decompress((tmp_path / (tag + ".grm.gz")).read_bytes()) if (tmp_path / (tag + ".grm.gz")).exists() else None
            res.append((p.stderr.replace(tag, "X"), out, [ln for ln in p.stdout.split("\n") if ln.startswith("Warning")]))
        assert res[0] == res[1], (k, c)


ID_FILES = {
    "extra_columns": "F0 I0 x y\nF1 I1 z\nF2 I2\n", "crlf": "F0\tI0\r\nF1\tI1\r\nF2\tI2\r\n", "blank_lines": "F0 I0\n\nF1 I1\n\n\nF2 I2\n",
    "no_final_newline": "F0 I0\nF1 I1\nF2 I2", "one_column_last": "F0 I0\nF1 I1\nF2", "one_column_last_nl": "F0 I0\nF1 I1\nF2\n",
    "too_many": "F0 I0\nF1 I1\nF2 I2\nF3 I3\n", "too_few": "F0 I0\nF1 I1\n", "duplicate": "F0 I0\nF0 I0\nF2 I2\n", "empty": "",
    "tabs_and_spaces": "  F0   I0  \n\tF1\tI1\t\nF2 I2\n",
}


@pytest.mark.parametrize("name", sorted(ID_FILES), ids=sorted(ID_FILES))
def test_unusual_grm_id_files(tmp_path, name):
    """.grm.id read like grm.cpp:392-402 (extra columns, CR LF, blank lines, a last line with one token, more or fewer
    IDs than the matrix has rows, duplicates), alone and followed by --keep and --grm-cutoff."""
    text = "".join("%d\t%d\t%d\t%.6e\n" % (i + 1, j + 1, 100 + i + j, 1.0 if i == j else 0.01 * (i - j)) for i in range(3) for j in range(i + 1))
    (tmp_path / "x.grm.gz").write_bytes(gzip.compress(text.encode()))
    (tmp_path / "x.grm.id").write_bytes(ID_FILES[name].encode())
    (tmp_path / "k.txt").write_text("F0 I0\nF2 I2\n")
    for extra in ([], ["--keep", "k.txt"], ["--grm-cutoff", "0.015"]):
        res = []
        for tag, binary in (("zzr", oracle.REF_BIN), ("zzm", BIN)):
            p = subprocess.run([binary, "--grm", "x"] + extra + ["--make-grm", "--out", tag], cwd=tmp_path, stdout=subprocess.PIPE,
                               stderr=subprocess.PIPE, text=True, timeout=60)
            files = []
            for ext in (".grm.gz", ".grm.id"):
                f = tmp_path / (tag + ext)
                files.append(None if not f.exists() else gzip.decompress(f.read_bytes()) if ext == ".grm.gz" else f.read_bytes())
                if f.exists():
                    os.remove(f)
            res.append((_body(p.stdout, tag), p.stderr.replace(tag, "X"), files))
        assert res[0] == res[1], (name, extra)


MGRM_LISTS = {
    "plain": "a\nb\n", "commas_semicolons": "a,\nb;\n", "two_tokens_on_a_line": "a b\n", "tabs_and_spaces": "\ta\t\n  b  \n", "crlf": "a\r\nb\r\n",
    "same_file_twice": "a\na\n", "single": "a\n", "empty": "", "blank_only": "\n\n", "missing_file": "a\nnothere\n", "no_final_newline": "a\nb",
    "comment_like": "a\nb\n#comment\n",
}


@pytest.mark.parametrize("name", sorted(MGRM_LISTS), ids=sorted(MGRM_LISTS))
def test_unusual_mgrm_lists(tmp_path, name):
    """The --mgrm file list is split like StrFunc::split_string (" ,\\t;\\n"); only lines with exactly one token count
    (grm.cpp:656-672)."""
    for nm, s in (("a", 0), ("b", 5)):
        text = "".join("%d\t%d\t%d\t%.6e\n" % (i + 1, j + 1, 100 + i + j + s, 1.0 if i == j else 0.01 * (i - j) + 0.001 * s)
                       for i in range(3) for j in range(i + 1))
        (tmp_path / (nm + ".grm.gz")).write_bytes(gzip.compress(text.encode()))
        (tmp_path / (nm + ".grm.id")).write_text("F0 I0\nF1 I1\nF2 I2\n")
    (tmp_path / "l.txt").write_bytes(MGRM_LISTS[name].encode())
    res = []
    for tag, binary in (("zzr", oracle.REF_BIN), ("zzm", BIN)):
        p = subprocess.run([binary, "--mgrm", "l.txt", "--make-grm", "--out", tag], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                           text=True, timeout=60)
        f = tmp_path / (tag + ".grm.gz")
        res.append((_body(p.stdout, tag), p.stderr.replace(tag, "X"), gzip.decompress(f.read_bytes()) if f.exists() else None))
    assert res[0] == res[1]


@pytest.mark.parametrize("seed", range(16))
def test_randomised_grm_manipulation(tmp_path, seed):
    """Random GRMs with many related pairs (so that --grm-cutoff has ties to break), random sizes down to n = 1, random
    keep lists, cutoffs, adjustment factors and merges of two or three files with shuffled / partial ID lists."""
    rng = np.random.default_rng(900 + seed)
    n = int(rng.integers(1, 45))
    ids = [("F%d" % i, "I%d" % i) for i in range(n)]

    def write(prefix, idl, scale):
        k = len(idl)
        with gzip.open(tmp_path / (prefix + ".grm.gz"), "wt") as f:
            for i in range(k):
                for j in range(i + 1):
                    a = 1.0 + 0.1 * rng.standard_normal() if i == j else scale * rng.standard_normal()
                    f.write("%d\t%d\t%d\t%.6e\n" % (i + 1, j + 1, int(rng.integers(0 if seed % 5 == 0 else 1, 4000)), np.float32(a)))
        (tmp_path / (prefix + ".grm.id")).write_text("".join("%s\t%s\n" % t for t in idl))

    write("a", ids, 0.2)
    args = ["--grm", "a"]
    if seed % 4 == 3 and n > 1:
        perm = [ids[i] for i in rng.permutation(n)]
        write("b", perm, 0.05)
        write("c", perm[: max(1, n - int(rng.integers(0, 3)))], 0.05)
        (tmp_path / "list.txt").write_text("a\nb\n" + ("c\n" if seed % 8 == 7 else ""))
        args = ["--mgrm", "list.txt"]
    else:
        if seed % 2 == 1 and n > 2:
            keep = sorted(rng.choice(n, size=int(rng.integers(1, n)), replace=False))
            (tmp_path / "k.txt").write_text("".join("F%d I%d\n" % (i, i) for i in keep))
            args += ["--keep" if seed % 3 else "--remove", "k.txt"]
        if seed % 3 != 1:
            args += ["--grm-cutoff", "%.3f" % float(rng.choice([0.0, 0.05, 0.15, 0.3]))]
        if seed % 4 == 2 and n > 3:
            args += ["--grm-adj", "%.2f" % float(rng.choice([0.0, 0.2, 1.0]))]
    args += ["--make-grm-bin" if seed % 6 == 4 else "--make-grm"]
    r = oracle.run_ref(args + ["--out", "zzr"], cwd=tmp_path, timeout=300)
    m = subprocess.run([BIN] + args + ["--out", "zzm"], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert _body(m.stdout, "zzm") == _body(r.stdout, "zzr"), " ".join(args)
    assert m.stderr.replace("zzm", "X") == r.stderr.replace("zzr", "X")
    for ext in (".grm.gz", ".grm.bin", ".grm.id"):
        rf, mf = tmp_path / ("zzr" + ext), tmp_path / ("zzm" + ext)
        assert rf.exists() == mf.exists(), ext
        if rf.exists():
            if ext == ".grm.gz":
                assert gzip.decompress(mf.read_bytes()) == gzip.decompress(rf.read_bytes()), " ".join(args)
            else:
                assert mf.read_bytes() == rf.read_bytes(), " ".join(args)


@pytest.mark.parametrize("name,args", CASES, ids=[c[0] for c in CASES])
def test_grm_manipulation_is_byte_identical_to_the_reference(data, name, args):
    rp, mp = "r_" + name, "m_" + name
    r = oracle.run_ref(args + ["--out", rp], cwd=data, timeout=300)
    m = subprocess.run([BIN] + args + ["--out", mp], cwd=data, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert m.returncode == 0 and r.returncode == 0
    assert _body(m.stdout, mp) == _body(r.stdout, rp)
    assert m.stderr.replace(mp, "X") == r.stderr.replace(rp, "X")
    produced = False
    for ext in (".grm.gz", ".grm.bin", ".grm.id"):
        rf, mf = data / (rp + ext), data / (mp + ext)
        assert rf.exists() == mf.exists(), ext
        if not rf.exists():
            continue
        produced = True
        if ext == ".grm.gz":
            assert gzip.decompress(mf.read_bytes()) == gzip.decompress(rf.read_bytes())
        else:
            assert mf.read_bytes() == rf.read_bytes()
    assert produced == (name not in ("mgsup2", "dcnosex", "adjrange", "missing", "badline", "short", "idxbad"))
