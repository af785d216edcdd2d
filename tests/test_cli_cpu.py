"""Host logic of the drop-in CLI (smithwaterman_b200/cli.py) on CPU: option parsing (OCL-main.rs:33-116), the list-file
reader (OCL-main.rs:268-288) and the batched clustering loop against a literal serial restatement of OCL-main.rs:149-243,
with the oracle standing in for the device alignments."""
import io

import numpy as np
import pytest

from oracle import oracle as O
from smithwaterman_b200 import cli
from smithwaterman_b200.engine import ALIGN_GLOBAL, ALIGN_LOCAL, SeqData

AA = "ARNDCQEGHILKMFPSTWYV"


def test_options():
    o = cli.Options(["-global", "a.fas", "b.fas"])
    assert (o.alignment_type, o.file1, o.file2, o.list, o.clustering) == (ALIGN_GLOBAL, "a.fas", "b.fas", False, False)
    o = cli.Options(["-list", "l.txt", "-out", "x"])
    assert o.list and o.file1 == "l.txt" and o.outfilename == "x" and o.alignment_type == ALIGN_LOCAL
    o = cli.Options(["-clustering", "-identity", "0.5", "-coverage_long", "0.7", "-out", "o", "s.fas"])
    assert o.clustering and o.c_identity == 0.5 and o.c_coverage_long == 0.7 and o.c_coverage_short is None
    for bad in (["a.fas"], ["-list", "-cluster", "x", "-out", "o"], ["-cluster", "x.fas"], ["-nope", "a", "b"], ["a", "b", "c"]):
        with pytest.raises(ValueError):
            cli.Options(bad)


def test_read_list(tmp_path):
    p = tmp_path / "l.txt"
    p.write_text("a.fas\tb.fas\nc.fas d.fas\nonly\nx\ty\tz\n")
    out = io.StringIO()
    assert cli.read_list(str(p), out) == [("a.fas", "b.fas"), ("c.fas", "d.fas")]
    assert "is ignoed" in out.getvalue() and "Only x y are used" in out.getvalue()


def _serial_cluster(seqs, align, identity, cshort, clong):
    """Literal restatement of OCL-main.rs:150-243, one alignment per call."""
    seq1 = sorted(seqs, key=lambda s: -len(s.seq))
    n = len(seq1)
    cluster_of, identical = list(range(n)), list(range(n))
    members, identical_members = [[] for _ in range(n)], [[] for _ in range(n)]
    for ii in range(n):
        if identical[ii] != ii:
            continue
        identical_members[ii].append(ii)
        si = "".join(seq1[ii].seq)
        for jj in range(ii + 1, n):
            if identical[jj] != jj:
                continue
            sj = "".join(seq1[jj].seq)
            if len(si) != len(sj):
                break
            if si == sj:
                identical[jj] = ii
                identical_members[ii].append(jj)
    for ii in range(n):
        if cluster_of[ii] != ii:
            continue
        if identical[ii] != ii:
            continue
        members[ii] += identical_members[ii]; identical_members[ii] = []
        si = "".join(seq1[ii].seq)
        for jj in range(ii + 1, n):
            if cluster_of[jj] != jj:
                continue
            if identical[jj] != jj:
                continue
            sj = "".join(seq1[jj].seq)
            if sj in si:
                if len(sj) / len(si) >= clong:
                    cluster_of[jj] = ii
                    members[ii] += identical_members[jj]; identical_members[jj] = []
            else:
                r0, r1 = align(seq1[ii], seq1[jj])
                alen = sum(1 for c in r0 if c != "-"); blen = sum(1 for c in r1 if c != "-")
                matchnum = sum(1 for x, y in zip(r0, r1) if x != "-" and y != "-" and x == y)
                lcov, scov = alen / len(seq1[ii].seq), blen / len(seq1[jj].seq)
                ident = matchnum / len(r0) if len(r0) else float("nan")
                if lcov >= clong and scov >= cshort and ident >= identity:
                    cluster_of[jj] = ii
                    members[ii] += identical_members[jj]; identical_members[jj] = []
    return seq1, cluster_of, members


@pytest.mark.parametrize("thr", [(0.8, 0.8, 0.8), (0.5, 0.6, 0.4)])
def test_cluster_batched_equals_serial(thr, tmp_path):
    rng = np.random.default_rng(11)
    base = ["".join(rng.choice(list(AA), int(k))) for k in rng.integers(40, 120, 6)]
    seqs = []
    for k in range(60):
        b = list(base[int(rng.integers(0, len(base)))])
        mode = rng.integers(0, 5)
        if mode == 1:                                   # point mutations
            for p in rng.integers(0, len(b), max(1, len(b) // int(rng.integers(4, 30)))):
                b[p] = AA[rng.integers(0, 20)]
        elif mode == 2:                                 # substring
            lo = int(rng.integers(0, len(b) // 3)); b = b[lo:lo + int(len(b) * rng.uniform(0.5, 1.0))]
        elif mode == 3:                                 # unrelated
            b = list(rng.choice(list(AA), int(rng.integers(30, 100))))
        seqs.append(SeqData("s%d" % k, "d%d" % k, b))
    oeng = O.Engine(O.Matrix.blosum62(), 10.0, 0.5, ALIGN_LOCAL)

    def one(a, b):
        r1, r2, _ = oeng.align("".join(a.seq), "".join(b.seq), False)
        return "".join(r1), "".join(r2)
    calls = []

    def many(rep, cands):
        calls.append(len(cands))
        return [one(rep, c) for c in cands]
    got = cli.cluster(seqs, many, *thr, out=io.StringIO())
    want = _serial_cluster(seqs, one, *thr)
    assert [s.name for s in got[0]] == [s.name for s in want[0]]
    assert got[1] == want[1] and got[2] == want[2]
    assert len(set(got[1])) < len(seqs) and max(calls) > 1          # something clustered, and batches were batches
    cli.write_clusters(str(tmp_path / "o.fas"), *got)
    reps = (tmp_path / "o.fas").read_text().split("\n")
    assert reps[0] == ">%s %s" % (got[0][0].name, got[0][0].desc) and reps[1] == "".join(got[0][0].seq)
    lines = (tmp_path / "o.fas.clstr").read_text().strip().split("\n")
    assert sorted(x for ln in lines for x in ln.split(" ")) == sorted(s.name for s in seqs)


def _swbcli():
    from smithwaterman_b200 import _lib
    _lib.build()
    return _lib.build_cli()


def test_cpp_host_mirror_fasta_and_formatting(tmp_path):
    """include/swb200.hpp (the C++ mirror of the reference's Rust API) reads FASTA and prints scores like the Python
    mirror (SA.rs:830-887 header rules; Rust `{}` of an f32, SA-main.rs:40).  No device needed."""
    import subprocess
    from smithwaterman_b200.engine import SeqData, rust_f32_display
    p = tmp_path / "odd.fas"
    p.write_text(">a1 first  description here\nACDE FGH\r\nIKL\n>b2\nMNPQ\n  > c3\ttabbed desc \nRST\n>\nVW\nno header > inside name x\nYY\n")
    out = subprocess.run([_swbcli(), "-dumpfasta", str(p)], capture_output=True, text=True, check=True).stdout.splitlines()
    want = ["name=%s|desc=%s|seq=%s" % (s.name, s.desc, "".join(s.seq)) for s in SeqData.load_fasta(str(p), False)]
    assert out == want and len(out) >= 4
    vals = ["485", "54.5", "-1770.5", "0", "0.1", "1234567.5", "-0.25", "3.4028235e38", "16777216", "0.3"]
    got = subprocess.run([_swbcli(), "-f32"] + vals, capture_output=True, text=True, check=True).stdout.split()
    assert got == [rust_f32_display(np.float32(v)) for v in vals]


def test_cpp_cli_option_errors():
    import subprocess
    exe = _swbcli()
    assert subprocess.run([exe], capture_output=True).returncode == 255                      # usage (SA-main.rs:12-15)
    for bad in (["-nope", "a", "b"], ["-cluster", "x.fas"], ["-list", "-cluster", "x", "-out", "o"], ["a", "b", "c"]):
        r = subprocess.run([exe] + bad, capture_output=True, text=True)
        assert r.returncode == 1 and "error" in r.stderr, bad


def test_perl_harness_rules_on_oracle_records(tmp_path):
    """test/check_results.pl (the reference's acceptance rules restated) accepts the records of the oracle for all
    100 pairs x 3 modes and rejects a corrupted one -- the CPU half of the harness test (the GPU half runs swbcli)."""
    import gzip
    import json
    import os
    import shutil
    import subprocess
    import sys
    from conftest import GOLDEN, ROOT
    if not shutil.which("perl"):
        pytest.skip("perl not installed")
    subprocess.run([sys.executable, os.path.join(ROOT, "test", "materialize.py"), str(tmp_path)], check=True)
    pairs = json.load(gzip.open(os.path.join(GOLDEN, "emboss.json.gz")))
    os.makedirs(tmp_path / "res_swb")
    m = O.Matrix.blosum62()
    for mode, tag in ((O.ALIGN_LOCAL, "local"), (O.ALIGN_GLOCAL, "glocal"), (O.ALIGN_GLOBAL, "global")):
        eng = O.Engine(m, 10.0, 0.5, mode)
        for p in pairs:
            r1, r2, sc = eng.align(p["seq1"], p["seq2"], True)
            (tmp_path / "res_swb" / ("seq%d.1.fas.res.%s" % (p["id"], tag))).write_text(O.cli_record("s1", "s2", r1, r2, sc, mode))
    chk = ["perl", os.path.join(ROOT, "test", "check_results.pl")]
    r = subprocess.run(chk, cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    assert r.stdout.splitlines()[-2:] == ["OK", "Checked 300 results"]
    f = tmp_path / "res_swb" / "seq7.1.fas.res.global"
    txt = f.read_text().split("\n")
    k = next(i for i, l in enumerate(txt) if l.startswith(">s2")) + 1
    pos = txt[k].index("-") if "-" in txt[k] else 5
    txt[k] = txt[k][:pos] + txt[k][pos + 1:] + "-"                 # move one gap to the end: still spells the input, differs from EMBOSS
    f.write_text("\n".join(txt))
    r = subprocess.run(chk, cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode != 0 and "differs from EMBOSS" in r.stderr
