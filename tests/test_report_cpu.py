"""srspair report writer (smithwaterman_b200/report.py) against the bodies of all 300 EMBOSS golden
reports of the reference (test/emboss_results/*.dat, committed as tests/golden/emboss_reports.json.gz by
tests/golden/make_emboss_reports.py).  Formatting only: runs without a GPU."""
import gzip
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import oracle as O
from smithwaterman_b200 import report
from smithwaterman_b200.engine import SubstitutionMatrix, ALIGN_LOCAL


@pytest.fixture(scope="module")
def reports():
    with gzip.open(os.path.join(GOLDEN, "emboss_reports.json.gz")) as f:
        return json.loads(f.read().decode())


def test_srspair_body_equals_all_golden_reports(emboss, reports):
    m = SubstitutionMatrix.get_blosum62_matrix()
    by_id = {p["id"]: p for p in emboss}
    n = 0
    for r in reports:
        g = by_id[r["id"]][r["kind"]]
        sa, sb = (r["first1"], r["first2"]) if r["kind"] == "local" else (1, 1)
        body = report.srspair_body(g["s1"], g["s2"], g["score"], m, 10.0, 0.5, sa, sb)
        assert body == r["body"], "%s %d" % (r["kind"], r["id"])
        n += 1
    assert n == 300


def test_starts_from_index_lists(emboss, reports):
    """The residue numbers water prints are the 1-based first aligned residues of the reference's own alignment."""
    m = SubstitutionMatrix.get_blosum62_matrix()
    om = O.Matrix(m.index_to_string, m.scores)
    eng = O.Engine(om, 10.0, 0.5, ALIGN_LOCAL)
    first = {r["id"]: (r["first1"], r["first2"]) for r in reports if r["kind"] == "local"}
    for p in emboss[:40]:
        xa, xb, sc, ei, ej = eng.raw(om.seq_to_index(p["seq1"]), om.seq_to_index(p["seq2"]))
        assert report.starts_of(xa, xb) == first[p["id"]]


def test_program_header_and_errors():
    m = SubstitutionMatrix.get_blosum62_matrix()
    txt = report.srspair_report("water", "inputs/seq0.1.fas", "inputs/seq0.2.fas", "out.dat", aln_a="AC-GT", aln_b="ACWGT",
                                score=12.5, matrix=m, gap_open=10.0, gap_extend=0.5)
    assert txt.startswith("########################################\n# Program: water\n")
    assert "# Gaps:           1/5 (20.0%)" in txt and "# Identity:       4/5 (80.0%)" in txt
    with pytest.raises(ValueError):
        report.srspair_body("AC", "A", 1.0, m, 10.0, 0.5)


def test_cpp_mirror_srspair_equals_all_golden_reports(emboss, reports, tmp_path):
    """include/swb200.hpp::srspair_body (the compiled host mirror) prints the same 300 bodies."""
    import subprocess
    from smithwaterman_b200 import _lib
    _lib.build()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "srspair")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-I", os.path.join(root, "include"), os.path.join(root, "tests", "cpp", "srspair.cpp"),
                           "-L", os.path.dirname(_lib.SO_PATH), "-lswb200", "-Wl,-rpath," + os.path.dirname(_lib.SO_PATH), "-o", exe])
    by_id = {p["id"]: p for p in emboss}
    rows = []
    for r in reports:
        g = by_id[r["id"]][r["kind"]]
        sa, sb = (r["first1"], r["first2"]) if r["kind"] == "local" else (1, 1)
        rows.append("\t".join([r["kind"], str(r["id"]), str(sa), str(sb), repr(g["score"]), g["s1"], g["s2"]]))
    (tmp_path / "cases.tsv").write_text("\n".join(rows) + "\n")
    out = subprocess.run([exe, str(tmp_path / "cases.tsv")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    got = {}
    for chunk in out.stdout.split("@@ ")[1:]:
        head, body = chunk.split("\n", 1)
        kind, k = head.split()
        got[(kind, int(k))] = body
    assert len(got) == 300
    for r in reports:
        assert got[(r["kind"], r["id"])] == r["body"], "%s %d" % (r["kind"], r["id"])
