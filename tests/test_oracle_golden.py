"""Pins the CPU oracle (oracle/) to the reference's own known answers:
Rust #[test] KATs (SA.rs:921-1221, OCL.rs:1292-1448), the 300 EMBOSS results of test/emboss_results,
the FASTA edge cases (SA.rs:1090-1130) and the matrix parser test (SA.rs:897-919).  CPU only."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, trim_terminal
from oracle import oracle as O


def make_matrix(spec):
    m = O.Matrix.blosum62() if spec["kind"] == "blosum62" else O.Matrix.match_mismatch(spec["match"], spec["mismatch"])
    if spec.get("scale", 1.0) != 1.0:
        m = O.Matrix(m.letters, m.scores * np.float32(spec["scale"]))
    return m


def test_rust_kats(rust_kats):
    assert len(rust_kats) == 26
    for c in rust_kats:
        m = make_matrix(c["matrix"])
        eng = O.Engine(m, c["go"], c["ge"], c["mode"])
        pssm = None
        if c["pssm"]:                                  # psm_test SA.rs:1057-1086: table[a + b*la]
            ia, ib = m.seq_to_index(c["seq1"]), m.seq_to_index(c["seq2"])
            pssm = m.scores[ia][:, ib].T.copy().ravel()
        r1, r2, sc = eng.align(c["seq1"], c["seq2"], c["retain_all"], pssm=pssm)
        if "score" in c:
            assert sc == c["score"], c["test"]
        if "r1" in c:
            assert "".join(r1) == c["r1"], c["test"]
            assert "".join(r2) == c["r2"], c["test"]


@pytest.mark.parametrize("mode,key", [(O.ALIGN_LOCAL, "local"), (O.ALIGN_GLOCAL, "glocal"), (O.ALIGN_GLOBAL, "global")])
def test_emboss_goldens(emboss, mode, key):
    assert len(emboss) == 100
    eng = O.Engine(O.Matrix.blosum62(), 10.0, 0.5, mode)
    for rec in emboss:
        r1, r2, sc = eng.align(rec["seq1"], rec["seq2"], True)
        s1, s2 = "".join(r1), "".join(r2)
        assert s1.replace("-", "") == rec["seq1"] and s2.replace("-", "") == rec["seq2"]   # check_results.pl:244-263
        if mode == O.ALIGN_LOCAL:
            s1, s2 = trim_terminal(s1, s2)
        assert s1 == rec[key]["s1"], (key, rec["id"])
        assert s2 == rec[key]["s2"], (key, rec["id"])
        assert sc == rec[key]["score"], (key, rec["id"])


def test_int_twin_equals_f32_on_goldens(emboss):
    m = O.Matrix.blosum62()
    mi = np.rint(m.scores * 2).astype(np.int32)
    seqs_a = [m.seq_to_index(r["seq1"]) for r in emboss]
    seqs_b = [m.seq_to_index(r["seq2"]) for r in emboss]
    a, oa = O.pack(seqs_a)
    b, ob = O.pack(seqs_b)
    for mode in (0, 1, 2):
        f, fi, fj = O.score_batch_f32(a, oa, b, ob, m.scores, 10.0, 0.5, mode, nthreads=4)
        s, si, sj = O.score_batch_int(a, oa, b, ob, mi, -20, -1, 2, mode, nthreads=4)
        assert np.array_equal(np.rint(f * 2).astype(np.int64), s)
        assert np.array_equal(fi, si) and np.array_equal(fj, sj)


def test_int_twin_fuzz():
    rng = np.random.default_rng(7)
    m = O.Matrix.blosum62()
    for trial in range(30):
        n = 40
        la = rng.integers(0, 60, n)
        lb = rng.integers(0, 60, n)
        la[0], lb[0] = 0, 5
        la[1], lb[1] = 5, 0
        alpha = 4 if trial % 2 else 20
        sa = [rng.integers(0, alpha, k).astype(np.int32) for k in la]
        sb = [rng.integers(0, alpha, k).astype(np.int32) for k in lb]
        a, oa = O.pack(sa)
        b, ob = O.pack(sb)
        go, ge = [(10, 0.5), (5, 2), (0.5, 0.5), (1, 0), (3, 1.5)][trial % 5]
        mi = np.rint(m.scores * 2).astype(np.int32)
        for mode in (0, 1, 2):
            f, fi, fj = O.score_batch_f32(a, oa, b, ob, m.scores, go, ge, mode)
            s, si, sj = O.score_batch_int(a, oa, b, ob, mi, -int(go * 2), -int(ge * 2), 2, mode)
            assert np.array_equal(np.rint(f * 2).astype(np.int64), s), (trial, mode)
            assert np.array_equal(fi, si) and np.array_equal(fj, sj), (trial, mode)


def test_fasta_edge_cases():
    with open(os.path.join(GOLDEN, "fasta_test1.json")) as f:
        expect = json.load(f)
    fas = O.load_fasta(os.path.join(GOLDEN, "fasta_test1.fas"))
    assert len(fas) == len(expect)
    for rec, (name, desc, seq) in zip(fas, expect):
        assert rec.name == name and rec.desc == desc and "".join(rec.seq) == seq


def test_matrix_parser():                               # SA.rs:897-919
    m = O.Matrix.load_from_lines(["  A T G C \n", "A 1 -3 -3 -3 ", "T -3  1 -3 -3 ", "G -3  -3  1 -3 ", "C -3  -3  -3  1"])
    for x in "ATGC":
        assert m.get_score_str(x, x) == 1.0
    for x, y in ("AC", "TG", "GA", "CT"):
        assert m.get_score_str(x, y) == -3.0


def test_rust_display():
    assert O.rust_f32_display(485.0) == "485"
    assert O.rust_f32_display(54.5) == "54.5"
    assert O.rust_f32_display(-1770.5) == "-1770.5"
    assert O.rust_f32_display(0.0) == "0"


def test_faithful_layout_equals_flat(emboss):
    """oracle/sw_oracle_faithful.c (the reference's own data structure: one heap-owning cell per DP cell, an indirect
    call per score lookup -- the `ref-faithful` CPU flavour of BASELINE.md section 3) shares the fill body with the flat
    port and gives the same scores and end cells on the golden pairs, all modes."""
    m = O.Matrix.blosum62()
    sa = [m.seq_to_index(r["seq1"]) for r in emboss[:40]]
    sb = [m.seq_to_index(r["seq2"]) for r in emboss[:40]]
    a, oa = O.pack(sa)
    b, ob = O.pack(sb)
    for mode in (O.ALIGN_GLOBAL, O.ALIGN_GLOCAL, O.ALIGN_LOCAL):
        flat = O.score_batch_f32(a, oa, b, ob, m.scores, 10.0, 0.5, mode, 4)
        faith = O.score_batch_f32(a, oa, b, ob, m.scores, 10.0, 0.5, mode, 4, layout="faithful")
        for x, y in zip(flat, faith):
            assert np.array_equal(x, y)
