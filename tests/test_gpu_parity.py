"""GPU parity tests: the CUDA engine (through the C ABI) against the CPU oracle and the committed
golden fixtures.  Bit-exact: scores, end cells, index lists and alignment strings."""
import numpy as np
import pytest

from conftest import trim_terminal
import smithwaterman_b200 as swb
from oracle import oracle as O

pytestmark = pytest.mark.gpu

AA = "ARNDCQEGHILKMFPSTWYV"


def _matrix(spec):
    m = swb.SubstitutionMatrix.get_blosum62_matrix() if spec["kind"] == "blosum62" else \
        swb.SubstitutionMatrix.get_mat_matrix(spec["match"], spec["mismatch"])
    if spec.get("scale", 1.0) != 1.0:
        m = swb.SubstitutionMatrix(m.index_to_string, m.scores * np.float32(spec["scale"]))
    return m


def _omatrix(m):
    return O.Matrix(m.index_to_string, m.scores)


def test_rust_kats(rust_kats):
    """SA.rs:921-1221 / OCL.rs:1292-1448 known answers, including the position-specific-matrix case."""
    done = 0; n_pssm = 0
    for c in rust_kats:
        if c["pssm"]:                                              # psm_test SA.rs:1057-1086: PositionSpecificMatrix filled from BLOSUM62
            sm = _matrix(c["matrix"])
            pm = swb.PositionSpecificMatrix()
            eng = swb.SequenceAlignment(pm, c["go"], c["ge"], c["mode"])
            eng.prepare(c["seq1"], c["seq2"])
            i1 = [sm.string_to_index.get(x, sm.string_to_index["X"]) for x in c["seq1"]]
            i2 = [sm.string_to_index.get(x, sm.string_to_index["X"]) for x in c["seq2"]]
            pm.scores[:len(i1) * len(i2)] = sm.scores[np.ix_(i2, i1)].reshape(-1)      # scores[a + b*a_length]
            assert pm.get_score(3, 5) == sm.scores[i1[3], i2[5]]
            r1, r2, sc = eng.align(c["seq1"], c["seq2"], c["retain_all"])
            assert (sc, "".join(r1), "".join(r2)) == (c["score"], c["r1"], c["r2"]), c["test"]
            eng.close()
            n_pssm += 1
            continue
        eng = swb.SequenceAlignment(_matrix(c["matrix"]), c["go"], c["ge"], c["mode"])
        r1, r2, sc = eng.align(swb.SeqData.create("", "", c["seq1"]), swb.SeqData.create("", "", c["seq2"]), c["retain_all"])
        if "score" in c:
            assert sc == c["score"], c["test"]
        if "r1" in c:
            assert "".join(r1) == c["r1"], c["test"]
            assert "".join(r2) == c["r2"], c["test"]
        done += 1
        eng.close()
    assert done == 25 and n_pssm == 1


@pytest.mark.parametrize("mode,key", [(swb.ALIGN_LOCAL, "local"), (swb.ALIGN_GLOCAL, "glocal"), (swb.ALIGN_GLOBAL, "global")])
def test_emboss_goldens(emboss, mode, key):
    """config[0]: the 100 test/inputs pairs vs test/emboss_results (strings + scores) and vs the oracle
    (index lists, end cells), through the batch-list call."""
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
    res = eng.align_list([r["seq1"] for r in emboss], [r["seq2"] for r in emboss], retain_all=True)
    oeng = O.Engine(_omatrix(m), 10.0, 0.5, mode)
    for rec, out in zip(emboss, res):
        s1, s2 = out["aln_a"], out["aln_b"]
        assert s1.replace("-", "") == rec["seq1"] and s2.replace("-", "") == rec["seq2"]
        if mode == swb.ALIGN_LOCAL:
            s1, s2 = trim_terminal(s1, s2)
        assert s1 == rec[key]["s1"] and s2 == rec[key]["s2"], (key, rec["id"])
        assert out["score"] == rec[key]["score"], (key, rec["id"])
        xa, xb, sc, ei, ej = oeng.raw(oeng.matrix.seq_to_index(rec["seq1"]), oeng.matrix.seq_to_index(rec["seq2"]))
        assert np.array_equal(out["idx_a"], xa) and np.array_equal(out["idx_b"], xb)
        assert (out["end_a"], out["end_b"]) == (ei, ej)
    # the score-only batch call agrees with the traceback call
    sc, ea, eb = eng.score_batch([r["seq1"] for r in emboss], [r["seq2"] for r in emboss])
    assert [float(x) for x in sc] == [o["score"] for o in res]
    assert [int(x) for x in ea] == [o["end_a"] for o in res] and [int(x) for x in eb] == [o["end_b"] for o in res]
    eng.close()


def _rand_seqs(rng, n, lo, hi, alphabet=AA, weights=None):
    ls = rng.integers(lo, hi + 1, n)
    letters = np.frombuffer(alphabet.encode(), np.uint8)
    return [bytes(letters[rng.choice(len(letters), k, p=weights)]) for k in ls]


def _oracle_scores(m, sa, sb, go, ge, mode, scale=2):
    om = _omatrix(m)
    a, oa = O.pack([om.seq_to_index(s.decode()) for s in sa])
    b, ob = O.pack([om.seq_to_index(s.decode()) for s in sb])
    mi = np.rint(m.scores * scale).astype(np.int32)
    return O.score_batch_int(a, oa, b, ob, mi, -int(round(go * scale)), -int(round(ge * scale)), scale, mode, nthreads=8)


@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL])
def test_score_batch_random(mode):
    """K1 (local, packed s16x2) / KW (other modes) vs the integer oracle: score and end cell, ragged shapes."""
    rng = np.random.default_rng(20242 + mode)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
    n = 3000 if mode == swb.ALIGN_LOCAL else 600
    sa = _rand_seqs(rng, n, 1, 200)
    sb = _rand_seqs(rng, n, 1, 200)
    # shapes around the 16/32-column lane boundaries and 16-row chunks; related pairs; repeats (tie-breaking)
    for la, lb in [(1, 1), (1, 33), (33, 1), (15, 16), (16, 16), (17, 17), (31, 32), (32, 32), (33, 33), (64, 64),
                   (100, 31), (100, 65), (300, 500), (500, 100), (129, 257)]:
        sa.append(_rand_seqs(rng, 1, la, la)[0]); sb.append(_rand_seqs(rng, 1, lb, lb)[0])
    for k in range(200):
        base = _rand_seqs(rng, 1, 50, 400)[0]
        mut = bytearray(base)
        for p in rng.integers(0, len(mut), max(1, len(mut) // 6)):
            mut[p] = ord(AA[rng.integers(0, 20)])
        cut = rng.integers(0, len(mut) // 2)
        sa.append(base); sb.append(bytes(mut[cut:]))
    for k in range(60):
        sa.append(b"K" * int(rng.integers(1, 120))); sb.append(b"K" * int(rng.integers(1, 120)))
        sa.append((b"AKL" * 50)[: int(rng.integers(3, 150))]); sb.append((b"KLA" * 50)[: int(rng.integers(3, 150))])
    sc, ea, eb = eng.score_batch(sa, sb)
    osc, oi, oj = _oracle_scores(m, sa, sb, 10.0, 0.5, mode)
    assert np.array_equal(np.rint(sc.astype(np.float64) * 2).astype(np.int64), osc)
    assert np.array_equal(ea, oi) and np.array_equal(eb, oj)
    eng.close()


def test_score_batch_edge_cases():
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    for mode in (swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL):
        eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
        sa = [b"", b"ACD", b"", b"W", b"AAAA", b"acdx", b"BZX*"]      # empty, ragged, lower-case -> X, special letters
        sb = [b"ACD", b"", b"", b"W", b"WWWW", b"ACDX", b"BZX*"]
        sc, ea, eb = eng.score_batch(sa, sb)
        osc, oi, oj = _oracle_scores(m, sa, sb, 10.0, 0.5, mode)
        assert np.array_equal(np.rint(sc.astype(np.float64) * 2).astype(np.int64), osc), mode
        assert np.array_equal(ea, oi) and np.array_equal(eb, oj), mode
        assert len(eng.score_batch([], [])[0]) == 0
        eng.close()


def test_s16_overflow_rerun():
    """Scores beyond the s16 lanes are re-run in int32 (overflow path must exist and be tested)."""
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    rng = np.random.default_rng(5)
    big = _rand_seqs(rng, 1, 3500, 3500)[0]
    sa = [b"W" * 2000, big, _rand_seqs(rng, 1, 300, 300)[0]]
    sb = [b"W" * 1800, big[100:], _rand_seqs(rng, 1, 200, 200)[0]]
    sc, ea, eb = eng.score_batch(sa, sb)
    assert eng.stats()["i32_reruns"] >= 2
    osc, oi, oj = _oracle_scores(m, sa, sb, 10.0, 0.5, swb.ALIGN_LOCAL)
    assert np.array_equal(np.rint(sc.astype(np.float64) * 2).astype(np.int64), osc)
    assert np.array_equal(ea, oi) and np.array_equal(eb, oj)
    assert sc[0] == 11.0 * 1800
    eng.close()


@pytest.mark.parametrize("go,ge,match,mismatch", [(5.0, 2.0, 2.0, -3.0), (10.0, 0.5, 5.0, -4.0), (0.5, 0.5, 1.0, -1.0), (2.0, 3.0, 2.0, -3.0)])
def test_dna_traceback_all_modes(go, ge, match, mismatch):
    """DNA-like inputs (+2/-3 of configs 4/5), including |open| < |extend| which only the 3-state kernel handles."""
    rng = np.random.default_rng(int(go * 10 + ge))
    m = swb.SubstitutionMatrix.get_mat_matrix(match, mismatch)
    for mode in (swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL):
        eng = swb.SequenceAlignment(m, go, ge, mode)
        oeng = O.Engine(_omatrix(m), go, ge, mode)
        sa, sb = [], []
        for k in range(40):
            base = bytearray(_rand_seqs(rng, 1, 20, 700, "ACGT")[0])
            out = bytearray()
            for ch in base:                                        # noisy copy: sub / ins / del
                r = rng.random()
                if r < 0.05:
                    out.append(ord("ACGT"[rng.integers(0, 4)]))
                elif r < 0.10:
                    out.append(ch); out.append(ord("ACGT"[rng.integers(0, 4)]))
                elif r < 0.15:
                    continue
                else:
                    out.append(ch)
            sa.append(bytes(base)); sb.append(bytes(out) if out else b"A")
        res = eng.align_list(sa, sb, retain_all=(mode != swb.ALIGN_LOCAL) or True)
        for a, b, out in zip(sa, sb, res):
            xa, xb, sc, ei, ej = oeng.raw(oeng.matrix.seq_to_index(a.decode()), oeng.matrix.seq_to_index(b.decode()))
            assert out["score"] == sc
            assert np.array_equal(out["idx_a"], xa) and np.array_equal(out["idx_b"], xb)
            r1, r2 = O.inflate(list(a.decode()), list(b.decode()), xa, xb, mode, True)
            assert out["aln_a"] == "".join(r1) and out["aln_b"] == "".join(r2)
        sc2, _, _ = eng.score_batch(sa, sb)
        assert [float(x) for x in sc2] == [o["score"] for o in res]
        eng.close()


def test_retain_all_false_and_empty_alignment():
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    oeng = O.Engine(_omatrix(m), 10.0, 0.5, swb.ALIGN_LOCAL)
    for a, b in [("WWWW", "PPPP"), ("AAAWCDEAAA", "GGWCDEGG"), ("", "ACD"), ("ACD", ""), ("MKV", "MKV")]:
        for ra in (True, False):
            r1, r2, sc = eng.align(a, b, ra)
            o1, o2, osc = oeng.align(a, b, ra)
            assert ("".join(r1), "".join(r2), sc) == ("".join(o1), "".join(o2), osc), (a, b, ra)
    eng.close()


def test_errors_instead_of_panics():
    m = swb.SubstitutionMatrix.get_mat_matrix(1.0, -1.0)
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_GLOBAL)
    with pytest.raises(swb.SwbError):
        eng.align("", "", True)                                    # reference underflows (SA.rs:360-363)
    sc, ea, eb = eng.score_batch([b""], [b""])                     # score_only of two empty sequences is 0 (SA.rs:451-462)
    assert sc[0] == 0.0
    eng.close()
    with pytest.raises(swb.SwbError):
        swb.SequenceAlignment(swb.SubstitutionMatrix.get_mat_matrix(1.0, -1.0), 10.0, 0.3337, swb.ALIGN_LOCAL)   # no integer grid


def test_unknown_letter_without_x():
    sm = swb.SubstitutionMatrix.load_from_lines(["  A T G C", "A 1 -3 -3 -3", "T -3 1 -3 -3", "G -3 -3 1 -3", "C -3 -3 -3 1"])
    eng = swb.SequenceAlignment(sm, 5.0, 2.0, swb.ALIGN_LOCAL)
    with pytest.raises(swb.SwbError):
        eng.score_batch([b"ACGN"], [b"ACGT"])                      # SA.rs:676
    with pytest.raises(swb.SwbError):
        eng.align("ACGN", "ACGT")
    sc, _, _ = eng.score_batch([b"ACGT"], [b"ACGT"])
    assert sc[0] == 4.0
    eng.close()


def test_symmetry_property_large():
    """Size-independent property at BASELINE config-2 shapes: score(a,b) == score(b,a) for a symmetric matrix
    (end cells transpose only when the maximum is unique, so only scores are compared)."""
    rng = np.random.default_rng(20242)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    n = 20000
    sa = _rand_seqs(rng, n, 100, 500)
    sb = _rand_seqs(rng, n, 100, 500)
    s1, ea, eb = eng.score_batch(sa, sb)
    s2, _, _ = eng.score_batch(sb, sa)
    assert np.array_equal(s1, s2)
    assert (s1 > 0).all() and (ea >= 1).all() and (eb >= 1).all()
    sub = rng.choice(n, 300, replace=False)
    osc, oi, oj = _oracle_scores(m, [sa[k] for k in sub], [sb[k] for k in sub], 10.0, 0.5, swb.ALIGN_LOCAL)
    assert np.array_equal(np.rint(s1[sub].astype(np.float64) * 2).astype(np.int64), osc)
    assert np.array_equal(ea[sub], oi) and np.array_equal(eb[sub], oj)
    eng.close()


def test_database_search_topk_with_traceback():
    """configs[2] in miniature: one query vs a database with planted homologs; top-k list (scores, order, end cells)
    and the hit alignments must equal the oracle's; the shared-query K1 path and the per-pair path must agree."""
    from smithwaterman_b200 import search as S
    rng = np.random.default_rng(20243)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    query = _rand_seqs(rng, 1, 700, 700)[0]
    seqs = _rand_seqs(rng, 3000, 30, 900)
    for k in range(0, 3000, 60):                                   # planted homologs: mutated windows of the query
        lo = int(rng.integers(0, 400)); w = bytearray(query[lo:lo + int(rng.integers(100, 300))])
        for p in rng.integers(0, len(w), len(w) // 4):
            w[p] = ord(AA[rng.integers(0, 20)])
        seqs[k] = bytes(w)
    seqs[61] = seqs[60]                                            # exact duplicate -> tie, lower index first
    db = S.Database.from_sequences(seqs)
    hits = S.search(eng, query, db, topk=40, traceback=True, retain_all=False)
    sc, eq, et = S.search_scores(eng, query, db)
    sc2, eq2, et2 = eng.score_batch([query] * len(seqs), seqs)
    assert np.array_equal(sc, sc2) and np.array_equal(eq, eq2) and np.array_equal(et, et2)
    osc, oi, oj = _oracle_scores(m, [query] * len(seqs), seqs, 10.0, 0.5, swb.ALIGN_LOCAL)
    assert np.array_equal(np.rint(sc.astype(np.float64) * 2).astype(np.int64), osc)
    order = np.lexsort((np.arange(len(seqs)), -osc))[:40]
    assert [h["idx"] for h in hits] == order.tolist()
    assert any(h["idx"] == 60 for h in hits) and [h["idx"] for h in hits].index(60) + 1 == [h["idx"] for h in hits].index(61)
    oeng = O.Engine(_omatrix(m), 10.0, 0.5, swb.ALIGN_LOCAL)
    for h in hits:
        r1, r2, s = oeng.align(query.decode(), seqs[h["idx"]].decode(), False)
        assert (h["aln_q"], h["aln_t"], h["score"]) == ("".join(r1), "".join(r2), s)
        assert (h["end_q"], h["end_t"]) == (oi[h["idx"]], oj[h["idx"]])
    eng.close()


@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL])
def test_long_pairs_team_wavefront(mode):
    """A few long pairs make the wavefront kernel sweep each pair with a team of warps (strips handed over through
    the global boundary ring): strings and scores must still equal the dense oracle."""
    rng = np.random.default_rng(77 + mode)
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    eng = swb.SequenceAlignment(m, 5.0, 2.0, mode)
    oeng = O.Engine(_omatrix(m), 5.0, 2.0, mode)
    sa, sb = [], []
    for la in (2500, 4000, 1800):
        base = _rand_seqs(rng, 1, la, la, "ACGT")[0]
        out = bytearray()
        for ch in base:
            r = rng.random()
            if r < 0.05:
                out.append(ord("ACGT"[rng.integers(0, 4)]))
            elif r < 0.08:
                out.append(ch); out.append(ord("ACGT"[rng.integers(0, 4)]))
            elif r < 0.11:
                continue
            else:
                out.append(ch)
        sa.append(base); sb.append(bytes(out))
    res = eng.align_list(sa, sb, True)
    for a, b, out in zip(sa, sb, res):
        r1, r2, sc = oeng.align(a.decode(), b.decode(), True)
        assert out["score"] == sc
        assert out["aln_a"] == "".join(r1) and out["aln_b"] == "".join(r2)
    sc2, ea, eb = eng.score_batch(sa, sb)                      # score-only path of the same kernel (non-local) / K1 (local)
    assert [float(x) for x in sc2] == [o["score"] for o in res]
    assert [int(x) for x in ea] == [o["end_a"] for o in res] and [int(x) for x in eb] == [o["end_b"] for o in res]
    eng.close()


def _mutate(rng, base, sub, ins, dele, alphabet=b"ACGT"):
    a = np.frombuffer(base, np.uint8)
    al = np.frombuffer(alphabet, np.uint8)
    b = a.copy()
    s = rng.random(len(a)) < sub
    b[s] = al[rng.integers(0, len(al), int(s.sum()))]
    b = b[rng.random(len(b)) >= dele]
    pos = np.nonzero(rng.random(len(b)) < ins)[0]
    return np.insert(b, pos, al[rng.integers(0, len(al), len(pos))]).tobytes()


def _rescore(aln_a, aln_b, match, mismatch, go, ge):
    """Score of an alignment given as strings under the reference's model (SA.rs:139-252): a column that opens a gap
    run -- also when the previous column was a gap in the OTHER sequence -- costs `go`, a column that continues a
    run in the same sequence costs `ge`."""
    a = np.frombuffer(aln_a.encode(), np.uint8); b = np.frombuffer(aln_b.encode(), np.uint8)
    ga, gb = a == ord("-"), b == ord("-")
    both = ~ga & ~gb
    sc = float(np.sum(np.where(a[both] == b[both], match, mismatch)))
    for g in (ga, gb):
        prev = np.concatenate(([False], g[:-1]))
        sc -= go * float(np.sum(g & ~prev)) + ge * float(np.sum(g & prev))
    return sc


@pytest.mark.parametrize("cols", ["2", "8"])
@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL])
def test_config4_dna_pair_dense_oracle(mode, cols, monkeypatch):
    """BASELINE configs[3] at 6 kb: +2/-3, open 5 / extend 2, full traceback; strings, score and end cell equal the
    dense oracle for both strip widths of the wavefront kernel."""
    monkeypatch.setenv("SWB_KW_COLS", cols)
    rng = np.random.default_rng(20244)
    a = _rand_seqs(rng, 1, 6000, 6000, "ACGT")[0]
    b = _mutate(rng, a, 0.08, 0.02, 0.02)
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    eng = swb.SequenceAlignment(m, 5.0, 2.0, mode)
    out = eng.align_raw(a, b, True)
    r1, r2, sc = O.Engine(_omatrix(m), 5.0, 2.0, mode).align(a.decode(), b.decode(), True)
    assert out["score"] == sc and out["aln_a"] == "".join(r1) and out["aln_b"] == "".join(r2)
    eng.close()


def test_config4_dna_pair_100kb_properties(monkeypatch):
    """BASELINE configs[3] at full size (100 kb x ~100 kb, 10^10 cells, 10 GB of direction bytes).  The dense oracle
    cannot hold this; checked instead: (i) the score equals the linear-memory integer oracle's on a 30 kb prefix pair,
    (ii) at full size the emitted strings re-score to the reported score, spell the inputs, and the global score is
    not above the local one, (iii) score-only and traceback paths agree, (iv) the checkpointed recompute (2 GB of
    direction bytes instead of 10 GB) returns the same strings, score and end cell as the direct store."""
    monkeypatch.delenv("SWB_CKPT_BYTES", raising=False)
    rng = np.random.default_rng(20244)
    a = _rand_seqs(rng, 1, 100000, 100000, "ACGT")[0]
    b = _mutate(rng, a, 0.08, 0.02, 0.02)
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    om = _omatrix(m)
    scores = {}
    for mode in (swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL):
        eng = swb.SequenceAlignment(m, 5.0, 2.0, mode)
        out = eng.align_raw(a, b, False)
        assert out["aln_a"].replace("-", "") in a.decode() and out["aln_b"].replace("-", "") in b.decode()
        if mode == swb.ALIGN_GLOBAL:
            assert out["aln_a"].replace("-", "") == a.decode() and out["aln_b"].replace("-", "") == b.decode()
        assert _rescore(out["aln_a"], out["aln_b"], 2.0, -3.0, 5.0, 2.0) == out["score"]
        sc, ea, eb = eng.score_batch([a], [b])
        assert float(sc[0]) == out["score"] and (int(ea[0]), int(eb[0])) == (out["end_a"], out["end_b"])
        scores[mode] = out["score"]
        monkeypatch.setenv("SWB_CKPT_BYTES", str(2 << 30))
        ck = eng.align_raw(a, b, False)
        monkeypatch.delenv("SWB_CKPT_BYTES", raising=False)
        assert eng.stats()["ckpt_pairs"] == 1
        assert (ck["score"], ck["end_a"], ck["end_b"], ck["aln_a"], ck["aln_b"]) == \
               (out["score"], out["end_a"], out["end_b"], out["aln_a"], out["aln_b"])
        pa, pb = a[:30000], b[:30000]
        sc, ea, eb = eng.score_batch([pa], [pb])
        ia, oa = O.pack([om.seq_to_index(pa.decode())]); ib, ob = O.pack([om.seq_to_index(pb.decode())])
        osc, oi, oj = O.score_batch_int(ia, oa, ib, ob, np.rint(m.scores).astype(np.int32), -5, -2, 1, mode, 1)
        assert (float(sc[0]), int(ea[0]), int(eb[0])) == (float(osc[0]), int(oi[0]), int(oj[0]))
        eng.close()
    assert scores[swb.ALIGN_GLOBAL] <= scores[swb.ALIGN_LOCAL]


@pytest.mark.timeout(1500)
def test_config4_100kb_score_and_end_cell_vs_integer_oracle_all_modes():
    """BASELINE configs[3] at FULL size (100 kb x ~100 kb, 10^10 cells) in all three modes: score and end cell of the
    score-only path against the linear-memory integer oracle (SA.rs:438-467 score_only scan); the three oracle runs
    (about a minute each on one host core) go in parallel threads."""
    import threading
    rng = np.random.default_rng(20244)
    a = _rand_seqs(rng, 1, 100000, 100000, "ACGT")[0]
    b = _mutate(rng, a, 0.08, 0.02, 0.02)
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    om = _omatrix(m)
    ia, oa = O.pack([om.seq_to_index(a.decode())]); ib, ob = O.pack([om.seq_to_index(b.decode())])
    mi = np.rint(m.scores).astype(np.int32)
    modes = (swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL, swb.ALIGN_GLOCAL)
    want = {}

    def run(mode):
        osc, oi, oj = O.score_batch_int(ia, oa, ib, ob, mi, -5, -2, 1, mode, 1)
        want[mode] = (float(osc[0]), int(oi[0]), int(oj[0]))
    th = [threading.Thread(target=run, args=(mode,)) for mode in modes]
    for t in th:
        t.start()
    got = {}
    for mode in modes:
        eng = swb.SequenceAlignment(m, 5.0, 2.0, mode)
        sc, ea, eb = eng.score_batch([a], [b])
        got[mode] = (float(sc[0]), int(ea[0]), int(eb[0]))
        eng.close()
    for t in th:
        t.join()
    assert got == want
    assert got[swb.ALIGN_GLOBAL][0] <= got[swb.ALIGN_GLOCAL][0] <= got[swb.ALIGN_LOCAL][0]


def test_glocal_batch_10kb_with_strings():
    """Glocal at the configs[4] size (10 kb noisy DNA pairs, one sequence a fragment of the other in half of them): scores
    against the integer oracle on all pairs, strings against the dense oracle on two, re-scoring of the core on all."""
    rng = np.random.default_rng(20246)
    n = 12
    sa = _rand_seqs(rng, n, 10000, 10000, "ACGT")
    sb = [_mutate(rng, s if k % 2 else s[1500:8200], 0.05, 0.04, 0.04) for k, s in enumerate(sa)]
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    om = _omatrix(m)
    eng = swb.SequenceAlignment(m, 5.0, 2.0, swb.ALIGN_GLOCAL)
    res = eng.align_list(sa, sb, False)
    sc, ea, eb = eng.score_batch(sa, sb)
    ia, oa = O.pack([om.seq_to_index(s.decode()) for s in sa]); ib, ob = O.pack([om.seq_to_index(s.decode()) for s in sb])
    osc, oi, oj = O.score_batch_int(ia, oa, ib, ob, np.rint(m.scores).astype(np.int32), -5, -2, 1, swb.ALIGN_GLOCAL, 8)
    assert [r["score"] for r in res] == [float(x) for x in osc] == [float(x) for x in sc]
    assert np.array_equal(ea, oi) and np.array_equal(eb, oj)
    oeng = O.Engine(om, 5.0, 2.0, swb.ALIGN_GLOCAL)
    for k in (0, 1):
        r1, r2, osc1 = oeng.align(sa[k].decode(), sb[k].decode(), False)
        assert (res[k]["aln_a"], res[k]["aln_b"], res[k]["score"]) == ("".join(r1), "".join(r2), osc1)
    eng.close()


def test_config5_nw_batch_10kb():
    """BASELINE configs[4]: Needleman-Wunsch (global) on a batch of 10 kb noisy-read-like DNA pairs with traceback
    strings: dense oracle on two pairs (strings), linear integer oracle on all (scores), re-scoring on all."""
    rng = np.random.default_rng(20245)
    n = 24
    sa = _rand_seqs(rng, n, 10000, 10000, "ACGT")
    sb = [_mutate(rng, s, 0.05, 0.05, 0.05) for s in sa]
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    om = _omatrix(m)
    eng = swb.SequenceAlignment(m, 5.0, 2.0, swb.ALIGN_GLOBAL)
    res = eng.align_list(sa, sb, True)
    ia, oa = O.pack([om.seq_to_index(s.decode()) for s in sa]); ib, ob = O.pack([om.seq_to_index(s.decode()) for s in sb])
    osc, _, _ = O.score_batch_int(ia, oa, ib, ob, np.rint(m.scores).astype(np.int32), -5, -2, 1, swb.ALIGN_GLOBAL, 8)
    assert [r["score"] for r in res] == [float(x) for x in osc]
    for a, b, r in zip(sa, sb, res):
        assert r["aln_a"].replace("-", "") == a.decode() and r["aln_b"].replace("-", "") == b.decode()
        assert _rescore(r["aln_a"], r["aln_b"], 2.0, -3.0, 5.0, 2.0) == r["score"]
    oeng = O.Engine(om, 5.0, 2.0, swb.ALIGN_GLOBAL)
    for k in (0, n - 1):
        r1, r2, sc = oeng.align(sa[k].decode(), sb[k].decode(), True)
        assert (res[k]["aln_a"], res[k]["aln_b"], res[k]["score"]) == ("".join(r1), "".join(r2), sc)
    eng.close()


def test_database_search_long_query_k2():
    """configs[2] with a long query: the K2 path (lanes of a warp split the columns; subjects bin-packed into warps,
    subjects of more than 1024 residues swept in several passes of the warp).  Scores and end cells equal the
    integer oracle; the per-pair K1 path and the top-k alignments agree."""
    from smithwaterman_b200 import search as S
    rng = np.random.default_rng(20246)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    query = _rand_seqs(rng, 1, 1500, 1500)[0]
    seqs = _rand_seqs(rng, 900, 20, 700) + _rand_seqs(rng, 12, 1025, 3300) + _rand_seqs(rng, 40, 1, 40) + [b"", b"A"]
    for k in range(0, 900, 45):                                    # planted homologs
        lo = int(rng.integers(0, 1000)); w = bytearray(query[lo:lo + int(rng.integers(100, 450))])
        for p in rng.integers(0, len(w), len(w) // 4):
            w[p] = ord(AA[rng.integers(0, 20)])
        seqs[k] = bytes(w)
    seqs[905] = seqs[905][:700] + query[200:1400] + seqs[905][700:]   # a long subject with a long exact match (> 1024 columns)
    seqs[46] = seqs[45]
    perm = rng.permutation(len(seqs))
    seqs = [seqs[k] for k in perm]
    db = S.Database.from_sequences(seqs)
    sc, eq, et = S.search_scores(eng, query, db)
    assert eng.stats()["launches"] >= 1
    osc, oi, oj = _oracle_scores(m, [query] * len(seqs), seqs, 10.0, 0.5, swb.ALIGN_LOCAL)
    assert np.array_equal(np.rint(sc.astype(np.float64) * 2).astype(np.int64), osc)
    nz = osc > 0
    assert np.array_equal(eq[nz], oi[nz]) and np.array_equal(et[nz], oj[nz])
    assert (eq[~nz] == -1).all() and (et[~nz] == -1).all()
    sc2, eq2, et2 = eng.score_batch([query] * len(seqs), seqs)
    assert np.array_equal(sc, sc2) and np.array_equal(eq, eq2) and np.array_equal(et, et2)
    hits = S.search(eng, query, db, topk=25, traceback=True, retain_all=False)
    order = np.lexsort((np.arange(len(seqs)), -osc))[:25]
    assert [h["idx"] for h in hits] == order.tolist()
    eng.close()


@pytest.mark.parametrize("flag,key", [("-local", "local"), ("-glocal", "glocal"), ("-global", "global")])
def test_cli_list_mode_vs_emboss(emboss, flag, key, tmp_path):
    """The reference's harness in one test (test/run_rust_list.pl + check_results.pl): `sa_opencl [-mode] -list L` over
    FASTA files, stdout records parsed and compared with test/emboss_results (strings; local after trimTerminal)."""
    import io
    from smithwaterman_b200 import cli
    recs = emboss[:40]
    lines = []
    for r in recs:
        f1, f2 = tmp_path / ("seq%s.1.fas" % r["id"]), tmp_path / ("seq%s.2.fas" % r["id"])
        f1.write_text(">a%s some description\n%s\n" % (r["id"], "\n".join(r["seq1"][k:k + 60] for k in range(0, len(r["seq1"]), 60))))
        f2.write_text(">b%s\n%s\n" % (r["id"], r["seq2"]))
        lines.append("%s\t%s" % (f1, f2))
    (tmp_path / "list.txt").write_text("\n".join(lines) + "\n")
    out = io.StringIO()
    assert cli.main([flag, "-list", str(tmp_path / "list.txt")], out) == 0
    blocks = out.getvalue().split("#score:")[1:]
    assert len(blocks) == len(recs)
    for r, blk in zip(recs, blocks):
        ln = blk.split("\n")
        assert ln[1] == "#type:%s alignment" % key.capitalize() and ln[2] == ">a%s" % r["id"] and ln[5] == ">b%s" % r["id"]
        s1, s2 = ln[3], ln[6]
        if key == "local":
            s1, s2 = trim_terminal(s1, s2)
        assert (s1, s2) == (r[key]["s1"], r[key]["s2"]), r["id"]
        assert float(ln[0]) == r[key]["score"]


@pytest.mark.parametrize("path", ["default", "wide"])
@pytest.mark.parametrize("seed", range(6))
def test_fuzz_random_models(seed, path, monkeypatch):
    """Random scoring models (random integer / half-integer matrices over a random alphabet with X, random penalties
    including open < extend, which K1 must hand to the exact 3-state kernel), all three modes, lengths 0..90 with
    low-complexity sequences (many ties): score batch and full alignments vs the oracle.  path "wide": the alignments
    are forced onto the int32 tile path (K3W + KT3) wherever it applies (local / global, open >= extend)."""
    if path == "wide":
        monkeypatch.setenv("SWB_NO_K3", "1"); monkeypatch.setenv("SWB_K3W_MIN", "1")
    rng = np.random.default_rng(9000 + seed)
    nlet = int(rng.integers(3, 12))
    letters = "X" + "".join(rng.choice(list("ACDEFGHIKLMNPQRSTVWY"), nlet - 1, replace=False))
    step = [1.0, 0.5][seed % 2]
    sym = rng.integers(-8, 9, (nlet, nlet)).astype(np.float32) * step
    sym = np.triu(sym) + np.triu(sym, 1).T
    sym[np.arange(nlet), np.arange(nlet)] = np.abs(sym[np.arange(nlet), np.arange(nlet)]) + step
    m = swb.SubstitutionMatrix(letters, sym)
    go = float(rng.integers(1, 24)) * step
    ge = float(rng.integers(0, 8)) * step if seed != 3 else go + 2 * step     # seed 3: extend dearer than open
    om = _omatrix(m)
    scale = 1 if step == 1.0 and float(go).is_integer() and float(ge).is_integer() else 2
    alpha = letters[1:] + "xb"                                              # unknown letters -> X
    for mode in (swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL):
        eng = swb.SequenceAlignment(m, go, ge, mode)
        assert eng.stats()["scale"] == scale
        sa = _rand_seqs(rng, 260, 0, 90, alpha) + [b"", b"", bytes(alpha[0], "ascii") * 40]
        sb = _rand_seqs(rng, 260, 0, 90, alpha[:3] + alpha[:2]) + [b"", bytes(alpha[1], "ascii") * 7, bytes(alpha[0], "ascii") * 33]
        keep = [k for k in range(len(sa)) if mode == swb.ALIGN_LOCAL or len(sa[k]) + len(sb[k]) > 0]
        sa, sb = [sa[k] for k in keep], [sb[k] for k in keep]
        sc, ea, eb = eng.score_batch(sa, sb)
        a, oa = O.pack([om.seq_to_index(s.decode()) for s in sa]); b, ob = O.pack([om.seq_to_index(s.decode()) for s in sb])
        osc, oi, oj = O.score_batch_int(a, oa, b, ob, np.rint(m.scores * scale).astype(np.int32), -int(round(go * scale)),
                                        -int(round(ge * scale)), scale, mode, 8)
        assert np.array_equal(np.rint(sc.astype(np.float64) * scale).astype(np.int64), osc), (seed, mode)
        assert np.array_equal(ea, oi) and np.array_equal(eb, oj), (seed, mode)
        oeng = O.Engine(om, go, ge, mode)
        sel = [k for k in range(0, len(sa), 4) if not (go + ge == 0 and (len(sa[k]) == 0 or len(sb[k]) == 0))]
        res = eng.align_list([sa[k] for k in sel], [sb[k] for k in sel], bool(seed & 1))
        for k, out in zip(sel, res):
            r1, r2, s = oeng.align(sa[k].decode(), sb[k].decode(), bool(seed & 1))
            assert (out["aln_a"], out["aln_b"], out["score"]) == ("".join(r1), "".join(r2), s), (seed, mode, k)
        eng.close()


def _string_counts(r):
    ra, rb = r["aln_a"], r["aln_b"]
    return (sum(1 for c in ra if c != "-"), sum(1 for c in rb if c != "-"),
            sum(1 for x, y in zip(ra, rb) if x != "-" and y != "-" and x == y), len(ra))


@pytest.mark.parametrize("retain_all", [False, True])
@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL])
def test_align_stats_equal_the_strings(mode, retain_all):
    """swb_align_stats_batch (counts taken during the traceback walk on the device) == the same counts read off the
    strings of swb_align_batch, which is what the clustering loop of the reference does (OCL-main.rs:205-233).
    Letters the matrix does not know (J, O, U -> scored as X) must not count as equal unless they ARE equal."""
    rng = np.random.default_rng(77 + mode * 2 + int(retain_all))
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
    base = _rand_seqs(rng, 30, 30, 400)
    sa, sb = [], []
    for s in base:
        for _ in range(3):
            sa.append(s); sb.append(_mutate(rng, s, 0.15, 0.03, 0.03, alphabet=AA.encode()))
    sa += _rand_seqs(rng, 40, 1, 300); sb += _rand_seqs(rng, 40, 1, 300)           # unrelated pairs, short ones included
    sa += [b"JJJJOUJOUWWWW", b"ACDEFGHIKL", b"", b"AAAA"]; sb += [b"OJJJOUJUUWWWW", b"ACDEFGHIKL", b"ACD", b""]
    st = eng.align_stats(sa, sb, retain_all)
    res = eng.align_list(sa, sb, retain_all)
    assert (st["status"] == 0).all()
    for k, r in enumerate(res):
        assert (int(st["a_letters"][k]), int(st["b_letters"][k]), int(st["matches"][k]), int(st["aln_len"][k])) == _string_counts(r), k
        assert float(st["score"][k]) == r["score"] and (int(st["end_a"][k]), int(st["end_b"][k])) == (r["end_a"], r["end_b"]), k
    eng.close()


def test_cluster_with_device_counts_equals_strings():
    """`-cluster` driven by swb_align_stats_batch makes the decisions of the string-reading loop (OCL-main.rs:149-243)."""
    import io
    from smithwaterman_b200 import cli
    rng = np.random.default_rng(4242)
    seqs = []
    for f, s in enumerate(_rand_seqs(rng, 12, 80, 300)):
        seqs.append(swb.SeqData.create("f%d_0" % f, "", s.decode()))
        for v in range(1, 6):
            t = _mutate(rng, s, 0.02 * v * v, 0.01 * v, 0.01 * v, alphabet=AA.encode())
            if v == 5:
                t = t[: len(t) // 2]                                 # a fragment: fails the coverage of the long sequence
            seqs.append(swb.SeqData.create("f%d_%d" % (f, v), "", t.decode()))
    seqs.append(swb.SeqData.create("dup", "", "".join(seqs[0].seq)))
    seqs = [seqs[k] for k in rng.permutation(len(seqs))]
    eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, swb.ALIGN_LOCAL)

    def align_many(rep, cands):
        return [(r["aln_a"], r["aln_b"]) for r in eng.align_list([rep] * len(cands), cands, False)]

    def stats_many(rep, cands):
        st = eng.align_stats([rep] * len(cands), cands, False)
        return list(zip(st["a_letters"].tolist(), st["b_letters"].tolist(), st["matches"].tolist(), st["aln_len"].tolist()))
    s1, c1, m1 = cli.cluster(seqs, align_many, 0.8, 0.8, 0.8, io.StringIO())
    s2, c2, m2 = cli.cluster(seqs, None, 0.8, 0.8, 0.8, io.StringIO(), stats_many=stats_many)
    assert [s.name for s in s1] == [s.name for s in s2] and c1 == c2 and m1 == m2
    n_clusters = sum(1 for k, c in enumerate(c1) if c == k)
    assert 12 <= n_clusters < len(seqs)
    # the loop run by the library (swb_cluster, sequences resident on the device), other thresholds as well
    for thr in ((0.8, 0.8, 0.8), (0.5, 0.3, 0.9), (0.95, 0.9, 0.9)):
        sp, cp, mp = cli.cluster(seqs, None, *thr, io.StringIO(), stats_many=stats_many)
        sn, cn, mn = cli.cluster_native(eng, seqs, *thr, io.StringIO())
        assert [s.name for s in sp] == [s.name for s in sn] and cp == cn and mp == mn, thr
    eng.close()


def test_cli_cluster_files(tmp_path):
    """`sa_opencl -cluster -out F x.fas` end to end: representative FASTA + .clstr member lists (OCL-main.rs:244-265)."""
    import io
    from smithwaterman_b200 import cli
    rng = np.random.default_rng(5)
    recs = []
    for f, s in enumerate(_rand_seqs(rng, 5, 120, 200)):
        for v in range(4):
            recs.append((">s%d_%d family %d" % (f, v, f), _mutate(rng, s, 0.03 * v, 0.0, 0.0, alphabet=AA.encode()).decode()))
    recs.append((">copy of s0_0", recs[0][1]))
    (tmp_path / "in.fas").write_text("".join("%s\n%s\n" % r for r in recs))
    assert cli.main(["-cluster", "-identity", "0.8", "-out", str(tmp_path / "o.fas"), str(tmp_path / "in.fas")], io.StringIO()) == 0
    reps = [ln[1:].split()[0] for ln in (tmp_path / "o.fas").read_text().splitlines() if ln.startswith(">")]
    lists = [ln.split() for ln in (tmp_path / "o.fas.clstr").read_text().splitlines()]
    assert sorted(x for l in lists for x in l) == sorted(r[0][1:].split()[0] for r in recs)       # every sequence in exactly one list
    assert len(lists) == 5 and all(len(l) >= 4 for l in lists)
    assert {l[0] for l in lists} <= set(reps)                                                    # a list starts with its representative


def _swbcli():
    from smithwaterman_b200 import _lib
    _lib.build()
    return _lib.build_cli()


@pytest.mark.parametrize("flag,key", [("-local", "local"), ("-glocal", "glocal"), ("-global", "global")])
def test_cpp_cli_list_mode_vs_emboss(emboss, flag, key, tmp_path):
    """The compiled command line (tools/swbcli.cpp over include/swb200.hpp, the C++ mirror of the reference's Rust API):
    `sa_opencl [-mode] -list L` against test/emboss_results, and byte-identical stdout with the Python mirror."""
    import io
    import subprocess
    from smithwaterman_b200 import cli
    recs = emboss[:30]
    lines = []
    for r in recs:
        f1, f2 = tmp_path / ("seq%s.1.fas" % r["id"]), tmp_path / ("seq%s.2.fas" % r["id"])
        f1.write_text(">a%s some description\n%s\n" % (r["id"], "\n".join(r["seq1"][k:k + 60] for k in range(0, len(r["seq1"]), 60))))
        f2.write_text(">b%s\n%s\n" % (r["id"], r["seq2"]))
        lines.append("%s\t%s" % (f1, f2))
    (tmp_path / "list.txt").write_text("\n".join(lines) + "\n")
    got = subprocess.run([_swbcli(), flag, "-list", str(tmp_path / "list.txt")], capture_output=True, text=True, check=True).stdout
    blocks = got.split("#score:")[1:]
    assert len(blocks) == len(recs)
    for r, blk in zip(recs, blocks):
        ln = blk.split("\n")
        s1, s2 = ln[3], ln[6]
        if key == "local":
            s1, s2 = trim_terminal(s1, s2)
        assert (s1, s2) == (r[key]["s1"], r[key]["s2"]), r["id"]
        assert float(ln[0]) == r[key]["score"]
    out = io.StringIO()
    assert cli.main([flag, "-list", str(tmp_path / "list.txt")], out) == 0
    assert out.getvalue() == got
    # pair mode and -out (the reference writes no newline after #score / #type there, OCL-main.rs:303-307)
    f1, f2 = lines[0].split("\t")
    pair = subprocess.run([_swbcli(), flag, f1, f2], capture_output=True, text=True, check=True).stdout
    assert pair == "#score:" + blocks[0]
    subprocess.run([_swbcli(), flag, "-list", str(tmp_path / "list.txt"), "-out", str(tmp_path / "cpp.out")], check=True, capture_output=True)
    assert cli.main([flag, "-list", str(tmp_path / "list.txt"), "-out", str(tmp_path / "py.out")], io.StringIO()) == 0
    assert (tmp_path / "cpp.out").read_text() == (tmp_path / "py.out").read_text()


def test_cpp_cli_cluster_equals_python(tmp_path):
    import io
    import subprocess
    from smithwaterman_b200 import cli
    rng = np.random.default_rng(11)
    recs = []
    for f, s in enumerate(_rand_seqs(rng, 8, 100, 260)):
        for v in range(5):
            recs.append((">q%d_%d fam %d  v%d" % (f, v, f, v), _mutate(rng, s, 0.04 * v, 0.01 * v, 0.0, alphabet=AA.encode()).decode()))
    recs.append((">twin", recs[3][1]))
    order = rng.permutation(len(recs))
    (tmp_path / "in.fas").write_text("".join("%s\n%s\n" % recs[k] for k in order))
    args = ["-cluster", "-identity", "0.7", "-coverage_short", "0.6", "-out"]
    r = subprocess.run([_swbcli()] + args + [str(tmp_path / "c.fas"), str(tmp_path / "in.fas")], capture_output=True, text=True, check=True)
    out = io.StringIO()
    assert cli.main(args + [str(tmp_path / "p.fas"), str(tmp_path / "in.fas")], out) == 0
    assert (tmp_path / "c.fas").read_text() == (tmp_path / "p.fas").read_text()
    assert (tmp_path / "c.fas.clstr").read_text() == (tmp_path / "p.fas.clstr").read_text()
    assert r.stdout == out.getvalue()
    assert 8 <= len((tmp_path / "c.fas.clstr").read_text().splitlines()) < len(recs)


def test_cpp_mirror_runs_the_reference_kats(rust_kats, tmp_path):
    """tests/cpp/kats.cpp: the reference's known-answer tests written against include/swb200.hpp the way the Rust tests
    are written against SequenceAlignment (SA.rs:921-1221)."""
    import os
    import subprocess
    from smithwaterman_b200 import _lib
    _lib.build()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "kats")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-I", os.path.join(root, "include"), os.path.join(root, "tests", "cpp", "kats.cpp"),
                           "-L", os.path.dirname(_lib.SO_PATH), "-lswb200", "-Wl,-rpath," + os.path.dirname(_lib.SO_PATH), "-o", exe])
    rows = []
    for c in rust_kats:
        if c["pssm"]:
            continue
        m = c["matrix"]
        rows.append("\t".join([c["test"], str(c["mode"]), repr(c["go"]), repr(c["ge"]), "1" if c["retain_all"] else "0", m["kind"],
                               repr(m.get("match", 0.0)), repr(m.get("mismatch", 0.0)), repr(m.get("scale", 1.0)), c["seq1"], c["seq2"],
                               swb.rust_f32_display(c["score"]) if "score" in c else "-", c.get("r1", "-"), c.get("r2", "-")]))
    (tmp_path / "cases.tsv").write_text("\n".join(rows) + "\n")
    r = subprocess.run([exe, str(tmp_path / "cases.tsv")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "kats: %d cases, 0 failures" % len(rows) in r.stdout


@pytest.mark.parametrize("mode", [swb.ALIGN_GLOBAL, swb.ALIGN_GLOCAL])
@pytest.mark.parametrize("go,ge", [(10.0, 0.5), (0.5, 0.5), (3.0, 1.0)])
def test_global_score_batch_packed_kernel(go, ge, mode):
    """Global and glocal mode through the packed s16x2 kernel (csrc/k1_global.cu): multi-pass, multi-block shapes, pairs
    near the s16 routing bound on both sides (longer ones go to the int32 wavefront kernel in the same call), self
    alignments (largest positive scores) and unrelated pairs of very different lengths (most negative scores).  Glocal:
    free last row / last column (SA.rs:166-170) as two tracked maxima over a right-aligned b."""
    rng = np.random.default_rng(int(go * 10 + ge * 100))
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, go, ge, mode)
    sa = _rand_seqs(rng, 2500, 1, 700); sb = _rand_seqs(rng, 2500, 1, 700)
    for la, lb in [(1, 700), (700, 1), (1400, 1400), (1450, 1460), (1500, 30), (30, 1500), (2500, 2500), (3000, 40), (8192, 5), (5, 8192),
                   (9000, 3), (31, 31), (32, 64), (96, 96), (97, 95)]:
        sa.append(_rand_seqs(rng, 1, la, la)[0]); sb.append(_rand_seqs(rng, 1, lb, lb)[0])
    for la in (200, 777, 1400, 1480, 1490, 1600):                   # self alignments: score = sum of the diagonal
        s = _rand_seqs(rng, 1, la, la)[0]
        sa.append(s); sb.append(s)
    for k in range(300):                                            # related pairs
        base = _rand_seqs(rng, 1, 50, 900)[0]
        sa.append(base); sb.append(_mutate(rng, base, 0.2, 0.05, 0.05, alphabet=AA.encode()))
    for k in range(200):                                            # glocal shapes: one sequence inside the other, overlaps
        base = _rand_seqs(rng, 1, 100, 900)[0]
        lo = int(rng.integers(0, len(base) // 2)); hi = int(rng.integers(lo + 20, len(base) + 1))
        sa.append(base); sb.append(_mutate(rng, base[lo:hi], 0.1, 0.02, 0.02, alphabet=AA.encode()) + (_rand_seqs(rng, 1, 0, 60)[0] if k % 2 else b""))
    sc, ea, eb = eng.score_batch(sa, sb)
    osc, oi, oj = _oracle_scores(m, sa, sb, go, ge, mode)
    assert np.array_equal(np.rint(sc.astype(np.float64) * 2).astype(np.int64), osc)
    assert np.array_equal(ea, oi) and np.array_equal(eb, oj)
    # one shared `a` (the database-search call) takes the same kernel
    from smithwaterman_b200 import search as S
    q = _rand_seqs(rng, 1, 350, 350)[0]
    db = S.Database.from_sequences(sb[:800])
    s2, _, _ = S.search_scores(eng, q, db)
    o2, _, _ = _oracle_scores(m, [q] * 800, sb[:800], go, ge, mode)
    assert np.array_equal(np.rint(s2.astype(np.float64) * 2).astype(np.int64), o2)
    eng.close()


@pytest.mark.parametrize("cols", ["2", "8"])
@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL])
def test_checkpointed_traceback_equals_direct(mode, cols, monkeypatch):
    """Checkpointed recompute (configs[3]): with a direction-byte budget far below the pairs' needs the traceback runs
    as forward sweep + row bands; index lists, strings, score and end cell must equal the oracle and the direct path.
    Pairs include related DNA (long diagonal path through every band), unrelated sequences (local: short path; global:
    path along the edges), very unequal lengths and |open| < |extend|."""
    monkeypatch.setenv("SWB_KW_COLS", cols)
    monkeypatch.setenv("SWB_NO_K3", "1")                           # this test is about the direction-byte path and its checkpointed form
    rng = np.random.default_rng(20250 + int(cols))
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    for go, ge in ((5.0, 2.0), (2.0, 3.0)):
        sa, sb = [], []
        for la in (900, 2500, 1700):
            a = _rand_seqs(rng, 1, la, la, "ACGT")[0]
            sa.append(a); sb.append(_mutate(rng, a, 0.08, 0.03, 0.03))
        sa.append(_rand_seqs(rng, 1, 1500, 1500, "ACGT")[0]); sb.append(_rand_seqs(rng, 1, 1300, 1300, "ACGT")[0])   # unrelated
        a = _rand_seqs(rng, 1, 3000, 3000, "ACGT")[0]
        sa.append(a); sb.append(a[1200:1500])                                                                        # tall
        sa.append(a[2000:2200]); sb.append(a)                                                                        # wide
        oeng = O.Engine(_omatrix(m), go, ge, mode)
        monkeypatch.delenv("SWB_CKPT_BYTES", raising=False)
        eng = swb.SequenceAlignment(m, go, ge, mode)
        direct = eng.align_list(sa, sb, True)
        assert eng.stats()["ckpt_pairs"] == 0
        for budget in ("65536", "300000"):
            monkeypatch.setenv("SWB_CKPT_BYTES", budget)
            res = eng.align_list(sa, sb, True)
            assert eng.stats()["ckpt_pairs"] >= 5
            for a_, b_, out, ref in zip(sa, sb, res, direct):
                xa, xb, sc, ei, ej = oeng.raw(oeng.matrix.seq_to_index(a_.decode()), oeng.matrix.seq_to_index(b_.decode()))
                assert out["score"] == sc == ref["score"]
                assert (out["end_a"], out["end_b"]) == (ref["end_a"], ref["end_b"])
                assert np.array_equal(out["idx_a"], xa) and np.array_equal(out["idx_b"], xb)
                assert out["aln_a"] == ref["aln_a"] and out["aln_b"] == ref["aln_b"]
        eng.close()


def test_checkpointed_traceback_30kb(monkeypatch):
    """A 30 kb x 30 kb DNA pair (+2/-3, 5/2) under a 64 MB direction-byte budget (the direct store needs 0.9 GB):
    same score, end cell and strings as the direct path; the strings re-score to the reported score."""
    rng = np.random.default_rng(20251)
    a = _rand_seqs(rng, 1, 30000, 30000, "ACGT")[0]
    b = _mutate(rng, a, 0.10, 0.02, 0.02)
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    for mode in (swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL):
        eng = swb.SequenceAlignment(m, 5.0, 2.0, mode)
        monkeypatch.delenv("SWB_CKPT_BYTES", raising=False)
        ref = eng.align_raw(a, b, mode != swb.ALIGN_LOCAL)
        monkeypatch.setenv("SWB_CKPT_BYTES", str(64 << 20))
        out = eng.align_raw(a, b, mode != swb.ALIGN_LOCAL)
        assert eng.stats()["ckpt_pairs"] == 1
        assert out["score"] == ref["score"] and out["aln_a"] == ref["aln_a"] and out["aln_b"] == ref["aln_b"]
        assert (out["end_a"], out["end_b"]) == (ref["end_a"], ref["end_b"])
        assert _rescore(out["aln_a"], out["aln_b"], 2.0, -3.0, 5.0, 2.0) == out["score"]
        eng.close()


@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL])
def test_checkpointed_counting_form(mode, monkeypatch):
    """swb_align_stats_batch on pairs over the direction-byte budget: the counts equal those of the direct path."""
    rng = np.random.default_rng(20252)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    sa = _rand_seqs(rng, 6, 600, 1500)
    sb = [bytes(_mutate(rng, a, 0.2, 0.03, 0.03, AA.encode())) for a in sa[:4]] + _rand_seqs(rng, 2, 700, 900)
    eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
    monkeypatch.delenv("SWB_CKPT_BYTES", raising=False)
    for retain_all in (False, True):
        ref = eng.align_stats(sa, sb, retain_all)
        monkeypatch.setenv("SWB_CKPT_BYTES", "100000")
        out = eng.align_stats(sa, sb, retain_all)
        assert eng.stats()["ckpt_pairs"] == 6
        monkeypatch.delenv("SWB_CKPT_BYTES", raising=False)
        assert out.tobytes() == ref.tobytes()
    eng.close()


def test_srspair_reports_from_the_engine(emboss):
    """8f-n3: the engine's alignments printed by the srspair writer equal the bodies of the reference's EMBOSS golden
    files byte for byte: all 100 water (local) reports, and every needle report whose terminal overhang EMBOSS places
    like the reference does (the rest differ only there, test/check_results.pl:486-508)."""
    import gzip, json, os
    from conftest import GOLDEN
    from smithwaterman_b200 import report
    with gzip.open(os.path.join(GOLDEN, "emboss_reports.json.gz")) as f:
        bodies = {(r["kind"], r["id"]): r["body"] for r in json.loads(f.read().decode())}
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    same = {}
    for mode, key in ((swb.ALIGN_LOCAL, "local"), (swb.ALIGN_GLOBAL, "global"), (swb.ALIGN_GLOCAL, "glocal")):
        eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
        res = eng.align_list([p["seq1"].encode() for p in emboss], [p["seq2"].encode() for p in emboss], mode != swb.ALIGN_LOCAL)
        same[key] = 0
        for p, r in zip(emboss, res):
            sa, sb = report.starts_of(r["idx_a"], r["idx_b"]) if mode == swb.ALIGN_LOCAL else (1, 1)
            body = report.srspair_body(r["aln_a"], r["aln_b"], r["score"], m, 10.0, 0.5, sa, sb)
            if mode == swb.ALIGN_LOCAL or (r["aln_a"], r["aln_b"]) == (p[key]["s1"], p[key]["s2"]):
                assert body == bodies[(key, p["id"])], "%s %d" % (key, p["id"])
                same[key] += 1
        eng.close()
    assert same["local"] == 100 and same["global"] > 0


def test_cpp_cli_emboss_reports(emboss, tmp_path):
    """`swbcli -emboss -local <list> <outdir>`: the compiled host mirror writes res<k>.dat files whose text equals the
    bodies of the reference's golden water reports (ids 0..39); the needle variants equal the Python writer's output."""
    import gzip, json, os, subprocess
    from conftest import GOLDEN
    from smithwaterman_b200 import report
    with gzip.open(os.path.join(GOLDEN, "emboss_reports.json.gz")) as f:
        bodies = {(r["kind"], r["id"]): r["body"] for r in json.loads(f.read().decode())}
    recs = emboss[:40]
    lines = []
    for r in recs:
        f1, f2 = tmp_path / ("seq%s.1.fas" % r["id"]), tmp_path / ("seq%s.2.fas" % r["id"])
        f1.write_text(">s1\n%s\n" % r["seq1"]); f2.write_text(">s2\n%s\n" % r["seq2"])
        lines.append("%s\t%s" % (f1, f2))
    (tmp_path / "list.txt").write_text("\n".join(lines) + "\n")
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    for flag, key, pre, mode in (("-local", "local", "res", swb.ALIGN_LOCAL), ("-global", "global", "needle_res", swb.ALIGN_GLOBAL)):
        out = tmp_path / key
        out.mkdir()
        subprocess.run([_swbcli(), "-emboss", flag, str(tmp_path / "list.txt"), str(out)], check=True, capture_output=True)
        eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
        res = eng.align_list([r["seq1"].encode() for r in recs], [r["seq2"].encode() for r in recs], mode != swb.ALIGN_LOCAL)
        eng.close()
        for k, (r, a) in enumerate(zip(recs, res)):
            txt = (out / ("%s%d.dat" % (pre, k))).read_text()
            if key == "local":
                assert txt == bodies[("local", r["id"])], r["id"]
            sa, sb = report.starts_of(a["idx_a"], a["idx_b"]) if key == "local" else (1, 1)
            assert txt == report.srspair_body(a["aln_a"], a["aln_b"], a["score"], m, 10.0, 0.5, sa, sb)


@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL])
def test_align_partial_region_pads_over_the_whole_sequences(mode):
    """align_partial(.., partial_region) (SA.rs:393-551): the DP runs on the prefixes, the retain_all padding loops run
    over s1.seq.len() / s2.seq.len() -- the tails behind the prefixes appear as overhang (Python mirror and C ABI)."""
    rng = np.random.default_rng(31 + mode)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
    oeng = O.Engine(_omatrix(m), 10.0, 0.5, mode)
    for _ in range(12):
        a, b = _rand_seqs(rng, 1, 30, 120)[0], _rand_seqs(rng, 1, 30, 120)[0]
        b = b[:10] + a[5:40] + b[10:]
        pr = (int(rng.integers(1, len(a) + 1)), int(rng.integers(1, len(b) + 1)))
        for ra in (True, False):
            r1, r2, s = eng.align_partial(a, b, ra, pr, False)
            o1, o2, os_ = oeng.align(a.decode(), b.decode(), ra, partial_region=pr)
            assert ("".join(r1), "".join(r2), s) == ("".join(o1), "".join(o2), os_), (pr, ra)
            if ra or mode != swb.ALIGN_LOCAL:
                assert "".join(r1).replace("-", "") == a.decode() and "".join(r2).replace("-", "") == b.decode()
        _, _, s2 = eng.align_partial(a, b, True, pr, True)
        assert s2 == oeng.align(a.decode(), b.decode(), True, score_only=True, partial_region=pr)[2]
    eng.close()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL])
def test_team_launches_from_two_streams_make_progress(mode, monkeypatch):
    """Forward progress of the team-scheduled wavefront: the pipelined host-buffer call alternates its chunks over two
    compute streams, and with 1024-pair chunks that each hold 2-3 pairs of more than 8192 residues every chunk launches a
    TEAM kernel (warps that wait on each other).  Team launches are cooperative and chained through one event, so they
    can neither oversubscribe the GPU nor interleave; the scores equal the oracle's."""
    monkeypatch.setenv("SWB_CHUNK_PAIRS", "1024")
    monkeypatch.setenv("SWB_CHUNK_FIRST", "1024")
    rng = np.random.default_rng(900 + mode)
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    eng = swb.SequenceAlignment(m, 5.0, 2.0, mode)
    n = 5 * 1024
    sa = _rand_seqs(rng, n, 20, 90, "ACGT"); sb = _rand_seqs(rng, n, 20, 90, "ACGT")
    longs = []
    for c in range(5):
        for t in range(2 + (c & 1)):
            k = c * 1024 + 100 + 300 * t
            base = _rand_seqs(rng, 1, 8300 + 400 * t, 8300 + 400 * t, "ACGT")[0]
            sa[k] = base; sb[k] = _mutate(rng, base, 0.06, 0.02, 0.02)
            longs.append(k)
    sc, ea, eb = eng.score_batch(sa, sb)
    chk = longs + list(rng.choice(n, 300, replace=False))
    osc, oi, oj = _oracle_scores(m, [sa[k] for k in chk], [sb[k] for k in chk], 5.0, 2.0, mode, scale=1)
    assert np.array_equal(np.rint(sc[chk].astype(np.float64)).astype(np.int64), osc)
    nz = osc > 0 if mode == swb.ALIGN_LOCAL else np.ones(len(chk), bool)
    assert np.array_equal(ea[chk][nz], oi[nz]) and np.array_equal(eb[chk][nz], oj[nz])
    eng.close()


def test_perl_harness_over_the_cli(tmp_path):
    """NS-5: the reference's acceptance harness (test/README.md:18-27) over the compiled drop-in CLI: materialise the data
    layout from the fixtures, run list mode for the three alignment types (test/run_swb.pl), compare with the EMBOSS
    results by the reference's rules (test/check_results.pl) -> "OK / Checked 300 results"."""
    import os
    import shutil
    import subprocess
    import sys
    from conftest import ROOT
    if not shutil.which("perl"):
        pytest.skip("perl not installed")
    cli = _swbcli()
    subprocess.run([sys.executable, os.path.join(ROOT, "test", "materialize.py"), str(tmp_path)], check=True)
    r = subprocess.run(["perl", os.path.join(ROOT, "test", "run_swb.pl"), cli], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    r = subprocess.run(["perl", os.path.join(ROOT, "test", "check_results.pl")], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    assert r.stdout.splitlines()[-2:] == ["OK", "Checked 300 results"]


@pytest.mark.parametrize("swap", [None, "0", "1"])
@pytest.mark.parametrize("go,ge,protein", [(5.0, 2.0, False), (10.0, 0.5, True), (2.0, 2.0, False)])
def test_packed_traceback_path_global(go, ge, protein, swap, monkeypatch):
    """The packed traceback path (csrc/k3_tile.cu: s16x2 fill with tile borders + tile-recompute walk) in global mode:
    strings, index lists and scores equal the dense oracle's on pairs of every shape -- single pass, several sweeps of
    1024 columns, fewer rows than one checkpoint, one-letter sequences, unrelated pairs -- and equal the direction-byte
    path on 6 kb pairs."""
    # swap: SWB_K3_SWAP forces every pair through the transposed form of the kernels (rows = b) or none; default: tall pairs only
    if swap is not None:
        monkeypatch.setenv("SWB_K3_SWAP", swap)
    rng = np.random.default_rng(int(go * 7 + ge * 3))
    m = swb.SubstitutionMatrix.get_blosum62_matrix() if protein else swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    al = AA if protein else "ACGT"
    eng = swb.SequenceAlignment(m, go, ge, swb.ALIGN_GLOBAL)
    oeng = O.Engine(_omatrix(m), go, ge, swb.ALIGN_GLOBAL)
    sa, sb = [], []
    for la, lb in [(1, 1), (1, 40), (40, 1), (5, 70), (63, 33), (64, 32), (65, 31), (130, 1100), (1100, 130), (300, 2100), (700, 700), (257, 96)]:
        if protein and max(la, lb) > 1400:
            la, lb = min(la, 1300), min(lb, 1300)
        sa.append(_rand_seqs(rng, 1, la, la, al)[0]); sb.append(_rand_seqs(rng, 1, lb, lb, al)[0])
    for k in range(25):
        base = _rand_seqs(rng, 1, 30, 1300, al)[0]
        sa.append(base); sb.append(_mutate(rng, base, 0.1, 0.04, 0.04, alphabet=al.encode()))
    res = eng.align_list(sa, sb, True)
    assert eng.stats()["k3_pairs"] == len(sa)
    for k, r in enumerate(res):
        r1, r2, s = oeng.align(sa[k].decode(), sb[k].decode(), True)
        assert (r["aln_a"], r["aln_b"], r["score"]) == ("".join(r1), "".join(r2), s), k
        xa, xb, _, ei, ej = oeng.raw(_omatrix(m).seq_to_index(sa[k].decode()), _omatrix(m).seq_to_index(sb[k].decode()))
        assert np.array_equal(r["idx_a"], xa) and np.array_equal(r["idx_b"], xb) and (r["end_a"], r["end_b"]) == (ei, ej)
    if not protein:                                               # 6 kb pairs: packed path == direction-byte path
        big_a = _rand_seqs(rng, 3, 6000, 6000, al)
        big_b = [_mutate(rng, x, 0.05, 0.05, 0.05) for x in big_a]
        r3 = eng.align_list(big_a, big_b, True)
        assert eng.stats()["k3_pairs"] == 3
        monkeypatch.setenv("SWB_NO_K3", "1")
        rw = eng.align_list(big_a, big_b, True)
        assert eng.stats()["k3_pairs"] == 0
        for x, y in zip(r3, rw):
            assert (x["aln_a"], x["aln_b"], x["score"]) == (y["aln_a"], y["aln_b"], y["score"])
    eng.close()


@pytest.mark.parametrize("go,ge", [(5.0, 2.0), (10.0, 0.5)])
def test_wide_tile_path_20kb_default_routing(go, ge):
    """A 20 kb x ~20 kb DNA pair (4e8 cells, score outside s16) takes the int32 tile path by default routing, with both gap
    models SURVEY 8(d) names for configs[3] (5/2 and 10/0.5, the latter on the half-integer grid): score and end cell
    against the linear-memory integer oracle, the strings re-score to the score and spell the inputs."""
    rng = np.random.default_rng(31 + int(go))
    a = _rand_seqs(rng, 1, 20500, 20500, "ACGT")[0]
    b = _mutate(rng, a, 0.08, 0.02, 0.02)
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    om = _omatrix(m)
    scale = 1 if float(ge).is_integer() else 2
    ia, oa = O.pack([om.seq_to_index(a.decode())]); ib, ob = O.pack([om.seq_to_index(b.decode())])
    for mode in (swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL):
        eng = swb.SequenceAlignment(m, go, ge, mode)
        out = eng.align_raw(a, b, False)
        assert eng.stats()["k3_pairs"] == 1
        osc, oi, oj = O.score_batch_int(ia, oa, ib, ob, np.rint(m.scores * scale).astype(np.int32), -int(round(go * scale)),
                                        -int(round(ge * scale)), scale, mode, 1)
        assert (out["score"], out["end_a"], out["end_b"]) == (float(osc[0]) / scale, int(oi[0]), int(oj[0]))
        assert _rescore(out["aln_a"], out["aln_b"], 2.0, -3.0, go, ge) == out["score"]
        if mode == swb.ALIGN_GLOBAL:
            assert out["aln_a"].replace("-", "") == a.decode() and out["aln_b"].replace("-", "") == b.decode()
        else:
            assert out["aln_a"].replace("-", "") in a.decode() and out["aln_b"].replace("-", "") in b.decode()
        eng.close()


@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL])
@pytest.mark.parametrize("go,ge,protein", [(5.0, 2.0, False), (10.0, 0.5, True), (2.0, 2.0, False)])
def test_wide_tile_path_small_pairs(go, ge, protein, mode, monkeypatch):
    """K3W + KT3 (the int32 tile path of single long pairs, csrc/k3_tile.cu) forced onto pairs of every shape -- one pass, several
    sweeps spread over several blocks (the cross-block progress words), fewer rows than one checkpoint, one-letter sequences,
    unrelated pairs (local score 0 -> empty alignment), the maximum in the last column / last row -- against the dense
    oracle: strings, index lists, score, end cell."""
    monkeypatch.setenv("SWB_NO_K3", "1"); monkeypatch.setenv("SWB_K3W_MIN", "1")
    rng = np.random.default_rng(int(go * 11 + ge * 5 + mode))
    m = swb.SubstitutionMatrix.get_blosum62_matrix() if protein else swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    al = AA if protein else "ACGT"
    eng = swb.SequenceAlignment(m, go, ge, mode)
    oeng = O.Engine(_omatrix(m), go, ge, mode)
    sa, sb = [], []
    for la, lb in [(1, 1), (1, 40), (40, 1), (5, 70), (63, 33), (64, 32), (65, 31), (130, 1100), (1100, 130), (300, 3100), (700, 700), (257, 96), (4, 2050)]:
        sa.append(_rand_seqs(rng, 1, la, la, al)[0]); sb.append(_rand_seqs(rng, 1, lb, lb, al)[0])
    for k in range(14):
        base = _rand_seqs(rng, 1, 30, 2600, al)[0]
        sa.append(base); sb.append(_mutate(rng, base, 0.1, 0.04, 0.04, alphabet=al.encode()))
    sa.append(_rand_seqs(rng, 1, 900, 900, al)[0]); sb.append(_rand_seqs(rng, 1, 50, 50, al)[0] + sa[-1][300:] )      # maximum in the last row and column
    if not protein:
        sa.append(b"AAAA"); sb.append(b"CCCCCC")                                                              # local: score 0
    res = eng.align_list(sa, sb, True)
    assert eng.stats()["k3_pairs"] >= 1
    for k, r in enumerate(res):
        r1, r2, s = oeng.align(sa[k].decode(), sb[k].decode(), True)
        assert (r["aln_a"], r["aln_b"], r["score"]) == ("".join(r1), "".join(r2), s), k
        xa, xb, _, ei, ej = oeng.raw(_omatrix(m).seq_to_index(sa[k].decode()), _omatrix(m).seq_to_index(sb[k].decode()))
        assert np.array_equal(r["idx_a"], xa) and np.array_equal(r["idx_b"], xb) and (r["end_a"], r["end_b"]) == (ei, ej), k
    eng.close()


@pytest.mark.parametrize("swap", [None, "0", "1"])
def test_packed_traceback_path_local_hits(swap, monkeypatch):
    """Local mode on the packed path: hits of a database search are traced inside the prefix box of their end cell with
    the scan's score as the s16 guarantee; strings equal the oracle's, with and without retain_all, for subjects of one
    pass up to several sweeps, and the direction-byte path gives the same."""
    from smithwaterman_b200 import search as S
    if swap is not None:
        monkeypatch.setenv("SWB_K3_SWAP", swap)
    rng = np.random.default_rng(99)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    oeng = O.Engine(_omatrix(m), 10.0, 0.5, swb.ALIGN_LOCAL)
    query = _rand_seqs(rng, 1, 1300, 1300)[0]
    seqs = _rand_seqs(rng, 200, 20, 400)
    for k in range(0, 200, 10):
        lo = int(rng.integers(0, 900)); w = query[lo:lo + int(rng.integers(40, 400))]
        seqs[k] = _rand_seqs(rng, 1, 0, 50)[0] + _mutate(rng, w, 0.2, 0.03, 0.03, alphabet=AA.encode()) + _rand_seqs(rng, 1, 0, 50)[0]
    seqs[5] = _rand_seqs(rng, 1, 1500, 1500)[0] + query[100:1200] + _rand_seqs(rng, 1, 300, 300)[0]     # hit far right in a long subject: 3 sweeps
    db = S.Database.from_sequences(seqs)
    ddb = S.DeviceDatabase(eng, db)
    hits = ddb.search(query, 30)
    for ra in (False, True):
        alns = ddb.align(query, hits, ra)
        assert eng.stats()["k3_pairs"] >= 20
        for k, h in enumerate(hits):
            r1, r2, s = oeng.align(query.decode(), seqs[int(h["index"])].decode(), ra)
            assert (alns[k]["aln_a"], alns[k]["aln_b"], alns[k]["score"]) == ("".join(r1), "".join(r2), s), (k, ra)
    ddb.free(); eng.close()
