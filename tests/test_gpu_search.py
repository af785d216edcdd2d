"""GPU tests of the native database-search driver (csrc/search.cu behind include/swb200.h): resident database, device-side
top-K, shard merge, per-owner traceback -- against the oracle and against the host-buffer path."""
import ctypes

import numpy as np
import pytest

import smithwaterman_b200 as swb
from smithwaterman_b200 import _lib
from smithwaterman_b200 import search as S
from oracle import oracle as O

pytestmark = pytest.mark.gpu

AA = "ARNDCQEGHILKMFPSTWYV"


def _rand(rng, n, lo, hi, alphabet=AA):
    al = np.frombuffer(alphabet.encode(), np.uint8)
    return [bytes(al[rng.integers(0, len(al), int(rng.integers(lo, hi + 1)))]) for _ in range(n)]


def _oracle_scores(m, sa, sb, go, ge, mode, scale=2, threads=8):
    om = O.Matrix(m.index_to_string, m.scores)
    a, oa = O.pack([om.seq_to_index(s.decode()) for s in sa])
    b, ob = O.pack([om.seq_to_index(s.decode()) for s in sb])
    return O.score_batch_int(a, oa, b, ob, np.rint(m.scores * scale).astype(np.int32), -int(round(go * scale)), -int(round(ge * scale)),
                             scale, mode, threads)


def _planted_db(rng, query, n, lo, hi, every):
    seqs = _rand(rng, n, lo, hi)
    for k in range(0, n, every):
        s0 = int(rng.integers(0, max(1, len(query) - 300))); w = bytearray(query[s0:s0 + int(rng.integers(100, 300))])
        for p in rng.integers(0, len(w), len(w) // 4):
            w[p] = ord(AA[rng.integers(0, 20)])
        seqs[k] = bytes(w)
    return seqs


@pytest.mark.parametrize("lq", [700, 1500])
def test_native_search_topk_equals_host_selection_and_oracle(lq):
    """swb_search (device top-K) == host selection over the full score vector == the oracle; ties by lower index;
    empty subjects; both the K1 (short query) and the K2 (long query) scan."""
    rng = np.random.default_rng(100 + lq)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    query = _rand(rng, 1, lq, lq)[0]
    seqs = _planted_db(rng, query, 2500, 20, 800, 50) + [b"", b"A"] + _rand(rng, 6, 1100, 2500)
    seqs[51] = seqs[50]; seqs[1201] = seqs[50]                    # exact copies of a hit: ties, lower index first
    db = S.Database.from_sequences(seqs)
    ddb = S.DeviceDatabase(eng, db)
    for topk in (40, 1, 3000):
        hits = ddb.search(query, topk)
        st = eng.stats()
        sc, eq, et = ddb.scores()
        sc2, eq2, et2 = S.search_scores(eng, query, db)           # host-buffer path
        assert np.array_equal(sc, sc2) and np.array_equal(eq, eq2) and np.array_equal(et, et2)
        ref = S.topk_records(sc, np.arange(db.n), eq, et, topk)
        assert len(hits) == min(topk, db.n)
        for f in ("score", "index", "end_q", "end_t"):
            assert np.array_equal(hits[f], ref[f]), (topk, f)
        assert st["d2h_bytes"] <= 32 * min(topk, db.n) + 64        # the score vector never crosses PCIe
    osc, oi, oj = _oracle_scores(m, [query] * len(seqs), seqs, 10.0, 0.5, swb.ALIGN_LOCAL)
    assert np.array_equal(np.rint(sc.astype(np.float64) * 2).astype(np.int64), osc)
    hits = ddb.search(query, 40)
    order = np.lexsort((np.arange(len(seqs)), -osc))[:40]
    assert hits["index"].tolist() == order.tolist()
    pos = hits["index"].tolist()
    assert pos.index(50) + 1 == pos.index(51) and pos.index(51) + 1 == pos.index(1201)
    # traceback of the hits from the resident residues, inside the prefix box of each hit
    oeng = O.Engine(O.Matrix(m.index_to_string, m.scores), 10.0, 0.5, swb.ALIGN_LOCAL)
    for ra in (False, True):
        alns = ddb.align(query, hits, retain_all=ra)
        assert sorted(alns) == list(range(40))
        for k, h in enumerate(hits):
            r1, r2, s = oeng.align(query.decode(), seqs[int(h["index"])].decode(), ra)
            assert (alns[k]["aln_a"], alns[k]["aln_b"], alns[k]["score"]) == ("".join(r1), "".join(r2), s), (k, ra)
            assert (alns[k]["end_a"], alns[k]["end_b"]) == (h["end_q"], h["end_t"])
    ddb.free(); eng.close()


@pytest.mark.parametrize("world", [2, 3, 8])
def test_native_search_shards_merge_to_the_single_list(world):
    """The shard path on one GPU: every shard resident in turn, per-shard device top-K, swb_search_merge == the
    single-database list; the per-shard score vectors scatter to the single vector; every hit has exactly one owner."""
    rng = np.random.default_rng(7 + world)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    query = _rand(rng, 1, 1200, 1200)[0]
    seqs = _planted_db(rng, query, 1800, 20, 700, 36)
    seqs[37] = seqs[36]
    db = S.Database.from_sequences(seqs)
    single = S.DeviceDatabase(eng, db)
    ref = single.search(query, 30).copy()
    full = single.scores()[0].copy()
    ref_alns = single.align(query, ref)
    parts, vec, owners, alns = [], np.full(db.n, -1.0, np.float32), np.zeros(30, int), {}
    for r in range(world):
        idx = S.native_shard_indices(db.offsets, r, world)
        assert np.array_equal(idx, S.shard_indices(db.lengths(), world)[r])
        sh = S.DeviceDatabase(eng, db, r, world)
        assert sh.n == len(idx)
        parts.append(sh.search(query, 30).copy())
        vec[idx] = sh.scores()[0]
        mine = sh.align(query, ref)
        for k, a in mine.items():
            owners[k] += 1; alns[k] = a
        sh.free()
    merged = S.merge_hits(parts, 30)
    assert merged.tobytes() == ref.tobytes()
    assert np.array_equal(vec, full)
    assert (owners == 1).all()
    for k in range(30):
        assert (alns[k]["aln_a"], alns[k]["aln_b"], alns[k]["score"]) == (ref_alns[k]["aln_a"], ref_alns[k]["aln_b"], ref_alns[k]["score"])
    single.free(); eng.close()


def test_search_sharded_one_process_two_engines():
    """swb_search_sharded: two engines (threads) on the same device, two resident shards, merged hits + alignments."""
    rng = np.random.default_rng(5)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    query = _rand(rng, 1, 1100, 1100)[0]
    seqs = _planted_db(rng, query, 1200, 20, 600, 40)
    db = S.Database.from_sequences(seqs)
    e0 = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    ref_db = S.DeviceDatabase(e0, db)
    ref = ref_db.search(query, 20).copy()
    ref_alns = ref_db.align(query, ref, True)
    engs = [swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL) for _ in range(2)]
    dbs = [S.DeviceDatabase(engs[g], db, g, 2) for g in range(2)]
    L = _lib.lib()
    eh = (ctypes.c_void_p * 2)(*[e._h for e in engs]); dh = (ctypes.c_void_p * 2)(*[d.h for d in dbs])
    hits = np.zeros(20, S.HIT_DTYPE); nh = ctypes.c_int64(0); outs = (_lib.Alignment * 20)()
    rc = L.swb_search_sharded(eh, dh, 2, query, len(query), 20, 1, hits.ctypes.data, ctypes.byref(nh), outs)
    assert rc == 0 and nh.value == 20
    for f in ("score", "index", "end_q", "end_t"):
        assert np.array_equal(hits[f], ref[f])
    for k in range(20):
        a = swb.SequenceAlignment._unpack(outs[k])
        assert (a["aln_a"], a["aln_b"]) == (ref_alns[k]["aln_a"], ref_alns[k]["aln_b"])
        L.swb_alignment_free(ctypes.byref(outs[k]))
    for d in dbs + [ref_db]:
        d.free()
    for e in engs + [e0]:
        e.close()


def test_search_overflow_rerun_and_other_modes():
    """A subject that contains the whole long query overflows the s16 lanes of the scan: re-run in int32 before the
    selection.  Global / glocal engines search through the same driver (scores of cell (lq, lt))."""
    rng = np.random.default_rng(11)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    query = _rand(rng, 1, 3600, 3600)[0]
    seqs = _rand(rng, 300, 50, 600)
    seqs[77] = seqs[77][:100] + query + seqs[77][100:200]         # ~3600 matches of ~5.5 on average, x scale 2 -> > 32767
    seqs[200] = query[500:3500]
    db = S.Database.from_sequences(seqs)
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    ddb = S.DeviceDatabase(eng, db)
    hits = ddb.search(query, 5)
    assert eng.stats()["i32_reruns"] >= 1
    assert hits["index"][0] == 77 and hits["score"][0] > 16384 and hits["index"][1] == 200
    osc, oi, oj = _oracle_scores(m, [query] * 2, [seqs[77], seqs[200]], 10.0, 0.5, swb.ALIGN_LOCAL)
    assert hits["score"][:2].tolist() == (osc / 2.0).tolist() and hits["end_q"][:2].tolist() == oi.tolist() and hits["end_t"][:2].tolist() == oj.tolist()
    al = ddb.align(query, hits[:2])
    assert al[0]["score"] == hits["score"][0] and al[0]["aln_a"].replace("-", "") in query.decode()
    ddb.free(); eng.close()
    q2, s2 = query[:400], [s[:500] for s in seqs[:120]] + [b""]
    for mode in (swb.ALIGN_GLOBAL, swb.ALIGN_GLOCAL):
        eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
        d2 = S.DeviceDatabase(eng, S.Database.from_sequences(s2))
        h = d2.search(q2, 10)
        sc = d2.scores()[0]
        osc, _, _ = _oracle_scores(m, [q2] * len(s2), s2, 10.0, 0.5, mode)
        assert np.array_equal(np.rint(sc.astype(np.float64) * 2).astype(np.int64), osc)
        order = np.lexsort((np.arange(len(s2)), -osc))[:10]
        assert h["index"].tolist() == order.tolist()
        d2.free(); eng.close()


def test_config2_named_size_against_the_oracle():
    """BASELINE configs[2] at the named size (5000-aa query, 2e8-residue database, 500 planted homologs): the top-100
    list and its alignment strings equal the oracle's over the planted set + a seeded 1 % sample of the database, and
    the full score vector of the two-shard path equals the single-database vector."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import search_bench as SB
    q, res, off = SB.make_db(5000, int(2e8))
    db = S.Database(res, off)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    ddb = S.DeviceDatabase(eng, db)
    hits = ddb.search(q, 100).copy()
    sc, eq, et = ddb.scores()
    sc = sc.copy(); eq = eq.copy(); et = et.copy()
    alns = ddb.align(q, hits)
    ddb.free()
    # two shards, one after the other on this GPU: merge == single list, scattered vectors == single vector
    vec = np.full(db.n, -1.0, np.float32); parts = []
    for r in range(2):
        sh = S.DeviceDatabase(eng, db, r, 2)
        parts.append(sh.search(q, 100).copy())
        vec[S.native_shard_indices(db.offsets, r, 2)] = sh.scores()[0]
        sh.free()
    assert np.array_equal(vec, sc)
    assert S.merge_hits(parts, 100).tobytes() == hits.tobytes()
    # oracle over every subject scoring >= the 100th hit's neighbourhood + a 1 % sample
    rng = np.random.default_rng(20243)
    sample = np.unique(np.concatenate([hits["index"], np.argsort(-sc, kind="stable")[:600], rng.choice(db.n, db.n // 100, replace=False)]))
    subs = [db.sequence(int(k)) for k in sample]
    osc, oi, oj = _oracle_scores(m, [q] * len(subs), subs, 10.0, 0.5, swb.ALIGN_LOCAL, threads=os.cpu_count() or 8)
    assert np.array_equal(np.rint(sc[sample].astype(np.float64) * 2).astype(np.int64), osc)
    nz = osc > 0
    assert np.array_equal(eq[sample][nz], oi[nz]) and np.array_equal(et[sample][nz], oj[nz])
    # the top-100 of the sample under the oracle's scores is the device list (every candidate for the list is in the sample)
    order = np.lexsort((sample, -osc))[:100]
    assert sample[order].tolist() == hits["index"].tolist()
    oeng = O.Engine(O.Matrix(m.index_to_string, m.scores), 10.0, 0.5, swb.ALIGN_LOCAL)
    for k in range(0, 100, 7):                                    # strings of a spread of the hits vs the dense oracle
        r1, r2, s = oeng.align(q.decode(), db.sequence(int(hits["index"][k])).decode(), False)
        assert (alns[k]["aln_a"], alns[k]["aln_b"], alns[k]["score"]) == ("".join(r1), "".join(r2), s)
    eng.close()
