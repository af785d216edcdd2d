#!/usr/bin/env python3
"""Small pass through every kernel of the library (K1, K2, KW with and without directions for both strip widths, the
team-scheduled long-pair path, KT, the PSSM fill, the staged batch API) in a few seconds: the run to put under a
profiler or a debugger (`python tools/exercise_kernels.py [k1 k2 kw team search k3 k3w glocal pssm staged static]`).  No oracle here -- the parity
tests check the values.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smithwaterman_b200 as swb                                   # noqa: E402
from smithwaterman_b200 import search as S                         # noqa: E402
from smithwaterman_b200.engine import pack_sequences               # noqa: E402

AA = np.frombuffer(b"ARNDCQEGHILKMFPSTWYV", np.uint8)


def rand(rng, n, lo, hi):
    return [AA[rng.integers(0, 20, int(rng.integers(lo, hi + 1)))].tobytes() for _ in range(n)]


def main():
    rng = np.random.default_rng(7)
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    only = set(sys.argv[1:])

    def want(k):
        return not only or k in only

    if want("k1"):                                                 # K1: ragged groups, empty / 1-letter pairs, several passes
        eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
        a = rand(rng, 700, 1, 150) + [b"", b"A", b"W" * 33]
        b = rand(rng, 700, 1, 150) + [b"AA", b"", b"W" * 65]
        sc, _, _ = eng.score_batch(a, b)
        print("k1", float(sc.sum()), eng.stats()["launches"])
        eng.close()
    if want("k2"):                                                 # K2: packed warps + one multi-sweep subject
        eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
        q = rand(rng, 1, 1100, 1100)[0]
        db = S.Database.from_sequences(rand(rng, 150, 1, 300) + rand(rng, 2, 1030, 1300) + [b"", b"A"])
        sc, _, _ = S.search_scores(eng, q, db)
        print("k2", float(sc.sum()))
        eng.close()
    for mode in (swb.ALIGN_LOCAL, swb.ALIGN_GLOCAL, swb.ALIGN_GLOBAL):
        if want("kw"):
            for cols in ("2", "8"):                                # KW + KT, both strip widths, several strips per pair
                os.environ["SWB_KW_COLS"] = cols
                eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
                a, b = rand(rng, 40, 1, 400), rand(rng, 40, 1, 400)
                res = eng.align_list([swb.SeqData.create("", "", x.decode()) for x in a],
                                     [swb.SeqData.create("", "", x.decode()) for x in b], True)
                print("kw dirs", mode, cols, sum(r["score"] for r in res))
                if mode != swb.ALIGN_LOCAL:
                    sc, _, _ = eng.score_batch(a, b)               # KW score-only (non-local batches)
                    print("kw score", mode, cols, float(sc.sum()))
                eng.close()
            os.environ.pop("SWB_KW_COLS", None)
        if want("team"):                                           # team-scheduled wavefront: one pair, many warps
            eng = swb.SequenceAlignment(m, 10.0, 0.5, mode)
            a, b = rand(rng, 1, 2500, 2500)[0], rand(rng, 1, 2100, 2100)[0]
            r = eng.align_raw(a, b, False)
            print("team", mode, r["score"])
            eng.close()
    if want("search"):                                             # resident database: K2 scan, device top-K, packed traceback of the hits (K3 + KT3)
        eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
        q = rand(rng, 1, 1100, 1100)[0]
        seqs = rand(rng, 300, 1, 300) + [q[100:900], b"", q[:500] + rand(rng, 1, 1200, 1200)[0]]
        ddb = S.DeviceDatabase(eng, S.Database.from_sequences(seqs))
        hits = ddb.search(q, 10)
        al = ddb.align(q, hits)
        print("search", float(hits["score"].sum()), len(al), eng.stats()["k3_pairs"])
        for w in range(2):                                         # two shards of the same database
            sh = S.DeviceDatabase(eng, S.Database.from_sequences(seqs), w, 2)
            print("shard", w, float(sh.search(q, 5)["score"].sum()))
            sh.free()
        ddb.free(); eng.close()
    if want("k3"):                                                 # packed traceback path, global mode: several sweeps and checkpoints
        eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_GLOBAL)
        a, b = rand(rng, 12, 1, 1300), rand(rng, 12, 1, 1300)
        res = eng.align_list(a, b, True)
        print("k3 global", sum(r["score"] for r in res), eng.stats()["k3_pairs"])
        eng.close()
    if want("glocal"):                                             # packed glocal score kernel (right-aligned b)
        eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_GLOCAL)
        sc, _, _ = eng.score_batch(rand(rng, 500, 1, 200), rand(rng, 500, 1, 200))
        print("glocal", float(sc.sum()))
        eng.close()
    if want("pssm"):
        pm = swb.PositionSpecificMatrix()
        eng = swb.SequenceAlignment(pm, 10.0, 0.5, swb.ALIGN_GLOBAL)
        s1, s2 = "ACDEFGHIKLMNPQRSTVWY" * 3, "ACDEFGHIKLNPQRSTVWYY" * 2
        eng.prepare(s1, s2)
        pm.scores[:len(s1) * len(s2)] = rng.integers(-4, 9, len(s1) * len(s2)).astype(np.float32)
        r1, r2, sc = eng.align(s1, s2, True)
        print("pssm", sc)
        eng.close()
    if want("staged"):                                             # staged API: upload once, run twice, fetch
        eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
        a, b = rand(rng, 300, 20, 120), rand(rng, 300, 20, 120)
        pa, oa = pack_sequences(a)
        pb, ob = pack_sequences(b)
        bt = eng.upload(pa, oa, pb, ob)
        bt.run(); bt.run_n(2)
        sc = bt.fetch()[0]
        print("staged", float(np.asarray(sc).sum()))
        bt.free()
        eng.close()
    if want("k3w"):                                                # int32 tile path of single long pairs, forced onto two small pairs
        os.environ["SWB_NO_K3"] = "1"; os.environ["SWB_K3W_MIN"] = "1"
        try:
            for mode in (swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL):
                eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0), 5.0, 2.0, mode)
                s1 = rand(rng, 1, 2500, 2500)[0]
                r = eng.align_list([s1, s1[:300]], [s1[100:2400], s1[:290]], True)
                print("k3w", mode, r[0]["score"], eng.stats()["k3_pairs"])
                eng.close()
        finally:
            os.environ.pop("SWB_NO_K3", None); os.environ.pop("SWB_K3W_MIN", None)
    if want("static"):                                             # allocation-free scorer: scratch in a static __device__ array
        import ctypes
        import torch
        eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
        a, b = rand(rng, 300, 0, 600), rand(rng, 300, 0, 600)
        pa, oa = pack_sequences(a)
        pb, ob = pack_sequences(b)
        ta, tb = torch.from_numpy(pa).cuda(), torch.from_numpy(pb).cuda()
        toa, tob = torch.from_numpy(oa).cuda(), torch.from_numpy(ob).cuda()
        tsc = torch.zeros(300, dtype=torch.float32, device="cuda"); tst = torch.zeros(1, dtype=torch.int32, device="cuda")
        for _ in range(2):
            eng._check(eng.L.swb_score_batch_static(eng._h, ta.data_ptr(), toa.data_ptr(), len(pa), tb.data_ptr(), tob.data_ptr(), len(pb),
                                                    300, tsc.data_ptr(), None, None, tst.data_ptr(), ctypes.c_void_p(0)))
        torch.cuda.synchronize()
        print("static", float(tsc.sum().item()), int(tst.item()))
        eng.close()
    print("exercise_kernels: done")


if __name__ == "__main__":
    main()
