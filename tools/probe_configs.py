#!/usr/bin/env python3
"""GPU probe of the non-headline configs (BASELINE.json configs[2..4]) at bounded sizes: times the C-ABI calls and
prints one JSON line per case.  Not a bench line: bench.py stays on configs[1]."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smithwaterman_b200 as swb  # noqa: E402
from smithwaterman_b200 import search as S  # noqa: E402

AA = np.frombuffer(b"ARNDCQEGHILKMFPSTWYV", np.uint8)
NT = np.frombuffer(b"ACGT", np.uint8)


def mutate(rng, a, sub, ins, dele, alphabet):
    out = []
    for c in a:
        u = rng.random()
        if u < dele:
            continue
        if u < dele + sub:
            out.append(int(alphabet[rng.integers(0, len(alphabet))]))
        else:
            out.append(int(c))
        if rng.random() < ins:
            out.append(int(alphabet[rng.integers(0, len(alphabet))]))
    return np.array(out, np.uint8)


def mutate_fast(rng, a, sub, ins, dele, alphabet):
    n = len(a)
    keep = rng.random(n) >= dele
    b = a.copy()
    s = rng.random(n) < sub
    b[s] = alphabet[rng.integers(0, len(alphabet), int(s.sum()))]
    b = b[keep]
    pos = np.nonzero(rng.random(len(b)) < ins)[0]
    b = np.insert(b, pos, alphabet[rng.integers(0, len(alphabet), len(pos))])
    return b


def case_search(lq, db_res, topk):
    rng = np.random.default_rng(20243)
    q = AA[rng.integers(0, 20, lq)]
    lens = np.clip(np.rint(np.exp(rng.normal(5.6, 0.6, int(db_res / 300)))), 30, 35000).astype(np.int64)
    lens = lens[np.cumsum(lens) <= db_res]
    off = np.zeros(len(lens) + 1, np.int64); off[1:] = np.cumsum(lens)
    res = AA[rng.integers(0, 20, int(off[-1]))]
    # planted homologs
    for k in rng.integers(0, len(lens), 50):
        w = int(min(lens[k], rng.integers(200, 2000), lq))
        s0 = int(rng.integers(0, lq - w + 1))
        m = mutate_fast(rng, q[s0:s0 + w], 0.3, 0.03, 0.03, AA)[:lens[k]]
        res[off[k]:off[k] + len(m)] = m
    db = S.Database(res, off)
    eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, swb.ALIGN_LOCAL)
    S.search_scores(eng, q.tobytes(), db)
    t0 = time.perf_counter()
    sc, eq, et = S.search_scores(eng, q.tobytes(), db)
    t1 = time.perf_counter()
    st = eng.stats()
    hits = S.search(eng, q.tobytes(), db, topk=topk, traceback=True)
    t2 = time.perf_counter()
    cells = float(lq) * float(off[-1])
    print(json.dumps({"case": "search", "lq": lq, "db_residues": int(off[-1]), "n_db": len(lens), "cells": cells,
                      "score_s": t1 - t0, "gcups_e2e": cells / (t1 - t0) / 1e9, "device_ms": st["kernel_ms"],
                      "gcups_device": cells / (st["kernel_ms"] * 1e-3) / 1e9, "reruns": st["i32_reruns"],
                      "search_topk_s": t2 - t1, "top_score": float(np.max(sc))}), flush=True)
    eng.close()
    return hits


def case_pairs(n, length, mode, alphabet, go, ge, match, mismatch, sub, indel, traceback, tag):
    rng = np.random.default_rng(20245)
    sa, sb = [], []
    for _ in range(n):
        a = alphabet[rng.integers(0, len(alphabet), length)]
        b = mutate_fast(rng, a, sub, indel, indel, alphabet)
        sa.append(a.tobytes()); sb.append(b.tobytes())
    m = swb.SubstitutionMatrix.get_mat_matrix(match, mismatch) if match is not None else swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, go, ge, mode)
    cells = float(sum(len(x) * len(y) for x, y in zip(sa, sb)))
    if traceback:                                             # warm-up: module load, scratch and direction-store growth
        eng.align_list(sa, sb, True)
    else:
        eng.score_batch(sa[:1], sb[:1])
    t0 = time.perf_counter()
    if traceback:
        out = eng.align_list(sa, sb, True)
        score = out[0]["score"]; ncols = len(out[0]["aln_a"])
    else:
        sc, _, _ = eng.score_batch(sa, sb)
        score = float(sc[0]); ncols = 0
    t1 = time.perf_counter()
    st = eng.stats()
    print(json.dumps({"case": tag, "n": n, "len": length, "mode": mode, "traceback": traceback, "cells": cells,
                      "wall_s": t1 - t0, "device_ms": st["kernel_ms"], "gcups_device": cells / max(st["kernel_ms"], 1e-6) / 1e6,
                      "gcups_wall": cells / (t1 - t0) / 1e9, "score0": score, "aln_len0": ncols}), flush=True)
    eng.close()


if __name__ == "__main__":
    which = sys.argv[1:] or ["search", "nw10k", "dna30k"]
    if "search" in which:
        case_search(5000, int(float(os.environ.get("DB_RES", "2e7"))), 100)
    if "nw10k" in which:
        case_pairs(int(os.environ.get("NW_PAIRS", "64")), 10000, swb.ALIGN_GLOBAL, NT, 5.0, 2.0, 2.0, -3.0, 0.05, 0.05, True, "nw10k")
        case_pairs(int(os.environ.get("NW_PAIRS", "64")), 10000, swb.ALIGN_GLOBAL, NT, 5.0, 2.0, 2.0, -3.0, 0.05, 0.05, False, "nw10k-score")
    if "dna30k" in which:
        L = int(os.environ.get("DNA_LEN", "30000"))
        case_pairs(1, L, swb.ALIGN_LOCAL, NT, 5.0, 2.0, 2.0, -3.0, 0.08, 0.02, True, "dna-long-local")
        case_pairs(1, L, swb.ALIGN_GLOBAL, NT, 5.0, 2.0, 2.0, -3.0, 0.08, 0.02, True, "dna-long-global")
        case_pairs(1, L, swb.ALIGN_LOCAL, NT, 5.0, 2.0, 2.0, -3.0, 0.08, 0.02, False, "dna-long-local-score")
