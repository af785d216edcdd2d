#!/usr/bin/env python3
"""GPU probe of the checkpointed traceback (BASELINE configs[3]): a 100 kb x 100 kb DNA pair through the direct
direction store and through checkpointed recompute (SWB_CKPT_BYTES), and one pair whose direction bytes exceed HBM
(default 450 kb x 450 kb = 203 GB) which only the checkpointed path can trace.  One JSON line per case."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smithwaterman_b200 as swb  # noqa: E402

NT = np.frombuffer(b"ACGT", np.uint8)


def mutate(rng, a, sub, ins, dele):
    b = a.copy()
    s = rng.random(len(a)) < sub
    b[s] = NT[rng.integers(0, 4, int(s.sum()))]
    b = b[rng.random(len(b)) >= dele]
    pos = np.nonzero(rng.random(len(b)) < ins)[0]
    return np.insert(b, pos, NT[rng.integers(0, 4, len(pos))])


def rescore(aln_a, aln_b, match, mismatch, go, ge):
    a = np.frombuffer(aln_a.encode(), np.uint8); b = np.frombuffer(aln_b.encode(), np.uint8)
    ga, gb = a == ord("-"), b == ord("-")
    both = ~ga & ~gb
    sc = float(np.sum(np.where(a[both] == b[both], match, mismatch)))
    for g in (ga, gb):
        prev = np.concatenate(([False], g[:-1]))
        sc -= go * float(np.sum(g & ~prev)) + ge * float(np.sum(g & prev))
    return sc


def run(eng, a, b, retain_all, budget):
    if budget is None:
        os.environ.pop("SWB_CKPT_BYTES", None)
    else:
        os.environ["SWB_CKPT_BYTES"] = str(budget)
    eng.align_raw(a[:2000].tobytes(), b[:2000].tobytes(), retain_all)     # warm
    t0 = time.perf_counter()
    out = eng.align_raw(a.tobytes(), b.tobytes(), retain_all)
    dt = time.perf_counter() - t0
    st = eng.stats()
    return out, dt, st


def main():
    big = int(float(sys.argv[1])) if len(sys.argv) > 1 else 450_000
    rng = np.random.default_rng(20260)
    m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
    a = NT[rng.integers(0, 4, 100_000)]
    b = mutate(rng, a, 0.10, 0.02, 0.02)
    for mode, name in ((swb.ALIGN_LOCAL, "local"), (swb.ALIGN_GLOBAL, "global")):
        eng = swb.SequenceAlignment(m, 5.0, 2.0, mode)
        ra = mode != swb.ALIGN_LOCAL
        ref, dt0, st0 = run(eng, a, b, ra, None)
        ref, dt0, st0 = run(eng, a, b, ra, None)                        # second call: direction store already allocated
        for budget in (4 << 30, 1 << 30, 256 << 20):
            out, dt, st = run(eng, a, b, ra, budget)
            same = out["score"] == ref["score"] and out["aln_a"] == ref["aln_a"] and out["aln_b"] == ref["aln_b"]
            print(json.dumps({"case": "configs[3] 100 kb x %d, %s, full traceback" % (len(b), name), "cells": float(len(a)) * len(b),
                              "direct_ms_wall": round(dt0 * 1e3, 1), "direct_ms_device": round(st0["kernel_ms"], 1),
                              "ckpt_budget_bytes": budget, "ckpt_ms_wall": round(dt * 1e3, 1), "ckpt_ms_device": round(st["kernel_ms"], 1),
                              "ckpt_launches": st["launches"], "identical_to_direct": bool(same), "score": out["score"]}), flush=True)
        eng.close()
    a = NT[rng.integers(0, 4, big)]
    b = mutate(rng, a, 0.10, 0.02, 0.02)
    for mode, name in ((swb.ALIGN_LOCAL, "local"), (swb.ALIGN_GLOBAL, "global")):
        eng = swb.SequenceAlignment(m, 5.0, 2.0, mode)
        out, dt, st = run(eng, a, b, mode != swb.ALIGN_LOCAL, None)      # default budget: 24 GB of direction bytes per band
        rs = rescore(out["aln_a"], out["aln_b"], 2.0, -3.0, 5.0, 2.0)
        spelled = out["aln_a"].replace("-", "") in a.tobytes().decode() if mode == swb.ALIGN_LOCAL else out["aln_a"].replace("-", "") == a.tobytes().decode()
        cells = float(len(a)) * len(b)
        print(json.dumps({"case": "%d x %d DNA pair, %s, checkpointed traceback (direct store would need %.0f GB)" % (len(a), len(b), name, cells / 1e9),
                          "cells": cells, "ms_wall": round(dt * 1e3, 1), "ms_device": round(st["kernel_ms"], 1), "gcups_wall": round(cells / dt / 1e9, 1),
                          "launches": st["launches"], "ckpt_pairs": st["ckpt_pairs"], "score": out["score"], "rescored": rs,
                          "rescore_equal": bool(rs == out["score"]), "spells_input": bool(spelled), "aln_len": len(out["aln_a"])}), flush=True)
        eng.close()


if __name__ == "__main__":
    main()
