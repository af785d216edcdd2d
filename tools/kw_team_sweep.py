#!/usr/bin/env python3
"""Tuning aid: one 100 kb x 100 kb DNA pair (configs[3]) under forced team sizes / strip widths (SWB_KW_TEAM, SWB_KW_COLS)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import smithwaterman_b200 as swb
from ckpt_probe import mutate, NT
rng = np.random.default_rng(20260)
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000
a = NT[rng.integers(0, 4, n)]; b = mutate(rng, a, 0.10, 0.02, 0.02)
m = swb.SubstitutionMatrix.get_mat_matrix(2.0, -3.0)
eng = swb.SequenceAlignment(m, 5.0, 2.0, swb.ALIGN_GLOBAL)
for cols, teams in (("2", (0, 300, 521, 782, 1042, 1563, 2300)), ("8", (0, 98, 131, 196, 391))):
    os.environ["SWB_KW_COLS"] = cols
    for t in teams:
        if t: os.environ["SWB_KW_TEAM"] = str(t)
        else: os.environ.pop("SWB_KW_TEAM", None)
        for trace in (False, True):
            best = 1e9
            for rep in range(3):
                if trace: eng.align_raw(a.tobytes(), b.tobytes(), True)
                else: eng.score_batch([a.tobytes()], [b.tobytes()])
                best = min(best, eng.stats()["kernel_ms"])
            print("cols %s team %5s %s: %.1f ms device" % (cols, t or "auto", "traceback" if trace else "score    ", best), flush=True)
