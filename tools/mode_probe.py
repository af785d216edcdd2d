import sys, time, numpy as np
sys.path.insert(0, ".")
import smithwaterman_b200 as swb
from bench import make_pairs
a, oa, b, ob = make_pairs(1000000, 20242)
cells = float(np.sum((oa[1:] - oa[:-1]) * (ob[1:] - ob[:-1])))
for mode, name in ((swb.ALIGN_GLOBAL, "global"), (swb.ALIGN_LOCAL, "local"), (swb.ALIGN_GLOCAL, "glocal")):
    n = 1000000 if mode != swb.ALIGN_GLOCAL else 100000
    c = float(np.sum((oa[1:n+1] - oa[:n]) * (ob[1:n+1] - ob[:n])))
    eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, mode)
    bt = eng.upload(a[:oa[n]], oa[:n+1], b[:ob[n]], ob[:n+1])
    bt.run_n(2)
    ms, nl = bt.run_n(3)
    print(name, "pairs", n, "ms/step", ms / 3, "GCUPS", c / (ms / 3 * 1e-3) / 1e9, "launches", nl)
    bt.free(); eng.close()
