"""Score-batch throughput of the three modes on the bench workload (configs[1] shapes): device time of the resident batch,
one JSON line per mode."""
import json, sys, numpy as np
sys.path.insert(0, ".")
import smithwaterman_b200 as swb
from bench import make_pairs
n = 1000000
a, oa, b, ob = make_pairs(n, 20242)
cells = float(np.sum((oa[1:] - oa[:-1]) * (ob[1:] - ob[:-1])))
for mode, name in ((swb.ALIGN_LOCAL, "local"), (swb.ALIGN_GLOBAL, "global"), (swb.ALIGN_GLOCAL, "glocal")):
    eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, mode)
    bt = eng.upload(a, oa, b, ob)
    bt.run_n(2)
    ms, nl = bt.run_n(5)
    st = eng.stats()
    print(json.dumps({"mode": name, "pairs": n, "ms_per_step": ms / 5, "gcups_device": cells / (ms / 5 * 1e-3) / 1e9, "launches_per_step": nl / 5,
                      "i32_reruns": st.get("i32_reruns")}))
    bt.free(); eng.close()
