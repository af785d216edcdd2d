import sys, os, gzip, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import smithwaterman_b200 as swb
from oracle import oracle as O
em = json.loads(gzip.open("tests/golden/emboss.json.gz").read().decode())
m = swb.SubstitutionMatrix.get_blosum62_matrix()
eng = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
sa = [r["seq1"] for r in em]; sb = [r["seq2"] for r in em]
sc, ea, eb = eng.score_batch(sa, sb)
bad = 0
for k, r in enumerate(em):
    exp = r["local"]["score"]
    if sc[k] != exp:
        bad += 1
        print("pair", k, "la", len(sa[k]), "lb", len(sb[k]), "got", sc[k], ea[k], eb[k], "exp", exp)
print("bad", bad, "reruns", eng.stats()["i32_reruns"])
# single-pair runs of the bad ones
for k, r in enumerate(em):
    if sc[k] != r["local"]["score"]:
        s1, a1, b1 = eng.score_batch([sa[k]], [sb[k]])
        print(" alone:", k, s1[0], a1[0], b1[0])
