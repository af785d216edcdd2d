#!/usr/bin/env python3
"""Aligns the pairs of a reference-style file list on the GPU and writes one EMBOSS srspair report per pair
(res<k>.dat / needle_res<k>.dat / needle_glocal_res<k>.dat), diffable against test/emboss_results of the reference.

  python tools/emboss_report.py -local|-global|-glocal -list file_list.txt -outdir DIR [-go 10 -ge 0.5]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smithwaterman_b200 as swb  # noqa: E402
from smithwaterman_b200 import report  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    g = ap.add_mutually_exclusive_group(required=True)
    g.add_argument("-local", action="store_true"); g.add_argument("-global", dest="glob", action="store_true")
    g.add_argument("-glocal", action="store_true")
    ap.add_argument("-list", required=True); ap.add_argument("-outdir", required=True)
    ap.add_argument("-go", type=float, default=10.0); ap.add_argument("-ge", type=float, default=0.5)
    a = ap.parse_args()
    mode, prog, pre = (swb.ALIGN_LOCAL, "water", "res") if a.local else \
        (swb.ALIGN_GLOBAL, "needle", "needle_res") if a.glob else (swb.ALIGN_GLOCAL, "needle", "needle_glocal_res")
    base = os.path.dirname(os.path.abspath(a.list))
    pairs = [ln.split() for ln in open(a.list) if len(ln.split()) >= 2]
    sa = [swb.SeqData.load_fasta(os.path.join(base, p[0]))[0] for p in pairs]
    sb = [swb.SeqData.load_fasta(os.path.join(base, p[1]))[0] for p in pairs]
    m = swb.SubstitutionMatrix.get_blosum62_matrix()
    eng = swb.SequenceAlignment(m, a.go, a.ge, mode)
    res = eng.align_list(["".join(s.seq).encode() for s in sa], ["".join(s.seq).encode() for s in sb], mode != swb.ALIGN_LOCAL)
    os.makedirs(a.outdir, exist_ok=True)
    for k, (p, r) in enumerate(zip(pairs, res)):
        s1, s2 = report.starts_of(r["idx_a"], r["idx_b"]) if mode == swb.ALIGN_LOCAL else (1, 1)
        out = os.path.join(a.outdir, "%s%d.dat" % (pre, k))
        with open(out, "w") as f:
            f.write(report.srspair_report(prog, p[0], p[1], out, aln_a=r["aln_a"], aln_b=r["aln_b"], score=r["score"], matrix=m,
                                          gap_open=a.go, gap_extend=a.ge, start_a=s1, start_b=s2))
    eng.close()
    print("wrote %d reports to %s" % (len(pairs), a.outdir))


if __name__ == "__main__":
    main()
