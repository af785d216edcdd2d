#!/usr/bin/env python3
"""`sa_opencl -cluster` (OCL-main.rs:149-265) over the CUDA engine -- greedy length-sorted clustering of a
synthetic protein set (families of mutated copies).  Times the loop driven by the device-side counts
(swb_align_stats_batch) and, on a smaller set, by the alignment strings (swb_align_batch), and checks that both give
the same clusters.  Prints one JSON line.

  python tools/cluster_bench.py [--families 400] [--members 8] [--strings-families 60]
"""
import argparse
import io
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smithwaterman_b200 as swb                                   # noqa: E402
from smithwaterman_b200 import cli                                 # noqa: E402

AA = np.frombuffer(b"ARNDCQEGHILKMFPSTWYV", np.uint8)


def make_set(rng, families, members):
    seqs = []
    for f in range(families):
        base = AA[rng.integers(0, 20, int(rng.integers(100, 501)))]
        for v in range(members):
            s = base.copy()
            sub = rng.random(len(s)) < 0.04 * v
            s[sub] = AA[rng.integers(0, 20, int(sub.sum()))]
            s = s[rng.random(len(s)) >= 0.01 * v]
            seqs.append(swb.SeqData.create("f%d_%d" % (f, v), "", s.tobytes().decode()))
    return [seqs[k] for k in rng.permutation(len(seqs))]


def run(eng, seqs, use_stats):
    n_aln = [0]; cells = [0.0]

    def stats_many(rep, cands):
        n_aln[0] += len(cands); cells[0] += float(len(rep)) * sum(map(len, cands))
        st = eng.align_stats(rep, cands, False)
        return list(zip(st["a_letters"].tolist(), st["b_letters"].tolist(), st["matches"].tolist(), st["aln_len"].tolist()))

    def align_many(rep, cands):
        n_aln[0] += len(cands); cells[0] += float(len(rep.seq)) * sum(len(c.seq) for c in cands)
        return [(r["aln_a"], r["aln_b"]) for r in eng.align_list([rep] * len(cands), cands, False)]
    t0 = time.perf_counter()
    s, c, m = cli.cluster(seqs, align_many, 0.8, 0.8, 0.8, io.StringIO(), stats_many=stats_many if use_stats else None)
    dt = time.perf_counter() - t0
    return dt, n_aln[0], cells[0], [x.name for x in s], c


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--families", type=int, default=400)
    ap.add_argument("--members", type=int, default=8)
    ap.add_argument("--strings-families", type=int, default=60)
    args = ap.parse_args()
    rng = np.random.default_rng(99)
    eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, swb.ALIGN_LOCAL)
    small = make_set(rng, args.strings_families, args.members)
    run(eng, small[:200], True)                                    # warm-up
    t_s, n_s, c_s, names_s, cl_s = run(eng, small, False)
    t_c, n_c, c_c, names_c, cl_c = run(eng, small, True)
    same = names_s == names_c and cl_s == cl_c
    big = make_set(rng, args.families, args.members)
    t_b, n_b, c_b, names_b, cl_b = run(eng, big, True)
    t0 = time.perf_counter()
    s_n, cl_n, _ = cli.cluster_native(eng, big, 0.8, 0.8, 0.8, io.StringIO())
    t_n = time.perf_counter() - t0
    native_same = [x.name for x in s_n] == names_b and cl_n == cl_b
    st = eng.stats()
    print(json.dumps({"case": "8f-n1 -cluster (identity 0.8, coverage 0.8/0.8, local, BLOSUM62 10/0.5)",
                      "small_set": {"sequences": len(small), "alignments": n_s, "strings_s": t_s, "counts_s": t_c,
                                    "same_clusters": bool(same), "gcups_strings": c_s / t_s / 1e9, "gcups_counts": c_c / t_c / 1e9},
                      "large_set": {"sequences": len(big), "clusters": int(sum(1 for k, c in enumerate(cl_b) if c == k)),
                                    "alignments": n_b, "cells": c_b, "seconds": t_b, "gcups": c_b / t_b / 1e9,
                                    "alignments_per_s": n_b / t_b,
                                    "native_loop_seconds": t_n, "native_gcups": c_b / t_n / 1e9, "native_alignments_per_s": n_b / t_n,
                                    "native_same_clusters": bool(native_same), "native_kernel_launches": st["launches"]}}), flush=True)
    eng.close()


if __name__ == "__main__":
    main()
