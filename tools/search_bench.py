#!/usr/bin/env python3
"""BASELINE.json configs[2]: 5000-aa query vs a synthetic protein database (default 2e8 residues), score + top-100
traceback, database sharded over the ranks (one process per GPU, torchrun).  Strong scaling: the database is fixed,
every rank scores its shard, ONE all-gather merges the per-rank top-k.  Prints one JSON line on rank 0.

  python tools/search_bench.py [--residues 2e8] [--steps 3]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/search_bench.py
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

AA = np.frombuffer(b"ARNDCQEGHILKMFPSTWYV", np.uint8)


def make_db(lq, residues, seed=20243, planted=500):
    rng = np.random.default_rng(seed)
    q = AA[rng.integers(0, 20, lq)]
    lens = np.clip(np.rint(np.exp(rng.normal(5.6, 0.6, int(residues / 250)))), 30, 35000).astype(np.int64)
    lens = lens[np.cumsum(lens) <= residues]
    off = np.zeros(len(lens) + 1, np.int64); off[1:] = np.cumsum(lens)
    res = AA[rng.integers(0, 20, int(off[-1]))]
    for k in rng.choice(len(lens), min(planted, len(lens)), replace=False):   # mutated windows of the query
        w = int(min(lens[k], rng.integers(200, 2000), lq))
        s0 = int(rng.integers(0, lq - w + 1))
        m = q[s0:s0 + w].copy()
        sub = rng.random(w) < 0.3
        m[sub] = AA[rng.integers(0, 20, int(sub.sum()))]
        m = m[rng.random(w) >= 0.03]
        res[off[k]:off[k] + len(m)] = m
    return q.tobytes(), res, off


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--residues", type=float, default=2e8)
    ap.add_argument("--query", type=int, default=5000)
    ap.add_argument("--topk", type=int, default=100)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--db", default=None, help="packed database file (Database.save); written first by rank 0 if it does not exist")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    import torch
    import torch.distributed as dist
    import smithwaterman_b200 as swb
    from smithwaterman_b200 import search as S
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    q, res, off = make_db(args.query, int(args.residues))
    db = S.Database(res, off)
    if args.db:                                                # the database as a resident service would hold it: memory-mapped file
        if rank == 0 and not os.path.exists(args.db):
            db.save(args.db)
        if world > 1:
            dist.barrier()
        db = S.Database.load(args.db)
    eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, swb.ALIGN_LOCAL, device=local_rank)
    dev = torch.device("cuda", local_rank)
    ddb = S.DeviceDatabase(eng, db, rank, world)               # this rank's shard, resident on its GPU (swb_db_upload_shard)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one(traceback):
        return S.search(eng, q, db, topk=args.topk, traceback=traceback, rank=rank, world_size=world, device=dev, ddb=ddb)

    hits = one(True)                                           # warm-up (module load, scratch growth)
    t_score, t_full, dev_ms = [], [], []
    for _ in range(args.steps):
        barrier(); t0 = time.perf_counter()
        one(False)
        barrier(); t1 = time.perf_counter()
        dev_ms.append(eng.stats()["kernel_ms"])
        hits = one(True)
        barrier(); t2 = time.perf_counter()
        t_score.append(t1 - t0); t_full.append(t2 - t1)
    t = torch.tensor([min(t_score), min(t_full), min(dev_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    cells = float(args.query) * float(off[-1])
    if rank == 0:
        print(json.dumps({"case": "configs[2] database search", "n_gpus": world, "query": args.query, "db_residues": int(off[-1]),
                          "db_sequences": int(db.n), "cells": cells, "topk": args.topk,
                          "score_stage_s": t[0].item(), "gcups_score_stage": cells / t[0].item() / 1e9,
                          "score_plus_traceback_s": t[1].item(), "gcups_with_traceback": cells / t[1].item() / 1e9,
                          "device_ms_max_rank": t[2].item(), "scaling": "strong",
                          "top_hit": {"idx": hits[0]["idx"], "score": hits[0]["score"]},
                          "checksum_top": float(sum(h["score"] for h in hits)), "idx_top": [h["idx"] for h in hits[:10]]}), flush=True)
    if world > 1:
        dist.destroy_process_group()
    eng.close()


if __name__ == "__main__":
    main()
