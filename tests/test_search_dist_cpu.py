"""Host logic of the sharded database search on CPU: world_size-2 gloo run (sharding, per-rank top-k, the single
all-gather, merge with deterministic tie-breaking).  Scoring is delegated to the oracle here (score_fn), so no GPU."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from smithwaterman_b200 import search as S
from oracle import oracle as O

AA = np.frombuffer(b"ARNDCQEGHILKMFPSTWYV", np.uint8)


def _make_db(seed=3, n=240):
    rng = np.random.default_rng(seed)
    query = bytes(AA[rng.integers(0, 20, 120)])
    seqs = []
    for k in range(n):
        if k % 17 == 0:                      # planted homologs (and exact duplicates -> score ties)
            w = bytearray(query[20:100])
            for p in rng.integers(0, len(w), 12):
                w[p] = AA[rng.integers(0, 20)]
            seqs.append(bytes(w))
        elif k % 17 == 1:
            seqs.append(seqs[-1])
        else:
            seqs.append(bytes(AA[rng.integers(0, 20, int(rng.integers(30, 200)))]))
    return query, S.Database.from_sequences(seqs)


def _oracle_score_fn(engine, query, db):
    m = O.Matrix.blosum62()
    qi = m.seq_to_index(query.decode())
    subs = [m.seq_to_index(db.sequence(k).decode()) for k in range(db.n)]
    a, oa = O.pack([qi] * db.n)
    b, ob = O.pack(subs)
    sc, ei, ej = O.score_batch_int(a, oa, b, ob, np.rint(m.scores * 2).astype(np.int32), -20, -1, 2, O.ALIGN_LOCAL, 2)
    return (sc / 2.0).astype(np.float32), ei, ej


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    query, db = _make_db()
    hits = S.search(None, query, db, topk=25, traceback=False, rank=rank, world_size=world, score_fn=_oracle_score_fn)
    out[rank] = [(h["idx"], h["score"], h["end_q"], h["end_t"]) for h in hits]
    dist.destroy_process_group()


def test_sharding_is_balanced_and_complete():
    rng = np.random.default_rng(0)
    ln = rng.integers(30, 3000, 1001)
    for w in (1, 2, 4, 8):
        parts = S.shard_indices(ln, w)
        assert sorted(np.concatenate(parts).tolist()) == list(range(1001))
        tot = [int(ln[p].sum()) for p in parts]
        assert max(tot) - min(tot) <= 3000 * 2           # residues balanced to within a couple of sequences


def test_world2_gloo_matches_single_process():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    query, db = _make_db()
    single = S.search(None, query, db, topk=25, traceback=False, score_fn=_oracle_score_fn)
    ref = [(h["idx"], h["score"], h["end_q"], h["end_t"]) for h in single]
    assert out[0] == ref and out[1] == ref
    # ties resolved by the lower database index; planted duplicates guarantee at least one tie
    scores = [r[1] for r in ref]
    assert any(scores[i] == scores[i + 1] and ref[i][0] < ref[i + 1][0] for i in range(len(ref) - 1))
    assert scores == sorted(scores, reverse=True)
