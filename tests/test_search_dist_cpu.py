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


def _worker(rank, world, port, out, db_path=None):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    query, db = _make_db()
    if db_path:                               # every rank maps the same packed database file (Database.save / load)
        db = S.Database.load(db_path)
    hits = S.search(None, query, db, topk=25, traceback=False, rank=rank, world_size=world, score_fn=_oracle_score_fn)
    out[rank] = [(h["idx"], h["score"], h["end_q"], h["end_t"]) for h in hits]
    # the second exchange: every rank's traced alignments on every rank (packed strings, two small collectives)
    mine = {k: {"aln_a": "AC-GT" * (k + 1), "aln_b": "A-TGT" * (k + 1), "score": float(k),
                "idx_a": np.where(np.tile([1, 1, 0, 1, 1], k + 1) > 0, 10 * k + np.cumsum(np.tile([1, 1, 0, 1, 1], k + 1)) - 1, -1),
                "idx_b": np.where(np.tile([1, 0, 1, 1, 1], k + 1) > 0, 3 + np.cumsum(np.tile([1, 0, 1, 1, 1], k + 1)) - 1, -1)}
            for k in range(rank, 7, world)}
    if rank == 1:
        mine[9] = {"aln_a": "", "aln_b": "", "score": 0.0, "idx_a": np.zeros(0, np.int64), "idx_b": np.zeros(0, np.int64)}
    got = S.allgather_alignments(mine)
    out["aln%d" % rank] = {k: (v["aln_a"], v["aln_b"], v["score"], v["idx_a"].tolist(), v["idx_b"].tolist()) for k, v in got.items()}
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
    assert out["aln0"] == out["aln1"] and sorted(out["aln0"]) == [0, 1, 2, 3, 4, 5, 6, 9]
    a3 = out["aln0"][3]
    assert a3[0] == "AC-GT" * 4 and a3[2] == 3.0 and a3[3][:5] == [30, 31, -1, 32, 33] and a3[4][:5] == [3, -1, 4, 5, 6]
    # ties resolved by the lower database index; planted duplicates guarantee at least one tie
    scores = [r[1] for r in ref]
    assert any(scores[i] == scores[i + 1] and ref[i][0] < ref[i + 1][0] for i in range(len(ref) - 1))
    assert scores == sorted(scores, reverse=True)


def test_world2_gloo_from_packed_database_file(tmp_path):
    """The same world-2 run with both ranks reading the memory-mapped database file and its stored shard order."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    query, db = _make_db()
    path = str(tmp_path / "db.swdb")
    db.save(path)
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out, path), nprocs=2, join=True)
    single = S.search(None, query, db, topk=25, traceback=False, score_fn=_oracle_score_fn)
    ref = [(h["idx"], h["score"], h["end_q"], h["end_t"]) for h in single]
    assert out[0] == ref and out[1] == ref
    assert out["aln0"] == out["aln1"] and sorted(out["aln0"]) == [0, 1, 2, 3, 4, 5, 6, 9]
    a3 = out["aln0"][3]
    assert a3[0] == "AC-GT" * 4 and a3[2] == 3.0 and a3[3][:5] == [30, 31, -1, 32, 33] and a3[4][:5] == [3, -1, 4, 5, 6]


def test_packed_database_file_round_trip(tmp_path):
    """8f-n3: the packed on-disk database (Database.save / load): memory-mapped load returns the same residues, offsets,
    names and shard deal as the in-memory object; FASTA ingest follows the reference's header rules; bad files raise."""
    import pytest
    rng = np.random.default_rng(7)
    aa = np.frombuffer(b"ARNDCQEGHILKMFPSTWYV", np.uint8)
    seqs = [bytes(aa[rng.integers(0, 20, int(k))]) for k in rng.integers(1, 400, 300)] + [b""]
    names = ["sp|%05d|x y" % k for k in range(len(seqs))]
    db = S.Database.from_sequences(seqs, names)
    path = str(tmp_path / "db.swdb")
    db.save(path)
    for mm in (True, False):
        d2 = S.Database.load(path, mmap=mm)
        assert d2.n == db.n and np.array_equal(d2.offsets, db.offsets) and np.array_equal(np.asarray(d2.residues), db.residues)
        assert d2.names == names and d2.sequence(17) == seqs[17] and d2.sequence(db.n - 1) == b""
        for w in (1, 2, 8):
            a = S.shard_indices(db.lengths(), w); b = S.shard_indices(d2.lengths(), w, d2.order)
            assert all(np.array_equal(x, y) for x, y in zip(a, b))
        sub = d2.subset(S.shard_indices(d2.lengths(), 4, d2.order)[1])
        ref = db.subset(S.shard_indices(db.lengths(), 4)[1])
        assert np.array_equal(sub.residues, ref.residues) and np.array_equal(sub.offsets, ref.offsets)
    # FASTA ingest
    fa = tmp_path / "x.fas"
    fa.write_text(">a first\nACD\nEF\n>b\nGHI\n")
    d3 = S.Database.from_fasta(str(fa))
    assert d3.n == 2 and d3.sequence(0) == b"ACDEF" and d3.sequence(1) == b"GHI" and d3.names == ["a", "b"]
    # bad files
    bad = tmp_path / "bad.swdb"
    bad.write_bytes(b"NOTADB" + bytes(5000))
    with pytest.raises(ValueError):
        S.Database.load(str(bad))
    raw = bytearray(open(path, "rb").read())
    open(str(bad), "wb").write(raw[:6000])
    with pytest.raises(ValueError):
        S.Database.load(str(bad))


def test_native_shard_and_merge_host_functions():
    """swb_db_shard / swb_search_merge are host-only entry points of the C ABI: the deal equals the numpy restatement,
    the merge equals topk over the concatenation with (score desc, index asc)."""
    rng = np.random.default_rng(3)
    ln = rng.integers(0, 900, 777)
    ln[5] = ln[6] = ln[700] = 450                                  # equal lengths: the stable order decides
    off = np.zeros(len(ln) + 1, np.int64); off[1:] = np.cumsum(ln)
    for w in (1, 2, 5, 8):
        for r in range(w):
            assert np.array_equal(S.native_shard_indices(off, r, w), S.shard_indices(ln, w)[r])
    parts = []
    for r in range(3):
        h = np.zeros(20, S.HIT_DTYPE)
        h["score"] = rng.integers(10, 30, 20).astype(np.float32) / 2
        h["index"] = rng.choice(1000, 20, replace=False) * 3 + r
        h["end_q"] = h["index"] + 1; h["end_t"] = h["index"] + 2
        parts.append(h)
    got = S.merge_hits(parts, 25)
    allh = np.concatenate(parts)
    order = np.lexsort((allh["index"], -allh["score"].astype(np.float64)))[:25]
    assert got.tobytes() == allh[order].tobytes()
    assert len(S.merge_hits(parts, 100)) == 60 and len(S.merge_hits([], 5)) == 0
