"""Database-search driver (BASELINE.json configs[2]): one query against a sequence database, top-K hits with
traceback, database sharded over the GPUs of one box.

The reference has no such driver; its closest paths are the all-vs-all loops of the CLIs (SA-main.rs:35-45,
OCL-main.rs:295-316: every `sw.align(ss1, ss2, true)` of one query file against all records of the other) and the
`-cluster` candidate scan (OCL-main.rs:190-243).  Semantics per (query, subject) pair are exactly those of
`align_partial(.., score_only=true)` (SA.rs:438-467) for the scan and `align(.., retain_all)` (SA.rs:389) for the hits.

Sharding: subjects are dealt round-robin over the length-sorted order, so every rank sees the same
length mix and nearly the same number of DP cells; shards are independent, the only exchange is one all-gather of the
per-rank top-K records (K x 24 bytes).  Ties are broken by (score desc, database index asc), which makes the result
independent of the number of ranks.
"""
import ctypes

import numpy as np

from .engine import SequenceAlignment, _ptr, pack_sequences


class Database:
    """Concatenated residues + offsets (+ optional names) of the subject sequences, host memory."""

    def __init__(self, residues, offsets, names=None):
        self.residues = np.ascontiguousarray(residues, np.uint8)
        self.offsets = np.ascontiguousarray(offsets, np.int64)
        self.names = names
        self.n = len(self.offsets) - 1

    @staticmethod
    def from_sequences(seqs, names=None):
        r, o = pack_sequences(seqs)
        return Database(r, o, names)

    def lengths(self):
        return self.offsets[1:] - self.offsets[:-1]

    def sequence(self, k):
        return self.residues[self.offsets[k]:self.offsets[k + 1]].tobytes()

    def subset(self, idx):
        """Contiguous copy of the subjects `idx` (a rank's shard)."""
        idx = np.asarray(idx, np.int64)
        ln = self.lengths()[idx]
        off = np.zeros(len(idx) + 1, np.int64)
        off[1:] = np.cumsum(ln)
        out = np.empty(int(off[-1]), np.uint8)
        src = self.offsets[idx]
        # vectorised gather: position of every output byte in the source
        if len(idx):
            rep = np.repeat(src - off[:-1], ln)
            out[:] = self.residues[np.arange(off[-1], dtype=np.int64) + rep]
        return Database(out, off, None)


def shard_indices(lengths, world_size):
    """Round-robin deal over the length-sorted order (stable, longest first): list of index arrays, one per rank."""
    order = np.argsort(-np.asarray(lengths, np.int64), kind="stable")
    return [np.sort(order[r::world_size]) for r in range(world_size)]


def topk_records(scores, idx, end_q, end_t, k):
    """Top-k by (score desc, database index asc) -> structured array."""
    scores = np.asarray(scores, np.float32)
    n = len(scores)
    if n > 4 * k > 0:                                         # linear-time selection first: everything scoring >= the k-th value
        kth = np.partition(scores, n - k)[n - k]
        sel = np.nonzero(scores >= kth)[0]
    else:
        sel = np.arange(n)
    rec = np.empty(len(sel), dtype=[("score", np.float32), ("idx", np.int64), ("end_q", np.int64), ("end_t", np.int64)])
    rec["score"], rec["idx"], rec["end_q"], rec["end_t"] = scores[sel], np.asarray(idx)[sel], np.asarray(end_q)[sel], np.asarray(end_t)[sel]
    order = np.lexsort((rec["idx"], -rec["score"].astype(np.float64)))
    return rec[order[:k]]


def merge_topk(parts, k):
    allr = np.concatenate(parts) if parts else np.empty(0, dtype=[("score", np.float32), ("idx", np.int64), ("end_q", np.int64), ("end_t", np.int64)])
    order = np.lexsort((allr["idx"], -allr["score"].astype(np.float64)))
    return allr[order[:k]]


def allgather_topk(local, k, device=None):
    """One collective: all-gather of the fixed-size per-rank top-k blocks (padded with score = -inf)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    buf = np.full((k, 4), -np.inf, np.float64)
    buf[:len(local), 0] = local["score"]
    buf[:len(local), 1] = local["idx"]
    buf[:len(local), 2] = local["end_q"]
    buf[:len(local), 3] = local["end_t"]
    t = torch.from_numpy(buf)
    if device is not None:
        t = t.to(device)
    outs = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(outs, t)
    parts = []
    for o in outs:
        a = o.cpu().numpy()
        a = a[np.isfinite(a[:, 0])]
        rec = np.empty(len(a), dtype=local.dtype)
        rec["score"], rec["idx"], rec["end_q"], rec["end_t"] = a[:, 0], a[:, 1].astype(np.int64), a[:, 2].astype(np.int64), a[:, 3].astype(np.int64)
        parts.append(rec)
    return parts


def search_scores(engine, query, db):
    """(scores f32, end_q, end_t) of the query against every subject of `db` (one pipelined device call)."""
    q = query if isinstance(query, (bytes, bytearray)) else query.encode("latin-1")
    n = db.n
    sc = np.empty(n, np.float32)
    eq = np.empty(n, np.int64)
    et = np.empty(n, np.int64)
    qb = ctypes.create_string_buffer(bytes(q), len(q))
    engine._check(engine.L.swb_search_scores(engine._h, ctypes.addressof(qb), len(q), _ptr(db.residues), _ptr(db.offsets), n,
                                             _ptr(sc), _ptr(eq), _ptr(et)))
    return sc, eq, et


def search(engine, query, db, topk=100, traceback=True, retain_all=False, rank=0, world_size=1, shard=None, device=None,
           score_fn=None):
    """Top-k hits of `query` in `db`.

    world_size > 1: every rank calls this with the full `db` (or its precomputed `shard` = (index array, Database));
    each scores its shard, one all-gather merges the per-rank top-k, every rank returns the same list.
    score_fn: replaces the device scoring call (used by the CPU multi-process tests of the host logic).
    Returns a list of dicts: idx, score, end_q, end_t (+ aln_q, aln_t, idx_q, idx_t when traceback).
    """
    if world_size > 1:
        if shard is None:
            idx = shard_indices(db.lengths(), world_size)[rank]
            shard = (idx, db.subset(idx))
        idx, sub = shard
    else:
        idx, sub = np.arange(db.n, dtype=np.int64), db
    sc, eq, et = (score_fn or search_scores)(engine, query, sub)
    local = topk_records(sc, idx, eq, et, topk)
    if world_size > 1:
        best = merge_topk(allgather_topk(local, topk, device), topk)
    else:
        best = local
    hits = [{"idx": int(r["idx"]), "score": float(r["score"]), "end_q": int(r["end_q"]), "end_t": int(r["end_t"])} for r in best]
    if traceback and hits and engine is not None:
        alns = engine.align_list([query] * len(hits), [db.sequence(h["idx"]) for h in hits], retain_all=retain_all)
        for h, a in zip(hits, alns):
            h.update(aln_q=a["aln_a"], aln_t=a["aln_b"], idx_q=a["idx_a"], idx_t=a["idx_b"])
            assert a["score"] == h["score"], "scan and traceback disagree"
    return hits
