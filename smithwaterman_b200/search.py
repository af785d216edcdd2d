"""Database-search driver (BASELINE.json configs[2]): one query against a sequence database, top-K hits with
traceback, database sharded over the GPUs of one box.

The reference has no such driver; its closest paths are the all-vs-all loops of the CLIs (SA-main.rs:35-45,
OCL-main.rs:295-316: every `sw.align(ss1, ss2, true)` of one query file against all records of the other) and the
`-cluster` candidate scan (OCL-main.rs:190-243).  Semantics per (query, subject) pair are exactly those of
`align_partial(.., score_only=true)` (SA.rs:438-467) for the scan and `align(.., retain_all)` (SA.rs:389) for the hits.

Sharding: subjects are dealt round-robin over the length-sorted order, so every rank sees the same
length mix and nearly the same number of DP cells; shards are independent, the only exchange is one all-gather of the
per-rank top-K records (K x 32 bytes).  Ties are broken by (score desc, database index asc), which makes the result
independent of the number of ranks.

The driver itself is native (csrc/search.cu behind include/swb200.h): `DeviceDatabase` is a shard resident on the GPU
(swb_db_upload_shard), `swb_search` scans, re-runs s16 overflows and selects the top-K on the device, `swb_search_merge`
merges the per-rank lists and `swb_search_align` traces the hits the shard owns.  This module is the thin caller that
adds the torch.distributed collectives; a C / Rust caller does the same with swb_search_sharded or its own gather.
"""
import ctypes
import struct
import os

import numpy as np

from . import _lib
from .engine import SequenceAlignment, SwbError, _ptr, pack_sequences


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


    # ---- packed on-disk form -------------------------------------------------------------------
    # One file, little endian, every section 4096-byte aligned so that it can be memory-mapped and handed to the
    # C ABI (which takes letters + int64 offsets) without a parse step:
    #   header (4096 B): magic "SWB200DB", u32 version, u32 flags, i64 n, i64 n_residues, i64 names_bytes,
    #                    i64 offset of {offsets, residues, order, names}
    #   offsets  int64[n+1]   residues  u8[n_residues] (letters as in the FASTA)   order  int64[n] (stable, longest first:
    #   what shard_indices deals round-robin, stored so that a rank does not sort 6*10^5 lengths per query)
    #   names    UTF-8, one per line
    MAGIC = b"SWB200DB"
    VERSION = 1

    def save(self, path):
        def pad(k):
            return (k + 4095) // 4096 * 4096
        names = ("\n".join(self.names) + "\n").encode() if self.names else b""
        order = np.argsort(-self.lengths(), kind="stable").astype(np.int64)
        o_off = 4096
        o_res = o_off + pad(self.offsets.nbytes)
        o_ord = o_res + pad(self.residues.nbytes)
        o_nam = o_ord + pad(order.nbytes)
        head = np.zeros(4096, np.uint8)
        head[:8] = np.frombuffer(self.MAGIC, np.uint8)
        head[8:16].view(np.uint32)[:] = (self.VERSION, 1 if self.names else 0)
        head[16:72].view(np.int64)[:] = (self.n, len(self.residues), len(names), o_off, o_res, o_ord, o_nam)
        with open(path, "wb") as f:
            f.write(head.tobytes())
            for off, blob in ((o_off, self.offsets.tobytes()), (o_res, self.residues.tobytes()), (o_ord, order.tobytes()), (o_nam, names)):
                f.seek(off)
                f.write(blob)
            f.truncate(o_nam + pad(len(names)))

    @staticmethod
    def load(path, mmap=True):
        """Memory-mapped (default) or in-memory Database from a file written by save()."""
        head = np.fromfile(path, np.uint8, 4096)
        if len(head) < 4096 or head[:8].tobytes() != Database.MAGIC:
            raise ValueError("%s: not a swb200 database file" % path)
        ver, flags = (int(v) for v in head[8:16].view(np.uint32))
        if ver != Database.VERSION:
            raise ValueError("%s: unsupported database version %d" % (path, ver))
        n, nres, nnames, o_off, o_res, o_ord, o_nam = (int(v) for v in head[16:72].view(np.int64))
        size = os.path.getsize(path)
        if n < 0 or nres < 0 or o_off + 8 * (n + 1) > size or o_res + nres > size or o_ord + 8 * n > size or o_nam + nnames > size:
            raise ValueError("%s: truncated database file" % path)
        rd = (lambda dt, off, cnt: np.memmap(path, dt, "r", off, (cnt,))) if mmap else \
             (lambda dt, off, cnt: np.fromfile(path, dt, cnt, offset=off))
        offsets = np.asarray(rd(np.int64, o_off, n + 1))
        if offsets[0] != 0 or offsets[-1] != nres or np.any(offsets[1:] < offsets[:-1]):
            raise ValueError("%s: corrupt offsets" % path)
        names = None
        if flags & 1:
            with open(path, "rb") as f:
                f.seek(o_nam)
                names = f.read(nnames).decode().split("\n")[:n]
        db = Database.__new__(Database)
        db.residues = rd(np.uint8, o_res, nres) if nres else np.zeros(0, np.uint8)
        db.offsets = np.ascontiguousarray(offsets)
        db.names = names
        db.n = n
        db.order = np.asarray(rd(np.int64, o_ord, n)) if n else np.zeros(0, np.int64)
        return db

    @staticmethod
    def from_fasta(path):
        """FASTA ingest with the reference's header rules (SeqData.load_fasta, SA.rs:830-887)."""
        from .engine import SeqData
        recs = SeqData.load_fasta(path)
        return Database.from_sequences([r.to_bytes() if hasattr(r, "to_bytes") else "".join(r.seq).encode() for r in recs],
                                       [r.name for r in recs])


def shard_indices(lengths, world_size, order=None):
    """Round-robin deal over the length-sorted order (stable, longest first): list of index arrays, one per rank.
    (The same deal as swb_db_shard in the library; tests/test_search_dist_cpu.py holds them equal.)"""
    if order is None:
        order = np.argsort(-np.asarray(lengths, np.int64), kind="stable")
    return [np.sort(order[r::world_size]) for r in range(world_size)]


def native_shard_indices(offsets, rank, world_size):
    """swb_db_shard: the index list of shard `rank` (host only, no device needed)."""
    off = np.ascontiguousarray(offsets, np.int64)
    n = len(off) - 1
    idx = np.empty(max(n, 1), np.int64)
    m = ctypes.c_int64(0)
    rc = _lib.lib().swb_db_shard(off.ctypes.data, n, int(rank), int(world_size), idx.ctypes.data, ctypes.byref(m))
    if rc:
        raise SwbError(rc, "swb_db_shard: bad argument")
    return idx[:m.value].copy()


HIT_DTYPE = np.dtype(_lib.HIT_DTYPE)


def merge_hits(parts, k):
    """swb_search_merge: global top-k of per-shard hit arrays (HIT_DTYPE); key (score desc, database index asc)."""
    allh = np.ascontiguousarray(np.concatenate(parts)) if len(parts) else np.zeros(0, HIT_DTYPE)
    out = np.zeros(max(int(k), 1), HIT_DTYPE)
    m = ctypes.c_int64(0)
    rc = _lib.lib().swb_search_merge(allh.ctypes.data, len(allh), int(k), out.ctypes.data, ctypes.byref(m))
    if rc:
        raise SwbError(rc, "swb_search_merge: bad argument")
    return out[:m.value].copy()


class DeviceDatabase:
    """A database shard resident on the engine's GPU (swb_db): encoded once, searched many times."""

    def __init__(self, engine, db, rank=0, world_size=1, global_index=None):
        self.engine, self.h = engine, ctypes.c_void_p()
        L = engine.L
        res = np.ascontiguousarray(db.residues)
        off = np.ascontiguousarray(db.offsets, np.int64)
        if global_index is not None:                          # `db` already is the shard; global_index ascending
            gi = np.ascontiguousarray(global_index, np.int64)
            engine._check(L.swb_db_upload(engine._h, _ptr(res), _ptr(off), db.n, _ptr(gi), ctypes.byref(self.h)))
        elif world_size > 1:
            engine._check(L.swb_db_upload_shard(engine._h, _ptr(res), _ptr(off), db.n, int(rank), int(world_size), ctypes.byref(self.h)))
        else:
            engine._check(L.swb_db_upload(engine._h, _ptr(res), _ptr(off), db.n, None, ctypes.byref(self.h)))
        self.n = L.swb_db_count(self.h)
        self.residues = L.swb_db_residues(self.h)

    def search(self, query, topk=100):
        """Local top-k of this shard -> numpy record array (HIT_DTYPE), sorted by (score desc, index asc)."""
        q = query if isinstance(query, (bytes, bytearray)) else query.encode("latin-1")
        hits = np.zeros(max(int(topk), 1), HIT_DTYPE)
        m = ctypes.c_int64(0)
        self.engine._check(self.engine.L.swb_search(self.engine._h, bytes(q), len(q), self.h, int(topk), hits.ctypes.data, ctypes.byref(m)))
        return hits[:m.value]

    def scores(self):
        """Full score vector of the last search, local subject order (tests / cross-checks)."""
        sc = np.empty(self.n, np.float32); eq = np.empty(self.n, np.int64); et = np.empty(self.n, np.int64)
        self.engine._check(self.engine.L.swb_db_scores(self.engine._h, self.h, _ptr(sc), _ptr(eq), _ptr(et)))
        return sc, eq, et

    def align(self, query, hits, retain_all=False):
        """Traceback of the hits this shard owns -> {position in `hits`: alignment dict}."""
        q = query if isinstance(query, (bytes, bytearray)) else query.encode("latin-1")
        hits = np.ascontiguousarray(hits, HIT_DTYPE)
        n = len(hits)
        outs = (_lib.Alignment * max(n, 1))()
        owned = np.zeros(max(n, 1), np.int32)
        import time
        t0 = time.perf_counter()
        self.engine._check(self.engine.L.swb_search_align(self.engine._h, bytes(q), len(q), self.h, hits.ctypes.data, n,
                                                          1 if retain_all else 0, outs, owned.ctypes.data))
        t1 = time.perf_counter()
        try:
            return {k: SequenceAlignment._unpack(outs[k]) for k in range(n) if owned[k]}
        finally:
            for k in range(n):
                if owned[k]:
                    self.engine.L.swb_alignment_free(ctypes.byref(outs[k]))
            self.last_align_s = (t1 - t0, time.perf_counter() - t1)      # (the C call, copying the results into Python objects)

    def free(self):
        if self.h is not None and self.h.value:
            self.engine.L.swb_db_free(self.engine._h, self.h)
            self.h = None

    def __del__(self):
        try:
            if self.engine._h is not None:
                self.free()
        except Exception:
            pass


def topk_records(scores, idx, end_q, end_t, k):
    """Top-k by (score desc, database index asc) -> structured array (host restatement of the device selection: the CPU
    tests of the multi-rank logic and the cross-checks of the GPU tests use it)."""
    scores = np.asarray(scores, np.float32)
    n = len(scores)
    if n > 4 * k > 0:                                         # linear-time selection first: everything scoring >= the k-th value
        kth = np.partition(scores, n - k)[n - k]
        sel = np.nonzero(scores >= kth)[0]
    else:
        sel = np.arange(n)
    rec = np.zeros(len(sel), HIT_DTYPE)
    rec["score"], rec["index"], rec["end_q"], rec["end_t"] = scores[sel], np.asarray(idx)[sel], np.asarray(end_q)[sel], np.asarray(end_t)[sel]
    order = np.lexsort((rec["index"], -rec["score"].astype(np.float64)))
    return rec[order[:k]]


def _allgather_blocks(block, device):
    """All-gather of one equally sized uint8 block per rank -> (world, len(block)) numpy array: one collective into one
    device tensor, one copy back."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    t = torch.from_numpy(block)
    if device is not None:
        t = t.to(device, non_blocking=True)
    out = torch.empty(world * t.numel(), dtype=torch.uint8, device=t.device)
    dist.all_gather_into_tensor(out, t)
    return out.cpu().numpy().reshape(world, -1)


def allgather_hits(local, k, device=None):
    """THE collective of the search: all-gather of the fixed-size per-rank hit blocks (k x 32 bytes, padded)."""
    buf = np.zeros(k, HIT_DTYPE)
    buf["index"] = -1
    buf[:len(local)] = local
    parts = []
    for row in _allgather_blocks(buf.view(np.uint8).copy(), device):
        a = row.view(HIT_DTYPE)
        parts.append(a[a["index"] >= 0])
    return parts


_ALN_HDR = struct.Struct("<qqdqq")          # position in the hit list, string length, score, first residue of a / of b


def _first_residue(idx):
    idx = np.asarray(idx)
    if not len(idx):
        return 0
    if idx[0] >= 0:
        return int(idx[0])
    nz = np.flatnonzero(idx >= 0)
    return int(idx[nz[0]]) if len(nz) else 0


def pack_alignments(mine):
    """{position: alignment dict} of core alignments (retain_all = False) -> one uint8 array: per hit a 40-byte header and
    the two strings.  The column index lists are not sent: they follow from the strings and the first residue index."""
    chunks = []
    for k, a in mine.items():
        chunks.append(_ALN_HDR.pack(k, len(a["aln_a"]), a["score"], _first_residue(a["idx_a"]), _first_residue(a["idx_b"])))
        chunks.append(a["aln_a"].encode("latin-1")); chunks.append(a["aln_b"].encode("latin-1"))
    return np.frombuffer(b"".join(chunks), np.uint8)


def column_indices(aln, start):
    """Residue index per alignment column (-1 = gap) of a core alignment string that starts at residue `start`."""
    g = np.frombuffer(aln.encode("latin-1"), np.uint8) != ord("-")
    return np.where(g, int(start) + np.cumsum(g) - 1, -1).astype(np.int64)


def unpack_alignments(buf, index_lists=True):
    out = {}
    p = 0
    raw = buf.tobytes()
    while p < len(raw):
        k, n, score, start_a, start_b = _ALN_HDR.unpack_from(raw, p); p += _ALN_HDR.size
        a = {"aln_a": raw[p:p + n].decode("latin-1"), "aln_b": raw[p + n:p + 2 * n].decode("latin-1"), "score": score,
             "start_a": start_a, "start_b": start_b}
        p += 2 * n
        if index_lists:
            a["idx_a"] = column_indices(a["aln_a"], start_a); a["idx_b"] = column_indices(a["aln_b"], start_b)
        out[k] = a
    return out


def allgather_alignments(mine, device=None, index_lists=True):
    """Every rank's traced hits on every rank: the sizes (8 bytes per rank), then the packed strings padded to the largest.
    index_lists = False leaves out the per-column index arrays of the received hits (column_indices rebuilds them)."""
    payload = pack_alignments(mine)
    sizes = _allgather_blocks(np.array([len(payload)], np.int64).view(np.uint8), device).view(np.int64).ravel()
    width = int(sizes.max())
    if width == 0:
        return {}
    block = np.zeros(width, np.uint8); block[:len(payload)] = payload
    rows = _allgather_blocks(block, device)
    out = {}
    for r, sz in zip(rows, sizes):
        out.update(unpack_alignments(r[:int(sz)], index_lists))
    return out


def search_scores(engine, query, db):
    """(scores f32, end_q, end_t) of the query against every subject of `db` (host buffers in and out, one pipelined call)."""
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
           score_fn=None, ddb=None, gather_alignments=True, timings=None, index_lists=True):
    """Top-k hits of `query` in `db`.

    ddb: the rank's resident shard (DeviceDatabase(engine, db, rank, world_size)); built here when absent -- callers that
    search more than once keep it.  world_size > 1: every rank scans its shard and selects on the device, ONE all-gather
    merges the per-rank lists (swb_search_merge), every rank traces the hits it owns and the alignments are gathered.
    score_fn: replaces the device path by a host scoring function over (engine, query, shard Database) -- the CPU
    multi-process tests of the rank logic.
    timings: a dict that receives the wall seconds of the stages (scan, gather_hits, align, gather_alignments).
    index_lists = False: hits traced by OTHER ranks come with aln_q / aln_t / start_q / start_t only (column_indices(aln, start)
    gives idx_q / idx_t on demand) -- rebuilding 2 x 100 index arrays in Python costs more than the collective.
    Returns a list of dicts: idx, score, end_q, end_t (+ aln_q, aln_t, idx_q, idx_t when traceback).
    """
    import time
    tick = [time.perf_counter()]

    def lap(name):
        if timings is not None:
            now = time.perf_counter(); timings[name] = now - tick[0]; tick[0] = now
    own_ddb = False
    if score_fn is not None:                                  # host restatement (tests): shard, score, select with numpy
        if world_size > 1:
            if shard is None:
                idx = shard_indices(db.lengths(), world_size, getattr(db, "order", None))[rank]
                shard = (idx, db.subset(idx))
            idx, sub = shard
        else:
            idx, sub = np.arange(db.n, dtype=np.int64), db
        sc, eq, et = score_fn(engine, query, sub)
        local = topk_records(sc, idx, eq, et, topk)
    else:
        own_ddb = ddb is None
        if own_ddb:
            ddb = DeviceDatabase(engine, db, rank, world_size)
        local = ddb.search(query, topk)
    lap("scan")
    best = merge_hits(allgather_hits(local, topk, device), topk) if world_size > 1 else local
    lap("gather_hits")
    hits = [{"idx": int(r["index"]), "score": float(r["score"]), "end_q": int(r["end_q"]), "end_t": int(r["end_t"])} for r in best]
    if traceback and hits and engine is not None:
        if score_fn is not None:
            mine = {k: a for k, a in enumerate(engine.align_list([query] * len(hits), [db.sequence(h["idx"]) for h in hits], retain_all=retain_all))}
        else:
            mine = ddb.align(query, best, retain_all)         # only the hits whose subject lives in this shard
        lap("align")
        if world_size > 1 and score_fn is None and gather_alignments:
            if not retain_all:                                # core alignments: strings only, two small collectives
                mine = allgather_alignments(mine, device, index_lists)
            else:
                import torch.distributed as dist
                parts = [None] * world_size
                dist.all_gather_object(parts, {k: (a["aln_a"], a["aln_b"], a["idx_a"], a["idx_b"], a["score"]) for k, a in mine.items()})
                mine = {k: {"aln_a": v[0], "aln_b": v[1], "idx_a": v[2], "idx_b": v[3], "score": v[4]} for p in parts for k, v in p.items()}
        lap("gather_alignments")
        for k, h in enumerate(hits):
            a = mine.get(k)
            if a is None:
                continue
            h.update(aln_q=a["aln_a"], aln_t=a["aln_b"])
            if "idx_a" in a:
                h.update(idx_q=a["idx_a"], idx_t=a["idx_b"])
            else:
                h.update(start_q=a["start_a"], start_t=a["start_b"])
            if a["score"] != h["score"]:
                raise SwbError(4, "scan and traceback disagree on subject %d" % h["idx"])
    if score_fn is None and own_ddb:
        ddb.free()
    return hits
