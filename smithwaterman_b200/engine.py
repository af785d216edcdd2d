"""Host-side mirror of the reference's alignment API over the C ABI (include/swb200.h).

Names and argument meaning follow rust/sequence_alignment/src/sequence_alignment.rs ("SA.rs"):

  SubstitutionMatrix.get_blosum62_matrix / get_mat_matrix / load_from_lines   SA.rs:681-794
  SeqData.create / load_fasta                                                  SA.rs:799-887
  SequenceAlignment(matrix, go, ge, alignment_type).align(s1, s2, retain_all)  SA.rs:26-45, 389-392
  SequenceAlignment.align_partial(.., partial_region, score_only)              SA.rs:393-467
  SequenceAlignment.align_list(pairs)    the `sa_opencl -list` loop            OCL-main.rs:295-316
  SequenceAlignment.score_batch(...)     score_only over a batch (one device call)

All DP work happens in the CUDA library; there is no Python or CPU compute fallback.  Errors that
are panics in the reference surface as SwbError.
"""
import ctypes

import numpy as np

from . import _lib

ALIGN_GLOBAL, ALIGN_GLOCAL, ALIGN_LOCAL = 0, 1, 2            # SA.rs:11-13
TYPE_NAMES = {ALIGN_GLOBAL: "Global alignment", ALIGN_GLOCAL: "Glocal alignment", ALIGN_LOCAL: "Local alignment"}


class SwbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("swb200 error %d: %s" % (code, msg))
        self.code = code


class SubstitutionMatrix:
    """letters (one byte each) + f32 scores[a][b]; SA.rs:626-631."""

    def __init__(self, letters, scores):
        self.index_to_string = list(letters)
        if any(len(s.encode("latin-1", "replace")) != 1 for s in self.index_to_string):
            raise ValueError("every matrix letter must be a single byte")
        self.string_to_index = {s: i for i, s in enumerate(self.index_to_string)}
        self.scores = np.ascontiguousarray(scores, dtype=np.float32).reshape(len(letters), len(letters))

    @staticmethod
    def get_blosum62_matrix():                                # SA.rs:697-733 (table lives in the library)
        letters = ctypes.create_string_buffer(24)
        sc = np.empty(24 * 24, np.float32)
        n = _lib.lib().swb_matrix_blosum62(ctypes.addressof(letters), sc.ctypes.data)
        return SubstitutionMatrix(letters.raw[:n].decode("latin-1"), sc)

    @staticmethod
    def get_mat_matrix(matchscore, mismatchscore):            # SA.rs:681-695
        letters = ctypes.create_string_buffer(26)
        sc = np.empty(26 * 26, np.float32)
        n = _lib.lib().swb_matrix_match_mismatch(matchscore, mismatchscore, ctypes.addressof(letters), sc.ctypes.data)
        return SubstitutionMatrix(letters.raw[:n].decode("latin-1"), sc)

    @staticmethod
    def load_from_lines(lines):                               # SA.rs:735-794
        cols, table, count = [], {}, -1
        for line in lines:
            tok = line.strip().split()
            if tok[0][0] == "#":
                continue
            count += 1
            if count == 0:
                for t in tok:
                    if t in cols:
                        raise ValueError("%s was already found." % t)
                    cols.append(t)
                continue
            if tok[0] not in cols:
                raise ValueError("%s was not found in the row name." % tok[0])
            for k, t in enumerate(tok[1:]):
                try:
                    v = float(t)
                except ValueError:
                    v = 0.0                                   # "can not be parsed! zero was assigned", SA.rs:769-772
                table[(tok[0][0], cols[k])] = v
        n = len(cols)
        sc = np.zeros((n, n), np.float32)
        for i, a in enumerate(cols):
            for j, b in enumerate(cols):
                if (a, b) not in table:
                    raise ValueError("score about %s %s is not defined." % (a, b))
                sc[i, j] = table[(a, b)]
        return SubstitutionMatrix(cols, sc)

    def get_score_str(self, a, b):                            # SA.rs:633-645
        if a not in self.string_to_index:
            raise KeyError("%s was not found in scoring matrix!" % a)
        if b not in self.string_to_index:
            raise KeyError("%s was not found in scoring matrix!" % b)
        return float(self.scores[self.string_to_index[a], self.string_to_index[b]])

    def get_score(self, a, b):                                # SA.rs:646-648
        return float(self.scores[a, b])


class PositionSpecificMatrix:
    """PositionSpecificMatrix of SA.rs:583-623: scores[a + b*a_length] for position a of seq1 and b of seq2."""

    def __init__(self):
        self.scores, self.a_length, self.b_length = np.zeros(0, np.float32), 0, 0

    def prepare(self, a, b):                                  # SA.rs:604-610
        self.a_length, self.b_length = len(_as_seq_bytes(a)), len(_as_seq_bytes(b))
        if self.a_length * self.b_length > len(self.scores):
            self.scores = np.zeros(self.a_length * self.b_length, np.float32)

    def get_score(self, a, b):                                # SA.rs:593-595
        return float(self.scores[a + b * self.a_length])

    def set_score(self, a, b, s):                             # SA.rs:601-603
        self.scores[a + b * self.a_length] = s

    def get_score_str(self, a, b):                            # SA.rs:590-592
        raise NotImplementedError("not implemented")


class SeqData:
    """name, desc, seq (list of one-letter strings); SA.rs:799-803."""

    def __init__(self, name="", desc="", seq=None):
        self.name, self.desc, self.seq = name, desc, list(seq or [])

    @staticmethod
    def create(n, d, s):                                      # SA.rs:808-810 (line_to_seq(s, true))
        return SeqData(n, d, [c for c in s if c not in "\r\n"])

    @staticmethod
    def line_to_seq(s, retainws):                             # SA.rs:811-821
        return [c for c in s if c not in "\r\n"] if retainws else [c for c in s if not c.isspace()]

    @staticmethod
    def load_fasta(filename, retainws=False):                 # SA.rs:830-887
        ret, buf, name, desc = [], [], "", ""
        with open(filename, "r") as f:
            for line in f.read().splitlines():
                if ">" in line:                               # any line holding '>' is a header (SA.rs:840-851)
                    if buf or name:
                        ret.append(SeqData(name, desc, buf))
                        buf = []
                    head = line.strip()
                    if head.startswith(">"):
                        head = head[1:]
                    parts = head.lstrip().split(None, 1) if head.strip() else []
                    name = parts[0] if parts else ""
                    desc = ""
                    if len(parts) > 1:
                        # everything after the first whitespace char that follows the name (SA.rs:861-870)
                        rest = head.lstrip()[len(name):]
                        desc = rest[1:] if rest else ""
                else:
                    buf.extend(SeqData.line_to_seq(line, retainws))
        if name or buf:
            ret.append(SeqData(name, desc, buf))
        return ret

    def as_bytes(self):
        return "".join(self.seq).encode("latin-1", "replace")


def _as_seq_bytes(s):
    if isinstance(s, SeqData):
        return s.as_bytes()
    if isinstance(s, str):
        return s.encode("latin-1", "replace")
    return bytes(s)


def pack_sequences(seqs):
    """list of str/bytes/SeqData -> (uint8 array, int64 offsets[n+1])."""
    bs = [s if type(s) is bytes else _as_seq_bytes(s) for s in seqs]
    off = np.zeros(len(bs) + 1, np.int64)
    if bs:
        off[1:] = np.cumsum([len(b) for b in bs])
    return np.frombuffer(b"".join(bs), dtype=np.uint8).copy() if bs else np.zeros(0, np.uint8), off


class SequenceAlignment:
    """SequenceAlignment of SA.rs:15-552 / OpenCLSequenceAlignment of OCL.rs:71-918 over the CUDA engine."""

    def __init__(self, ss, go, ge, alignment_type, device=0):
        self._h = None
        self.L = _lib.lib()
        self.scoring_matrix, self.alignment_type = ss, int(alignment_type)
        self._pssm = isinstance(ss, PositionSpecificMatrix)
        if self._pssm:                                        # the engine only carries penalties and mode; a 1-letter dummy matrix
            letters, scores = b"X", np.zeros(1, np.float32)
        else:
            letters, scores = "".join(ss.index_to_string).encode("latin-1"), ss.scores
        self._letters = ctypes.create_string_buffer(letters, len(letters))
        self._scores = np.ascontiguousarray(scores, np.float32)
        m = _lib.Matrix(len(letters), ctypes.cast(self._letters, ctypes.c_char_p),
                        self._scores.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
        h = ctypes.c_void_p()
        rc = self.L.swb_engine_create(ctypes.byref(m), go, ge, self.alignment_type, device, ctypes.byref(h))
        if rc:
            raise SwbError(rc, (self.L.swb_last_error(None) or b"").decode())
        self._h = h

    def close(self):
        if self._h is not None:
            self.L.swb_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc:
            raise SwbError(rc, (self.L.swb_last_error(self._h) or b"").decode())

    def stats(self):
        st = _lib.Stats()
        self._check(self.L.swb_get_stats(self._h, ctypes.byref(st)))
        return {k: getattr(st, k) for k, _ in st._fields_}

    def prepare(self, s1, s2):                                # SA.rs:47-50: sizes the position-specific matrix
        if self._pssm:
            self.scoring_matrix.prepare(s1, s2)

    # -- single pair, SA.rs:389-551 -------------------------------------------------------------
    def align(self, s1, s2, retain_all=True):
        return self.align_partial(s1, s2, retain_all, None, False)

    def align_partial(self, s1, s2, retain_all, partial_region=None, score_only=False):
        a, b = _as_seq_bytes(s1), _as_seq_bytes(s2)
        if self._pssm or (partial_region is None and not score_only):
            if partial_region is not None:
                a, b = a[:partial_region[0]], b[:partial_region[1]]
            res = self.align_raw(a, b, retain_all)
            return ([], [], res["score"]) if score_only else (list(res["aln_a"]), list(res["aln_b"]), res["score"])
        # the DP runs on the prefixes (SA.rs:396-405), the padding spells the whole sequences (SA.rs:512-549)
        pa, pb = (len(a), len(b)) if partial_region is None else (int(partial_region[0]), int(partial_region[1]))
        out = _lib.Alignment()
        self._check(self.L.swb_align_partial(self._h, a, len(a), b, len(b), pa, pb, int(bool(retain_all)), int(bool(score_only)),
                                             ctypes.byref(out)))
        try:
            res = self._unpack(out)
        finally:
            self.L.swb_alignment_free(ctypes.byref(out))
        if score_only:                                        # SA.rs:438-467
            return [], [], res["score"]
        return list(res["aln_a"]), list(res["aln_b"]), res["score"]

    def align_raw(self, a, b, retain_all=True):
        a, b = _as_seq_bytes(a), _as_seq_bytes(b)
        out = _lib.Alignment()
        if self._pssm:
            pm = self.scoring_matrix
            if (pm.a_length, pm.b_length) != (len(a), len(b)):
                raise SwbError(1, "PositionSpecificMatrix was prepared for other lengths (SA.rs:604-610)")
            sc = np.ascontiguousarray(pm.scores[:len(a) * len(b)], np.float32)
            self._check(self.L.swb_align_pssm(self._h, sc.ctypes.data, a, len(a), b, len(b), int(bool(retain_all)), ctypes.byref(out)))
        else:
            self._check(self.L.swb_align(self._h, a, len(a), b, len(b), int(bool(retain_all)), ctypes.byref(out)))
        try:
            return self._unpack(out)
        finally:
            self.L.swb_alignment_free(ctypes.byref(out))

    @staticmethod
    def _unpack(out):
        n = out.n_cols
        # (one string_at per list: np.ctypeslib.as_array builds a ctypes array type per call, ~10x slower for short lists)
        return {"score": float(out.score), "end_a": out.end_a, "end_b": out.end_b,
                "idx_a": np.frombuffer(ctypes.string_at(out.idx_a, 8 * n), np.int64) if n else np.zeros(0, np.int64),
                "idx_b": np.frombuffer(ctypes.string_at(out.idx_b, 8 * n), np.int64) if n else np.zeros(0, np.int64),
                "aln_a": ctypes.string_at(out.aln_a, out.aln_len).decode("latin-1"),
                "aln_b": ctypes.string_at(out.aln_b, out.aln_len).decode("latin-1")}

    # -- batch-list path, OCL-main.rs:295-316 ------------------------------------------------------
    def align_list(self, seqs_a, seqs_b, retain_all=True):
        """n pairs with traceback in one device pass -> list of dicts (see align_raw)."""
        a, oa = pack_sequences(seqs_a)
        b, ob = pack_sequences(seqs_b)
        n = len(oa) - 1
        outs = (_lib.Alignment * max(n, 1))()
        self._check(self.L.swb_align_batch(self._h, a.ctypes.data, oa.ctypes.data, b.ctypes.data, ob.ctypes.data, n,
                                           1 if retain_all else 0, outs))
        try:
            return [self._unpack(outs[k]) for k in range(n)]
        finally:
            for k in range(n):
                self.L.swb_alignment_free(ctypes.byref(outs[k]))

    def align_stats(self, seqs_a, seqs_b, retain_all=False):
        """align(a, b, retain_all) of n pairs (seqs_a may be ONE sequence to be aligned with every b) reduced on the device to the counts the clustering loop reads off the
        strings (OCL-main.rs:205-233) -> numpy record array (score, status, end_a, end_b, aln_len, a_letters,
        b_letters, matches); no strings are built and 48 bytes per pair come back."""
        b, ob = pack_sequences(seqs_b)
        if isinstance(seqs_a, (bytes, str, SeqData)):         # one sequence against all of seqs_b
            one = np.frombuffer(_as_seq_bytes(seqs_a), np.uint8)
            a, oa = np.tile(one, len(ob) - 1), np.arange(len(ob), dtype=np.int64) * len(one)
        else:
            a, oa = pack_sequences(seqs_a)
        n = len(oa) - 1
        out = np.zeros(max(n, 1), dtype=_lib.ALN_STATS_DTYPE)
        self._check(self.L.swb_align_stats_batch(self._h, a.ctypes.data, oa.ctypes.data, b.ctypes.data, ob.ctypes.data, n,
                                                 1 if retain_all else 0, out.ctypes.data))
        return out[:n]

    def cluster(self, seqs, identity=0.8, coverage_short=0.8, coverage_long=0.8):
        """The clustering loop of `sa_opencl -cluster` (OCL-main.rs:149-243) with the sequences resident on the device
        -> (order, cluster_of, identical, n_alignments), the arrays indexed by sorted position (see swb_cluster)."""
        s, off = pack_sequences(seqs)
        n = len(off) - 1
        order = np.zeros(max(n, 1), np.int64); cl = np.zeros(max(n, 1), np.int64); ident = np.zeros(max(n, 1), np.int64)
        nal = ctypes.c_int64(0)
        self._check(self.L.swb_cluster(self._h, s.ctypes.data, off.ctypes.data, n, float(identity), float(coverage_short),
                                       float(coverage_long), order.ctypes.data, cl.ctypes.data, ident.ctypes.data, ctypes.byref(nal)))
        return order[:n], cl[:n], ident[:n], nal.value

    def score_batch(self, seqs_a, seqs_b):
        a, oa = pack_sequences(seqs_a)
        b, ob = pack_sequences(seqs_b)
        return self.score_packed(a, oa, b, ob)

    def score_packed(self, a, oa, b, ob, out=None):
        """Host buffers in, host buffers out (H2D + kernels + D2H inside the call).  `out` = (scores f32[n], end_a
        i64[n], end_b i64[n]) to fill, e.g. pinned arrays: the device then writes the results straight into them."""
        n = len(oa) - 1
        if out is not None:
            sc, ea, eb = out
        else:
            sc = np.empty(n, np.float32)
            ea = np.empty(n, np.int64)
            eb = np.empty(n, np.int64)
        self._check(self.L.swb_score_batch(self._h, _ptr(a), _ptr(oa), _ptr(b), _ptr(ob), n, _ptr(sc), _ptr(ea), _ptr(eb)))
        return sc, ea, eb

    # -- staged batch (inputs resident in HBM between runs) ---------------------------------------
    def upload(self, a, oa, b, ob):
        h = ctypes.c_void_p()
        self._check(self.L.swb_batch_upload(self._h, _ptr(a), _ptr(oa), _ptr(b), _ptr(ob), len(oa) - 1, ctypes.byref(h)))
        return Batch(self, h, len(oa) - 1)


def _ptr(x):
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):                                # torch tensor (pinned host memory in bench.py)
        return x.data_ptr()
    return x.ctypes.data


class Batch:
    def __init__(self, eng, h, n):
        self.eng, self.h, self.n = eng, h, n

    def run(self):
        self.eng._check(self.eng.L.swb_batch_run(self.eng._h, self.h))

    def last_run_ms(self):
        ms, nl = ctypes.c_float(), ctypes.c_int64()
        self.eng._check(self.eng.L.swb_batch_last_run_ms(self.eng._h, self.h, ctypes.byref(ms), ctypes.byref(nl)))
        return ms.value, nl.value

    def run_n(self, steps):
        """`steps` back-to-back runs timed with CUDA events on the engine's stream -> (total ms, kernels launched)."""
        ms, nl = ctypes.c_float(), ctypes.c_int64()
        self.eng._check(self.eng.L.swb_batch_run_n(self.eng._h, self.h, int(steps), ctypes.byref(ms), ctypes.byref(nl)))
        return ms.value, nl.value

    def fetch(self):
        sc = np.empty(self.n, np.float32)
        ea = np.empty(self.n, np.int64)
        eb = np.empty(self.n, np.int64)
        self.eng._check(self.eng.L.swb_batch_fetch(self.eng._h, self.h, _ptr(sc), _ptr(ea), _ptr(eb)))
        return sc, ea, eb

    def free(self):
        if self.h is not None:
            self.eng.L.swb_batch_free(self.eng._h, self.h)
            self.h = None


def rust_f32_display(x):
    """Rust `{}` formatting of an f32 score (SA-main.rs:40): shortest round-trip, no trailing .0."""
    v = np.float32(x)
    if np.isfinite(v) and v == np.floor(v) and abs(float(v)) < 1e16:
        return "%d" % int(v)
    return np.format_float_positional(v, unique=True, trim="-")


def cli_record(name1, name2, r1, r2, score, alignment_type):
    """One stdout record of `seqalign` / `sa_opencl` (SA-main.rs:40-43)."""
    return "#score:%s\n#type:%s\n>%s\n%s\n\n>%s\n%s\n\n" % (
        rust_f32_display(score), TYPE_NAMES[alignment_type], name1, "".join(r1), name2, "".join(r2))


def dpx_peak(which, device=0):
    """K0 micro-benchmark: thread-instructions/s of instruction mix `which` (see include/swb200.h)."""
    v = ctypes.c_double()
    rc = _lib.lib().swb_dpx_peak(device, which, ctypes.byref(v))
    if rc:
        raise SwbError(rc, "swb_dpx_peak failed")
    return v.value
