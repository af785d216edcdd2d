"""CPU ORACLE -- Python side (test infrastructure, NOT product code).

ctypes binding of oracle/liboracle.so plus a restatement of the small host-side
pieces of the reference around the DP (all `SA.rs` = rust/sequence_alignment/src/
sequence_alignment.rs of the reference):

  * matrices          SA.rs:681-733 (get_mat_matrix, get_blosum62_matrix), 735-794 (load_from_lines)
  * seq_to_index      SA.rs:649-679 (unknown letter -> index of "X", else error)
  * FASTA loader      SA.rs:811-887
  * string inflation  SA.rs:470-550 (retain_all padding rules)
  * CLI record format SA-main.rs:40-43 (Rust `{}` formatting of the f32 score)

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference)
may import this module.  Parity: pinned by tests/test_oracle_golden.py.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

ALIGN_GLOBAL, ALIGN_GLOCAL, ALIGN_LOCAL = 0, 1, 2   # SA.rs:11-13
MODE_NAMES = {ALIGN_GLOBAL: "Global alignment", ALIGN_GLOCAL: "Glocal alignment",
              ALIGN_LOCAL: "Local alignment"}        # SA-main.rs:8-31


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("sw_oracle.c", "sw_oracle_int.c", "sw_oracle_faithful.c", "sw_oracle_fill.inc")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        c = ctypes
        p = c.c_void_p
        L.swo_create.restype = p
        L.swo_destroy.argtypes = [p]
        L.swo_align.restype = c.c_int64
        L.swo_align.argtypes = [p, p, c.c_int64, p, c.c_int64, p, c.c_int64, c.c_int, c.c_float, c.c_float,
                                c.c_int, c.c_int, p, p, p, p, p]
        L.swo_cell.argtypes = [p, c.c_int64, c.c_int64, c.c_int64, p, p]
        L.swo_score_batch.restype = c.c_int
        L.swo_score_batch.argtypes = [p, p, p, p, c.c_int64, p, c.c_int64, c.c_float, c.c_float, c.c_int,
                                      p, p, p, c.c_int]
        L.swf_score_batch.restype = c.c_int
        L.swf_score_batch.argtypes = L.swo_score_batch.argtypes
        L.swo_int_score.restype = c.c_int
        L.swo_int_score.argtypes = [p, c.c_int64, p, c.c_int64, p, c.c_int64, c.c_int32, c.c_int32, c.c_int32,
                                    c.c_int, p, p, p]
        L.swo_int_score_batch.restype = c.c_int
        L.swo_int_score_batch.argtypes = [p, p, p, p, c.c_int64, p, c.c_int64, c.c_int32, c.c_int32, c.c_int32,
                                          c.c_int, p, p, p, c.c_int]
        _LIB = L
    return _LIB


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


# ----------------------------------------------------------------------------- matrices
BLOSUM62_LINES = """   A  R  N  D  C  Q  E  G  H  I  L  K  M  F  P  S  T  W  Y  V  B  Z  X  *
A  4 -1 -2 -2  0 -1 -1  0 -2 -1 -1 -1 -1 -2 -1  1  0 -3 -2  0 -2 -1  0 -4
R -1  5  0 -2 -3  1  0 -2  0 -3 -2  2 -1 -3 -2 -1 -1 -3 -2 -3 -1  0 -1 -4
N -2  0  6  1 -3  0  0  0  1 -3 -3  0 -2 -3 -2  1  0 -4 -2 -3  3  0 -1 -4
D -2 -2  1  6 -3  0  2 -1 -1 -3 -4 -1 -3 -3 -1  0 -1 -4 -3 -3  4  1 -1 -4
C  0 -3 -3 -3  9 -3 -4 -3 -3 -1 -1 -3 -1 -2 -3 -1 -1 -2 -2 -1 -3 -3 -2 -4
Q -1  1  0  0 -3  5  2 -2  0 -3 -2  1  0 -3 -1  0 -1 -2 -1 -2  0  3 -1 -4
E -1  0  0  2 -4  2  5 -2  0 -3 -3  1 -2 -3 -1  0 -1 -3 -2 -2  1  4 -1 -4
G  0 -2  0 -1 -3 -2 -2  6 -2 -4 -4 -2 -3 -3 -2  0 -2 -2 -3 -3 -1 -2 -1 -4
H -2  0  1 -1 -3  0  0 -2  8 -3 -3 -1 -2 -1 -2 -1 -2 -2  2 -3  0  0 -1 -4
I -1 -3 -3 -3 -1 -3 -3 -4 -3  4  2 -3  1  0 -3 -2 -1 -3 -1  3 -3 -3 -1 -4
L -1 -2 -3 -4 -1 -2 -3 -4 -3  2  4 -2  2  0 -3 -2 -1 -2 -1  1 -4 -3 -1 -4
K -1  2  0 -1 -3  1  1 -2 -1 -3 -2  5 -1 -3 -1  0 -1 -3 -2 -2  0  1 -1 -4
M -1 -1 -2 -3 -1  0 -2 -3 -2  1  2 -1  5  0 -2 -1 -1 -1 -1  1 -3 -1 -1 -4
F -2 -3 -3 -3 -2 -3 -3 -3 -1  0  0 -3  0  6 -4 -2 -2  1  3 -1 -3 -3 -1 -4
P -1 -2 -2 -1 -3 -1 -1 -2 -2 -3 -3 -1 -2 -4  7 -1 -1 -4 -3 -2 -2 -1 -2 -4
S  1 -1  1  0 -1  0  0  0 -1 -2 -2  0 -1 -2 -1  4  1 -3 -2 -2  0  0  0 -4
T  0 -1  0 -1 -1 -1 -1 -2 -2 -1 -1 -1 -1 -2 -1  1  5 -2 -2  0 -1 -1  0 -4
W -3 -3 -4 -4 -2 -2 -3 -2 -2 -3 -2 -3 -1  1 -4 -3 -2 11  2 -3 -4 -3 -2 -4
Y -2 -2 -2 -3 -2 -1 -2 -3  2 -1 -1 -2 -1  3 -3 -2 -2  2  7 -1 -3 -2 -1 -4
V  0 -3 -3 -3 -1 -2 -2 -3 -3  3  1 -2  1 -1 -2 -2  0 -3 -1  4 -3 -2 -1 -4
B -2 -1  3  4 -3  0  1 -1  0 -3 -4  0 -3 -3 -2  0 -1 -4 -3 -3  4  1 -1 -4
Z -1  0  0  1 -3  3  4 -2  0 -3 -3  1 -1 -3 -1  0 -1 -3 -2 -2  1  4 -1 -4
X  0 -1 -1 -1 -2 -1 -1 -1 -1 -1 -1 -1 -1 -1 -2  0  0 -2 -1 -1 -1 -1 -1 -4
* -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4 -4  1""".split("\n")


class Matrix:
    """SubstitutionMatrix of SA.rs:626-795: letters, letter->index map, f32 scores[a][b]."""

    def __init__(self, letters, scores):
        self.letters = list(letters)
        self.index = {s: i for i, s in enumerate(self.letters)}
        self.scores = np.ascontiguousarray(scores, dtype=np.float32)

    @staticmethod
    def load_from_lines(lines):                      # SA.rs:735-794
        cols, hm, count = [], {}, -1
        for line in lines:
            ptt = line.strip().split()
            if not ptt:
                raise ValueError("empty matrix line")    # reference: index panic on ptt[0]
            if ptt[0][0] == "#":
                continue
            count += 1
            if count == 0:
                for pp in ptt:
                    if pp in cols:
                        raise ValueError("%s was already found." % pp)
                    cols.append(pp)
            else:
                if ptt[0] not in cols:
                    raise ValueError("%s was not found in the row name." % ptt[0])
                for ll in range(1, len(ptt)):
                    try:
                        v = float(np.float32(float(ptt[ll])))
                    except ValueError:
                        v = 0.0                          # SA.rs:769-772
                    hm[(ptt[0][0], cols[ll - 1])] = v
        n = len(cols)
        sc = np.zeros((n, n), np.float32)
        for i in range(n):
            for j in range(n):
                if (cols[i], cols[j]) not in hm:
                    raise ValueError("score about %s %s is not defined." % (cols[i], cols[j]))
                sc[i, j] = hm[(cols[i], cols[j])]
        return Matrix(cols, sc)

    @staticmethod
    def blosum62():                                  # SA.rs:697-733
        return Matrix.load_from_lines(BLOSUM62_LINES)

    @staticmethod
    def match_mismatch(match, mismatch):             # SA.rs:681-695
        letters = list("ABCDEFGHIJKLMNOPQRSTUVWXYZ")
        sc = np.full((26, 26), mismatch, np.float32)
        np.fill_diagonal(sc, match)
        return Matrix(letters, sc)

    def get_score_str(self, a, b):                   # SA.rs:633-645
        return float(self.scores[self.index[a], self.index[b]])

    def seq_to_index(self, seq):                     # SA.rs:649-679
        out = np.empty(len(seq), np.int32)
        x = self.index.get("X")
        for k, ch in enumerate(seq):
            i = self.index.get(ch)
            if i is None:
                if x is None:
                    raise ValueError("unknown letter %s. please set X to allow scoring for undefined letter pair." % ch)
                i = x
            out[k] = i
        return out


# ----------------------------------------------------------------------------- FASTA (SA.rs:799-887)
class SeqData:
    def __init__(self, name="", desc="", seq=""):
        self.name, self.desc = name, desc
        self.seq = [c for c in seq if c not in "\r\n"]   # SeqData::create -> line_to_seq(s, true), SA.rs:808-816


def load_fasta_text(text, retainws=False):
    """SeqData::load_fasta (SA.rs:830-887) over the lines of `text`."""
    ret, seqbuff, name, desc = [], [], "", ""
    for line in text.splitlines():
        x = line.find(">")
        if x >= 0:
            if seqbuff or name:
                ret.append(_mk(name, desc, seqbuff))
                seqbuff = []
            s = line.strip()
            nameflag, nb, db = True, [], []
            for sii, ch in enumerate(s):
                if nameflag:
                    if sii == 0 and ch == ">":
                        continue
                    if ch.isspace():
                        if nb:
                            nameflag = False
                        continue
                    nb.append(ch)
                else:
                    db.append(ch)
            name, desc = "".join(nb), "".join(db)
        else:
            if retainws:
                seqbuff.extend(c for c in line if c not in "\r\n")
            else:
                seqbuff.extend(c for c in line if not c.isspace())
    if name or seqbuff:
        ret.append(_mk(name, desc, seqbuff))
    return ret


def _mk(name, desc, seq):
    s = SeqData(name, desc)
    s.seq = list(seq)
    return s


def load_fasta(path, retainws=False):
    with open(path, "r") as f:
        return load_fasta_text(f.read(), retainws)


# ----------------------------------------------------------------------------- align
class Engine:
    """SequenceAlignment of SA.rs:15-552 (matrix, go, ge, mode) over the C oracle."""

    def __init__(self, matrix, go, ge, mode):
        self.matrix, self.go, self.ge, self.mode = matrix, float(go), float(ge), int(mode)
        self._h = lib().swo_create()

    def __del__(self):
        try:
            lib().swo_destroy(self._h)
        except Exception:
            pass

    def raw(self, ia, ib, score_only=False, pssm=None):
        """fill + backtrack on index vectors -> (idx_a, idx_b, score, end_i, end_j)."""
        ia = np.ascontiguousarray(ia, np.int32)
        ib = np.ascontiguousarray(ib, np.int32)
        la, lb = len(ia), len(ib)
        if pssm is None:
            mat, n, is_p = self.matrix.scores, self.matrix.scores.shape[0], 0
        else:                                   # PositionSpecificMatrix: scores[a + b*a_length], SA.rs:593-595
            mat, n, is_p = np.ascontiguousarray(pssm, np.float32), la, 1
        xa = np.empty(la + lb + 1, np.int64)
        xb = np.empty(la + lb + 1, np.int64)
        sc = ctypes.c_float()
        ei, ej = ctypes.c_int64(), ctypes.c_int64()
        n_cols = lib().swo_align(self._h, _ptr(ia), la, _ptr(ib), lb, _ptr(mat), n, is_p, self.go, self.ge,
                                 self.mode, int(score_only), _ptr(xa), _ptr(xb), ctypes.byref(sc),
                                 ctypes.byref(ei), ctypes.byref(ej))
        if n_cols == -1:
            raise MemoryError("oracle allocation failed")
        if n_cols == -2:
            raise RuntimeError("reference panics on this input (SA.rs:360-371)")
        return xa[:n_cols].copy(), xb[:n_cols].copy(), float(sc.value), ei.value, ej.value

    def cell(self, lb, i, j):
        sc = np.empty(3, np.float32)
        pv = np.empty(3, np.uint8)
        lib().swo_cell(self._h, lb, i, j, _ptr(sc), _ptr(pv))
        return sc, pv

    def align(self, s1, s2, retain_all=True, score_only=False, pssm=None, partial_region=None):
        """align_partial (SA.rs:393-551) -> (list[str], list[str], score).  partial_region: the DP sees the prefixes
        (seq_to_index(.., Some(x)), SA.rs:396-405, 649-656), the padding loops run over the whole sequences (SA.rs:512-549)."""
        if isinstance(s1, str):
            s1 = SeqData(seq=s1)
        if isinstance(s2, str):
            s2 = SeqData(seq=s2)
        if pssm is None:
            p1, p2 = (len(s1.seq), len(s2.seq)) if partial_region is None else partial_region
            ia, ib = self.matrix.seq_to_index(s1.seq[:p1]), self.matrix.seq_to_index(s2.seq[:p2])
        else:
            ia, ib = np.arange(len(s1.seq), dtype=np.int32), np.arange(len(s2.seq), dtype=np.int32)
        xa, xb, score, _, _ = self.raw(ia, ib, score_only, pssm)
        if score_only:
            return [], [], score
        r1, r2 = inflate(s1.seq, s2.seq, xa, xb, self.mode, retain_all)
        return r1, r2, score


def inflate(seq1, seq2, xa, xb, mode, retain_all):
    """String re-inflation + retain_all padding, SA.rs:470-550."""
    ret1, ret2 = [], []
    start1 = start2 = end1 = end2 = -1
    for ii in xa:
        if ii > -1:
            if start1 < 0:
                start1 = int(ii)
            ret1.append(seq1[ii])
            end1 = int(ii)
        else:
            ret1.append("-")
    for ii in xb:
        if ii > -1:
            if start2 < 0:
                start2 = int(ii)
            ret2.append(seq2[ii])
            end2 = int(ii)
        else:
            ret2.append("-")
    if mode == ALIGN_LOCAL and not retain_all:
        return ret1, ret2
    r1, r2 = [], []
    if start1 < 0 or start2 < 0:                       # SA.rs:512-524
        for c in seq1:
            r1.append(c); r2.append("-")
        for c in seq2:
            r1.append("-"); r2.append(c)
        return r1, r2
    for c in seq1[:start1]:
        r1.append(c); r2.append("-")
    for c in seq2[:start2]:
        r1.append("-"); r2.append(c)
    r1 += ret1
    r2 += ret2
    for c in seq1[end1 + 1:]:
        r1.append(c); r2.append("-")
    for c in seq2[end2 + 1:]:
        r1.append("-"); r2.append(c)
    return r1, r2


def rust_f32_display(x):
    """Rust `{}` of an f32 (SA-main.rs:40): shortest round-trip decimal, no exponent, no trailing `.0`."""
    v = np.float32(x)
    if v == np.floor(v) and abs(float(v)) < 1e16:
        s = "%d" % int(v)
        return "-0" if (s == "0" and np.signbit(v)) else s
    return np.format_float_positional(v, unique=True, trim="-")


def cli_record(name1, name2, r1, r2, score, mode):
    """One stdout record of seqalign, SA-main.rs:40-43."""
    return "#score:%s\n#type:%s\n>%s\n%s\n\n>%s\n%s\n\n" % (
        rust_f32_display(score), MODE_NAMES[mode], name1, "".join(r1), name2, "".join(r2))


# ----------------------------------------------------------------------------- batches
def pack(seqs_idx):
    off = np.zeros(len(seqs_idx) + 1, np.int64)
    off[1:] = np.cumsum([len(s) for s in seqs_idx])
    allv = np.concatenate(seqs_idx).astype(np.int32) if len(seqs_idx) else np.zeros(0, np.int32)
    return np.ascontiguousarray(allv), off


def score_batch_f32(a_all, off_a, b_all, off_b, matrix_f32, go, ge, mode, nthreads=1, layout="flat"):
    """Dense f32 fill + score_only scan per pair (the reference's CPU path), OpenMP over pairs.
    layout "flat": contiguous arrays; "faithful": the reference's own data structure, one heap-owning cell per DP cell
    and an indirect call per score lookup (sw_oracle_faithful.c, SA.rs:556-572)."""
    n = len(off_a) - 1
    sc = np.empty(n, np.float32)
    ei = np.empty(n, np.int64)
    ej = np.empty(n, np.int64)
    m = np.ascontiguousarray(matrix_f32, np.float32)
    fn = lib().swf_score_batch if layout == "faithful" else lib().swo_score_batch
    rc = fn(_ptr(a_all), _ptr(off_a), _ptr(b_all), _ptr(off_b), n, _ptr(m), m.shape[0],
                               float(go), float(ge), int(mode), _ptr(sc), _ptr(ei), _ptr(ej), int(nthreads))
    if rc:
        raise MemoryError
    return sc, ei, ej


def score_batch_int(a_all, off_a, b_all, off_b, matrix_int, o, e, unit, mode, nthreads=1):
    """Linear-memory integer twin (sw_oracle_int.c); o,e are the scaled negative penalties."""
    n = len(off_a) - 1
    sc = np.empty(n, np.int64)
    ei = np.empty(n, np.int64)
    ej = np.empty(n, np.int64)
    m = np.ascontiguousarray(matrix_int, np.int32)
    rc = lib().swo_int_score_batch(_ptr(a_all), _ptr(off_a), _ptr(b_all), _ptr(off_b), n, _ptr(m), m.shape[0],
                                   int(o), int(e), int(unit), int(mode), _ptr(sc), _ptr(ei), _ptr(ej), int(nthreads))
    if rc:
        raise MemoryError
    return sc, ei, ej
