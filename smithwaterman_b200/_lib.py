"""Loader / builder of the C-ABI shared library (include/swb200.h).

The library is built in-tree (smithwaterman_b200/libswb200.so) with nvcc for sm_100a.  There is no
fallback: if the library cannot be built or loaded, importing callers get an exception.
"""
import ctypes
import glob
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
SO_PATH = os.path.join(_HERE, "libswb200.so")
HEADER = os.path.join(os.path.dirname(_HERE), "include", "swb200.h")

NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-shared"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build():
    if not os.path.exists(SO_PATH):
        return True
    t = os.path.getmtime(SO_PATH)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh")) + [HEADER]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force=False, verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -> libswb200.so (cross-compiles without a GPU)."""
    if not force and not needs_build():
        return SO_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO_PATH] + sources()
    subprocess.check_call(cmd)
    return SO_PATH


CLI_SRC = os.path.join(os.path.dirname(HEADER), "..", "tools", "swbcli.cpp")
CLI_PATH = os.path.join(os.path.dirname(SO_PATH), "swbcli")


def build_cli(force=False):
    """g++ tools/swbcli.cpp (the C++ host mirror include/swb200.hpp + the reference's command lines) -> swbcli."""
    src = os.path.normpath(CLI_SRC)
    hpp = os.path.join(os.path.dirname(HEADER), "swb200.hpp")
    if not force and os.path.exists(CLI_PATH) and all(os.path.getmtime(CLI_PATH) >= os.path.getmtime(p) for p in (src, hpp, HEADER, SO_PATH)):
        return CLI_PATH
    subprocess.check_call([os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", "-I", os.path.dirname(HEADER), src,
                           "-L", os.path.dirname(SO_PATH), "-lswb200", "-Wl,-rpath,$ORIGIN", "-o", CLI_PATH])
    return CLI_PATH


class Matrix(ctypes.Structure):
    _fields_ = [("n_letters", ctypes.c_int32), ("letters", ctypes.c_char_p), ("scores", ctypes.POINTER(ctypes.c_float))]


class Alignment(ctypes.Structure):
    _fields_ = [("score", ctypes.c_float), ("end_a", ctypes.c_int64), ("end_b", ctypes.c_int64),
                ("n_cols", ctypes.c_int64), ("idx_a", ctypes.POINTER(ctypes.c_int64)),
                ("idx_b", ctypes.POINTER(ctypes.c_int64)), ("aln_len", ctypes.c_int64),
                ("aln_a", ctypes.c_void_p), ("aln_b", ctypes.c_void_p)]


# swb_aln_stats as a numpy record (48 bytes, the C layout)
ALN_STATS_DTYPE = [("score", "<f4"), ("status", "<i4"), ("end_a", "<i8"), ("end_b", "<i8"), ("aln_len", "<i8"),
                   ("a_letters", "<i8"), ("b_letters", "<i8"), ("matches", "<i8")]


class Hit(ctypes.Structure):            # swb_hit, 32 bytes
    _fields_ = [("score", ctypes.c_float), ("shard", ctypes.c_int32), ("index", ctypes.c_int64),
                ("end_q", ctypes.c_int64), ("end_t", ctypes.c_int64)]


HIT_DTYPE = [("score", "<f4"), ("shard", "<i4"), ("index", "<i8"), ("end_q", "<i8"), ("end_t", "<i8")]


class Stats(ctypes.Structure):
    _fields_ = [("cells", ctypes.c_double), ("kernel_ms", ctypes.c_double), ("h2d_bytes", ctypes.c_double),
                ("d2h_bytes", ctypes.c_double), ("launches", ctypes.c_int64), ("i32_reruns", ctypes.c_int64),
                ("scale", ctypes.c_int32), ("exact_scale", ctypes.c_int32), ("ckpt_pairs", ctypes.c_int64), ("k3_pairs", ctypes.c_int64)]


_lib = None


def lib():
    """ctypes handle with argtypes set for every symbol declared in include/swb200.h."""
    global _lib
    if _lib is not None:
        return _lib
    so = os.environ.get("SWB_LIB")          # tuning aid: A/B a differently compiled build of the same sources
    if not so:
        if needs_build():
            build()
        so = SO_PATH
    L = ctypes.CDLL(so)
    c = ctypes
    vp, i64, i32, f32 = c.c_void_p, c.c_int64, c.c_int32, c.c_float
    sig = {
        "swb_matrix_blosum62": (i32, [vp, vp]),
        "swb_matrix_match_mismatch": (i32, [f32, f32, vp, vp]),
        "swb_engine_create": (c.c_int, [c.POINTER(Matrix), f32, f32, c.c_int, c.c_int, c.POINTER(vp)]),
        "swb_engine_destroy": (None, [vp]),
        "swb_last_error": (c.c_char_p, [vp]),
        "swb_get_stats": (c.c_int, [vp, c.POINTER(Stats)]),
        "swb_device_count": (c.c_int, []),
        "swb_align": (c.c_int, [vp, vp, i64, vp, i64, c.c_int, c.POINTER(Alignment)]),
        "swb_align_pssm": (c.c_int, [vp, vp, vp, i64, vp, i64, c.c_int, c.POINTER(Alignment)]),
        "swb_align_batch": (c.c_int, [vp, vp, vp, vp, vp, i64, c.c_uint, c.POINTER(Alignment)]),
        "swb_alignment_free": (None, [c.POINTER(Alignment)]),
        "swb_align_stats_batch": (c.c_int, [vp, vp, vp, vp, vp, i64, c.c_uint, vp]),
        "swb_cluster": (c.c_int, [vp, vp, vp, i64, c.c_double, c.c_double, c.c_double, vp, vp, vp, vp]),
        "swb_score_batch": (c.c_int, [vp, vp, vp, vp, vp, i64, vp, vp, vp]),
        "swb_search_scores": (c.c_int, [vp, vp, i64, vp, vp, i64, vp, vp, vp]),
        "swb_score_batch_static": (c.c_int, [vp, vp, vp, i64, vp, vp, i64, i64, vp, vp, vp, vp, vp]),
        "swb_batch_upload": (c.c_int, [vp, vp, vp, vp, vp, i64, c.POINTER(vp)]),
        "swb_batch_run": (c.c_int, [vp, vp]),
        "swb_batch_fetch": (c.c_int, [vp, vp, vp, vp, vp]),
        "swb_batch_free": (None, [vp, vp]),
        "swb_batch_last_run_ms": (c.c_int, [vp, vp, c.POINTER(f32), c.POINTER(i64)]),
        "swb_batch_run_n": (c.c_int, [vp, vp, c.c_int, c.POINTER(f32), c.POINTER(i64)]),
        "swb_dpx_peak": (c.c_int, [c.c_int, c.c_int, c.POINTER(c.c_double)]),
        "swb_align_partial": (c.c_int, [vp, vp, i64, vp, i64, i64, i64, c.c_int, c.c_int, c.POINTER(Alignment)]),
        "swb_db_shard": (c.c_int, [vp, i64, c.c_int, c.c_int, vp, c.POINTER(i64)]),
        "swb_db_upload": (c.c_int, [vp, vp, vp, i64, vp, c.POINTER(vp)]),
        "swb_db_upload_shard": (c.c_int, [vp, vp, vp, i64, c.c_int, c.c_int, c.POINTER(vp)]),
        "swb_db_free": (None, [vp, vp]),
        "swb_db_count": (i64, [vp]),
        "swb_db_residues": (i64, [vp]),
        "swb_search": (c.c_int, [vp, vp, i64, vp, c.c_int, vp, c.POINTER(i64)]),
        "swb_db_scores": (c.c_int, [vp, vp, vp, vp, vp]),
        "swb_search_merge": (c.c_int, [vp, i64, c.c_int, vp, c.POINTER(i64)]),
        "swb_search_align": (c.c_int, [vp, vp, i64, vp, vp, i64, c.c_uint, vp, vp]),
        "swb_search_sharded": (c.c_int, [vp, vp, c.c_int, vp, i64, c.c_int, c.c_uint, vp, c.POINTER(i64), vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)          # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    L._swb_symbols = sorted(sig)
    _lib = L
    return L
