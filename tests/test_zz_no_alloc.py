"""GPU test of swb_score_batch_static (csrc/static_scratch.cu): the allocation-free variant of the pair scorer -- device
buffers in and out, kernel scratch in a static __device__ array -- against the existing entry point swb_score_batch on the
same pairs, bit-exact (integer arithmetic; the f32 scores are exact multiples of 1/scale), on two calls in a row that
share the scratch.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest

import smithwaterman_b200 as swb
from smithwaterman_b200.engine import pack_sequences

pytestmark = pytest.mark.gpu

AA = np.frombuffer(b"ARNDCQEGHILKMFPSTWYV", np.uint8)
MAX_PAIRS, MAX_LEN, MAX_RES = 4096, 1024, 1 << 20          # SWB_STATIC_MAX_* of include/swb200.h
SWB_ERR_ARG, SWB_ERR_LETTER = 1, 3


def _pairs(rng, n, lo, hi):
    sa = [bytes(AA[rng.integers(0, 20, int(rng.integers(lo, hi + 1)))]) for _ in range(n)]
    sb = [bytes(AA[rng.integers(0, 20, int(rng.integers(lo, hi + 1)))]) for _ in range(n)]
    for k in range(0, n, 7):                                       # related pairs: long local alignments, gaps on the path
        w = bytearray(sa[k])
        for p in rng.integers(0, max(1, len(w)), len(w) // 5):
            if len(w):
                w[p] = AA[rng.integers(0, 20)]
        del w[len(w) // 3:len(w) // 3 + int(rng.integers(0, 12))]
        sb[k] = bytes(w)
    return sa, sb


def _static_call(eng, a, oa, b, ob, with_ends=True, check_free=False):
    """One call through the C ABI with torch-owned device buffers (allocated before the call, by the caller)."""
    import torch
    dev = torch.device("cuda:0")
    n = len(oa) - 1
    t_a = torch.from_numpy(np.ascontiguousarray(a)).to(dev) if len(a) else torch.zeros(1, dtype=torch.uint8, device=dev)
    t_b = torch.from_numpy(np.ascontiguousarray(b)).to(dev) if len(b) else torch.zeros(1, dtype=torch.uint8, device=dev)
    t_oa = torch.from_numpy(np.ascontiguousarray(oa, dtype=np.int64)).to(dev)
    t_ob = torch.from_numpy(np.ascontiguousarray(ob, dtype=np.int64)).to(dev)
    t_sc = torch.full((max(n, 1),), -12345.0, dtype=torch.float32, device=dev)
    t_ea = torch.full((max(n, 1),), -777, dtype=torch.int64, device=dev)
    t_eb = torch.full((max(n, 1),), -777, dtype=torch.int64, device=dev)
    t_st = torch.full((1,), 99, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    before, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    rc = eng.L.swb_score_batch_static(eng._h, t_a.data_ptr(), t_oa.data_ptr(), len(a), t_b.data_ptr(), t_ob.data_ptr(), len(b), n,
                                      t_sc.data_ptr(), t_ea.data_ptr() if with_ends else None, t_eb.data_ptr() if with_ends else None,
                                      t_st.data_ptr(), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    eng._check(rc)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    if check_free:                                                 # a warm call: the device's free memory does not shrink either
        assert torch.cuda.mem_get_info()[0] >= free_before
    return t_sc.cpu().numpy()[:n], t_ea.cpu().numpy()[:n], t_eb.cpu().numpy()[:n], int(t_st.item())


@pytest.mark.parametrize("mode", [swb.ALIGN_LOCAL, swb.ALIGN_GLOBAL, swb.ALIGN_GLOCAL])
def test_static_scratch_scores_equal_score_batch_on_two_calls_in_a_row(mode):
    rng = np.random.default_rng(4242 + mode)
    eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, mode)
    # call 1: many short pairs (several pairs per ring, empty sequences, one maximal pair); call 2: fewer, longer pairs of
    # other shapes -- it runs on the scratch the first call left behind
    for call, (n, lo, hi) in enumerate([(700, 0, 300), (300, 200, MAX_LEN)]):
        sa, sb = _pairs(rng, n, lo, hi)
        sa[3] = b""; sb[5] = b""; sa[9] = b""; sb[9] = b"" if mode == swb.ALIGN_LOCAL else b"AC"
        sa[11] = bytes(AA[rng.integers(0, 20, MAX_LEN)]); sb[11] = sa[11][:700] + bytes(AA[rng.integers(0, 20, MAX_LEN - 700)])
        a, oa = pack_sequences(sa); b, ob = pack_sequences(sb)
        ref_sc, ref_ea, ref_eb = eng.score_packed(a, oa, b, ob)          # the existing entry point (host buffers)
        sc, ea, eb, status = _static_call(eng, a, oa, b, ob)
        assert status == 0, call
        assert np.array_equal(sc, ref_sc), (call, np.nonzero(sc != ref_sc)[0][:10])
        assert np.array_equal(ea, ref_ea) and np.array_equal(eb, ref_eb), call
        sc2, ea2, eb2, _ = _static_call(eng, a, oa, b, ob, with_ends=False, check_free=True)   # end arrays are optional and stay untouched
        assert np.array_equal(sc2, ref_sc) and (ea2 == -777).all() and (eb2 == -777).all()
    eng.close()


def test_static_scratch_limits_and_status():
    rng = np.random.default_rng(7)
    eng = swb.SequenceAlignment(swb.SubstitutionMatrix.get_blosum62_matrix(), 10.0, 0.5, swb.ALIGN_LOCAL)
    L = eng.L
    # beyond the stated maxima: rejected on the host with a clear message, nothing is enqueued
    rc = L.swb_score_batch_static(eng._h, None, None, 0, None, None, 0, MAX_PAIRS + 1, None, None, None, None, None)
    assert rc == SWB_ERR_ARG and b"at most 4096" in L.swb_last_error(eng._h)
    rc = L.swb_score_batch_static(eng._h, None, None, MAX_RES, None, None, 1, 1, None, None, None, None, None)
    assert rc == SWB_ERR_ARG and b"at most 1048576" in L.swb_last_error(eng._h)
    # a sequence longer than the maximum: flagged in the status word, the pair is scored as empty; the others are right
    sa, sb = _pairs(rng, 40, 10, 200)
    sa[4] = bytes(AA[rng.integers(0, 20, MAX_LEN + 1)])
    a, oa = pack_sequences(sa); b, ob = pack_sequences(sb)
    ref_sc, _, _ = eng.score_packed(a, oa, b, ob)
    sc, ea, eb, status = _static_call(eng, a, oa, b, ob)
    assert status == SWB_ERR_ARG and sc[4] == 0.0 and ea[4] == -1
    keep = np.arange(40) != 4
    assert np.array_equal(sc[keep], ref_sc[keep])
    # unknown letter with a matrix that has no X: SWB_ERR_LETTER in the status word (SA.rs:676); the next call is clean
    m = swb.SubstitutionMatrix.load_from_lines(["  A T G C", "A 5 -4 -4 -4", "T -4 5 -4 -4", "G -4 -4 5 -4", "C -4 -4 -4 5"])
    e2 = swb.SequenceAlignment(m, 10.0, 0.5, swb.ALIGN_LOCAL)
    a, oa = pack_sequences([b"ACGT", b"AC#T"]); b, ob = pack_sequences([b"ACGT", b"ACGT"])
    _, _, _, status = _static_call(e2, a, oa, b, ob)
    assert status == SWB_ERR_LETTER
    a, oa = pack_sequences([b"ACGT", b"ACTT"])
    sc, _, _, status = _static_call(e2, a, oa, b, ob)
    assert status == 0 and list(sc) == [20.0, 11.0]
    # no pairs: status only
    sc, _, _, status = _static_call(e2, np.zeros(0, np.uint8), np.zeros(1, np.int64), np.zeros(0, np.uint8), np.zeros(1, np.int64))
    assert status == 0 and len(sc) == 0
    eng.close(); e2.close()
