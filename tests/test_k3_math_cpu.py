"""The arithmetic of the packed traceback path (csrc/k3_tile.cu) restated on the CPU and checked cell by cell against
the oracle's three-state matrix: from the values K3 stores on its border grid (Gotoh H / raw E / raw F / raw M of the
last column of a pass; H of rows r-1 and r and raw F of row r at a row checkpoint) the walker's reconstruction
   M(r, j) = H(r-1, j-1) + s,   X(r, j) = row recurrence from the stored E of the left border,   Y(r, j) = F(r, j)
followed by the exact three-state cell reproduces every state value AND every predecessor code of the reference inside a
tile -- for local and global mode, including |open| == |extend| and tie-rich sequences.  (No GPU: this pins the math the
kernels implement; the GPU tests pin the kernels.)"""
import numpy as np
import pytest

from oracle import oracle as O

NEG = -(1 << 29)


def am(m, x, y):
    return 0 if (m >= x and m >= y) else (1 if x >= y else 2)


def kw_cell(mode, hd, amd, t, l, s, o, e):
    """Port of kw_cell (csrc/kw_cell.cuh): returns (m, x, y, (pm, px, py)) with code 3 = local stop."""
    tM, tX, tY = t
    lM, lX, lY = l
    a, b, c = tM + o, tX + e, tY + o
    f, g, h = lM + o, lY + e, lX + o
    pm = amd
    m = hd + s
    if mode != O.ALIGN_LOCAL:
        x = max(a, b, c); px = 1 if (b >= a and b >= c) else (0 if a >= c else 2)
        y = max(f, g, h); py = 2 if (g >= f and g >= h) else (0 if f >= h else 1)
    else:
        x = max(a, b, c, 0); px = 2 if (c >= a and c >= b) else (0 if a >= b else 1)
        y = max(f, g, h, 0); py = 1 if (h >= f and h >= g) else (0 if f >= g else 2)
        m = max(m, 0)
        if m <= 0: pm = 3
        if x <= 0: px = 3
        if y <= 0: py = 3
    return m, x, y, (pm, px, py)


def gotoh(mode, a, b, mat, o, e):
    """What K3 computes: H, raw E, raw F, raw M (local: H clamped, the others not), with the reference's boundaries."""
    la, lb = len(a), len(b)
    so, se = (o, e) if mode == O.ALIGN_GLOBAL else (0, 0)
    sent = 10 * o + 10 * e
    H = np.zeros((la + 1, lb + 1), np.int64); E = np.full((la + 1, lb + 1), NEG, np.int64)
    F = np.full((la + 1, lb + 1), NEG, np.int64); M = np.full((la + 1, lb + 1), NEG, np.int64)
    for j in range(1, lb + 1):
        ls = j * se + (so - se); H[0, j] = ls; E[0, j] = ls; F[0, j] = ls + sent; M[0, j] = ls + sent
    for i in range(1, la + 1):
        ls = i * se + (so - se); H[i, 0] = ls; F[i, 0] = ls; E[i, 0] = ls + sent; M[i, 0] = ls + sent
    E[0, 0] = -1; F[0, 0] = -1; M[0, 0] = 0
    for i in range(1, la + 1):
        for j in range(1, lb + 1):
            E[i, j] = max(E[i, j - 1] + e, H[i, j - 1] + o)
            F[i, j] = max(F[i - 1, j] + e, H[i - 1, j] + o)
            M[i, j] = H[i - 1, j - 1] + mat[a[i - 1], b[j - 1]]
            h = max(M[i, j], E[i, j], F[i, j])
            H[i, j] = max(h, 0) if mode == O.ALIGN_LOCAL else h
    return H, E, F, M


def border_states(mode, H, E, F, M, i, j):
    """(M, X, Y) of border cell (i, j) as the walker takes them from the store (kt3_border_cell)."""
    if i == 0 or j == 0:
        return int(M[i, j]), int(E[i, j]), int(F[i, j])            # matrix boundary: analytic, never clamped
    v = (int(M[i, j]), int(E[i, j]), int(F[i, j]))
    return tuple(max(x, 0) for x in v) if mode == O.ALIGN_LOCAL else v


@pytest.mark.parametrize("mode", [O.ALIGN_LOCAL, O.ALIGN_GLOBAL])
@pytest.mark.parametrize("o,e,scale", [(-20, -1, 2), (-2, -2, 2), (-5, -2, 1), (-1, 0, 2)])
def test_tile_reconstruction_reproduces_every_state_and_code(mode, o, e, scale):
    rng = np.random.default_rng(abs(o) * 100 + abs(e) * 10 + mode)
    m = O.Matrix.blosum62() if scale == 2 else O.Matrix.match_mismatch(2.0, -3.0)
    mat = np.rint(m.scores * scale).astype(np.int64)
    nlet = 20 if scale == 2 else 4
    oeng = O.Engine(m, -o / scale, -e / scale, mode)
    for trial in range(6):
        la, lb = int(rng.integers(20, 60)), int(rng.integers(20, 60))
        a = rng.integers(0, nlet, la).astype(np.int32)
        b = rng.integers(0, nlet, lb).astype(np.int32)
        if trial % 2:                                           # related pairs: long diagonals, many ties with o == e
            b = np.concatenate([a[: la // 2], rng.integers(0, nlet, 5).astype(np.int32), a[la // 2:]])[:lb] if lb <= la + 5 else b
            lb = len(b)
        oeng.raw(a, b)
        H, E, F, M = gotoh(mode, a, b, mat, o, e)
        # the Gotoh values equal the three-state values (|open| >= |extend|)
        for i, j in ((la, lb), (la // 2, lb // 3), (1, 1), (la, 1), (1, lb)):
            sc, _ = oeng.cell(lb, i, j)
            ref = np.rint(sc.astype(np.float64) * scale).astype(np.int64)
            got = border_states(mode, H, E, F, M, i, j)
            assert (int(ref[0]), int(ref[1]), int(ref[2])) == got
        R, W = 8, 7                                             # tile height / width of the emulation
        for rtop in range(0, la, R):
            for c0 in range(0, lb, W):
                cx, cy = min(la, rtop + R), min(lb, c0 + W)
                ncols = cy - c0
                # --- what the walker reconstructs on the top border (row rtop) ---
                top = []
                for l in range(ncols):
                    j = c0 + 1 + l
                    if rtop == 0:
                        top.append(border_states(mode, H, E, F, M, 0, j))
                        continue
                    Mu = int(H[rtop - 1, j - 1]) + int(mat[a[rtop - 1], b[j - 1]])
                    ev = int(E[rtop, c0]) + (l + 1) * e
                    for k in range(l + 1):
                        ev = max(ev, int(H[rtop, c0 + k]) + o + (l - k) * e)
                    Yu = int(F[rtop, j])
                    v = (Mu, ev, Yu)
                    top.append(tuple(max(x, 0) for x in v) if mode == O.ALIGN_LOCAL else v)
                left = [border_states(mode, H, E, F, M, rtop + r, c0) for r in range(cx - rtop + 1)]
                # --- the exact cell over the tile, row by row ---
                up = list(top)
                for r in range(1, cx - rtop + 1):
                    i = rtop + r
                    t = left[r]
                    d = left[r - 1]
                    for l in range(ncols):
                        j = c0 + 1 + l
                        hd, ad = max(d), am(*d)
                        mm, xx, yy, codes = kw_cell(mode, hd, ad, t, up[l], int(mat[a[i - 1], b[j - 1]]), o, e)
                        sc, pv = oeng.cell(lb, i, j)
                        ref = np.rint(sc.astype(np.float64) * scale).astype(np.int64)
                        assert (mm, xx, yy) == (int(ref[0]), int(ref[1]), int(ref[2])), (mode, i, j)
                        for s_ in range(3):
                            if codes[s_] != 3:
                                assert codes[s_] == int(pv[s_]), (mode, i, j, s_)
                            else:
                                assert (mm, xx, yy)[s_] == 0
                        d = up[l]
                        up[l] = (mm, xx, yy)
                        t = (mm, xx, yy)
