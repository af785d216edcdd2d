// The exact three-state cell of the reference (fill_matrix SA.rs:139-268) with its tie-breaking, int32, branch-free:
// shared by the wavefront kernel (kw_trace.cu) and the tile walker of the packed traceback path (k3_tile.cu).
#pragma once
#include "swb_internal.h"

#define NEG_INF_I32 (-(1 << 29))

// Priority argmax of a cell's three states, M > X > Y (the MATCH predecessor rule, SA.rs:139-158).
// (Decisions compare the candidates with each other, never a maximum with its own operands: nvcc 12.9 folds
// `x = max(a, b); x == b` into the predicate outputs of VIMNMX[.RELU], which report the selected operand, not equality,
// and the tie cases then come out wrong.)
__device__ __forceinline__ unsigned kw_am(int m, int x, int y) { return (m >= x && m >= y) ? 0u : ((x >= y) ? 1u : 2u); }

// One cell, branch-free.  hd / amd: max and priority argmax of the diagonal cell's states; (tM,tX,tY) the left cell,
// (lM,lX,lY) the upper cell.  The predecessor of a gap state is read off the maximum by equality tests in priority
// order, which is the same decision as the reference's nested comparisons on the sums (SA.rs:164-252).
template <int MODE>
__device__ __forceinline__ void kw_cell(int hd, unsigned amd, int tM, int tX, int tY, int lM, int lX, int lY, int s,
                                        int po, int pe, int qo, int qe, int &m, int &x, int &y, unsigned &code)
{
    const int a = tM + po, b = tX + pe, c = tY + po;
    const int f = lM + qo, g = lY + qe, h = lX + qo;
    unsigned pm = amd, px, py;
    m = hd + s;
    if (MODE != SWB_LOCAL) {
        x = __vimax3_s32(a, b, c); px = (b >= a && b >= c) ? 4u : ((a >= c) ? 0u : 8u);    // X > M > Y  (SA.rs:172-189)
        y = __vimax3_s32(f, g, h); py = (g >= f && g >= h) ? 32u : ((f >= h) ? 0u : 16u);  // Y > M > X  (SA.rs:192-211)
    } else {
        x = __vimax3_s32_relu(a, b, c); px = (c >= a && c >= b) ? 8u : ((a >= b) ? 0u : 4u);   // Y > M > X  (SA.rs:214-231)
        y = __vimax3_s32_relu(f, g, h); py = (h >= f && h >= g) ? 16u : ((f >= g) ? 0u : 32u); // X > M > Y  (SA.rs:233-252)
        m = max(m, 0);                                                               // clamp after the choice (SA.rs:256-260)
        pm = (m <= 0) ? 3u : pm; px = (x <= 0) ? 12u : px; py = (y <= 0) ? 48u : py; // code 3: "score is 0", the walk's stop test
    }
    code = pm + px + py;
}

