// K1, global (Needleman-Wunsch) form: score of cell (la, lb) for a batch of independent pairs, packed s16x2.
//
// Same work layout, packed recurrence and memory layouts as the local kernel in k1_score.cu (one thread = one pair, 32
// columns per pass, low s16 lane = columns [32q, 32q+16) at row i, high lane = columns [32q+16, 32q+32) at row i-1);
// what differs is what the reference's global mode differs in (fill_matrix SA.rs:55-272 with alignment_type GLOBAL):
//   * boundary row / column (SA.rs:88-118): H(0,0) = 0, H(0,j) = open + (j-1) ext (reached through GAPINX), H(i,0)
//     likewise through GAPINY; the other two states of a boundary cell hold the sentinel 10 open + 10 ext below it;
//   * no clamp at 0;
//   * the result is the maximum of the three states at (la, lb) (SA.rs:296-309), i.e. H(la, lb); no running maximum.
// The Gotoh form used here equals the reference's 3-state form in value when |open| >= |ext| (SURVEY.md 8a note 3);
// global mode charges the same penalties on the last row and column as elsewhere (so = open, se = ext, SA.rs:166-170),
// so the uniform packed cell applies as it is (glocal, whose last row / column are free, needs the GLOCAL form below).
// The s16 range is guaranteed by the host's routing (batch.cu k1_take: |scores| bounded from the lengths), not by flags.
// Cells beyond (la, lb) are computed on PAD residues; they may wrap, but nothing flows from them to a real cell.
//
// GLOCAL = true: the reference's glocal mode (SA.rs:74-82: so = se = 0) -- boundary row and column cost nothing, and so do
// gaps in the last row (SA.rs:166-167) and the last column (SA.rs:169-170).  A free move only ever copies a value along
// the last row / column, so the three-state maximum at (la, lb) is
//     max(0, max_i H(i, lb), max_j H(la, j))
// of the matrix filled with the INTERIOR penalties everywhere (M and the gap state that enters the last row / column are
// computed from interior cells with interior penalties; the free state is the running maximum of the cells before it).
// The kernel therefore runs the uniform cell with an all-zero boundary and tracks two maxima.  To keep the last column
// at a fixed register (word 15, high half, of the last pass) the packer RIGHT-aligns b: the first pass starts with PAD
// columns, and in this kernel PAD scores 0 -- a zero-score column next to the zero boundary reproduces the boundary
// exactly (H = max(0 + 0, E, F) = 0 with E, F <= open), so left padding is invisible to the real columns.
#include <cstdlib>

#include "swb_internal.h"
#include "k1_common.cuh"

#define K1G_CELL(k)                                                                                  \
    {                                                                                                \
        unsigned s_lo = lds32(cb_lo[k] + roff_lo);                                                   \
        unsigned s_hi = lds32(cb_hi[k] + roff_hi);                                                   \
        unsigned s2 = s_hi * 65536u + s_lo;                                                          \
        unsigned Fk = __viaddmax_s16x2(F[k], ext2, Ho[k]);                                           \
        unsigned G = __viaddmax_s16x2(HoD, s2, Fk);                                                  \
        E = __viaddmax_s16x2(E, ext2, GoP);                                                          \
        GoP = __vadd2(G, open2);                                                                     \
        HoD = Ho[k];                                                                                 \
        Ho[k] = __viaddmax_s16x2(E, open2, GoP);                                                     \
        F[k] = Fk;                                                                                   \
    }

template <bool GLOCAL>
__global__ void __maxnreg__(120)
k1_global_kernel(const uint8_t *__restrict__ a_pk, const uint8_t *__restrict__ b_pk,
                 const K1Group *__restrict__ groups, int n_groups, const int32_t *__restrict__ la_arr,
                 const int32_t *__restrict__ lb_arr, const int32_t *__restrict__ mat, int nl, int open, int ext,
                 uint4 *__restrict__ bnd, int *counter, const int32_t *__restrict__ order, K1Result *__restrict__ res)
{
    extern __shared__ unsigned T_sh[];               // [nl*nl][32 lanes]: (s - open) & 0xFFFF
    for (int idx = threadIdx.x; idx < nl * nl * 32; idx += blockDim.x) {
        const int cell = idx >> 5;
        const bool is_pad = (cell % nl == nl - 1) || (cell / nl == nl - 1);
        T_sh[idx] = (unsigned)((GLOCAL && is_pad ? 0 : mat[cell]) - open) & 0xFFFFu;      // glocal: PAD scores 0 (see above)
    }
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const int wslot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    uint4 *mybnd = bnd + (size_t)wslot * 32 + lane;              // [row block][resident warp][lane], as the local kernel
    const size_t bstride = (size_t)gridDim.x * (blockDim.x >> 5) * 32;
    const unsigned tbase = (unsigned)__cvta_generic_to_shared(T_sh) + lane * 4;
    const unsigned rowstride = (unsigned)nl * 128u;
    const unsigned open2 = pack2(open, open), ext2 = pack2(ext, ext);
    const unsigned long long pol = keep_policy();
    const int sent = 10 * open + 10 * ext;           // the "not reachable" states of a boundary cell sit this far below it
    // Ho of boundary cell (0, j) or (j, 0), j >= 1:  H + open = 2 open + (j-1) ext;  (0,0): open
    auto ho_bnd = [=](int j) { return (GLOCAL || j == 0) ? open : 2 * open + (j - 1) * ext; };     // glocal: H = 0 on the whole boundary

    for (;;) {
        int g = 0;
        if (lane == 0) g = atomicAdd(counter, 1);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= n_groups) break;
        const K1Group G_ = groups[g];
        const int p = G_.first_pair + lane;
        const bool valid = lane < G_.n_pairs;
        const int la = valid ? la_arr[p] : 0, lb = valid ? lb_arr[p] : 0;
        int score = 0;
        int best_row = open;                            // glocal: maximum of Ho over row la
        unsigned cm = open2;                            // glocal: running maximum of the last column (high half of word 15, last pass)
        const unsigned *ap = reinterpret_cast<const unsigned *>(a_pk + G_.a_off) + lane;  // rows 4B+1..4B+4 at ap[B*32]
        const uint4 *bp = reinterpret_cast<const uint4 *>(b_pk + G_.b_off) + lane * 2;    // pass q at bp[q*64]
        const int nblk = (G_.la_max + 1 + 3) >> 2;      // row steps run in blocks of 4

        for (int q = 0; q < G_.n_pass; ++q) {
            const int jbase = q * K1_STRIP;
            unsigned cb_lo[K1_HALF], cb_hi[K1_HALF], Ho[K1_HALF], F[K1_HALF];
            const unsigned fdelta = pack2(sent - open, sent - open);     // F(0,j) = H(0,j) + sent = Ho - open + sent
            {
                const uint4 c0 = bp[q * 64], c1 = bp[q * 64 + 1];
                const unsigned w[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
                for (int k = 0; k < K1_HALF; ++k) {
                    cb_lo[k] = tbase + ((w[k >> 2] >> ((k & 3) * 8)) & 0xFFu) * 128u;
                    cb_hi[k] = tbase + ((w[4 + (k >> 2)] >> ((k & 3) * 8)) & 0xFFu) * 128u;
                    Ho[k] = pack2(ho_bnd(jbase + k + 1), ho_bnd(jbase + K1_HALF + k + 1));      // row 0 (SA.rs:100-108)
                    F[k] = __vadd2(Ho[k], fdelta);
                }
            }
            const bool first = (q == 0), last = (q == G_.n_pass - 1);
            unsigned HoD = pack2(ho_bnd(jbase), open);                      // diag of the low lane's first column: cell (0, jbase)
            unsigned r_prev = (unsigned)(nl - 1);                           // PAD row for the high lane's "row 0" step (overwritten below)
            unsigned lo15_Ho = pack2(ho_bnd(jbase + K1_HALF), open);        // cell (0, jbase+16): the high lane's diagonal at row 1
            unsigned lo15_E = open2;
            // the cell that holds the result: column lb of row la, in the low half (step i == la) or the high half (step i == la + 1)
            const int tcol = lb - 1 - jbase;
            const int trow = (tcol >= 0 && tcol < K1_STRIP) ? (tcol < K1_HALF ? la : la + 1) : -1;
            const uint4 open4 = make_uint4(open2, open2, open2, open2);
            uint4 bq_next = first ? open4 : ld_keep(mybnd, pol);
            unsigned word_next = ap[0];
            unsigned c1 = open2, c2 = open2, c3 = open2; // boundary words of rows i0-3..i0-1 waiting for their 4th

#pragma unroll 1
            for (int blk = 0; blk < nblk; ++blk) {
                const int i0 = blk * 4;                           // rows i0+1 .. i0+4 (low lane)
                const unsigned word = word_next;
                const uint4 bq4 = bq_next;
                word_next = ap[(blk + 1) * 32];                   // a_pk / bnd have one block of slack
                if (!first) bq_next = ld_keep(mybnd + (size_t)(blk + 1) * bstride, pol);
                unsigned bq[4] = {bq4.x, bq4.y, bq4.z, bq4.w};
                if (first) {                                      // column 0 (SA.rs:109-118): Ho = H + open, E = H + sent
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const int hb = ho_bnd(i0 + t + 1);
                        bq[t] = pack2(hb, hb - open + sent);
                    }
                }
                unsigned wout[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int i = i0 + t + 1;                    // low lane row; high lane row = i-1
                    const unsigned r = (word >> (t * 8)) & 0xFFu;
                    const unsigned roff_lo = r * rowstride, roff_hi = r_prev * rowstride;
                    r_prev = r;
                    // packed left boundary: low lane <- previous pass (row i), high lane <- low lane's column 15 (row i-1)
                    unsigned GoP = __byte_perm(bq[t], lo15_Ho, 0x5410);
                    unsigned E = __byte_perm(bq[t], lo15_E, 0x5432);
                    const unsigned HoD_next = GoP;
                    K1G_CELL(0) K1G_CELL(1) K1G_CELL(2) K1G_CELL(3) K1G_CELL(4) K1G_CELL(5) K1G_CELL(6) K1G_CELL(7)
                    K1G_CELL(8) K1G_CELL(9) K1G_CELL(10) K1G_CELL(11) K1G_CELL(12) K1G_CELL(13) K1G_CELL(14) K1G_CELL(15)
                    if (t == 0 && blk == 0) {
                        // the high lane has just run its "row 0" step on PAD residues: put the boundary row there instead
#pragma unroll
                        for (int k = 0; k < K1_HALF; ++k) {
                            const int hb = ho_bnd(jbase + K1_HALF + k + 1);
                            Ho[k] = (Ho[k] & 0xFFFFu) | ((unsigned)hb << 16);
                            F[k] = (F[k] & 0xFFFFu) | ((unsigned)(hb - open + sent) << 16);
                        }
                        E = (E & 0xFFFFu) | ((unsigned)(ho_bnd(jbase + K1_STRIP) - open) << 16);   // X(0, jbase+32) = H(0, .)
                    }
                    // pass boundary: high lanes of (Ho[15], E) are column 32q+31 at row i-1
                    wout[t] = __byte_perm(Ho[15], E, 0x7632);
                    if (t == 0 && !last && blk > 0) st_keep(mybnd + (size_t)(blk - 1) * bstride, make_uint4(c1, c2, c3, wout[0]), pol);
                    lo15_Ho = Ho[15]; lo15_E = E;
                    HoD = HoD_next;
                    if (GLOCAL) {
                        if (last && i <= la + 1) cm = __vmaxs2(cm, Ho[15]);       // cell (i-1, lb); row 0 holds the boundary 0
                        if (i == la || i == la + 1) {                                  // row la: low halves now, high halves one step later
                            unsigned rm = __vimax3_s16x2(Ho[0], Ho[1], Ho[2]);
                            rm = __vimax3_s16x2(rm, Ho[3], Ho[4]); rm = __vimax3_s16x2(rm, Ho[5], Ho[6]);
                            rm = __vimax3_s16x2(rm, Ho[7], Ho[8]); rm = __vimax3_s16x2(rm, Ho[9], Ho[10]);
                            rm = __vimax3_s16x2(rm, Ho[11], Ho[12]); rm = __vimax3_s16x2(rm, Ho[13], Ho[14]);
                            rm = __vmaxs2(rm, Ho[15]);
                            best_row = max(best_row, i == la ? lo16(rm) : hi16(rm));
                        }
                    } else if (i == trow) {
                        int v = 0;
#pragma unroll
                        for (int k = 0; k < K1_HALF; ++k) {
                            if (k == tcol) v = lo16(Ho[k]);
                            if (k + K1_HALF == tcol) v = hi16(Ho[k]);
                        }
                        score = v - open;
                    }
                }
                c1 = wout[1]; c2 = wout[2]; c3 = wout[3];
            }
            if (!last) st_keep(mybnd + (size_t)(nblk - 1) * bstride, make_uint4(c1, c2, c3, open2), pol);   // rows 4(nblk-1)+1..+3
        }
        if (GLOCAL) score = max(max(best_row, hi16(cm)), open) - open;      // max(0, last row, last column)
        if (valid) {
            K1Result r;
            r.score = score; r.end_i = la; r.end_j = lb; r.flags = 0;       // SA.rs:296-309: the walk starts at (la, lb)
            res[order[p]] = r;
        }
    }
}

cudaError_t k1_global_launch(const swb_engine *e, const swb_batch *b, int n_blocks, void *d_bnd, cudaStream_t st)
{
    const int nl = e->n + 1;
    const size_t smem = (size_t)nl * nl * 32 * sizeof(unsigned);
    const bool glocal = e->mode == SWB_GLOCAL;
    cudaError_t err = glocal ? cudaFuncSetAttribute(k1_global_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                             : cudaFuncSetAttribute(k1_global_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    err = cudaMemsetAsync(b->d_counter, 0, sizeof(int32_t), st);
    if (err != cudaSuccess) return err;
    if (glocal)
        k1_global_kernel<true><<<n_blocks, K1_THREADS * 2, smem, st>>>(
            b->d_a, b->d_b, b->d_groups, b->n_groups, b->d_la, b->d_lb, e->d_mat, nl, e->o, e->e,
            reinterpret_cast<uint4 *>(d_bnd), b->d_counter, b->d_order, b->d_res);
    else
        k1_global_kernel<false><<<n_blocks, K1_THREADS * 2, smem, st>>>(
            b->d_a, b->d_b, b->d_groups, b->n_groups, b->d_la, b->d_lb, e->d_mat, nl, e->o, e->e,
            reinterpret_cast<uint4 *>(d_bnd), b->d_counter, b->d_order, b->d_res);
    return cudaGetLastError();
}
