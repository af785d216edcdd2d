// K1 -- inter-sequence batch kernel: local (Smith-Waterman) score + first-maximum cell.
//
// Replaces the fill + score_only scan of the reference for a batch of independent pairs
// (fill_matrix SA.rs:55-272, score scan SA.rs:438-467 / start-cell scan SA.rs:282-295; the serial
// per-pair loop of `sa_opencl -list`, OCL-main.rs:295-316).
//
// Layout of the work: ONE THREAD = ONE PAIR, 32 pairs of similar shape = one warp-sized work item
// fetched from a global counter.  A thread sweeps its DP matrix in passes of 32 columns.  The two
// s16 lanes of every packed register belong to the SAME pair: the low lane holds columns
// [32q, 32q+16) at row i, the high lane columns [32q+16, 32q+32) at row i-1 (skewed by one row so
// that the high lane's left neighbour -- the low lane's column 15 -- is already known).  Every DPX
// instruction (VIADDMNMX.S16x2[.RELU], VIMNMX.S16x2, VIMNMX3.S16x2, VIADD.16x2) therefore advances
// two cells.
//
// Recurrence (Gotoh form, equal in value to the reference's 3-state form when |open| >= |extend|,
// SURVEY.md 8a note 3), all registers hold Ho = H + open:
//     E  = max(E + ext, Ho_left)           VIADDMNMX.S16x2
//     F  = max(F + ext, Ho_up)             VIADDMNMX.S16x2
//     H  = max(Ho_diag + (s - open), E, F, 0)   VIMNMX.S16x2 + VIADDMNMX.S16x2.RELU
//     Ho = H + open                        VIADD.16x2
// Substitution scores (s - open) come from a shared-memory table replicated per lane
// (bank == lane, conflict-free): one LDS per s16 lane, merged with one IMAD.
//
// End cell = first strict maximum of H in (row, column) order == the reference's first strict
// maximum of the MATCH state (SA.rs:286-294): a maximal H cell reached through a gap state always
// has an equal-valued cell before it in that order.  Per row step only a packed row maximum is
// kept (VIMNMX3, half an instruction per cell); the column is recovered by a register scan in
// the (rare) steps where the running best improves.
#include "swb_internal.h"

__device__ __forceinline__ unsigned lds32(unsigned addr)
{
    unsigned v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ int lo16(unsigned v) { return (int)(short)(v & 0xFFFFu); }
__device__ __forceinline__ int hi16(unsigned v) { return ((int)v) >> 16; }
__device__ __forceinline__ unsigned pack2(int lo, int hi) { return ((unsigned)lo & 0xFFFFu) | ((unsigned)hi << 16); }

#define K1_CELL(k)                                                                                   \
    {                                                                                                \
        unsigned s_lo = lds32(cb_lo[k] + roff_lo);                                                   \
        unsigned s_hi = lds32(cb_hi[k] + roff_hi);                                                   \
        unsigned s2 = s_hi * 65536u + s_lo;                                                          \
        E = __viaddmax_s16x2(E, ext2, HoL);                                                          \
        unsigned Fk = __viaddmax_s16x2(F[k], ext2, Ho[k]);                                           \
        unsigned T = __vmaxs2(E, Fk);                                                           \
        unsigned H = __viaddmax_s16x2_relu(HoD, s2, T);                                              \
        HoD = Ho[k];                                                                                 \
        HoL = __vadd2(H, open2);                                                                     \
        Ho[k] = HoL;                                                                                 \
        F[k] = Fk;                                                                                   \
    }

__global__ void __launch_bounds__(K1_THREADS * 2, 1)
k1_local_kernel(const uint8_t *__restrict__ a_pk, const uint8_t *__restrict__ b_pk,
                const K1Group *__restrict__ groups, int n_groups, const int32_t *__restrict__ la_arr,
                const int32_t *__restrict__ lb_arr, const int32_t *__restrict__ mat, int nl, int open, int ext,
                int limit_ho, uint2 *__restrict__ bnd, int bnd_rows, int *counter, K1Result *__restrict__ res)
{
    extern __shared__ unsigned T_sh[];               // [nl*nl][32 lanes]: (s - open) & 0xFFFF
    for (int idx = threadIdx.x; idx < nl * nl * 32; idx += blockDim.x)
        T_sh[idx] = (unsigned)(mat[idx >> 5] - open) & 0xFFFFu;
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const int wslot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    uint2 *mybnd = bnd + (size_t)wslot * (size_t)bnd_rows * 32 + lane;      // [row][lane]
    const unsigned tbase = (unsigned)__cvta_generic_to_shared(T_sh) + lane * 4;
    const unsigned rowstride = (unsigned)nl * 128u;
    const unsigned open2 = pack2(open, open), ext2 = pack2(ext, ext);

    for (;;) {
        int g = 0;
        if (lane == 0) g = atomicAdd(counter, 1);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= n_groups) break;
        const K1Group G = groups[g];
        const int p = G.first_pair + lane;
        const bool valid = lane < G.n_pairs;
        const int la = valid ? la_arr[p] : 0;
        int bestHo = open, bi = -1, bj = -1, ovf = 0;
        const uint4 *ap = reinterpret_cast<const uint4 *>(a_pk + G.a_off) + lane;       // chunk c at ap[c*32]
        const uint4 *bp = reinterpret_cast<const uint4 *>(b_pk + G.b_off) + lane * 2;   // pass q at bp[q*64]
        const int nsteps = G.la_max + 1;

        for (int q = 0; q < G.n_pass; ++q) {
            unsigned cb_lo[K1_HALF], cb_hi[K1_HALF], Ho[K1_HALF], F[K1_HALF];
            {
                const uint4 c0 = bp[q * 64], c1 = bp[q * 64 + 1];
                const unsigned w[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
#pragma unroll
                for (int k = 0; k < K1_HALF; ++k) {
                    cb_lo[k] = tbase + ((w[k >> 2] >> ((k & 3) * 8)) & 0xFFu) * 128u;
                    cb_hi[k] = tbase + ((w[4 + (k >> 2)] >> ((k & 3) * 8)) & 0xFFu) * 128u;
                    Ho[k] = open2;          // H(0,j) = 0
                    F[k] = 0u;              // any value <= 0 is equivalent under the clamp
                }
            }
            const bool first = (q == 0), last = (q == G.n_pass - 1);
            unsigned HoD = open2;                        // diag of the left boundary, row 0
            unsigned HoL_next, E_next;                   // packed left boundary for the coming step
            {
                uint2 b0 = first ? make_uint2(open2, open2) : mybnd[32];   // row 1, high lanes hold the values
                HoL_next = __byte_perm(b0.x, open2, 0x5432);
                E_next = __byte_perm(b0.y, open2, 0x5432);
            }
            unsigned r_prev = (unsigned)(nl - 1);        // PAD row for the high lane's "row 0"
            const int jbase = q * K1_STRIP;

            for (int c = 0; c * K1_ROWCHUNK < nsteps; ++c) {
                uint4 ar = ap[c * 32];
#pragma unroll 1
                for (int w4 = 0; w4 < 4; ++w4) {
                    const int i0 = c * K1_ROWCHUNK + w4 * 4;      // rows i0+1 .. i0+4 (low lane)
                    if (i0 >= nsteps) break;
                    const unsigned word = ar.x;
                    ar.x = ar.y; ar.y = ar.z; ar.z = ar.w;
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const int i = i0 + t + 1;                // low lane row; high lane row = i-1
                        const unsigned r = (word >> (t * 8)) & 0xFFu;
                        const unsigned roff_lo = r * rowstride, roff_hi = r_prev * rowstride;
                        r_prev = r;
                        unsigned HoL = HoL_next, E = E_next;
                        const unsigned HoD_next = HoL;
                        // prefetch next row's left boundary (previous pass wrote rows 1..nsteps-1; beyond = PAD rows)
                        uint2 bn = make_uint2(open2, open2);
                        if (!first && i + 1 < nsteps) bn = mybnd[(size_t)(i + 1) * 32];
                        K1_CELL(0) K1_CELL(1) K1_CELL(2) K1_CELL(3) K1_CELL(4) K1_CELL(5) K1_CELL(6) K1_CELL(7)
                        K1_CELL(8) K1_CELL(9) K1_CELL(10) K1_CELL(11) K1_CELL(12) K1_CELL(13) K1_CELL(14) K1_CELL(15)
                        // pass boundary: high lanes of (Ho[15], E) are column 32q+31 at row i-1
                        if (!last && i >= 2) mybnd[(size_t)(i - 1) * 32] = make_uint2(Ho[15], E);
                        HoL_next = __byte_perm(bn.x, Ho[15], 0x5432);
                        E_next = __byte_perm(bn.y, E, 0x5432);
                        HoD = HoD_next;
                        // packed row maximum of Ho (low lane: row i, high lane: row i-1)
                        unsigned rm = __vimax3_s16x2(Ho[0], Ho[1], Ho[2]);
                        rm = __vimax3_s16x2(rm, Ho[3], Ho[4]);
                        rm = __vimax3_s16x2(rm, Ho[5], Ho[6]);
                        rm = __vimax3_s16x2(rm, Ho[7], Ho[8]);
                        rm = __vimax3_s16x2(rm, Ho[9], Ho[10]);
                        rm = __vimax3_s16x2(rm, Ho[11], Ho[12]);
                        rm = __vimax3_s16x2(rm, Ho[13], Ho[14]);
                        rm = __vmaxs2(rm, Ho[15]);
                        const int rlo = lo16(rm), rhi = hi16(rm);
                        const int mm = max(rlo, rhi);
                        if (mm > bestHo || (mm == bestHo && mm > open && i - 1 < bi)) {
                            if (mm >= limit_ho) ovf = 1;
                            // low lane: row i
                            if (i <= la && (rlo > bestHo || (rlo == bestHo && rlo > open && i < bi))) {
                                int kk = 0;
#pragma unroll
                                for (int k = K1_HALF - 1; k >= 0; --k)
                                    if (lo16(Ho[k]) == rlo) kk = k;
                                bestHo = rlo; bi = i; bj = jbase + kk + 1;
                            }
                            // high lane: row i-1
                            if (i >= 2 && i - 1 <= la && (rhi > bestHo || (rhi == bestHo && rhi > open && i - 1 < bi))) {
                                int kk = 0;
#pragma unroll
                                for (int k = K1_HALF - 1; k >= 0; --k)
                                    if (hi16(Ho[k]) == rhi) kk = k;
                                bestHo = rhi; bi = i - 1; bj = jbase + K1_HALF + kk + 1;
                            }
                        }
                    }
                }
            }
        }
        if (valid) {
            K1Result r;
            r.score = bestHo - open;
            r.end_i = (bestHo > open) ? bi : -1;
            r.end_j = (bestHo > open) ? bj : -1;
            r.flags = ovf;
            res[p] = r;
        }
    }
}

size_t k1_bnd_bytes(int n_blocks, int la_max)
{
    return (size_t)n_blocks * (K1_THREADS * 2 / 32) * (size_t)(la_max + 2) * 32 * sizeof(uint2);
}

cudaError_t k1_launch(const swb_engine *e, const swb_batch *b, int n_blocks, cudaStream_t st)
{
    const int nl = e->n + 1;
    const size_t smem = (size_t)nl * nl * 32 * sizeof(unsigned);
    cudaError_t err = cudaFuncSetAttribute(k1_local_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    err = cudaMemsetAsync(b->d_counter, 0, sizeof(int32_t), st);
    if (err != cudaSuccess) return err;
    // Ho must stay below 32767 - smax so that Ho_diag + (s - open) cannot wrap.
    const int limit_ho = 32767 - 2 * e->smax + e->o - 1;
    k1_local_kernel<<<n_blocks, K1_THREADS * 2, smem, st>>>(
        b->d_a, b->d_b, b->d_groups, (int)b->groups.size(), b->d_la, b->d_lb, e->d_mat, nl, e->o, e->e, limit_ho,
        reinterpret_cast<uint2 *>(e->d_bnd), b->la_max + 2, b->d_counter, b->d_res);
    return cudaGetLastError();
}
