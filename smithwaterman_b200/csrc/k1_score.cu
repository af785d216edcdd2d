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
//     F  = max(F + ext, Ho_up)                  VIADDMNMX.S16x2
//     G  = max(Ho_diag + (s - open), F, 0)      VIADDMNMX.S16x2.RELU
//     Go = G + open                             VIADD.16x2
//     E  = max(E + ext, Go_left)                VIADDMNMX.S16x2   (the only op on the row's dependency chain)
//     Ho = max(E + open, Go)  (= H + open)      VIADDMNMX.S16x2
// Substitution scores (s - open) come from a shared-memory table replicated per lane
// (bank == lane, conflict-free): one LDS per s16 lane, merged with one IMAD.
//
// End cell = first strict maximum of H in (row, column) order == the reference's first strict
// maximum of the MATCH state (SA.rs:286-294): a maximal H cell reached through a gap state always
// has an equal-valued cell before it in that order.  Per row step only a packed row maximum is
// kept (VIMNMX3, half an instruction per cell, seeded with the current threshold so that one compare
// tells whether anything improved).  In the (rare) steps where the running best of a lane improves,
// only the row and the 16-column half are decided; the step's 16 packed Ho words are parked in a
// per-thread shared-memory slot (16 STS; K2: 4 STS.128) and the column is read off that slot once per pair.
// (K1_NOTRACK=1 builds a timing-only variant without the end cell: the per-row test, its branch and
// the rare path together cost 10 % of the kernel, see DESIGN.md.)
#include <cstdlib>

#include "swb_internal.h"
#include "k1_common.cuh"

// K1_BLK_UNROLL=2 unrolls two 4-row blocks (8 rows) per loop trip: compiles to 119 registers without spills and about
// half the back-edge register moves per row; not measured yet (tools/ab_k1.sh with -DK1_BLK_UNROLL=2).
#ifndef K1_BLK_UNROLL
#define K1_BLK_UNROLL 1
#endif
constexpr int kBlkUnroll = K1_BLK_UNROLL;
#ifndef K1_NOTRACK
#define K1_NOTRACK 0
#endif
// K1_VOTE=1: the per-row improvement test becomes a warp-uniform branch (VOTE.ANY + BRA.U, no BSSY/BSYNC pair per row).
#ifndef K1_VOTE
#define K1_VOTE 0
#endif

// One packed cell-vector.  The only loop-carried dependency along k is the single VIADDMNMX of E:
//   G  = max(Hdiag + s, F, 0)                 (independent of the row chain)
//   E  = max(E + ext, Go_prev)                == max(E + ext, Ho_prev) because open <= ext
//   Ho = max(Go, E + open)                    == max(G, E) + open
#define K1_CELL(k)                                                                                   \
    {                                                                                                \
        unsigned s_lo = lds32(cb_lo[k] + roff_lo);                                                   \
        unsigned s_hi = lds32(cb_hi[k] + roff_hi);                                                   \
        unsigned s2 = s_hi * 65536u + s_lo;                                                          \
        unsigned Fk = __viaddmax_s16x2(F[k], ext2, Ho[k]);                                           \
        unsigned G = __viaddmax_s16x2_relu(HoD, s2, Fk);                                             \
        E = __viaddmax_s16x2(E, ext2, GoP);                                                          \
        GoP = __vadd2(G, open2);                                                                     \
        HoD = Ho[k];                                                                                 \
        Ho[k] = __viaddmax_s16x2(E, open2, GoP);                                                     \
        F[k] = Fk;                                                                                   \
    }

// (120 registers, not the 128 that 512 threads would allow: the 4096 left over hold one 128-thread helper block of the
// upload / result path per SM, see batch.cu SWB_HELPER; the kernel needs 119)
__global__ void __maxnreg__(120)
k1_local_kernel(const uint8_t *__restrict__ a_pk, const uint8_t *__restrict__ b_pk,
                const K1Group *__restrict__ groups, int n_groups, const int32_t *__restrict__ la_arr,
                const int32_t *__restrict__ lb_arr, const int32_t *__restrict__ mat, int nl, int open, int ext,
                int limit_ho, uint4 *__restrict__ bnd, int bnd_blocks, int *counter, const int32_t *__restrict__ order,
                K1Result *__restrict__ res)
{
    extern __shared__ __align__(16) unsigned T_sh[];               // [nl*nl][32 lanes]: (s - open) & 0xFFFF
    for (int idx = threadIdx.x; idx < nl * nl * 32; idx += blockDim.x)
        T_sh[idx] = (unsigned)(mat[idx >> 5] - open) & 0xFFFFu;
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const int wslot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    // Pass-boundary column, one word per row (low s16 = Ho, high s16 = E), stored as [row block of 4][lane] uint4:
    // storage block B holds rows 4B+1 .. 4B+4, i.e. exactly what the 4 row steps i0 = 4B need.
    // row-block major ([row block][resident warp][lane]): the blocks every pair touches (short rows) form one dense
    // region at the start of the buffer, which the launcher pins in L2 with a persisting access window
    uint4 *mybnd = bnd + (size_t)wslot * 32 + lane;
    const size_t bstride = (size_t)gridDim.x * (blockDim.x >> 5) * 32;
    const unsigned tbase = (unsigned)__cvta_generic_to_shared(T_sh) + lane * 4;
    const unsigned rowstride = (unsigned)nl * 128u;
    const unsigned open2 = pack2(open, open), ext2 = pack2(ext, ext);
    const unsigned long long pol = keep_policy();
    // per-thread snapshot of the 16 packed Ho words of the row step that holds the running best: [16][thread]
    unsigned *snap = T_sh + nl * nl * 32 + threadIdx.x;

    for (;;) {
        int g = 0;
        if (lane == 0) g = atomicAdd(counter, 1);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= n_groups) break;
        const K1Group G_ = groups[g];
        const int p = G_.first_pair + lane;
        const bool valid = lane < G_.n_pairs;
        const int la = valid ? la_arr[p] : 0;
        int bestHo = open, bi = -1, bj = -1;
        const unsigned *ap = reinterpret_cast<const unsigned *>(a_pk + G_.a_off) + lane;  // rows 4B+1..4B+4 at ap[B*32]
        const uint4 *bp = reinterpret_cast<const uint4 *>(b_pk + G_.b_off) + lane * 2;    // pass q at bp[q*64]
        const int nblk = (G_.la_max + 1 + 3) >> 2;      // row steps run in blocks of 4

        for (int q = 0; q < G_.n_pass; ++q) {
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
            const bool first = (q == 0), last = (q == G_.n_pass - 1);
            unsigned HoD = open2;                        // diag of the left boundary, row 0
            unsigned r_prev = (unsigned)(nl - 1);        // PAD row for the high lane's "row 0"
            unsigned lo15_Ho = open2, lo15_E = open2;    // low lane's column 15 at the previous row (low s16 used)
            const int jbase = q * K1_STRIP;
            // Improvement test: a half of the packed row maximum beats thr.  In passes q > 0 a tie with the best can
            // still win (smaller row), so ties enter the exact path as well.
            unsigned thr2 = pack2(bestHo - (first ? 0 : 1), bestHo - (first ? 0 : 1));
            const uint4 open4 = make_uint4(open2, open2, open2, open2);
            uint4 bq_next = first ? open4 : ld_keep(mybnd, pol);
            unsigned word_next = ap[0];
            unsigned c1 = open2, c2 = open2, c3 = open2; // boundary words of rows i0-3..i0-1 waiting for their 4th

#pragma unroll kBlkUnroll
            for (int blk = 0; blk < nblk; ++blk) {
                const int i0 = blk * 4;                           // rows i0+1 .. i0+4 (low lane)
                const unsigned word = word_next;
                const uint4 bq4 = bq_next;
                word_next = ap[(blk + 1) * 32];                   // a_pk / bnd have one block of slack
                if (!first) bq_next = ld_keep(mybnd + (size_t)(blk + 1) * bstride, pol);
                const unsigned bq[4] = {bq4.x, bq4.y, bq4.z, bq4.w};
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
                    K1_CELL(0) K1_CELL(1) K1_CELL(2) K1_CELL(3) K1_CELL(4) K1_CELL(5) K1_CELL(6) K1_CELL(7)
                    K1_CELL(8) K1_CELL(9) K1_CELL(10) K1_CELL(11) K1_CELL(12) K1_CELL(13) K1_CELL(14) K1_CELL(15)
                    // pass boundary: high lanes of (Ho[15], E) are column 32q+31 at row i-1
                    wout[t] = __byte_perm(Ho[15], E, 0x7632);
                    if (t == 0 && !last && blk > 0) st_keep(mybnd + (size_t)(blk - 1) * bstride, make_uint4(c1, c2, c3, wout[0]), pol);
                    lo15_Ho = Ho[15]; lo15_E = E;
                    HoD = HoD_next;
                    // packed row maximum of Ho (low lane: row i, high lane: row i-1)
                    // (thr2 rides in the chain: 8 max3 for 17 operands, and "rm != thr2" is the improvement test)
                    unsigned rm = __vimax3_s16x2(thr2, Ho[0], Ho[1]);
                    rm = __vimax3_s16x2(rm, Ho[2], Ho[3]);
                    rm = __vimax3_s16x2(rm, Ho[4], Ho[5]);
                    rm = __vimax3_s16x2(rm, Ho[6], Ho[7]);
                    rm = __vimax3_s16x2(rm, Ho[8], Ho[9]);
                    rm = __vimax3_s16x2(rm, Ho[10], Ho[11]);
                    rm = __vimax3_s16x2(rm, Ho[12], Ho[13]);
                    rm = __vimax3_s16x2(rm, Ho[14], Ho[15]);
                    // (the predicate outputs of __vibmax_s16x2 gave wrong answers here with nvcc 12.9; compare the max instead)
#if K1_NOTRACK == 1
                    thr2 = rm;      // timing experiment only: no end cell
#else
#if K1_VOTE
                    if (__any_sync(0xffffffffu, rm != thr2)) {
#else
                    if (rm != thr2) {
#endif
                        // Rare path (the running best of some lane improves).  Only the row and the half are decided
                        // here; the 16 packed Ho words of the step are parked in shared memory (16 STS of the lanes
                        // that improved; scalar stores: vector stores made the compiler shuffle Ho into aligned quads in the hot path) and the column is read off them once, after the last pass.
                        const int rlo = lo16(rm), rhi = hi16(rm);
                        const bool ulo = i <= la && (rlo > bestHo || (rlo == bestHo && rlo > open && i < bi));
                        if (ulo) { bestHo = rlo; bi = i; bj = jbase; }
                        const bool uhi = i >= 2 && i - 1 <= la && (rhi > bestHo || (rhi == bestHo && rhi > open && i - 1 < bi));
                        if (uhi) { bestHo = rhi; bi = i - 1; bj = jbase + K1_HALF; }
                        if (ulo || uhi) {
#pragma unroll
                            for (int k = 0; k < K1_HALF; ++k) snap[k * K1_THREADS * 2] = Ho[k];
                            thr2 = pack2(bestHo - (first ? 0 : 1), bestHo - (first ? 0 : 1));
                        }
                    }
#endif
                }
                c1 = wout[1]; c2 = wout[2]; c3 = wout[3];
            }
            if (!last) st_keep(mybnd + (size_t)(nblk - 1) * bstride, make_uint4(c1, c2, c3, open2), pol);   // rows 4(nblk-1)+1..+3
#if K1_NOTRACK
            bestHo = max(bestHo, max(lo16(thr2), hi16(thr2)));
#endif
        }
        if (bestHo > open) {                        // column of the best cell: first word of its half that holds bestHo
            const bool hi = (bj & K1_HALF) != 0;    // jbase is a multiple of 32
            int kk = 0;
#pragma unroll
            for (int k = K1_HALF - 1; k >= 0; --k)
                { const unsigned w = snap[k * K1_THREADS * 2]; if ((hi ? hi16(w) : lo16(w)) == bestHo) kk = k; }
            bj += kk + 1;
        }
        if (valid) {
            K1Result r;
            r.score = bestHo - open;
            r.end_i = (bestHo > open) ? bi : -1;
            r.end_j = (bestHo > open) ? bj : -1;
            r.flags = (bestHo >= limit_ho) ? 1 : 0;      // the running best is monotone: one overflow test per pair
            res[order[p]] = r;                      // scatter to the caller's order: the host pass over the results is linear
        }
    }
}

static int k1_bnd_blocks(int la_max) { return (la_max + 1 + 3) / 4 + 2; }

size_t k1_bnd_bytes(int n_blocks, int la_max)
{
    return (size_t)n_blocks * (K1_THREADS * 2 / 32) * (size_t)k1_bnd_blocks(la_max) * 32 * sizeof(uint4);
}

cudaError_t k1_launch(const swb_engine *e, const swb_batch *b, int n_blocks, void *d_bnd, cudaStream_t st)
{
    const int nl = e->n + 1;
    const size_t smem = (size_t)nl * nl * 32 * sizeof(unsigned) + (size_t)K1_THREADS * 2 * 4 * sizeof(uint4);
    cudaError_t err = cudaFuncSetAttribute(k1_local_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    err = cudaMemsetAsync(b->d_counter, 0, sizeof(int32_t), st);
    if (err != cudaSuccess) return err;
    // Ho must stay below 32767 - smax so that Ho_diag + (s - open) cannot wrap.
    const int limit_ho = 32767 - 2 * e->smax + e->o - 1;
    k1_local_kernel<<<n_blocks, K1_THREADS * 2, smem, st>>>(
        b->d_a, b->d_b, b->d_groups, b->n_groups, b->d_la, b->d_lb, e->d_mat, nl, e->o, e->e, limit_ho,
        reinterpret_cast<uint4 *>(d_bnd), k1_bnd_blocks(b->la_max), b->d_counter, b->d_order, b->d_res);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------------------------
// K2 -- the same packed recurrence with the LANES OF A WARP SPLITTING THE COLUMNS of a pair: database search with a
// long shared query (SA-main.rs:35-45 all-vs-all loop with one fixed `a`; per-pair semantics of align_partial(..,
// score_only), SA.rs:438-467).
//
// With one thread per pair (K1) a work item lasts la * ceil(lb/32) row steps: 5000 x 157 for a 5000-aa query against a
// 5000-aa subject, i.e. hundreds of milliseconds for one warp, which leaves most of the GPU (and most of a multi-GPU
// box) waiting for the longest subjects.  Here every lane owns ONE 32-column pass of some subject; the passes of a
// subject sit in consecutive lanes (a "segment") and run skewed by one 4-row block per lane, so that the last column
// of a pass reaches the next lane by four __shfl_up per row block instead of through memory.  Subjects of up to 32
// passes are bin-packed into warps by the host; every item then lasts (la + 4*31) row steps regardless of the subject
// lengths.  A subject of more than 32 passes takes a whole warp for several sweeps of 1024 columns, handing the last
// column of a sweep to lane 0 of the next one through a small global buffer.  Rows before a lane's start and after its
// end are PAD rows, which leave the DP state equivalent to the initial one (H = 0; E, F <= 0), so all lanes run the same
// instruction stream.
__global__ void __launch_bounds__(K1_THREADS * 2, 1)
k2_search_kernel(const unsigned *__restrict__ a_words, int la, const uint8_t *__restrict__ codes_b,
                 const int64_t *__restrict__ off_b, long long base_b, const int32_t *__restrict__ item_pairs,
                 const int32_t *__restrict__ item_sweeps, int n_items, const int32_t *__restrict__ mat, int nl, int open,
                 int ext, int limit_ho, uint4 *__restrict__ gbnd, int gb_blocks, int *counter, K1Result *__restrict__ res)
{
    extern __shared__ __align__(16) unsigned T_sh[];               // [nl*nl][32 lanes]: (s - open) & 0xFFFF
    __shared__ int cand_sh[K1_THREADS * 2 / 32][32][4];
    for (int idx = threadIdx.x; idx < nl * nl * 32; idx += blockDim.x)
        T_sh[idx] = (unsigned)(mat[idx >> 5] - open) & 0xFFFFu;
    __syncthreads();

    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int wslot = blockIdx.x * (blockDim.x >> 5) + wib;
    uint4 *mybnd = gbnd + (size_t)wslot * (size_t)gb_blocks;          // one uint4 per row block (a single lane uses it)
    const unsigned tbase = (unsigned)__cvta_generic_to_shared(T_sh) + lane * 4;
    const unsigned rowstride = (unsigned)nl * 128u;
    const unsigned open2 = pack2(open, open), ext2 = pack2(ext, ext);
    const unsigned pad = (unsigned)(nl - 1), pad4 = pad * 0x01010101u;
    uint4 *snap = reinterpret_cast<uint4 *>(T_sh + nl * nl * 32) + threadIdx.x;    // [4][thread] uint4 (K2 keeps the vector form: measured faster here)
    const int nblk = (la + 1 + 3) >> 2;              // row blocks of a pass (one step past la finishes the high half)
    const int nwords = (la + 3) >> 2;                // query words that hold real rows

    for (;;) {
        int g = 0;
        if (lane == 0) g = atomicAdd(counter, 1);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= n_items) break;
        const int pair = item_pairs[(size_t)g * 32 + lane];           // caller index of this lane's subject, -1 = idle
        const int sweeps = item_sweeps[g];
        const int prev_pair = __shfl_up_sync(0xffffffffu, pair, 1);
        const bool head = (lane == 0) || (pair != prev_pair);
        const unsigned heads = __ballot_sync(0xffffffffu, head);
        const int head_lane = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));
        const int pos = lane - head_lane;                             // pass index of this lane within its subject
        const unsigned next_heads = heads & ~((2u << lane) - 1u);     // heads above this lane
        const int seg_end = next_heads ? (__ffs(next_heads) - 1) : 32;   // one past the last lane of this segment
        long long ob = 0; int lb = 0;
        if (pair >= 0) { ob = off_b[pair] - base_b; lb = (int)(off_b[pair + 1] - off_b[pair]); }
        int pmax = pos;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) pmax = max(pmax, __shfl_xor_sync(0xffffffffu, pmax, d));
        const int n_gb = nblk + pmax;                                 // global row blocks: every lane finishes its rows
        int bestHo = open, bi = -1, bj = -1;
        int tb = open;                                                // best of the SUBJECT known to this lane: the exact path is
                                                                      // entered only by rows that reach it (ties included)

        for (int sw = 0; sw < sweeps; ++sw) {
            const int jbase = (sw * 32 + pos) * K1_STRIP;
            unsigned cb_lo[K1_HALF], cb_hi[K1_HALF], Ho[K1_HALF], F[K1_HALF];
#pragma unroll
            for (int k = 0; k < K1_HALF; ++k) {
                const int j1 = jbase + k, j2 = jbase + K1_HALF + k;
                const unsigned c1 = (j1 < lb) ? (unsigned)codes_b[ob + j1] : pad;
                const unsigned c2 = (j2 < lb) ? (unsigned)codes_b[ob + j2] : pad;
                cb_lo[k] = tbase + c1 * 128u; cb_hi[k] = tbase + c2 * 128u;
                Ho[k] = open2; F[k] = 0u;
            }
            const bool first = (pos == 0 && sw == 0);                 // left boundary = column 0: H = 0
            const bool from_g = (pos == 0 && sw > 0);                 // left boundary = last column of the previous sweep
            const bool to_g = (lane == 31 && sw + 1 < sweeps);
            unsigned HoD = open2, r_prev = pad, lo15_Ho = open2, lo15_E = open2;
            unsigned thr2 = pack2(tb - 1, tb - 1);
            unsigned c1 = open2, c2 = open2, c3 = open2;              // this lane's boundary words of the previous block
            int bl = -pos;                                            // this lane's row block
            unsigned word_next = (bl >= 0 && bl < nwords) ? a_words[bl] : pad4;
            uint4 g_next = make_uint4(open2, open2, open2, open2);
            if (from_g) g_next = mybnd[0];

#pragma unroll 1
            for (int gbk = 0; gbk < n_gb; ++gbk, ++bl) {
                const int i0 = bl * 4;                            // low-half rows i0+1 .. i0+4 (negative before the start)
                const unsigned word = word_next;
                const uint4 g4 = g_next;
                word_next = (bl + 1 >= 0 && bl + 1 < nwords) ? a_words[bl + 1] : pad4;
                if (from_g && bl + 1 < nblk) g_next = mybnd[bl + 1];
                // rows i0+1..i0+3 of the left boundary: the left lane's words of ITS previous block (it runs one block ahead)
                const unsigned s1 = __shfl_up_sync(0xffffffffu, c1, 1), s2 = __shfl_up_sync(0xffffffffu, c2, 1),
                               s3 = __shfl_up_sync(0xffffffffu, c3, 1);
                unsigned bq[4];
                bq[0] = first ? open2 : (from_g ? g4.x : s1);
                bq[1] = first ? open2 : (from_g ? g4.y : s2);
                bq[2] = first ? open2 : (from_g ? g4.z : s3);
                bq[3] = open2;
                unsigned wout[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int i = i0 + t + 1;                    // low lane row; high lane row = i-1
                    const unsigned r = (word >> (t * 8)) & 0xFFu;
                    const unsigned roff_lo = r * rowstride, roff_hi = r_prev * rowstride;
                    r_prev = r;
                    unsigned GoP = __byte_perm(bq[t], lo15_Ho, 0x5410);
                    unsigned E = __byte_perm(bq[t], lo15_E, 0x5432);
                    const unsigned HoD_next = GoP;
                    K1_CELL(0) K1_CELL(1) K1_CELL(2) K1_CELL(3) K1_CELL(4) K1_CELL(5) K1_CELL(6) K1_CELL(7)
                    K1_CELL(8) K1_CELL(9) K1_CELL(10) K1_CELL(11) K1_CELL(12) K1_CELL(13) K1_CELL(14) K1_CELL(15)
                    wout[t] = __byte_perm(Ho[15], E, 0x7632);    // column 32q+31 at row i-1
                    if (t == 0) {
                        // row i0+4 of the left boundary is the left lane's first word of THIS block
                        const unsigned s0 = __shfl_up_sync(0xffffffffu, wout[0], 1);
                        bq[3] = first ? open2 : (from_g ? g4.w : s0);
                        if (to_g && bl > 0) mybnd[bl - 1] = make_uint4(c1, c2, c3, wout[0]);
                    }
                    lo15_Ho = Ho[15]; lo15_E = E;
                    HoD = HoD_next;
                    unsigned rm = __vimax3_s16x2(thr2, Ho[0], Ho[1]);
                    rm = __vimax3_s16x2(rm, Ho[2], Ho[3]);
                    rm = __vimax3_s16x2(rm, Ho[4], Ho[5]);
                    rm = __vimax3_s16x2(rm, Ho[6], Ho[7]);
                    rm = __vimax3_s16x2(rm, Ho[8], Ho[9]);
                    rm = __vimax3_s16x2(rm, Ho[10], Ho[11]);
                    rm = __vimax3_s16x2(rm, Ho[12], Ho[13]);
                    rm = __vimax3_s16x2(rm, Ho[14], Ho[15]);
                    if (rm != thr2) {                           // rare path as in K1: row + half here, column from the parked words
                        const int rlo = lo16(rm), rhi = hi16(rm);
                        const bool ulo = i >= 1 && i <= la && (rlo > bestHo || (rlo == bestHo && rlo > open && i < bi));
                        if (ulo) { bestHo = rlo; bi = i; bj = jbase; }
                        const bool uhi = i >= 2 && i - 1 <= la && (rhi > bestHo || (rhi == bestHo && rhi > open && i - 1 < bi));
                        if (uhi) { bestHo = rhi; bi = i - 1; bj = jbase + K1_HALF; }
                        if (ulo || uhi) {
                            snap[0] = make_uint4(Ho[0], Ho[1], Ho[2], Ho[3]);
                            snap[K1_THREADS * 2] = make_uint4(Ho[4], Ho[5], Ho[6], Ho[7]);
                            snap[K1_THREADS * 4] = make_uint4(Ho[8], Ho[9], Ho[10], Ho[11]);
                            snap[K1_THREADS * 6] = make_uint4(Ho[12], Ho[13], Ho[14], Ho[15]);
                        }
                        tb = max(tb, bestHo);
                        thr2 = pack2(tb - 1, tb - 1);
                    }
                }
                c1 = wout[1]; c2 = wout[2]; c3 = wout[3];
                {   // the subject's best diffuses along the segment, one lane per row block in both directions
                    const int up = __shfl_up_sync(0xffffffffu, tb, 1), dn = __shfl_down_sync(0xffffffffu, tb, 1);
                    int nb = tb;
                    if (!head) nb = max(nb, up);
                    if (lane + 1 < seg_end) nb = max(nb, dn);
                    if (nb != tb) { tb = nb; thr2 = pack2(tb - 1, tb - 1); }
                }
            }
            if (to_g) mybnd[nblk - 1] = make_uint4(c1, c2, c3, open2);
            __syncwarp();
        }

        if (bestHo > open) {                        // column of this lane's best cell (see K1)
            const uint4 s0 = snap[0], s1 = snap[K1_THREADS * 2], s2 = snap[K1_THREADS * 4], s3 = snap[K1_THREADS * 6];
            const unsigned w[K1_HALF] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w, s2.x, s2.y, s2.z, s2.w, s3.x, s3.y, s3.z, s3.w};
            const bool hi = (bj & K1_HALF) != 0;
            int kk = 0;
#pragma unroll
            for (int k = K1_HALF - 1; k >= 0; --k)
                if ((hi ? hi16(w[k]) : lo16(w[k])) == bestHo) kk = k;
            bj += kk + 1;
        }
        // per-lane candidates -> the segment's head lane: score desc, row asc, column asc (SA.rs:286-294 scan order)
        cand_sh[wib][lane][0] = bestHo; cand_sh[wib][lane][1] = bi; cand_sh[wib][lane][2] = bj; cand_sh[wib][lane][3] = (bestHo >= limit_ho) ? 1 : 0;
        __syncwarp();
        if (head && pair >= 0) {
            int ovf = (bestHo >= limit_ho) ? 1 : 0;
            for (int l2 = lane + 1; l2 < seg_end; ++l2) {
                const int ob_ = cand_sh[wib][l2][0], oi = cand_sh[wib][l2][1], oj = cand_sh[wib][l2][2];
                ovf |= cand_sh[wib][l2][3];
                if (ob_ > bestHo || (ob_ == bestHo && ob_ > open && (oi < bi || (oi == bi && oj < bj)))) { bestHo = ob_; bi = oi; bj = oj; }
            }
            K1Result r;
            r.score = bestHo - open;
            r.end_i = (bestHo > open) ? bi : -1;
            r.end_j = (bestHo > open) ? bj : -1;
            r.flags = ovf;
            res[pair] = r;
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------------------------
// K2P -- K2 with a QUERY PROFILE.  The query is the same for every subject of a search, so the substitution scores are
// laid out once per block as bytes  QP[column residue c][row x] = s(query[x], c) - open  (0 <= x < la; 0 in the margins
// and for the PAD residue, i.e. s = open there: a PAD row keeps the initial state, a PAD column only ever holds
// (real H) + open < best).  A cell-vector then fetches its two scores with two LDS.U8 whose row offset is an IMMEDIATE
// (t = 0..3 inside a 4-row block; the per-column pointers advance by 4 once per block) and merges them with one IMAD:
// no row-residue extraction, no per-cell address arithmetic -- 9 issue slots per cell-vector instead of 14.  The
// recurrence, the lane skew, the boundary exchange and the end-cell logic are K2's.  Needs every s - open in [0, 127]
// (|open| >= |smin|: BLOSUM62 with 10/0.5 gives 12..42) and a query whose profile fits shared memory; k2_launch falls
// back to K2 otherwise.
template <int IMM>
__device__ __forceinline__ unsigned ldsu8i(unsigned addr)
{
    unsigned v;
    asm("ld.shared.u8 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(IMM));
    return v;
}

#define K2P_MARGIN 128                 // profile bytes before row 1: lane l starts l row blocks early (the high half one row earlier still)
#define K2P_TAIL 144                   // and after row la: 31 blocks of skew + the block that finishes the high half

#define K2P_CELL(k, T)                                                                               \
    {                                                                                                \
        unsigned s_lo = ldsu8i<T>(plo[k]);                                                           \
        unsigned s_hi = ldsu8i<T>(phi[k]);                                                           \
        unsigned s2 = s_hi * 65536u + s_lo;                                                          \
        unsigned Fk = __viaddmax_s16x2(F[k], ext2, Ho[k]);                                           \
        unsigned G = __viaddmax_s16x2_relu(HoD, s2, Fk);                                             \
        E = __viaddmax_s16x2(E, ext2, GoP);                                                          \
        GoP = __vadd2(G, open2);                                                                     \
        HoD = Ho[k];                                                                                 \
        Ho[k] = __viaddmax_s16x2(E, open2, GoP);                                                     \
        F[k] = Fk;                                                                                   \
    }

#define K2P_ROW(T)                                                                                                     \
    {                                                                                                                  \
        const int i = i0 + T + 1;                    /* low lane row; high lane row = i-1 */                           \
        unsigned GoP = __byte_perm(bq[T], lo15_Ho, 0x5410);                                                            \
        unsigned E = __byte_perm(bq[T], lo15_E, 0x5432);                                                               \
        const unsigned HoD_next = GoP;                                                                                 \
        K2P_CELL(0, T) K2P_CELL(1, T) K2P_CELL(2, T) K2P_CELL(3, T) K2P_CELL(4, T) K2P_CELL(5, T) K2P_CELL(6, T) K2P_CELL(7, T) \
        K2P_CELL(8, T) K2P_CELL(9, T) K2P_CELL(10, T) K2P_CELL(11, T) K2P_CELL(12, T) K2P_CELL(13, T) K2P_CELL(14, T) K2P_CELL(15, T) \
        wout[T] = __byte_perm(Ho[15], E, 0x7632);    /* column 32q+31 at row i-1 */                                    \
        if (T == 0) {                                                                                                  \
            const unsigned s0 = __shfl_up_sync(0xffffffffu, wout[0], 1);                                               \
            bq[3] = first ? open2 : (from_g ? g4.w : s0);                                                              \
            if (to_g && bl > 0) mybnd[bl - 1] = make_uint4(c1, c2, c3, wout[0]);                                       \
        }                                                                                                              \
        lo15_Ho = Ho[15]; lo15_E = E;                                                                                  \
        HoD = HoD_next;                                                                                                \
        unsigned rm = __vimax3_s16x2(thr2, Ho[0], Ho[1]);                                                              \
        rm = __vimax3_s16x2(rm, Ho[2], Ho[3]);                                                                         \
        rm = __vimax3_s16x2(rm, Ho[4], Ho[5]);                                                                         \
        rm = __vimax3_s16x2(rm, Ho[6], Ho[7]);                                                                         \
        rm = __vimax3_s16x2(rm, Ho[8], Ho[9]);                                                                         \
        rm = __vimax3_s16x2(rm, Ho[10], Ho[11]);                                                                       \
        rm = __vimax3_s16x2(rm, Ho[12], Ho[13]);                                                                       \
        rm = __vimax3_s16x2(rm, Ho[14], Ho[15]);                                                                       \
        if (rm != thr2) {                            /* rare path as in K2 */                                          \
            const int rlo = lo16(rm), rhi = hi16(rm);                                                                  \
            const bool ulo = i >= 1 && i <= la && (rlo > bestHo || (rlo == bestHo && rlo > open && i < bi));           \
            if (ulo) { bestHo = rlo; bi = i; bj = jbase; }                                                             \
            const bool uhi = i >= 2 && i - 1 <= la && (rhi > bestHo || (rhi == bestHo && rhi > open && i - 1 < bi));   \
            if (uhi) { bestHo = rhi; bi = i - 1; bj = jbase + K1_HALF; }                                               \
            if (ulo || uhi) {                                                                                          \
                snap[0] = make_uint4(Ho[0], Ho[1], Ho[2], Ho[3]);                                                      \
                snap[K1_THREADS * 2] = make_uint4(Ho[4], Ho[5], Ho[6], Ho[7]);                                         \
                snap[K1_THREADS * 4] = make_uint4(Ho[8], Ho[9], Ho[10], Ho[11]);                                       \
                snap[K1_THREADS * 6] = make_uint4(Ho[12], Ho[13], Ho[14], Ho[15]);                                     \
            }                                                                                                          \
            tb = max(tb, bestHo);                                                                                      \
            thr2 = pack2(tb - 1, tb - 1);                                                                              \
        }                                                                                                              \
    }

__global__ void __launch_bounds__(K1_THREADS * 2, 1)
k2p_search_kernel(const uint8_t *__restrict__ a_codes, int la, const uint8_t *__restrict__ codes_b,
                  const int64_t *__restrict__ off_b, long long base_b, const int32_t *__restrict__ item_pairs,
                  const int32_t *__restrict__ item_sweeps, int n_items, const int32_t *__restrict__ mat, int nl, int open,
                  int ext, int limit_ho, int pitch, uint4 *__restrict__ gbnd, int gb_blocks, int *counter, K1Result *__restrict__ res,
                  unsigned one)
{
    extern __shared__ __align__(16) unsigned char QP_sh[];          // [nl][pitch] bytes, then the snapshot slots
    __shared__ int cand_sh[K1_THREADS * 2 / 32][32][4];
    for (int idx = threadIdx.x; idx < nl * pitch; idx += blockDim.x) {
        const int c = idx / pitch, x = idx - c * pitch - K2P_MARGIN;
        int v = 0;                                                   // margins and the PAD column: s = open
        if (x >= 0 && x < la && c != nl - 1) v = mat[(int)a_codes[x] * nl + c] - open;
        QP_sh[idx] = (unsigned char)v;
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int wslot = blockIdx.x * (blockDim.x >> 5) + wib;
    uint4 *mybnd = gbnd + (size_t)wslot * (size_t)gb_blocks;
    const unsigned qbase = (unsigned)__cvta_generic_to_shared(QP_sh) + K2P_MARGIN;
    const unsigned open2 = pack2(open, open), ext2 = pack2(ext, ext);
    const unsigned pad = (unsigned)(nl - 1);
    uint4 *snap = reinterpret_cast<uint4 *>(QP_sh + (((size_t)nl * pitch + 15) & ~(size_t)15)) + threadIdx.x;   // [4][thread] uint4
    const int nblk = (la + 1 + 3) >> 2;              // row blocks of a pass (one step past la finishes the high half)

    for (;;) {
        int g = 0;
        if (lane == 0) g = atomicAdd(counter, 1);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= n_items) break;
        const int pair = item_pairs[(size_t)g * 32 + lane];           // caller index of this lane's subject, -1 = idle
        const int sweeps = item_sweeps[g];
        const int prev_pair = __shfl_up_sync(0xffffffffu, pair, 1);
        const bool head = (lane == 0) || (pair != prev_pair);
        const unsigned heads = __ballot_sync(0xffffffffu, head);
        const int head_lane = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));
        const int pos = lane - head_lane;                             // pass index of this lane within its subject
        const unsigned next_heads = heads & ~((2u << lane) - 1u);     // heads above this lane
        const int seg_end = next_heads ? (__ffs(next_heads) - 1) : 32;
        long long ob = 0; int lb = 0;
        if (pair >= 0) { ob = off_b[pair] - base_b; lb = (int)(off_b[pair + 1] - off_b[pair]); }
        // Lanes are skewed by their LANE index (K2: by the position inside the subject): lane l is at row block g - l, so the 32
        // profile words of one LDS lie in 32 different banks whatever the residues are (the profile rows of all residues
        // start in the same bank: pitch is a multiple of 128 bytes).  A subject that starts at lane h idles h blocks first.
        const int last_lane = 31 - __clz(__ballot_sync(0xffffffffu, pair >= 0) | 1u);
        const int n_gb = nblk + last_lane;
        int bestHo = open, bi = -1, bj = -1;
        int tb = open;

        for (int sw = 0; sw < sweeps; ++sw) {
            const int jbase = (sw * 32 + pos) * K1_STRIP;
            unsigned plo[K1_HALF], phi[K1_HALF], Ho[K1_HALF], F[K1_HALF];
            // profile pointers of this lane's 32 columns at its first row block (block -pos: rows -4 pos + 1 ..); the high
            // half is one row behind
            const unsigned rowoff = qbase + (unsigned)(-4 * lane);
#pragma unroll
            for (int k = 0; k < K1_HALF; ++k) {
                const int j1 = jbase + k, j2 = jbase + K1_HALF + k;
                const unsigned c1 = (j1 < lb) ? (unsigned)codes_b[ob + j1] : pad;
                const unsigned c2 = (j2 < lb) ? (unsigned)codes_b[ob + j2] : pad;
                plo[k] = rowoff + c1 * (unsigned)pitch; phi[k] = rowoff + c2 * (unsigned)pitch - 1u;
                Ho[k] = open2; F[k] = 0u;
            }
            const bool first = (pos == 0 && sw == 0);
            const bool from_g = (pos == 0 && sw > 0);
            const bool to_g = (lane == 31 && sw + 1 < sweeps);
            unsigned HoD = open2, lo15_Ho = open2, lo15_E = open2;
            unsigned thr2 = pack2(tb - 1, tb - 1);
            unsigned c1 = open2, c2 = open2, c3 = open2;
            int bl = -lane;
            uint4 g_next = make_uint4(open2, open2, open2, open2);
            if (from_g) g_next = mybnd[0];

#pragma unroll 1
            for (int gbk = 0; gbk < n_gb; ++gbk, ++bl) {
                const int i0 = bl * 4;
                const uint4 g4 = g_next;
                if (from_g && bl + 1 < nblk) g_next = mybnd[bl + 1];
                const unsigned s1 = __shfl_up_sync(0xffffffffu, c1, 1), s2_ = __shfl_up_sync(0xffffffffu, c2, 1),
                               s3 = __shfl_up_sync(0xffffffffu, c3, 1);
                unsigned bq[4];
                bq[0] = first ? open2 : (from_g ? g4.x : s1);
                bq[1] = first ? open2 : (from_g ? g4.y : s2_);
                bq[2] = first ? open2 : (from_g ? g4.z : s3);
                bq[3] = open2;
                unsigned wout[4];
                K2P_ROW(0) K2P_ROW(1) K2P_ROW(2) K2P_ROW(3)
                // the 32 pointer bumps go to the FMA pipe (x * one + 4 with `one` a kernel argument the compiler cannot fold): the
                // ALU pipe, which the DPX instructions saturate, is the kernel's limit, the FMA pipe is 78 % idle
#pragma unroll
                for (int k = 0; k < K1_HALF; ++k) {
                    asm("mad.lo.u32 %0, %0, %1, 4;" : "+r"(plo[k]) : "r"(one));
                    asm("mad.lo.u32 %0, %0, %1, 4;" : "+r"(phi[k]) : "r"(one));
                }
                c1 = wout[1]; c2 = wout[2]; c3 = wout[3];
                {
                    const int up = __shfl_up_sync(0xffffffffu, tb, 1), dn = __shfl_down_sync(0xffffffffu, tb, 1);
                    int nb = tb;
                    if (!head) nb = max(nb, up);
                    if (lane + 1 < seg_end) nb = max(nb, dn);
                    if (nb != tb) { tb = nb; thr2 = pack2(tb - 1, tb - 1); }
                }
            }
            if (to_g) mybnd[nblk - 1] = make_uint4(c1, c2, c3, open2);
            __syncwarp();
        }

        if (bestHo > open) {
            const uint4 s0 = snap[0], s1 = snap[K1_THREADS * 2], s2 = snap[K1_THREADS * 4], s3 = snap[K1_THREADS * 6];
            const unsigned w[K1_HALF] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w, s2.x, s2.y, s2.z, s2.w, s3.x, s3.y, s3.z, s3.w};
            const bool hi = (bj & K1_HALF) != 0;
            int kk = 0;
#pragma unroll
            for (int k = K1_HALF - 1; k >= 0; --k)
                if ((hi ? hi16(w[k]) : lo16(w[k])) == bestHo) kk = k;
            bj += kk + 1;
        }
        cand_sh[wib][lane][0] = bestHo; cand_sh[wib][lane][1] = bi; cand_sh[wib][lane][2] = bj; cand_sh[wib][lane][3] = (bestHo >= limit_ho) ? 1 : 0;
        __syncwarp();
        if (head && pair >= 0) {
            int ovf = (bestHo >= limit_ho) ? 1 : 0;
            for (int l2 = lane + 1; l2 < seg_end; ++l2) {
                const int ob_ = cand_sh[wib][l2][0], oi = cand_sh[wib][l2][1], oj = cand_sh[wib][l2][2];
                ovf |= cand_sh[wib][l2][3];
                if (ob_ > bestHo || (ob_ == bestHo && ob_ > open && (oi < bi || (oi == bi && oj < bj)))) { bestHo = ob_; bi = oi; bj = oj; }
            }
            K1Result r;
            r.score = bestHo - open;
            r.end_i = (bestHo > open) ? bi : -1;
            r.end_j = (bestHo > open) ? bj : -1;
            r.flags = ovf;
            res[pair] = r;
        }
        __syncwarp();
    }
}

// Profile form available?  Every s - open must be a non-negative byte, the profile must fit next to the snapshot slots.
static int k2p_pitch(const swb_engine *e, int la)
{
    if (getenv("SWB_NO_K2P")) return 0;
    if (e->smin - e->o < 0 || e->smax - e->o > 127) return 0;
    int pitch = K2P_MARGIN + la + K2P_TAIL;
    pitch = (pitch + 127) / 128 * 128;                             // every residue's row starts in bank 0 (see the lane skew in the kernel)
    const int nl = e->n + 1;
    const size_t smem = (((size_t)nl * pitch + 15) & ~(size_t)15) + (size_t)K1_THREADS * 2 * 4 * sizeof(uint4);
    return smem <= 220 * 1024 ? pitch : 0;
}

size_t k2_bnd_bytes(int n_blocks, int la) { return (size_t)n_blocks * (K1_THREADS * 2 / 32) * (size_t)((la + 1 + 3) / 4 + 2) * sizeof(uint4); }

cudaError_t k2_launch(const swb_engine *e, const uint8_t *d_query, int la, const uint8_t *d_codes_b, const int64_t *d_off_b,
                      int64_t base_b, const int32_t *d_item_pairs, const int32_t *d_item_sweeps, int n_items, void *d_gbnd,
                      int *d_counter, K1Result *d_res, int n_blocks, cudaStream_t st)
{
    const int nl = e->n + 1;
    const size_t smem = (size_t)nl * nl * 32 * sizeof(unsigned) + (size_t)K1_THREADS * 2 * 4 * sizeof(uint4);
    cudaError_t err = cudaFuncSetAttribute(k2_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    err = cudaMemsetAsync(d_counter, 0, sizeof(int32_t), st);
    if (err != cudaSuccess) return err;
    const int limit_ho = 32767 - 2 * e->smax + e->o - 1;
    if (const int pitch = k2p_pitch(e, la)) {
        const size_t psmem = (((size_t)nl * pitch + 15) & ~(size_t)15) + (size_t)K1_THREADS * 2 * 4 * sizeof(uint4);
        err = cudaFuncSetAttribute(k2p_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psmem);
        if (err != cudaSuccess) return err;
        k2p_search_kernel<<<n_blocks, K1_THREADS * 2, psmem, st>>>(
            d_query, la, d_codes_b, d_off_b, (long long)base_b, d_item_pairs, d_item_sweeps, n_items, e->d_mat, nl, e->o, e->e, limit_ho,
            pitch, reinterpret_cast<uint4 *>(d_gbnd), (la + 1 + 3) / 4 + 2, d_counter, d_res, 1u);
        return cudaGetLastError();
    }
    k2_search_kernel<<<n_blocks, K1_THREADS * 2, smem, st>>>(
        reinterpret_cast<const unsigned *>(d_query), la, d_codes_b, d_off_b, (long long)base_b, d_item_pairs, d_item_sweeps,
        n_items, e->d_mat, nl, e->o, e->e, limit_ho, reinterpret_cast<uint4 *>(d_gbnd), (la + 1 + 3) / 4 + 2, d_counter, d_res);
    return cudaGetLastError();
}
