// K3 + KT3 -- the packed traceback path: score-speed fill with sparse tile borders, traceback by recomputing only the
// tiles the path crosses.
//
// Replaces fill_matrix + backtrack of the reference (SA.rs:55-387) for pairs whose scores provably fit s16 lanes and
// whose penalties satisfy |open| >= |extend| (so that the two-value Gotoh form equals the three-state form in VALUE,
// SURVEY.md 8a note 3).  The direction-byte path (kw_trace.cu) pays ~42 int32 instructions and one byte of HBM per
// cell; here
//   * K3 fills the matrix with the packed s16x2 cell of K1/K2 (5.5 DPX-pipe operations per TWO cells) and stores only
//     the cells on a grid: the last column of every 32-column pass (H, E, M, F: 8 bytes per row) and, every 64 rows,
//     three snapshots of the pass's H registers and two of F (320 bytes per pass);
//   * KT3 walks the path tile by tile (32 columns x <= 64 rows): a warp re-computes the part of the tile above-left of
//     the cursor with the EXACT three-state int32 cell (kw_cell.cuh: the reference's five priority orders, clamp after
//     the choice) from the stored border values -- M(r,j) = H(r-1,j-1) + s, X(r,j) by the row recurrence from the
//     stored E of the pass's left border, Y = F -- writes the predecessor codes to shared memory and follows them.
// Every border cell holds the value it has in the full matrix, so every recomputed cell and every tie decision is the
// reference's.  Local mode is given the end cell by the scan (K1/K2) and fills only the prefix box that ends there;
// global mode reads the start state off the recomputed cell (la, lb).  Glocal stays on the direction-byte path.
#include <algorithm>
#include <cstdlib>
#include <climits>

#include <cuda/atomic>
#include <cstdio>

#include "swb_internal.h"
#include "k1_common.cuh"
#include "kw_cell.cuh"

#define K3_R 64                       // rows between row checkpoints (a multiple of 4)
#define K3_ROWB_WORDS 80              // per (checkpoint, pass): Ho after steps t=1,2,3 (3 x 16 words) + F after t=2,3 (2 x 16)
#define K3_WARPS 12                   // 384 threads: up to 168 registers per thread, no spills in the global form

template <int MODE>
struct K3Cell {
    static __device__ __forceinline__ unsigned g(unsigned hod, unsigned s2, unsigned fk)
    {
        return MODE == SWB_LOCAL ? __viaddmax_s16x2_relu(hod, s2, fk) : __viaddmax_s16x2(hod, s2, fk);
    }
};

#define K3_CELL(k)                                                                                   \
    {                                                                                                \
        unsigned s_lo = lds32(cb_lo[k] + roff_lo);                                                   \
        unsigned s_hi = lds32(cb_hi[k] + roff_hi);                                                   \
        unsigned s2 = s_hi * 65536u + s_lo;                                                          \
        unsigned Fk = __viaddmax_s16x2(F[k], ext2, Ho[k]);                                           \
        unsigned G = K3Cell<MODE>::g(HoD, s2, Fk);                                                   \
        if (k == 15) m15 = __vadd2(HoD, s2);                                                         \
        E = __viaddmax_s16x2(E, ext2, GoP);                                                          \
        GoP = __vadd2(G, open2);                                                                     \
        HoD = Ho[k];                                                                                 \
        Ho[k] = __viaddmax_s16x2(E, open2, GoP);                                                     \
        F[k] = Fk;                                                                                   \
    }

// ---- K3: fill ---------------------------------------------------------------------------------------------------------
// A TEAM of `team` warps of one block = one pair; warp w takes the sweeps w, w + team, ... (a sweep = 32 passes = 1024
// columns); lane l of a sweep owns the 32-column pass q = 32 sw + l and runs l row blocks (of 4 rows) behind lane 0, so the
// last column of a pass reaches the next lane by shuffles (as in K2); the last lane's column reaches lane 0 of the next
// sweep -- another warp of the team, about 33 row blocks behind -- through the pair's own border store, which the walker
// reads anyway, guarded by a progress word in shared memory (published row blocks, monotone over the sweeps).  The warps
// of a team sit in one block, so they are resident together by construction: no inter-block waiting anywhere.
// (A lone warp runs this cell at an IPC of ~0.25 -- 119 registers leave no room to software-pipeline the table loads --
// so a pair needs several warps as soon as there are fewer pairs than warp slots.)
__device__ __forceinline__ void k3_team_bar(int slot, int team)
{
    if (team == 1) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(slot + 1), "r"(team * 32) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(K3_WARPS * 32, 1)
k3_fill_kernel(const uint8_t *__restrict__ a_all, const uint8_t *__restrict__ b_all, const K3Pair *__restrict__ pairs, int n_pairs,
               const int32_t *__restrict__ mat, int nl, int open, int ext, uint4 *colb, unsigned *rowb, int *counter, int team)
{
    extern __shared__ __align__(16) unsigned T_sh[];               // [nl*nl][32 lanes]: (s - open) & 0xFFFF
    __shared__ volatile int prog_sh[K3_WARPS];                     // per warp: sweep * (nblk + 2) + row blocks of its last pass published
    __shared__ int pair_sh[K3_WARPS];
    for (int idx = threadIdx.x; idx < nl * nl * 32; idx += blockDim.x)
        T_sh[idx] = (unsigned)(mat[idx >> 5] - open) & 0xFFFFu;
    __syncthreads();

    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int slot = wib / team, w = wib - slot * team;             // team slot of the block, warp within the team
    const int pw = slot * team + (w + team - 1) % team;             // the warp that owns the sweep before this warp's
    const unsigned tbase = (unsigned)__cvta_generic_to_shared(T_sh) + lane * 4;
    const unsigned rowstride = (unsigned)nl * 128u;
    const unsigned open2 = pack2(open, open), ext2 = pack2(ext, ext);
    const unsigned pad = (unsigned)(nl - 1);
    const int sent = 10 * open + 10 * ext;
    // Ho = H + open of boundary cell (0, j) / (j, 0): global 2 open + (j-1) ext (SA.rs:100-118), local 0 + open
    auto ho_bnd = [=](int j) { return (MODE == SWB_LOCAL || j == 0) ? open : 2 * open + (j - 1) * ext; };

    for (;;) {
        k3_team_bar(slot, team);                                   // every warp of the team is done with the previous pair
        if (lane == 0) prog_sh[wib] = -1;
        if (w == 0 && lane == 0) pair_sh[slot] = atomicAdd(counter, 1);
        k3_team_bar(slot, team);
        const int g = pair_sh[slot];
        if (g >= n_pairs) break;
        const K3Pair P = pairs[g];
        const uint8_t *a = P.swap ? b_all + P.b_off : a_all + P.a_off, *b = P.swap ? a_all + P.a_off : b_all + P.b_off;
        const int la = P.la, lb = P.lb;
        const int n_pass = (lb + K1_STRIP - 1) / K1_STRIP;
        const int nblk = (la + 1 + 3) >> 2;                          // row blocks: steps 1 .. la+1 (the high half trails by one row)
        const int sweeps = (n_pass + 31) >> 5;
        const int n_ck = la / K3_R;                                  // checkpoint rows R m - 1, m = 1 .. n_ck
        uint4 *cb = colb + P.colb_off;                               // [pass][nblk + 1][2]
        unsigned *rb = rowb + P.rowb_off;                            // [m - 1][pass][80]

        const int pbase = nblk + 2;
        for (int sw = w; sw < sweeps; sw += team) {
            const int q = sw * 32 + lane;
            const bool has_pass = q < n_pass;
            const int jbase = q * K1_STRIP;
            const int lanes_here = min(32, n_pass - sw * 32);
            const int n_gb = nblk + lanes_here - 1;
            unsigned cb_lo[K1_HALF], cb_hi[K1_HALF], Ho[K1_HALF], F[K1_HALF];
            const unsigned fdelta = pack2(sent - open, sent - open);
#pragma unroll
            for (int k = 0; k < K1_HALF; ++k) {
                const int j1 = jbase + k, j2 = jbase + K1_HALF + k;
                const unsigned c1 = (has_pass && j1 < lb) ? (unsigned)b[j1] : pad;
                const unsigned c2 = (has_pass && j2 < lb) ? (unsigned)b[j2] : pad;
                cb_lo[k] = tbase + c1 * 128u; cb_hi[k] = tbase + c2 * 128u;
                if (MODE == SWB_LOCAL) { Ho[k] = open2; F[k] = 0u; }
                else { Ho[k] = pack2(ho_bnd(j1 + 1), ho_bnd(j2 + 1)); F[k] = __vadd2(Ho[k], fdelta); }
            }
            const bool first = (q == 0);                              // left boundary = column 0 of the matrix
            const bool from_g = (lane == 0 && sw > 0);                // left boundary = the stored last column of pass q - 1
            unsigned HoD = MODE == SWB_LOCAL ? open2 : pack2(ho_bnd(jbase), open);
            unsigned r_prev = pad;
            unsigned lo15_Ho = MODE == SWB_LOCAL ? open2 : pack2(ho_bnd(jbase + K1_HALF), open), lo15_E = open2;
            unsigned c1 = open2, c2 = open2, c3 = open2;              // this lane's (Ho, E) words of rows 1..3 of its previous block
            const uint4 *gin = cb + ((size_t)(q > 0 ? q - 1 : 0) * (nblk + 1)) * 2;
            uint4 g_cur = make_uint4(open2, open2, open2, open2), g_next = g_cur;
            // lane 0 of a later sweep: wait until the row blocks it is about to read have been published by the sweep before
            auto wait_pub = [&](int need) {
                if (team > 1) {
                    const int tgt = (sw - 1) * pbase + min(need, nblk);
                    while (prog_sh[pw] < tgt) __nanosleep(20);
                    __threadfence_block();
                }
            };
            // (the wait is executed by the WHOLE warp -- its condition is warp-uniform -- and only the two loads are lane 0's:
            //  a spin inside lane-0-only control flow left the warp split into {lane 0} and {lanes 1..31} for the whole
            //  sweep, each half issuing every instruction of the loop: measured, 2 x the warp instructions)
            if (sw > 0) {
                wait_pub(2);
                if (lane == 0) { g_cur = __ldcg(gin); g_next = __ldcg(gin + 2); }
            }
            uint4 *gout = cb + ((size_t)q * (nblk + 1)) * 2;
            auto load_rows = [&](int blk) -> unsigned {                 // residues of rows 4 blk + 1 .. 4 blk + 4, PAD beyond la
                unsigned wd = 0;
                if (has_pass && blk >= 0 && blk < nblk) {
#pragma unroll
                    for (int t = 0; t < 4; ++t) wd |= ((4 * blk + t < la) ? (unsigned)a[4 * blk + t] : pad) << (8 * t);
                }
                return wd;
            };
            unsigned word_next = load_rows(-lane);

#pragma unroll 1
            for (int gbk = 0; gbk < n_gb; ++gbk) {
                const int bl = gbk - lane;                            // this lane's row block
                const bool active = has_pass && bl >= 0 && bl < nblk;
                const int i0 = bl * 4;
                // rows i0+1..i0+3 of the left boundary: the left lane's words 1..3 of ITS previous block (it runs one block ahead)
                const unsigned s1 = __shfl_up_sync(0xffffffffu, c1, 1), s2_ = __shfl_up_sync(0xffffffffu, c2, 1),
                               s3 = __shfl_up_sync(0xffffffffu, c3, 1);
                unsigned bq[4];
                if (first) {
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const int hb = ho_bnd(i0 + t + 1);
                        bq[t] = MODE == SWB_LOCAL ? open2 : pack2(hb, hb - open + sent);      // column 0 (SA.rs:109-118)
                    }
                } else if (from_g) { bq[0] = g_cur.y; bq[1] = g_cur.z; bq[2] = g_cur.w; bq[3] = g_next.x; }
                else { bq[0] = s1; bq[1] = s2_; bq[2] = s3; bq[3] = open2; }
                const unsigned word = word_next;                       // fetched one block ahead: no load latency on the block's critical path
                word_next = load_rows(bl + 1);
                const bool snap = active && ((bl + 1) % (K3_R / 4) == 0) && ((bl + 1) / (K3_R / 4) <= n_ck);
                unsigned *rdst = rb + ((size_t)(snap ? (bl + 1) / (K3_R / 4) - 1 : 0) * n_pass + q) * K3_ROWB_WORDS;
                unsigned wout[4] = {open2, open2, open2, open2}, mf[4] = {0u, 0u, 0u, 0u};
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    if (t == 1) {
                        // row i0+4 of the left boundary is the left lane's first word of THIS block
                        const unsigned s0 = __shfl_up_sync(0xffffffffu, wout[0], 1);
                        if (!first && !from_g) bq[3] = s0;
                    }
                    if (active) {
                        const unsigned r = (word >> (t * 8)) & 0xFFu;
                        const unsigned roff_lo = r * rowstride, roff_hi = r_prev * rowstride;
                        r_prev = r;
                        unsigned GoP = __byte_perm(bq[t], lo15_Ho, 0x5410);
                        unsigned E = __byte_perm(bq[t], lo15_E, 0x5432);
                        const unsigned HoD_next = GoP;
                        unsigned m15 = 0;
                        K3_CELL(0) K3_CELL(1) K3_CELL(2) K3_CELL(3) K3_CELL(4) K3_CELL(5) K3_CELL(6) K3_CELL(7)
                        K3_CELL(8) K3_CELL(9) K3_CELL(10) K3_CELL(11) K3_CELL(12) K3_CELL(13) K3_CELL(14) K3_CELL(15)
                        if (MODE != SWB_LOCAL && t == 0 && bl == 0) {
                            // the high half has just run its "row 0" step on PAD residues: put the boundary row there instead
#pragma unroll
                            for (int k = 0; k < K1_HALF; ++k) {
                                const int hb = ho_bnd(jbase + K1_HALF + k + 1);
                                Ho[k] = (Ho[k] & 0xFFFFu) | ((unsigned)hb << 16);
                                F[k] = (F[k] & 0xFFFFu) | ((unsigned)(hb - open + sent) << 16);
                            }
                            E = (E & 0xFFFFu) | ((unsigned)(ho_bnd(jbase + K1_STRIP) - open) << 16);   // X(0, jbase+32) = H(0, .)
                            m15 = (m15 & 0xFFFFu) | ((unsigned)(ho_bnd(jbase + K1_STRIP) - open + sent) << 16);
                        }
                        wout[t] = __byte_perm(Ho[15], E, 0x7632);        // (Ho, E) of column 32q+32 at row i0 + t
                        mf[t] = __byte_perm(m15, F[15], 0x7632);         // (M, F) of the same cell
                        lo15_Ho = Ho[15]; lo15_E = E;
                        HoD = HoD_next;
                        if (snap && t >= 1) {                            // row checkpoint: see the header comment / KT3
                            uint4 *d4 = reinterpret_cast<uint4 *>(rdst + (t - 1) * 16);
                            d4[0] = make_uint4(Ho[0], Ho[1], Ho[2], Ho[3]); d4[1] = make_uint4(Ho[4], Ho[5], Ho[6], Ho[7]);
                            d4[2] = make_uint4(Ho[8], Ho[9], Ho[10], Ho[11]); d4[3] = make_uint4(Ho[12], Ho[13], Ho[14], Ho[15]);
                            if (t >= 2) {
                                uint4 *f4 = reinterpret_cast<uint4 *>(rdst + 48 + (t - 2) * 16);
                                f4[0] = make_uint4(F[0], F[1], F[2], F[3]); f4[1] = make_uint4(F[4], F[5], F[6], F[7]);
                                f4[2] = make_uint4(F[8], F[9], F[10], F[11]); f4[3] = make_uint4(F[12], F[13], F[14], F[15]);
                            }
                        }
                    }
                }
                if (active) {
                    gout[(size_t)bl * 2] = make_uint4(wout[0], wout[1], wout[2], wout[3]);         // rows i0 .. i0+3 of column 32q+32
                    gout[(size_t)bl * 2 + 1] = make_uint4(mf[0], mf[1], mf[2], mf[3]);
                    c1 = wout[1]; c2 = wout[2]; c3 = wout[3];
                    if (team > 1 && lane == lanes_here - 1) { __threadfence_block(); prog_sh[wib] = sw * pbase + bl + 1; }
                }
                if (sw > 0 && gbk < nblk) {                             // lane 0 has finished its block gbk: its next two border entries
                    const bool more = gbk + 2 < nblk;
                    if (more) wait_pub(gbk + 3);
                    if (lane == 0) { g_cur = g_next; if (more) g_next = __ldcg(gin + (size_t)(gbk + 2) * 2); }
                }
            }
            __syncwarp();                                             // the last lane's column is read by lane 0 in the next sweep
        }
    }
}

// ---- K3W: the int32 form of the fill for ONE long pair -----------------------------------------------------------------------
// A pair whose scores leave the s16 range (a 100 kb x 100 kb DNA pair scores 148 609) cannot take the packed cell; on the
// direction-byte path it costs 98 ms of KW + 38 ms of KT.  K3W is K3 with one cell per instruction: a lane owns a 32-column
// pass in int32 registers (Ho, F, table pointers), runs 4-row blocks, hands its last column to the next lane by shuffles and
// stores it (Ho, E, M, F: 16 bytes per row and pass = 0.5 B/cell) for the walker; every 64 rows the pass's H of rows r-1, r
// and F of row r (r = 64 m - 1) are saved.  The sweeps (1024 columns each) of the ONE pair are dealt to all warps of a
// cooperative grid, consecutive sweeps to different SMs; a sweep follows the one before it ~33 row blocks behind, through
// the border store, guarded by a progress word per sweep in global memory (release store after the rows are written,
// acquire load before they are read; waits only ever point to a lower sweep, every warp is resident: no deadlock).
// Local mode tracks the first maximum itself (per-lane candidates, merged by k3w_pick_kernel in the reference's scan
// order, SA.rs:286-294); the walk then starts there inside the full-matrix border store.
#define K3W_WARPS 4
#ifndef K3W_COLS
#define K3W_COLS 16                   // columns per lane: half a 32-column pass (measured 32 / 16 / 8: 48.5 / 40.8 / 43.6 ms for the 100 kb global pair)
#endif
#define K3W_PARTS (K1_STRIP / K3W_COLS)
#define K3W_AHEAD 10                  // a sweep polls the progress of the sweep before it for this many row blocks at once
#define K3W_ROWB_WORDS 96             // per (checkpoint, pass): Ho of row r-1, Ho of row r, F of row r (32 words each)

typedef cuda::atomic_ref<int, cuda::thread_scope_device> k3w_flag;

template <int MODE>
__global__ void __launch_bounds__(K3W_WARPS * 32, 1)
k3w_fill_kernel(const uint8_t *__restrict__ a, const uint8_t *__restrict__ b, int la, int lb, const int32_t *__restrict__ mat, int nl,
                int open, int ext, uint4 *colb, int *rowb, int *prog, int4 *cand)
{
    extern __shared__ __align__(16) int TW_sh[];                     // [nl*nl][32 lanes]: s - open
    for (int idx = threadIdx.x; idx < nl * nl * 32; idx += blockDim.x) TW_sh[idx] = mat[idx >> 5] - open;
    __syncthreads();
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int team = gridDim.x * K3W_WARPS;
    const int gw = wib * gridDim.x + blockIdx.x;                     // consecutive sweeps sit on different SMs
    const unsigned tbase = (unsigned)__cvta_generic_to_shared(TW_sh) + lane * 4;
    const unsigned rowstride = (unsigned)nl * 128u;
    const unsigned pad = (unsigned)(nl - 1);
    const int sent = 10 * open + 10 * ext;
    const int n_pass = (lb + K1_STRIP - 1) / K1_STRIP;
    const int n_half = K3W_PARTS * n_pass;                                  // a lane owns HALF a pass (K3W_COLS = 16 columns): twice the sweeps,
    const int nblk = (la + 3) >> 2;                                  // half the work per step -- a lone warp per SM sub-partition is latency-bound
    const int sweeps = (n_half + 31) >> 5;
    const int n_ck = la / K3_R;
    // Ho = H + open of boundary cell (0, j) / (j, 0): global 2 open + (j-1) ext (SA.rs:100-118), local 0 + open
    auto ho_bnd = [=](int j) { return (MODE == SWB_LOCAL || j == 0) ? open : 2 * open + (j - 1) * ext; };
    int bestHo = open, bi = -1, bj = -1;                             // local: this lane's first maximum over all its sweeps

    for (int sw = gw; sw < sweeps; sw += team) {
        const int hq = sw * 32 + lane;                               // half-pass index; pass q = hq / 2, right half = hq & 1
        const int q = hq / K3W_PARTS, part = hq - q * K3W_PARTS;
        const bool right = (part == K3W_PARTS - 1);                  // the part that holds the pass's last column
        const bool has_pass = hq < n_half;
        const int jbase = hq * K3W_COLS;
        const int lanes_here = min(32, n_half - sw * 32);            // even: the last lane of a sweep is a right half
        const int n_gb = nblk + lanes_here - 1;
        int Ho[K3W_COLS], F[K3W_COLS];
        unsigned cb[K3W_COLS];
#pragma unroll
        for (int k = 0; k < K3W_COLS; ++k) {
            const int j = jbase + k;
            const unsigned c = (has_pass && j < lb) ? (unsigned)b[j] : pad;
            cb[k] = tbase + c * 128u;
            if (MODE == SWB_LOCAL) { Ho[k] = open; F[k] = 0; }
            else { Ho[k] = ho_bnd(j + 1); F[k] = Ho[k] - open + sent; }       // Y(0, j) = H(0, j) + sentinel (SA.rs:100-104)
        }
        const bool first = (hq == 0);
        const bool from_g = (lane == 0 && sw > 0);
        int HoD = ho_bnd(jbase);                                     // Ho of (0, jbase): the diagonal of the first cell
        int o0 = open, o1 = open, o2 = open, o3 = open;              // this lane's last column of its previous block: Ho of rows 1..4
        int e0 = open, e1 = open, e2 = open, e3 = open;              //                                                  E
        const uint4 *gin = colb + ((size_t)(q > 0 ? q - 1 : 0) * (nblk + 1)) * 4;
        uint4 *gout = colb + ((size_t)q * (nblk + 1)) * 4;
        int seen = 0;                                                // row blocks of the sweep before this one known to be published
        // lane 0 of a later sweep reads the last column of the sweep before it from the border store.  Neither the progress
        // poll nor the data load may sit on a step's critical path (two L2 round trips, ~1200 cycles, against ~2000 of
        // arithmetic): the poll waits for K3W_AHEAD blocks at once, the data of block k + 1 is loaded during block k.
        auto wait_pub = [&](int need) {
            if (seen < need) {
                const int want = min(need + K3W_AHEAD, nblk);
                int v = 0;
                do {
                    if (lane == 0) v = k3w_flag(prog[sw - 1]).load(cuda::memory_order_acquire);
                    v = __shfl_sync(0xffffffffu, v, 0);
                    if (v < want) __nanosleep(200);
                } while (v < want);
                seen = v;
            }
        };
        auto load_rows = [&](int blk) -> unsigned {                  // residues of rows 4 blk + 1 .. 4 blk + 4, PAD beyond la
            unsigned wd = 0;
            if (has_pass && blk >= 0 && blk < nblk) {
#pragma unroll
                for (int t = 0; t < 4; ++t) wd |= ((4 * blk + t < la) ? (unsigned)a[4 * blk + t] : pad) << (8 * t);
            }
            return wd;
        };
        unsigned word_next = load_rows(-lane);
        uint4 pf0 = make_uint4(0, 0, 0, 0), pf1 = pf0, pf2 = pf0, pf3 = pf0;
        if (sw > 0) {
            wait_pub(1);
            if (lane == 0) { pf0 = __ldcg(gin + 0); pf1 = __ldcg(gin + 1); pf2 = __ldcg(gin + 2); pf3 = __ldcg(gin + 3); }
        }

#pragma unroll 1
        for (int gbk = 0; gbk < n_gb; ++gbk) {
            const int bl = gbk - lane;
            const bool active = has_pass && bl >= 0 && bl < nblk;
            const int i0 = bl * 4;
            // left boundary of this block's rows: the left lane's last column of ITS previous block (it runs one block ahead)
            int lo[4], le[4];
            lo[0] = __shfl_up_sync(0xffffffffu, o0, 1); lo[1] = __shfl_up_sync(0xffffffffu, o1, 1);
            lo[2] = __shfl_up_sync(0xffffffffu, o2, 1); lo[3] = __shfl_up_sync(0xffffffffu, o3, 1);
            le[0] = __shfl_up_sync(0xffffffffu, e0, 1); le[1] = __shfl_up_sync(0xffffffffu, e1, 1);
            le[2] = __shfl_up_sync(0xffffffffu, e2, 1); le[3] = __shfl_up_sync(0xffffffffu, e3, 1);
            if (sw > 0 && gbk < nblk) {                              // lane 0: block gbk of the sweep before (loaded one step ago)
                if (lane == 0) {
                    lo[0] = (int)pf0.x; le[0] = (int)pf0.y; lo[1] = (int)pf1.x; le[1] = (int)pf1.y;
                    lo[2] = (int)pf2.x; le[2] = (int)pf2.y; lo[3] = (int)pf3.x; le[3] = (int)pf3.y;
                }
                if (gbk + 1 < nblk) {
                    wait_pub(gbk + 2);
                    if (lane == 0) {
                        const uint4 *gp = gin + (size_t)(gbk + 1) * 4;
                        pf0 = __ldcg(gp); pf1 = __ldcg(gp + 1); pf2 = __ldcg(gp + 2); pf3 = __ldcg(gp + 3);
                    }
                }
            }
            if (first) {
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int hb = ho_bnd(i0 + t + 1);
                    lo[t] = MODE == SWB_LOCAL ? open : hb;           // column 0 (SA.rs:109-118): X(i, 0) = H(i, 0) + sentinel
                    le[t] = MODE == SWB_LOCAL ? open : hb - open + sent;
                }
            }
            (void)from_g;
            const unsigned word = word_next;                         // fetched one block ahead: no load latency on the block's critical path
            word_next = load_rows(bl + 1);
            if (active) {
                const bool snap = ((bl + 1) % (K3_R / 4) == 0) && ((bl + 1) / (K3_R / 4) <= n_ck);
                int *rdst = rowb + ((size_t)(snap ? (bl + 1) / (K3_R / 4) - 1 : 0) * n_pass + q) * K3W_ROWB_WORDS + part * K3W_COLS;
                int oo[4], ee[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int i = i0 + t + 1;
                    const unsigned roff = ((word >> (8 * t)) & 0xFFu) * rowstride;
                    int GoP = lo[t], E = le[t];
                    const int HoD_next = GoP;
                    int m31 = 0;
#pragma unroll
                    for (int k = 0; k < K3W_COLS; ++k) {
                        const int s = (int)lds32(cb[k] + roff);
                        const int Fk = __viaddmax_s32(F[k], ext, Ho[k]);
                        const int G = MODE == SWB_LOCAL ? __viaddmax_s32_relu(HoD, s, Fk) : __viaddmax_s32(HoD, s, Fk);
                        if (k == K3W_COLS - 1) m31 = HoD + s;
                        E = __viaddmax_s32(E, ext, GoP);
                        GoP = G + open;
                        HoD = Ho[k];
                        Ho[k] = __viaddmax_s32(E, open, GoP);
                        F[k] = Fk;
                    }
                    HoD = HoD_next;
                    oo[t] = Ho[K3W_COLS - 1]; ee[t] = E;
                    if (right) gout[(size_t)bl * 4 + t] = make_uint4((unsigned)Ho[K3W_COLS - 1], (unsigned)E, (unsigned)m31, (unsigned)F[K3W_COLS - 1]);
                    if (MODE == SWB_LOCAL && i <= la) {              // first strict maximum of this lane (rows ascend; first column inside the row)
                        // (a tree, not a chain: independent maxima fill the bubbles of the row's E chain)
                        int rmx[K3W_COLS / 2];
#pragma unroll
                        for (int k = 0; k < K3W_COLS / 2; ++k) rmx[k] = max(Ho[2 * k], Ho[2 * k + 1]);
#pragma unroll
                        for (int w2 = K3W_COLS / 4; w2 >= 1; w2 >>= 1) {
#pragma unroll
                            for (int k = 0; k < w2; ++k) rmx[k] = max(rmx[k], rmx[k + w2]);
                        }
                        const int rm = max(rmx[0], bestHo);
                        // only the row and the pass are kept (two predicated moves): along a long alignment the best improves in
                        // almost every row, and finding the column here (32 compares) cost 13 of the fill's 50 ms.  The walker
                        // reads the column off the first tile it re-computes (KT3Smem::lastM).
                        if (rm > bestHo) { bestHo = rm; bi = i; bj = min((q + 1) * K1_STRIP, lb); }
                    }
                    if (snap && t == 1) {
#pragma unroll
                        for (int k = 0; k < K3W_COLS; k += 4) *reinterpret_cast<int4 *>(rdst + k) = make_int4(Ho[k], Ho[k + 1], Ho[k + 2], Ho[k + 3]);
                    }
                    if (snap && t == 2) {
#pragma unroll
                        for (int k = 0; k < K3W_COLS; k += 4) {
                            *reinterpret_cast<int4 *>(rdst + 32 + k) = make_int4(Ho[k], Ho[k + 1], Ho[k + 2], Ho[k + 3]);
                            *reinterpret_cast<int4 *>(rdst + 64 + k) = make_int4(F[k], F[k + 1], F[k + 2], F[k + 3]);
                        }
                    }
                }
                o0 = oo[0]; o1 = oo[1]; o2 = oo[2]; o3 = oo[3];
                e0 = ee[0]; e1 = ee[1]; e2 = ee[2]; e3 = ee[3];
                // publish this sweep's last column every 8 row blocks (the release store waits for the rows to be visible:
                // an L2 round trip that must not be paid in every step)
                if (lane == lanes_here - 1 && sw + 1 < sweeps && (((bl + 1) & 7) == 0 || bl + 1 == nblk))
                    k3w_flag(prog[sw]).store(bl + 1, cuda::memory_order_release);
            }
        }
        __syncwarp();
    }
    if (MODE == SWB_LOCAL) cand[(size_t)gw * 32 + lane] = make_int4(bestHo, bi, bj, 0);
}

// Local mode: the end cell = the best candidate in the reference's scan order (score desc, row asc, column asc -- the passes
// are ordered by column, the column inside the pass is resolved by the walker); written into the pair record the walker
// reads (cx0, cy0, score).  One block.
__global__ void __launch_bounds__(256)
k3w_pick_kernel(const int4 *__restrict__ cand, int n_cand, int open, K3Pair *pair)
{
    __shared__ int4 best_sh[256];
    int4 best = make_int4(open, -1, -1, 0);
    for (int k = threadIdx.x; k < n_cand; k += blockDim.x) {
        const int4 c = cand[k];
        if (c.x > best.x || (c.x == best.x && c.x > open && (c.y < best.y || (c.y == best.y && c.z < best.z)))) best = c;
    }
    best_sh[threadIdx.x] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < (int)blockDim.x; ++k) {
            const int4 c = best_sh[k];
            if (c.x > best.x || (c.x == best.x && c.x > open && (c.y < best.y || (c.y == best.y && c.z < best.z)))) best = c;
        }
        pair->score = best.x - open;
        pair->cx0 = best.x > open ? best.y : -1;
        pair->cy0 = best.x > open ? best.z : -1;                     // the LAST column of the pass that holds the end cell:
        pair->find_col = 1;                                              // the walker finds the column itself (first MATCH == score in that row)
    }
}

// ---- KT3: the walk ------------------------------------------------------------------------------------------------------
// One BLOCK of KT3_WARPS warps per pair.  The walk is serial, but re-computing a tile is not part of the dependency: a tile's
// codes follow from the stored borders alone.  Each round, warp w re-computes the w-th tile that a straight diagonal from the
// cursor would visit (warp 0: the cursor's own tile, only the part above-left of the cursor; the others: whole tiles), then
// warp 0 walks through as many of those tiles as the path really visits and the block starts the next round at the first
// tile that was not predicted.  A walk alone spends ~85 % of its time re-computing; with 8 warps a round costs one
// re-compute plus the walk through several tiles.
#define KT3_WARPS 8                   // default; SWB_KT3_WARPS (1..16) overrides at launch
#define KT3_MAX_WARPS 16

struct KT3Smem {
    uint8_t dir[K3_R][32];          // predecessor codes of the recomputed part of the tile
    int4 left[K3_R + 2];            // (M, X, Y, raw E) of column c0, rows rtop .. cx; local mode: M, X, Y clamped as the reference's states
    int hm1[33], h0[33];            // H(rtop - 1, c0 + k), H(rtop, c0 + k), k = 0 .. 32
    int arow[K3_R + 2];             // score-table row offset of the residue of row rtop + r (no global load inside the wavefront)
    int rtop, c0, tcx, tcy;         // the tile held: rows (rtop, tcx], columns (c0, tcy]; tcx < 0 = none
    int fin[3];                     // (M, X, Y) of the tile's last cell (tcx, tcy): the start state of a global walk
    int lastM[32];                  // MATCH of the tile's last row, per column: the wide local form finds its start column here
};

// (M, X, Y, raw E, H) of a stored border cell (row, column 32 q + 32) -- or of a matrix boundary cell.
template <int MODE, bool WIDE>
__device__ __forceinline__ void kt3_border_cell(const uint4 *cb, int nblk, int q_left, int row, int col0, int open, int o, int e, int &M, int &X, int &Y, int &Eraw, int &H)
{
    const int so = (MODE == SWB_GLOBAL) ? o : 0, se = (MODE == SWB_GLOBAL) ? e : 0, sent = 10 * o + 10 * e;
    if (q_left < 0) {                                   // column 0 (SA.rs:88-93, 109-118)
        if (row == 0) { M = 0; X = -1; Y = -1; Eraw = -1; H = 0; return; }
        const int ls = row * se + (so - se);
        M = ls + sent; X = ls + sent; Y = ls; Eraw = X; H = ls; return;
    }
    if (row == 0) {                                     // row 0 (SA.rs:100-108)
        const int ls = col0 * se + (so - se);
        M = ls + sent; X = ls; Y = ls + sent; Eraw = X; H = ls; return;
    }
    if (WIDE) {                                         // K3W: one uint4 (Ho, E, M, F) per row, rows 4 b + 1 .. 4 b + 4 in block b
        const uint4 g = cb[((size_t)q_left * (nblk + 1) + ((row - 1) >> 2)) * 4 + ((row - 1) & 3)];
        H = (int)g.x - open; Eraw = (int)g.y; M = (int)g.z; Y = (int)g.w; X = Eraw;
        if (MODE == SWB_LOCAL) { M = max(M, 0); X = max(X, 0); Y = max(Y, 0); }
        return;
    }
    const uint4 *ent = cb + ((size_t)q_left * (nblk + 1) + (row >> 2)) * 2;
    const uint4 w0 = ent[0], w1 = ent[1];
    const int t = row & 3;
    const unsigned a = t == 0 ? w0.x : (t == 1 ? w0.y : (t == 2 ? w0.z : w0.w));
    const unsigned b = t == 0 ? w1.x : (t == 1 ? w1.y : (t == 2 ? w1.z : w1.w));
    H = lo16(a) - open; Eraw = hi16(a); M = lo16(b); Y = hi16(b); X = Eraw;
    if (MODE == SWB_LOCAL) { M = max(M, 0); X = max(X, 0); Y = max(Y, 0); }
}

struct KT3Ctx {
    const uint8_t *a, *b;
    int la, lb, nblk, n_pass, nl, o, e;
    const uint4 *cb;
    const unsigned *rb;
    const int *mat_sh;
    bool swp;
};

// Re-compute rows (rtop, hi_row] x columns (c0, hi_col] of the tile of (hi_row, hi_col) into S (one warp).
template <int MODE, bool WIDE>
__device__ __forceinline__ void kt3_tile(KT3Smem &S, const KT3Ctx &C, int lane, int hi_row, int hi_col)
{
    const int o = C.o, e = C.e, open = C.o, nl = C.nl, lb = C.lb;
    const int so = (MODE == SWB_GLOBAL) ? o : 0, se = (MODE == SWB_GLOBAL) ? e : 0, sent = 10 * o + 10 * e;
    const bool swp = C.swp;
    auto am = [swp](int m, int x, int y) -> unsigned { return (m >= x && m >= y) ? 0u : (swp ? ((y >= x) ? 2u : 1u) : ((x >= y) ? 1u : 2u)); };
    const int q = (hi_col - 1) >> 5;
    const int c0 = q << 5;
    const int m = hi_row / K3_R;
    const int rtop = m ? K3_R * m - 1 : 0;
    const int nrows = hi_row - rtop, ncols = hi_col - c0;
    __syncwarp();
    if (lane == 0) { S.rtop = rtop; S.c0 = c0; S.tcx = hi_row; S.tcy = hi_col; }
    for (int r = lane; r <= nrows; r += 32) {
        int M, X, Y, Er, H;
        kt3_border_cell<MODE, WIDE>(C.cb, C.nblk, q - 1, rtop + r, c0, open, o, e, M, X, Y, Er, H);
        S.left[r] = make_int4(M, X, Y, Er);
        if (r == 0) S.h0[0] = H;
        S.arow[r] = (rtop + r >= 1) ? (int)C.a[rtop + r - 1] * nl : 0;
    }
    if (lane == 0) {
        int M, X, Y, Er, H = 0;
        if (rtop > 0) kt3_border_cell<MODE, WIDE>(C.cb, C.nblk, q - 1, rtop - 1, c0, open, o, e, M, X, Y, Er, H);
        S.hm1[0] = H;
    }
    // top border row rtop of this lane's column j = c0 + 1 + lane
    const int j = c0 + 1 + lane;
    int Mu, Xu, Yu;
    if (rtop == 0) {
        const int ls = j * se + (so - se);
        Mu = ls + sent; Xu = ls; Yu = ls + sent;
        S.h0[lane + 1] = ls; S.hm1[lane + 1] = 0;
    } else {
        int hm1, h0, f0;
        if (WIDE) {
            const int *w = reinterpret_cast<const int *>(C.rb) + ((size_t)(m - 1) * C.n_pass + q) * K3W_ROWB_WORDS;
            hm1 = w[lane] - open; h0 = w[32 + lane] - open; f0 = w[64 + lane];
        } else {
            const unsigned *w = C.rb + ((size_t)(m - 1) * C.n_pass + q) * K3_ROWB_WORDS;
            const int k = lane & 15;
            if (lane < 16) { hm1 = lo16(w[k]) - open; h0 = lo16(w[16 + k]) - open; f0 = lo16(w[48 + k]); }
            else { hm1 = hi16(w[16 + k]) - open; h0 = hi16(w[32 + k]) - open; f0 = hi16(w[64 + k]); }
        }
        S.hm1[lane + 1] = hm1; S.h0[lane + 1] = h0;
        Yu = f0; Mu = 0; Xu = 0;
    }
    __syncwarp();
    if (rtop > 0) {
        const int sc = (j <= lb) ? C.mat_sh[(int)C.a[rtop - 1] * nl + (int)C.b[j - 1]] : 0;
        Mu = S.hm1[lane] + sc;                               // M(rtop, j) = H(rtop-1, j-1) + s
        int ev = S.left[0].w + (lane + 1) * e;               // X(rtop, j): the row recurrence from the stored E of column c0
        for (int k = 0; k <= lane; ++k) ev = max(ev, S.h0[k] + o + (lane - k) * e);
        Xu = ev;
        if (MODE == SWB_LOCAL) { Mu = max(Mu, 0); Xu = max(Xu, 0); Yu = max(Yu, 0); }
    }
    // diagonal of this lane's first cell: cell (rtop, j - 1)
    int dM = __shfl_up_sync(0xffffffffu, Mu, 1), dX = __shfl_up_sync(0xffffffffu, Xu, 1), dY = __shfl_up_sync(0xffffffffu, Yu, 1);
    if (lane == 0) { const int4 v = S.left[0]; dM = v.x; dX = v.y; dY = v.z; }
    int hd0 = __vimax3_s32(dM, dX, dY); unsigned ad0 = am(dM, dX, dY);
    int lastM = 0, lastX = 0, lastY = 0;
    const int bcode = (j <= lb) ? (int)C.b[j - 1] : 0;
    const int n_steps = nrows + ncols - 1;
    auto score_of = [&](int r) { return (lane < ncols && r >= 1 && r <= nrows) ? C.mat_sh[S.arow[r] + bcode] : 0; };
    int s_nxt = score_of(1 - lane);
    for (int t = 0; t < n_steps; ++t) {
        const int r = t - lane + 1;                          // row of this lane in this step, 1 .. nrows
        int tM = __shfl_up_sync(0xffffffffu, lastM, 1), tX = __shfl_up_sync(0xffffffffu, lastX, 1), tY = __shfl_up_sync(0xffffffffu, lastY, 1);
        const int sc = s_nxt;
        s_nxt = score_of(r + 1);                             // the next step's score: two shared-memory loads off the critical path
        if (lane < ncols && r >= 1 && r <= nrows) {
            if (lane == 0) { const int4 v = S.left[r]; tM = v.x; tX = v.y; tY = v.z; }
            const int hd = hd0; const unsigned ad = ad0;
            hd0 = __vimax3_s32(tM, tX, tY); ad0 = am(tM, tX, tY);
            int mm, xx, yy; unsigned code;
            kw_cell<MODE>(hd, ad, tM, tX, tY, Mu, Xu, Yu, sc, o, e, o, e, mm, xx, yy, code);
            Mu = mm; Xu = xx; Yu = yy;
            lastM = mm; lastX = xx; lastY = yy;
            S.dir[r - 1][lane] = (uint8_t)code;
        }
    }
    if (lane == ncols - 1) { S.fin[0] = Mu; S.fin[1] = Xu; S.fin[2] = Yu; }
    S.lastM[lane] = (lane < ncols) ? Mu : INT_MIN;
    __syncwarp();
}

struct KT3Walk { int cx, cy, st, n, score, done, sy; };

template <int MODE, bool WIDE>
__global__ void __launch_bounds__(KT3_MAX_WARPS * 32, 2)
kt3_walk_kernel(const uint8_t *__restrict__ a_all, const uint8_t *__restrict__ b_all, const K3Pair *__restrict__ pairs, int n_pairs,
                const int32_t *__restrict__ mat, int nl, int o, int e, const uint4 *__restrict__ colb, const unsigned *__restrict__ rowb,
                int32_t *__restrict__ idx_a, int32_t *__restrict__ idx_b, int32_t *__restrict__ ncols_out, KWResult *__restrict__ res)
{
    extern __shared__ __align__(16) int mat_sh[];                    // nl * nl scores (rounded up to 16 bytes), then one KT3Smem per warp
    __shared__ KT3Walk W;
    const int n_warps = blockDim.x >> 5;
    KT3Smem *sm_all = reinterpret_cast<KT3Smem *>(mat_sh + ((nl * nl + 3) & ~3));
    for (int k = threadIdx.x; k < nl * nl; k += blockDim.x) mat_sh[k] = mat[k];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int p = blockIdx.x;
    const K3Pair P = pairs[p];
    // Transposed pairs (P.swap): rows = b, columns = a.  The gap-state rules of the reference mirror each other (GAPINX prefers
    // X > M > Y where GAPINY prefers Y > M > X, SA.rs:172-211; local Y > M > X / X > M > Y, SA.rs:214-252) and so do the
    // boundaries (SA.rs:100-118), so the cell is the same with the state names X and Y exchanged; what is NOT symmetric is
    // the MATCH predecessor priority M > X > Y (SA.rs:139-158) and the start-state order M, X, Y (SA.rs:302-308): in the
    // transposed names they read M > Y > X.  The host exchanges the two column lists and the end cell back.
    KT3Ctx C;
    C.swp = P.swap != 0;
    C.a = C.swp ? b_all + P.b_off : a_all + P.a_off; C.b = C.swp ? a_all + P.a_off : b_all + P.b_off;
    C.la = P.la; C.lb = P.lb; C.nblk = WIDE ? (P.la + 3) >> 2 : (P.la + 1 + 3) >> 2; C.n_pass = (P.lb + K1_STRIP - 1) / K1_STRIP;
    C.nl = nl; C.o = o; C.e = e; C.cb = colb + P.colb_off; C.rb = rowb + P.rowb_off; C.mat_sh = mat_sh;
    const int la = P.la, lb = P.lb;
    const int sent = 10 * o + 10 * e;
    int32_t *xa = idx_a + P.out_off, *xb = idx_b + P.out_off;
    const int sx = P.cx0 ? P.cx0 : la, sy = P.cy0 ? P.cy0 : lb;     // the cell the walk starts at (wide local form: found by the fill)
    if (threadIdx.x == 0) { W.cx = max(sx, 0); W.cy = max(sy, 0); W.st = P.start_state; W.n = 0; W.score = P.score; W.done = (sx < 0) ? 1 : 0; W.sy = sy; }
    bool need_col = WIDE && MODE == SWB_LOCAL && P.find_col == 1 && sx >= 1;      // sy is the end of the start cell's pass, not yet its column
    if (lane == 0) sm_all[wib].tcx = -1;
    bool need_start = (MODE != SWB_LOCAL);
    int va = 0, vb = 0;                                              // warp 0: lane (n & 31) holds column n of the lists (one coalesced store per 32 columns)
    __syncthreads();

    for (; !W.done;) {
        // ---- every warp: the w-th tile on the straight diagonal from the cursor
        {
            int x = W.cx, y = W.cy;
            bool have = (x >= 1 && y >= 1);
            for (int k = 0; k < wib && have; ++k) {
                const int rt = (x / K3_R) ? K3_R * (x / K3_R) - 1 : 0, cc = ((y - 1) >> 5) << 5;
                const int d = min(x - rt, y - cc);
                x -= d; y -= d;
                have = (x >= 1 && y >= 1);
            }
            if (have) {
                int hi_row = x, hi_col = y;
                if (wib > 0) {                                       // a predicted tile: all of it (the entry point is not known yet)
                    const int m = x / K3_R;
                    hi_row = min(m ? K3_R * m + K3_R - 1 : K3_R - 1, la); hi_col = min((((y - 1) >> 5) << 5) + 32, lb);
                }
                kt3_tile<MODE, WIDE>(sm_all[wib], C, lane, hi_row, hi_col);
            } else if (lane == 0) sm_all[wib].tcx = -1;
        }
        __syncthreads();
        // ---- warp 0: walk through the tiles at hand (backtrack, SA.rs:346-383); the same on every lane, lane 0 writes
        if (wib == 0) {
            int cx = W.cx, cy = W.cy, st = W.st, n = W.n, score = W.score, done = 0;
            if (need_col) {                                          // first MATCH == score in the start row (SA.rs:286-294: columns ascend)
                const unsigned hit = __ballot_sync(0xffffffffu, sm_all[0].lastM[lane] == score);
                if (hit) { cy = sm_all[0].c0 + __ffs(hit); if (lane == 0) W.sy = cy; }
                else { n = -2; done = 1; }                           // cannot happen: the fill saw this value in this row of this pass
            }
            if (need_start && cx >= 1 && cy >= 1) {                  // cell (la, lb): first strict maximum in the order M, X, Y (SA.rs:302-308)
                const int fm = sm_all[0].fin[0], fx = sm_all[0].fin[1], fy = sm_all[0].fin[2];
                score = fm; st = 0;
                if (!C.swp) { if (fx > score) { score = fx; st = 1; } if (fy > score) { score = fy; st = 2; } }
                else { if (fy > score) { score = fy; st = 2; } if (fx > score) { score = fx; st = 1; } }
            }
            // Inside a tile the walk is a tight loop: one code byte per step at a linear position, no boundary, panic or end
            // tests (a tile's cells have cx, cy >= 1); everything else -- changing tiles, the matrix boundary, the end of the
            // walk -- is the slow path around it.
            for (; !done;) {
                if (cx >= 1 && cy >= 1) {
                    int found = -1;
                    for (int k = 0; k < n_warps; ++k) {
                        const KT3Smem *Q = &sm_all[k];
                        if (found < 0 && Q->tcx >= 0 && cx > Q->rtop && cx <= Q->tcx && cy > Q->c0 && cy <= Q->tcy) found = k;
                    }
                    if (found < 0) break;                            // not predicted: next round
                    const KT3Smem *T = &sm_all[found];
                    const int t_rtop = T->rtop, t_c0 = T->c0;
                    const uint8_t *t_dir = &T->dir[0][0];
                    int pos = (cx - t_rtop - 1) * 32 + (cy - t_c0 - 1);
                    bool stop = false;
                    do {
                        const unsigned prev = ((unsigned)t_dir[pos] >> (2 * st)) & 3u;
                        if (MODE == SWB_LOCAL && prev == 3) { stop = true; break; }
                        const int dx = (st != 1), dy = (st != 2);
                        if (lane == (n & 31)) { va = dx ? cx - 1 : -1; vb = dy ? cy - 1 : -1; }
                        ++n;
                        if ((n & 31) == 0) { xa[n - 32 + lane] = va; xb[n - 32 + lane] = vb; }
                        cx -= dx; cy -= dy; pos -= dx * 32 + dy;
                        st = (int)prev;
                    } while (cx > t_rtop && cy > t_c0);
                    if (stop) { done = 1; break; }
                    if (cx == 0 && cy == 0) { done = 1; break; }
                    continue;
                }
                // on the matrix boundary (SA.rs:105-116): row 0 -> GAPINX, column 0 -> GAPINY for all states
                unsigned prev;
                if (cx == 0 && cy == 0) prev = 0;
                else if (cx == 0) { prev = 1; if (MODE == SWB_LOCAL && (st == 1 || sent == 0)) prev = 3; }
                else { prev = 2; if (MODE == SWB_LOCAL && (st == 2 || sent == 0)) prev = 3; }
                if (MODE == SWB_LOCAL && prev == 3) { done = 1; break; }
                if ((st != 2 && cy == 0) || (st != 1 && cx == 0)) { n = -2; done = 1; break; }   // the reference's panic cases (SA.rs:360-371)
                if (lane == (n & 31)) { va = (st == 1) ? -1 : cx - 1; vb = (st == 2) ? -1 : cy - 1; }
                ++n;
                if ((n & 31) == 0) { xa[n - 32 + lane] = va; xb[n - 32 + lane] = vb; }
                cx -= (st != 1); cy -= (st != 2);
                if (cx == 0 && cy == 0) { done = 1; break; }
                st = (int)prev;
            }
            if (n > 0 && lane < (n & 31)) { xa[(n & ~31) + lane] = va; xb[(n & ~31) + lane] = vb; }   // the columns of the last, partial group
            __syncwarp();
            if (lane == 0) { W.cx = cx; W.cy = cy; W.st = st; W.n = n; W.score = score; W.done = done; }
        }
        need_start = false; need_col = false;
        __syncthreads();
        if (W.done) break;
        __syncthreads();                                             // everybody has read the cursor before warp 0 may write it again
    }
    if (threadIdx.x == 0) {
        ncols_out[p] = W.n;
        KWResult R; R.score = W.score; R.end_i = sx; R.end_j = (sx >= 1) ? W.sy : sy; R.state = P.start_state;
        res[p] = R;
    }
}

// ---- host side ----------------------------------------------------------------------------------------------------------
// Which pairs may take the packed path.  |open| >= |extend| makes the Gotoh values equal the three-state values.  Local
// mode: the caller knows the pair's score from the scan and has checked it against the s16 limit (every H is <= it).
// Global mode has no clamp: the lengths must bound every value.  Upwards min(la,lb) best substitutions; downwards a cell
// (i, j) is at least the path "min(i,j) worst substitutions, then one gap of |i-j|", whose minimum over the matrix sits in
// a corner, and the states / temporaries of a cell lie at most 3 openings + one substitution + the sentinel below that.
bool k3_eligible(const swb_engine *e, int64_t la, int64_t lb)
{
    if (e->o > e->e || la < 1 || lb < 1 || la > 0x3FFFFFF0 || lb > 0x3FFFFFF0) return false;
    if (getenv("SWB_NO_K3")) return false;
    if (e->mode == SWB_LOCAL) return true;
    if (e->mode != SWB_GLOBAL) return false;
    const int64_t mn = std::min(la, lb), mx = std::max(la, lb);
    const int64_t ao = -(int64_t)e->o, ae = -(int64_t)e->e, sp = std::max(e->smax, 0), sn = std::max(-e->smin, 0);
    const int64_t up = mn * sp + ao + sp + 64;
    const int64_t corner = std::max(sn * mn + ae * (mx - mn), ae * mx) + ao;
    const int64_t down = corner + 3 * ao + sn + (10 * ao + 10 * ae) + 64;
    return up <= 32000 && down <= 32000;
}

void k3_sizes(int32_t la, int32_t lb, int64_t *colb_uint4, int64_t *rowb_words)
{
    const int64_t n_pass = (lb + K1_STRIP - 1) / K1_STRIP, nblk = (la + 1 + 3) >> 2, n_ck = la / K3_R;
    *colb_uint4 = n_pass * (nblk + 1) * 2;
    *rowb_words = n_ck * n_pass * K3_ROWB_WORDS;
}

cudaError_t k3_launch(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const K3Pair *d_pairs, int n_pairs,
                      uint4 *d_colb, unsigned *d_rowb, int *d_counter, int32_t *d_idx_a, int32_t *d_idx_b, int32_t *d_ncols,
                      KWResult *d_res, cudaStream_t st, int max_sweeps)
{
    if (n_pairs <= 0) return cudaSuccess;
    const int nl = e->n + 1;
    const size_t smem = (size_t)nl * nl * 32 * sizeof(unsigned);
    cudaError_t err = cudaMemsetAsync(d_counter, 0, sizeof(int), st);
    if (err != cudaSuccess) return err;
    // team size: a divisor of the block's warps, enough to fill the warp slots of the GPU when pairs are few, no more
    // than the sweeps of the widest pair (`max_sweeps`, from the caller)
    int team = 1;
    {
        const int want = (e->sm_count * K3_WARPS + n_pairs - 1) / n_pairs;
        const int divs[] = {1, 2, 3, 4, 6, 12};
        for (int d : divs) if (d <= std::max(1, max_sweeps) && (team < want)) team = d;
        if (const char *f = getenv("SWB_K3_TEAM")) { const int v = atoi(f); for (int d : divs) if (d == v) team = v; }
    }
    const int slots = K3_WARPS / team;
    const int fill_blocks = std::min(e->sm_count, (n_pairs + slots - 1) / slots);
    const int walk_blocks = n_pairs;                                   // one block of KT3_WARPS warps per pair
    int kt3_warps = KT3_WARPS;
    if (const char *f = getenv("SWB_KT3_WARPS")) { const int v = atoi(f); if (v >= 1 && v <= KT3_MAX_WARPS) kt3_warps = v; }
    const size_t wsmem = (size_t)((nl * nl + 3) & ~3) * sizeof(int) + (size_t)kt3_warps * sizeof(KT3Smem);
    if (e->mode == SWB_LOCAL) {
        if ((err = cudaFuncSetAttribute(k3_fill_kernel<SWB_LOCAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return err;
        if ((err = cudaFuncSetAttribute(kt3_walk_kernel<SWB_LOCAL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem)) != cudaSuccess) return err;
        k3_fill_kernel<SWB_LOCAL><<<fill_blocks, K3_WARPS * 32, smem, st>>>(d_a, d_b, d_pairs, n_pairs, e->d_mat, nl, e->o, e->e, d_colb, d_rowb, d_counter, team);
        kt3_walk_kernel<SWB_LOCAL, false><<<walk_blocks, kt3_warps * 32, wsmem, st>>>(d_a, d_b, d_pairs, n_pairs, e->d_mat, nl, e->o, e->e, d_colb, d_rowb,
                                                                              d_idx_a, d_idx_b, d_ncols, d_res);
    } else {
        if ((err = cudaFuncSetAttribute(k3_fill_kernel<SWB_GLOBAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return err;
        if ((err = cudaFuncSetAttribute(kt3_walk_kernel<SWB_GLOBAL, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem)) != cudaSuccess) return err;
        k3_fill_kernel<SWB_GLOBAL><<<fill_blocks, K3_WARPS * 32, smem, st>>>(d_a, d_b, d_pairs, n_pairs, e->d_mat, nl, e->o, e->e, d_colb, d_rowb, d_counter, team);
        kt3_walk_kernel<SWB_GLOBAL, false><<<walk_blocks, kt3_warps * 32, wsmem, st>>>(d_a, d_b, d_pairs, n_pairs, e->d_mat, nl, e->o, e->e, d_colb, d_rowb,
                                                                               d_idx_a, d_idx_b, d_ncols, d_res);
    }
    return cudaGetLastError();
}

// ---- K3W host side ------------------------------------------------------------------------------------------------------
static int k3w_team_cap(const swb_engine *e) { return e->sm_count * K3W_WARPS; }

void k3w_sizes(int32_t la, int32_t lb, int64_t *colb_uint4, int64_t *rowb_words, int64_t *scratch_ints)
{
    const int64_t n_pass = (lb + K1_STRIP - 1) / K1_STRIP, nblk = (la + 3) >> 2, n_ck = la / K3_R, sweeps = (K3W_PARTS * n_pass + 31) / 32;
    *colb_uint4 = n_pass * (nblk + 1) * 4;
    *rowb_words = std::max<int64_t>(n_ck * n_pass * K3W_ROWB_WORDS, 4);
    *scratch_ints = ((sweeps + 3) / 4 * 4) + 4 * 32 * (int64_t)(148 * 2 * K3W_WARPS);      // progress words, then the candidates (int4 per lane)
}

cudaError_t k3w_launch(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, K3Pair *d_pair, const K3Pair &hp,
                       uint4 *d_colb, unsigned *d_rowb, int *d_scratch, int32_t *d_idx_a, int32_t *d_idx_b, int32_t *d_ncols,
                       KWResult *d_res, cudaStream_t st)
{
    const int nl = e->n + 1;
    const size_t smem = (size_t)nl * nl * 32 * sizeof(int);
    const int n_pass = (hp.lb + K1_STRIP - 1) / K1_STRIP, sweeps = (K3W_PARTS * n_pass + 31) / 32;    // a sweep = 32 lanes of K3W_COLS columns
    const int blocks = std::max(1, std::min(e->sm_count, sweeps));
    const int team = blocks * K3W_WARPS;
    if (team > k3w_team_cap(e) * 2) return cudaErrorInvalidValue;
    int *d_prog = d_scratch;
    int4 *d_cand = reinterpret_cast<int4 *>(d_scratch + (sweeps + 3) / 4 * 4);
    cudaError_t err = cudaMemsetAsync(d_prog, 0, (size_t)((sweeps + 3) / 4 * 4) * sizeof(int), st);
    if (err != cudaSuccess) return err;
    const uint8_t *pa = hp.swap ? d_b + hp.b_off : d_a + hp.a_off, *pb = hp.swap ? d_a + hp.a_off : d_b + hp.b_off;   // transposed: rows = b
    int la = hp.la, lb = hp.lb, nl_ = nl, open = e->o, ext = e->e;
    const int32_t *mat = e->d_mat;
    int *rowb_i = reinterpret_cast<int *>(d_rowb);
    void *args[] = {(void *)&pa, (void *)&pb, (void *)&la, (void *)&lb, (void *)&mat, (void *)&nl_, (void *)&open, (void *)&ext,
                    (void *)&d_colb, (void *)&rowb_i, (void *)&d_prog, (void *)&d_cand};
    const void *fn = e->mode == SWB_LOCAL ? reinterpret_cast<const void *>(&k3w_fill_kernel<SWB_LOCAL>) : reinterpret_cast<const void *>(&k3w_fill_kernel<SWB_GLOBAL>);
    if ((err = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return err;
    // the sweeps wait for each other across blocks: the grid must be resident at once (one block per SM by launch bounds), and
    // like KW's team launches it is chained through the engine's team event so that two such grids never share the GPU
    swb_engine *me = const_cast<swb_engine *>(e);
    if (!me->ev_team) { if ((err = cudaEventCreateWithFlags(&me->ev_team, cudaEventDisableTiming)) != cudaSuccess) return err; }
    else if ((err = cudaStreamWaitEvent(st, me->ev_team, 0)) != cudaSuccess) return err;
    if ((err = cudaLaunchCooperativeKernel(fn, dim3(blocks), dim3(K3W_WARPS * 32), args, smem, st)) != cudaSuccess) return err;
    if ((err = cudaEventRecord(me->ev_team, st)) != cudaSuccess) return err;
    if (e->mode == SWB_LOCAL) k3w_pick_kernel<<<1, 256, 0, st>>>(d_cand, team * 32, e->o, d_pair);
    int kt3_warps = KT3_MAX_WARPS;                                   // one pair, the GPU is otherwise idle: the deepest prediction (measured 8 / 12 / 16 warps: 68.8 / 67.2 / 66.8 ms)
    if (const char *f = getenv("SWB_KT3_WARPS")) { const int v = atoi(f); if (v >= 1 && v <= KT3_MAX_WARPS) kt3_warps = v; }
    const size_t wsmem = (size_t)((nl * nl + 3) & ~3) * sizeof(int) + (size_t)kt3_warps * sizeof(KT3Smem);
    if (e->mode == SWB_LOCAL) {
        if ((err = cudaFuncSetAttribute(kt3_walk_kernel<SWB_LOCAL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem)) != cudaSuccess) return err;
        kt3_walk_kernel<SWB_LOCAL, true><<<1, kt3_warps * 32, wsmem, st>>>(d_a, d_b, d_pair, 1, e->d_mat, nl, e->o, e->e, d_colb, d_rowb, d_idx_a, d_idx_b, d_ncols, d_res);
    } else {
        if ((err = cudaFuncSetAttribute(kt3_walk_kernel<SWB_GLOBAL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem)) != cudaSuccess) return err;
        kt3_walk_kernel<SWB_GLOBAL, true><<<1, kt3_warps * 32, wsmem, st>>>(d_a, d_b, d_pair, 1, e->d_mat, nl, e->o, e->e, d_colb, d_rowb, d_idx_a, d_idx_b, d_ncols, d_res);
    }
    return cudaGetLastError();
}

