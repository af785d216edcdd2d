// KW -- intra-sequence anti-diagonal wavefront kernel, one warp per pair, exact 3-state int32
// recurrence with the reference's tie-breaking, optional direction bytes for the traceback.
// KT -- pointer walk over the direction bytes.
//
// Replaces fill_matrix (SA.rs:55-272) and backtrack (SA.rs:273-387) of the reference (OpenCL twins
// OCL.rs:354-524 and OCL.rs:180-323).  Semantics reproduced cell for cell:
//   MATCH  = max(d.M, d.X, d.Y) + s            priority M > X > Y                 (SA.rs:139-158)
//   GAPINX = max(t.M+po, t.X+pe, t.Y+po)       local  Y > M > X  (SA.rs:214-231)
//                                              other  X > M > Y  (SA.rs:172-189), (po,pe)=(so,se) on row la
//   GAPINY = max(l.M+qo, l.Y+qe, l.X+qo)       local  X > M > Y  (SA.rs:233-252)
//                                              other  Y > M > X  (SA.rs:192-211), (qo,qe)=(so,se) on column lb
//   local: clamp to 0 after the predecessor is chosen (SA.rs:256-260)
// with d=(i-1,j-1), t=(i,j-1), l=(i-1,j); boundaries as SA.rs:88-118.
//
// Work layout: lane l of the warp owns KW_C consecutive columns of a 32*KW_C wide strip and walks
// down the rows skewed by one step per lane (lane l is at row t-l+1 in step t), so the cells of
// one step lie on an anti-diagonal of column blocks.  The left neighbour's last column travels by
// __shfl_up; the last column of a strip goes through a small global buffer to lane 0 of the next
// strip.  Direction bytes (three 2-bit predecessor codes; code 3 = "local score is 0, stop here",
// SA.rs:354-357) are stored step-major: the warp writes 32*KW_C contiguous bytes per step.
#include <algorithm>

#include "swb_internal.h"

#define KW_C 8
#define KW_SW (32 * KW_C)
#define KW_WARPS 8
#define NEG_INF_I32 (-(1 << 29))

int64_t kw_dir_bytes(int32_t la, int32_t lb)
{
    int64_t strips = (lb + KW_SW - 1) / KW_SW;
    return strips * (int64_t)(la + 31) * KW_SW;
}

template <int MODE>
__device__ __forceinline__ void kw_cell(int dM, int dX, int dY, int tM, int tX, int tY, int lM, int lX, int lY, int s,
                                        int po, int pe, int qo, int qe, int &m, int &x, int &y, unsigned &code)
{
    int pm, px, py;
    if (dM >= dX) { if (dM >= dY) { m = dM; pm = 0; } else { m = dY; pm = 2; } }
    else          { if (dX >= dY) { m = dX; pm = 1; } else { m = dY; pm = 2; } }
    m += s;
    const int a = tM + po, b = tX + pe, c = tY + po;
    const int f = lM + qo, g = lY + qe, h = lX + qo;
    if (MODE != SWB_LOCAL) {
        if (a > b) { if (a >= c) { x = a; px = 0; } else { x = c; px = 2; } }
        else       { if (b >= c) { x = b; px = 1; } else { x = c; px = 2; } }
        if (f > g) { if (f >= h) { y = f; py = 0; } else { y = h; py = 1; } }
        else       { if (g >= h) { y = g; py = 2; } else { y = h; py = 1; } }
    } else {
        if (a >= b) { if (a > c) { x = a; px = 0; } else { x = c; px = 2; } }
        else        { if (b > c) { x = b; px = 1; } else { x = c; px = 2; } }
        if (f >= g) { if (f > h) { y = f; py = 0; } else { y = h; py = 1; } }
        else        { if (g > h) { y = g; py = 2; } else { y = h; py = 1; } }
        if (m <= 0) { m = 0; pm = 3; }
        if (x <= 0) { x = 0; px = 3; }
        if (y <= 0) { y = 0; py = 3; }
    }
    code = (unsigned)pm | ((unsigned)px << 2) | ((unsigned)py << 4);
}

// Team scheduling: a pair is swept by a TEAM of T warps (T = 1 for batches of many pairs; up to several hundred for a
// few long pairs).  Warp w of the team takes the column strips w, w+T, w+2T, ...; strip s consumes the last column of
// strip s-1 row by row through a ring of T+1 column buffers in global memory (L2 resident) guarded by a per-strip
// progress counter (rows published, release/acquire through __threadfence + volatile accesses).  Waits only ever point
// to a lower strip, so there is no cycle; the launcher sizes the grid so that every warp of every team is resident.
__device__ __forceinline__ int ld_volatile(const int *p) { return *reinterpret_cast<const volatile int *>(p); }

template <int MODE, bool DIRS>
__global__ void __launch_bounds__(KW_WARPS * 32)
kw_kernel(const uint8_t *__restrict__ a_all, const uint8_t *__restrict__ b_all, const KWPair *__restrict__ pairs,
          int n_pairs, int team_warps, int n_teams, const int32_t *__restrict__ mat, int nl, int o, int e, int unit,
          uint8_t *__restrict__ dir, int4 *__restrict__ bnd, int *__restrict__ scratch, KWResult *__restrict__ res)
{
    extern __shared__ int mat_sh[];
    for (int k = threadIdx.x; k < nl * nl; k += blockDim.x) mat_sh[k] = mat[k];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int gw = blockIdx.x * KW_WARPS + (threadIdx.x >> 5);
    const int team = gw / team_warps, w = gw - team * team_warps;
    if (team >= n_teams) return;
    const int so = (MODE == SWB_GLOBAL) ? o : 0, se = (MODE == SWB_GLOBAL) ? e : 0;
    const int sent = 10 * o + 10 * e;

    for (int pidx = team; pidx < n_pairs; pidx += n_teams) {
        const KWPair P = pairs[pidx];
        const uint8_t *a = a_all + P.a_off, *b = b_all + P.b_off;
        const int la = P.la, lb = P.lb;
        const int n_strips = (lb + KW_SW - 1) / KW_SW;
        const int nbuf = min(n_strips, team_warps + 1);
        int *counter = scratch + P.scr_off;                    // [0] arrivals, [1..n_strips] rows published per strip,
        int *prog = counter + 1;                               // then 4 ints per team warp (local-mode candidates)
        int *cand = prog + n_strips;
        int best = 0, bi = -1, bj = -1;                        // local: first strict maximum of MATCH
        int fin_score = 0, fin_state = 0;                      // non-local: cell (la, lb)
        bool have_fin = false;

        for (int sidx = w; sidx < n_strips; sidx += team_warps) {
            const int j0 = sidx * KW_SW + lane * KW_C;         // 0-based first column of this lane; DP column = j0+c+1
            int4 *bnd_out = bnd + P.bnd_off + (int64_t)(sidx % nbuf) * (la + 2);
            const int4 *bnd_in = bnd + P.bnd_off + (int64_t)((sidx + nbuf - 1) % nbuf) * (la + 2);
            const bool wait_prev = (team_warps > 1 && sidx > 0);
            int Mu[KW_C], Xu[KW_C], Yu[KW_C], brow[KW_C];
#pragma unroll
            for (int c = 0; c < KW_C; ++c) {
                const int j = j0 + c + 1;
                const int ls = j * se + (so - se);             // SA.rs:100-104
                Mu[c] = ls + sent; Xu[c] = ls; Yu[c] = ls + sent;
                brow[c] = (j <= lb) ? (int)b[j - 1] : (nl - 1);
            }
            int dM0, dX0, dY0;                                 // diag of column j0 at row 0: cell (0, j0)
            if (j0 == 0) { dM0 = 0; dX0 = -unit; dY0 = -unit; }            // SA.rs:88-93
            else { const int ls = j0 * se + (so - se); dM0 = ls + sent; dX0 = ls; dY0 = ls + sent; }
            int lastM = 0, lastX = 0, lastY = 0;               // this lane's last column, row of the previous step
            uint8_t *dstrip = DIRS ? dir + P.dir_off + (int64_t)sidx * (la + 31) * KW_SW + lane * KW_C : nullptr;

            const int n_steps = la + 31;
            for (int t = 0; t < n_steps; ++t) {
                if (wait_prev && (t & 31) == 0 && t < la) {    // rows t+1 .. t+32 are consumed by lane 0 in the next 32 steps
                    const int need = min(la, t + 32);
                    if (lane == 0) {
                        while (ld_volatile(prog + sidx - 1) < need) __nanosleep(32);
                        __threadfence();
                    }
                    __syncwarp();
                }
                const int i = t - lane + 1;                    // DP row of this lane in this step
                int tM = __shfl_up_sync(0xffffffffu, lastM, 1);
                int tX = __shfl_up_sync(0xffffffffu, lastX, 1);
                int tY = __shfl_up_sync(0xffffffffu, lastY, 1);
                const bool active = (i >= 1 && i <= la);
                if (lane == 0 && active) {
                    if (sidx == 0) { const int ls = i * se + (so - se); tM = ls + sent; tX = ls + sent; tY = ls; }   // SA.rs:109-113
                    else { const int4 v = __ldcg(bnd_in + i); tM = v.x; tX = v.y; tY = v.z; }
                }
                if (active) {
                    const int ra = (int)a[i - 1] * nl;
                    const int po = (MODE != SWB_LOCAL && i == la) ? so : o;      // SA.rs:166-167
                    const int pe = (MODE != SWB_LOCAL && i == la) ? se : e;
                    int dM = dM0, dX = dX0, dY = dY0;
                    dM0 = tM; dX0 = tX; dY0 = tY;                                 // next row's diag for column 0
                    unsigned long long codes = 0ull;
#pragma unroll
                    for (int c = 0; c < KW_C; ++c) {
                        const int j = j0 + c + 1;
                        const int qo = (MODE != SWB_LOCAL && j == lb) ? so : o;  // SA.rs:169-170
                        const int qe = (MODE != SWB_LOCAL && j == lb) ? se : e;
                        const int s = mat_sh[ra + brow[c]];
                        int m, x, y; unsigned code;
                        kw_cell<MODE>(dM, dX, dY, tM, tX, tY, Mu[c], Xu[c], Yu[c], s, po, pe, qo, qe, m, x, y, code);
                        dM = Mu[c]; dX = Xu[c]; dY = Yu[c];
                        Mu[c] = m; Xu[c] = x; Yu[c] = y;
                        tM = m; tX = x; tY = y;
                        codes |= (unsigned long long)code << (8 * c);
                        if (j <= lb) {
                            if (MODE == SWB_LOCAL) {
                                if (m > best || (m == best && m > 0 && i < bi)) { best = m; bi = i; bj = j; }
                            } else if (i == la && j == lb) {
                                fin_score = m; fin_state = 0;                     // SA.rs:302-308: first strict max, order M,X,Y
                                if (x > fin_score) { fin_score = x; fin_state = 1; }
                                if (y > fin_score) { fin_score = y; fin_state = 2; }
                                have_fin = true;
                            }
                        }
                    }
                    lastM = Mu[KW_C - 1]; lastX = Xu[KW_C - 1]; lastY = Yu[KW_C - 1];
                    if (lane == 31 && sidx + 1 < n_strips) {
                        bnd_out[i] = make_int4(lastM, lastX, lastY, 0);
                        if (team_warps > 1 && ((i & 31) == 0 || i == la)) {       // publish rows 1..i of this strip's last column
                            __threadfence();
                            *reinterpret_cast<volatile int *>(prog + sidx) = i;
                        }
                    }
                    if (DIRS) *reinterpret_cast<unsigned long long *>(dstrip + (int64_t)t * KW_SW) = codes;
                }
            }
            __syncwarp();
        }

        if (MODE == SWB_LOCAL) {
            // warp argmax with the reference's scan order: score desc, row asc, column asc
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) {
                const int ob = __shfl_xor_sync(0xffffffffu, best, d);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, d);
                const int oj = __shfl_xor_sync(0xffffffffu, bj, d);
                const bool take = (ob > best) || (ob == best && ob > 0 && (oi < bi || (oi == bi && oj < bj)));
                if (take) { best = ob; bi = oi; bj = oj; }
            }
            if (team_warps == 1) {
                if (lane == 0) { KWResult R; R.score = best; R.end_i = (best > 0) ? bi : -1; R.end_j = (best > 0) ? bj : -1; R.state = 0; res[pidx] = R; }
            } else {
                int arrived = 0;
                if (lane == 0) {
                    cand[w * 4 + 0] = best; cand[w * 4 + 1] = bi; cand[w * 4 + 2] = bj;
                    __threadfence();
                    arrived = atomicAdd(counter, 1);
                }
                arrived = __shfl_sync(0xffffffffu, arrived, 0);
                if (arrived == team_warps - 1) {                                  // last warp of the team: merge the candidates
                    __threadfence();
                    int cb = 0, ci = -1, cj = -1;
                    for (int k = lane; k < team_warps; k += 32) {
                        const int ob = ld_volatile(cand + k * 4), oi = ld_volatile(cand + k * 4 + 1), oj = ld_volatile(cand + k * 4 + 2);
                        if (ob > cb || (ob == cb && ob > 0 && (oi < ci || (oi == ci && oj < cj)))) { cb = ob; ci = oi; cj = oj; }
                    }
#pragma unroll
                    for (int d = 16; d >= 1; d >>= 1) {
                        const int ob = __shfl_xor_sync(0xffffffffu, cb, d);
                        const int oi = __shfl_xor_sync(0xffffffffu, ci, d);
                        const int oj = __shfl_xor_sync(0xffffffffu, cj, d);
                        if (ob > cb || (ob == cb && ob > 0 && (oi < ci || (oi == ci && oj < cj)))) { cb = ob; ci = oi; cj = oj; }
                    }
                    if (lane == 0) { KWResult R; R.score = cb; R.end_i = (cb > 0) ? ci : -1; R.end_j = (cb > 0) ? cj : -1; R.state = 0; res[pidx] = R; }
                }
            }
        } else {
            const unsigned who = __ballot_sync(0xffffffffu, have_fin);
            if (who) {                                                            // exactly one lane of one team warp owns (la, lb)
                const int src = __ffs(who) - 1;
                KWResult R;
                R.score = __shfl_sync(0xffffffffu, fin_score, src);
                R.state = __shfl_sync(0xffffffffu, fin_state, src);
                R.end_i = la; R.end_j = lb;
                if (lane == 0) res[pidx] = R;
            }
        }
    }
}

// Host-side plan: team size, grid, scratch offsets.  Fills bnd_off / scr_off of the pairs.
// Resident blocks per SM, minimum over all instantiations (teams spin on each other, so every block of a team launch
// must be resident at once; the value is queried from the runtime, not assumed).
static int kw_blocks_per_sm(const swb_engine *e)
{
    static int cached[64] = {0};
    const int nl = e->n + 1;
    if (cached[nl]) return cached[nl];
    const size_t smem = (size_t)nl * nl * sizeof(int);
    int best = 1 << 30, v = 0;
#define KW_OCC(M, D) if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, kw_kernel<M, D>, KW_WARPS * 32, smem) == cudaSuccess) best = std::min(best, v); else best = 0;
    KW_OCC(SWB_GLOBAL, true) KW_OCC(SWB_GLOBAL, false) KW_OCC(SWB_GLOCAL, true) KW_OCC(SWB_GLOCAL, false)
    KW_OCC(SWB_LOCAL, true) KW_OCC(SWB_LOCAL, false)
#undef KW_OCC
    cached[nl] = std::max(1, std::min(best, 8));
    return cached[nl];
}

void kw_plan(const swb_engine *e, KWPair *pairs, int n_pairs, KWPlan *plan)
{
    const int capacity = e->sm_count * kw_blocks_per_sm(e) * KW_WARPS;   // warps that are guaranteed to be co-resident
    int T = 1;
    if (n_pairs > 0 && n_pairs < capacity / 4) {
        int useful = 1;
        for (int k = 0; k < n_pairs; ++k) {
            const int strips = (pairs[k].lb + KW_SW - 1) / KW_SW;
            useful = std::max(useful, std::min(strips, std::max(1, pairs[k].la / 128)));
        }
        T = std::max(1, std::min(capacity / n_pairs, useful));
    }
    plan->team_warps = T;
    plan->n_teams = (T == 1) ? n_pairs : std::max(1, std::min(n_pairs, capacity / T));
    const int64_t warps = (int64_t)plan->n_teams * T;
    plan->blocks = (int)((warps + KW_WARPS - 1) / KW_WARPS);
    int64_t bnd = 0, scr = 0;
    for (int k = 0; k < n_pairs; ++k) {
        const int strips = (pairs[k].lb + KW_SW - 1) / KW_SW;
        pairs[k].bnd_off = bnd; pairs[k].scr_off = scr;
        bnd += (int64_t)std::min(strips, T + 1) * (pairs[k].la + 2);
        scr += 1 + strips + 4 * (int64_t)T;
    }
    plan->bnd_elems = std::max<int64_t>(bnd, 1); plan->scratch_ints = std::max<int64_t>(scr, 1);
}

template <int MODE>
static cudaError_t kw_launch_mode(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pairs,
                                  int n_pairs, const KWPlan &pl, uint8_t *d_dir, int4 *d_bnd, int *d_scratch,
                                  KWResult *d_res, cudaStream_t st)
{
    const int nl = e->n + 1;
    const size_t smem = (size_t)nl * nl * sizeof(int);
    cudaError_t err = cudaMemsetAsync(d_scratch, 0, (size_t)pl.scratch_ints * sizeof(int), st);
    if (err != cudaSuccess) return err;
    if (d_dir)
        kw_kernel<MODE, true><<<pl.blocks, KW_WARPS * 32, smem, st>>>(d_a, d_b, d_pairs, n_pairs, pl.team_warps, pl.n_teams,
                                                                      e->d_mat, nl, e->o, e->e, e->scale, d_dir, d_bnd,
                                                                      d_scratch, d_res);
    else
        kw_kernel<MODE, false><<<pl.blocks, KW_WARPS * 32, smem, st>>>(d_a, d_b, d_pairs, n_pairs, pl.team_warps, pl.n_teams,
                                                                       e->d_mat, nl, e->o, e->e, e->scale, d_dir, d_bnd,
                                                                       d_scratch, d_res);
    return cudaGetLastError();
}

cudaError_t kw_launch(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pairs, int n_pairs,
                      const KWPlan &plan, uint8_t *d_dir, int4 *d_bnd, int *d_scratch, KWResult *d_res, cudaStream_t st)
{
    if (n_pairs <= 0) return cudaSuccess;
    switch (e->mode) {
    case SWB_GLOBAL: return kw_launch_mode<SWB_GLOBAL>(e, d_a, d_b, d_pairs, n_pairs, plan, d_dir, d_bnd, d_scratch, d_res, st);
    case SWB_GLOCAL: return kw_launch_mode<SWB_GLOCAL>(e, d_a, d_b, d_pairs, n_pairs, plan, d_dir, d_bnd, d_scratch, d_res, st);
    default: return kw_launch_mode<SWB_LOCAL>(e, d_a, d_b, d_pairs, n_pairs, plan, d_dir, d_bnd, d_scratch, d_res, st);
    }
}

// ------------------------------------------------------------------------------------------------
// KT: backtrack (SA.rs:346-383).  One thread per pair follows the 2-bit predecessor codes from the
// start cell and emits the column lists in walk order (end -> start); the host reverses them
// (SA.rs:384-385).  idx value -1 = gap.  n_cols < 0 reports the reference's panic cases.
__device__ __forceinline__ unsigned kt_code(const uint8_t *dir, int64_t dir_off, int la, int i, int j, int state)
{
    const int jj = j - 1;
    const int sidx = jj / KW_SW, within = jj - sidx * KW_SW;
    const int t = (i - 1) + within / KW_C;
    const uint8_t byte = dir[dir_off + ((int64_t)sidx * (la + 31) + t) * KW_SW + within];
    return (byte >> (2 * state)) & 3u;
}

template <int MODE>
__global__ void kt_kernel(const KWPair *__restrict__ pairs, const KWResult *__restrict__ res, int n_pairs,
                          const uint8_t *__restrict__ dir, const int64_t *__restrict__ out_off,
                          int32_t *__restrict__ idx_a, int32_t *__restrict__ idx_b, int32_t *__restrict__ ncols, int sent)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    const KWPair P = pairs[p];
    const KWResult R = res[p];
    int32_t *xa = idx_a + out_off[p], *xb = idx_b + out_off[p];
    int n = 0;
    if (R.end_i < 0) { ncols[p] = 0; return; }                     // SA.rs:346-348
    int cx = R.end_i, cy = R.end_j, st = R.state;
    for (;;) {
        unsigned prev;
        if (cx >= 1 && cy >= 1) prev = kt_code(dir, P.dir_off, P.la, cx, cy, st);
        else if (cx == 0 && cy == 0) prev = 0;                     // only reachable when la == lb == 0
        else if (cx == 0) {                                        // row 0: prev = GAPINX for all states (SA.rs:105-107)
            prev = 1;
            if (MODE == SWB_LOCAL && (st == 1 || sent == 0)) prev = 3;   // X(0,j) = 0 in local mode -> stop (SA.rs:354-357)
        } else {                                                   // column 0: prev = GAPINY (SA.rs:114-116)
            prev = 2;
            if (MODE == SWB_LOCAL && (st == 2 || sent == 0)) prev = 3;
        }
        if (MODE == SWB_LOCAL && prev == 3) break;
        if (st == 0) {
            if (cx == 0 || cy == 0) { n = -2; break; }             // reference underflows / panics
            xa[n] = cx - 1; xb[n] = cy - 1; ++n; --cx; --cy;
        } else if (st == 1) {
            if (cy == 0) { n = -2; break; }
            xa[n] = -1; xb[n] = cy - 1; ++n; --cy;
        } else {
            if (cx == 0) { n = -2; break; }                        // SA.rs:369-371
            xa[n] = cx - 1; xb[n] = -1; ++n; --cx;
        }
        if (cx == 0 && cy == 0) break;
        st = (int)prev;
    }
    ncols[p] = n;
}

cudaError_t kt_launch(const swb_engine *e, const KWPair *d_pairs, const KWResult *d_res, int n_pairs,
                      const uint8_t *d_dir, const int64_t *d_out_off, int32_t *d_idx_a, int32_t *d_idx_b,
                      int32_t *d_ncols, cudaStream_t st)
{
    if (n_pairs <= 0) return cudaSuccess;
    const int threads = 32, blocks = (n_pairs + threads - 1) / threads;
    const int sent = 10 * e->o + 10 * e->e;
    switch (e->mode) {
    case SWB_GLOBAL: kt_kernel<SWB_GLOBAL><<<blocks, threads, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, d_out_off, d_idx_a, d_idx_b, d_ncols, sent); break;
    case SWB_GLOCAL: kt_kernel<SWB_GLOCAL><<<blocks, threads, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, d_out_off, d_idx_a, d_idx_b, d_ncols, sent); break;
    default: kt_kernel<SWB_LOCAL><<<blocks, threads, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, d_out_off, d_idx_a, d_idx_b, d_ncols, sent); break;
    }
    return cudaGetLastError();
}
