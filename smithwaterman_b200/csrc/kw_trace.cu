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
#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>

#include <cuda/atomic>

#include "swb_internal.h"

#define KT_AHEAD 16

int64_t kw_dir_bytes(int32_t la, int32_t lb, int cols)
{
    const int sw = 32 * cols;
    int64_t strips = (lb + sw - 1) / sw;
    return strips * (int64_t)(la + 31) * sw;
}

#include "kw_cell.cuh"

// Team scheduling: a pair is swept by a TEAM of T warps (T = 1 for batches of many pairs; up to several hundred for a
// few long pairs).  Warp w of the team takes the column strips w, w+T, w+2T, ...; strip s consumes the last column of
// strip s-1 row by row through a ring of T+1 column buffers in global memory (L2 resident) guarded by a per-strip
// progress counter (rows published: a release store of cuda::atomic_ref<int, thread_scope_device> after the rows are
// written, an acquire load before they are read -- not the fence-less spinning of the reference, OCL.rs:374-376).  Waits
// only ever point to a lower strip, so there is no cycle, and every warp of every team is resident: team launches are
// COOPERATIVE launches (an oversubscribed grid is an error, not a hang), sized from the occupancy of the exact
// instantiation that is launched, and serialised among themselves through the engine's team event.
// C = columns per lane (strip width 32*C): 8 for batches (fewest instructions per cell), 2 when a few long pairs must be
// spread over many warps (the per-step dependency chain, not the issue rate, bounds a lone warp).
typedef cuda::atomic_ref<int, cuda::thread_scope_device> kw_flag;

template <int C> struct KWCodes;
template <> struct KWCodes<8> { typedef unsigned long long T; };
template <> struct KWCodes<4> { typedef unsigned T; };
template <> struct KWCodes<2> { typedef unsigned short T; };

// BAND: the (single) pair is a row band of a larger matrix (KWBand, swb_internal.h): absolute rows = band.row0 + i, the
// top row comes from a checkpoint, checkpoint rows are written on the way (checkpointed traceback, align.cu).
template <int MODE, bool DIRS, int C, bool BAND>
__global__ void __launch_bounds__(KW_WARPS * 32)
kw_kernel(const uint8_t *__restrict__ a_all, const uint8_t *__restrict__ b_all, const KWPair *__restrict__ pairs,
          int n_pairs, int team_warps, int n_teams, const int32_t *__restrict__ mat, int nl, int o, int e, int unit,
          uint8_t *__restrict__ dir, int4 *__restrict__ bnd, int *__restrict__ scratch, KWResult *__restrict__ res,
          const int32_t *__restrict__ pssm_all, const KWBand band)
{
    constexpr int SW = 32 * C;
    typedef typename KWCodes<C>::T codes_t;
    extern __shared__ int mat_sh[];
    __shared__ int4 bnd_sh[KW_WARPS][32];
    __shared__ uint8_t a_sh[KW_WARPS][64];
    // local mode: PAD row/column scores are made so negative that a padded column's MATCH is always clamped to 0 and can
    // never become a strict maximum (the int32 lanes have the room; K1's s16 table keeps the small PAD of the engine)
    for (int k = threadIdx.x; k < nl * nl; k += blockDim.x)
        mat_sh[k] = (MODE == SWB_LOCAL && (k % nl == nl - 1 || k / nl == nl - 1)) ? NEG_INF_I32 : mat[k];
    __syncthreads();
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int gw = blockIdx.x * KW_WARPS + wib;
    const int team = gw / team_warps, w = gw - team * team_warps;
    if (team >= n_teams) return;
    const int so = (MODE == SWB_GLOBAL) ? o : 0, se = (MODE == SWB_GLOBAL) ? e : 0;
    const int sent = 10 * o + 10 * e;

    for (int pidx = team; pidx < n_pairs; pidx += n_teams) {
        const KWPair P = pairs[pidx];
        const uint8_t *a = a_all + P.a_off, *b = b_all + P.b_off;
        const int la = P.la, lb = P.lb;
        const int row0 = BAND ? band.row0 : 0;                 // absolute row = row0 + i
        const int la_abs = BAND ? band.la_full : la, lb_abs = BAND ? band.lb_full : lb;
        // PositionSpecificMatrix (SA.rs:583-623): score of (row i, column j) = scores[(i-1) + (j-1)*la]
        const int32_t *pssm = (pssm_all != nullptr && P.pssm_off >= 0) ? pssm_all + P.pssm_off : nullptr;
        const int n_strips = (lb + SW - 1) / SW;
        const int nbuf = min(n_strips, team_warps + 1);
        int *counter = scratch + P.scr_off;                    // [0] arrivals, [1..n_strips] rows published per strip,
        int *prog = counter + 1;                               // then 4 ints per team warp (local-mode candidates)
        int *cand = prog + n_strips;
        int best = 0, bi = -1, bj = -1;                        // local: first strict maximum of MATCH
        int fin_score = 0, fin_state = 0;                      // non-local: cell (la, lb)
        bool have_fin = false;

        for (int sidx = w; sidx < n_strips; sidx += team_warps) {
            const int j0 = sidx * SW + lane * C;               // 0-based first column of this lane; DP column = j0+c+1
            int4 *bnd_out = bnd + P.bnd_off + (int64_t)(sidx % nbuf) * (la + 2);
            const int4 *bnd_in = bnd + P.bnd_off + (int64_t)((sidx + nbuf - 1) % nbuf) * (la + 2);
            const bool wait_prev = (team_warps > 1 && sidx > 0);
            // per column: the three states of the cell above, their max / priority argmax (the next row's diagonal),
            // the column's offset into the score table
            int Mu[C], Xu[C], Yu[C], Hu[C], boff[C];
            unsigned Au[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const int j = j0 + c + 1;
                const int ls = j * se + (so - se);             // SA.rs:100-104: X = ls, M = Y = ls + sentinel, prev = GAPINX
                Mu[c] = ls + sent; Xu[c] = ls; Yu[c] = ls + sent;
                if (BAND && band.ckpt_in != nullptr && j <= lb) { const int4 v = band.ckpt_in[j]; Mu[c] = v.x; Xu[c] = v.y; Yu[c] = v.z; }
                Hu[c] = __vimax3_s32(Mu[c], Xu[c], Yu[c]); Au[c] = kw_am(Mu[c], Xu[c], Yu[c]);
                boff[c] = (j <= lb) ? (pssm ? 0 : (int)b[j - 1]) : (nl - 1);
            }
            int hd0; unsigned ad0;                             // diagonal of column 0 at the next row: cell (row-1, j0)
            if (j0 == 0) { hd0 = 0; ad0 = 0u; }                // (0,0): M = 0, X = Y = -1 (SA.rs:88-93)
            else { const int ls = j0 * se + (so - se); hd0 = __vimax3_s32(ls + sent, ls, ls + sent); ad0 = kw_am(ls + sent, ls, ls + sent); }
            if (BAND && row0 > 0) {                            // the diagonal of the band's first row is cell (row0, j0)
                if (j0 == 0) { const int ls = row0 * se + (so - se); hd0 = __vimax3_s32(ls + sent, ls + sent, ls); ad0 = kw_am(ls + sent, ls + sent, ls); }   // SA.rs:109-113
                else if (j0 <= lb) { const int4 v = band.ckpt_in[j0]; hd0 = __vimax3_s32(v.x, v.y, v.z); ad0 = kw_am(v.x, v.y, v.z); }
            }
            int lastM = 0, lastX = 0, lastY = 0;               // this lane's last column, row of the previous step
            uint8_t *dstrip = DIRS ? dir + P.dir_off + (int64_t)sidx * (la + 31) * SW + lane * C : nullptr;
            int sb = 0, si = -1, sj = -1;                      // local: this strip's first strict maximum in this lane
            const int own = lb_abs - 1 - j0;                   // register index of column lb when it is in this lane

            const int n_steps = la + 31;
            for (int t = 0; t < n_steps; ++t) {
                if ((t & 31) == 0) {
                    // Once per 32 steps the warp stages what the next 32 steps consume, so that no global-memory latency
                    // sits on the per-step critical path: the 32 rows of the previous strip's last column that lane 0 will
                    // read (one coalesced 512-byte load) and the next 32 residues of `a` (ring of 64: lane l reads index t-l).
                    if (wait_prev && t < la) {                 // rows t+1 .. t+32 are consumed by lane 0 in the next 32 steps
                        const int need = min(la, t + 32);
                        if (lane == 0) {
                            kw_flag f(prog[sidx - 1]);
                            while (f.load(cuda::memory_order_acquire) < need) __nanosleep(32);
                        }
                    }
                    __syncwarp();
                    if (sidx > 0 && t + 1 + lane <= la) bnd_sh[wib][lane] = __ldcg(bnd_in + t + 1 + lane);
                    a_sh[wib][(t + lane) & 63] = (t + lane < la) ? a[t + lane] : (uint8_t)(nl - 1);
                    __syncwarp();
                }
                const int i = t - lane + 1;                    // DP row of this lane in this step
                int tM = __shfl_up_sync(0xffffffffu, lastM, 1);
                int tX = __shfl_up_sync(0xffffffffu, lastX, 1);
                int tY = __shfl_up_sync(0xffffffffu, lastY, 1);
                const bool active = (i >= 1 && i <= la);
                if (lane == 0 && active) {
                    if (sidx == 0) { const int ls = (row0 + i) * se + (so - se); tM = ls + sent; tX = ls + sent; tY = ls; }   // SA.rs:109-113
                    else { const int4 v = bnd_sh[wib][t & 31]; tM = v.x; tX = v.y; tY = v.z; }
                }
                if (active) {
                    const int *mrow = mat_sh + (int)a_sh[wib][(i - 1) & 63] * nl;
                    const int po = (MODE == SWB_GLOCAL && row0 + i == la_abs) ? so : o;     // SA.rs:166-167 (global: so == o)
                    const int pe = (MODE == SWB_GLOCAL && row0 + i == la_abs) ? se : e;
                    int hd = hd0; unsigned ad = ad0;
                    hd0 = __vimax3_s32(tM, tX, tY); ad0 = kw_am(tM, tX, tY);     // next row's diagonal for column 0
                    codes_t codes = 0;
                    const int sb_in = sb;
                    int sv[C];
                    if (pssm) {
#pragma unroll
                        for (int c = 0; c < C; ++c)
                            sv[c] = (boff[c] != nl - 1) ? __ldg(pssm + (i - 1) + (int64_t)(j0 + c) * la) : NEG_INF_I32;
                    } else {
#pragma unroll
                        for (int c = 0; c < C; ++c) sv[c] = mrow[boff[c]];
                    }
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        const int qo = (MODE == SWB_GLOCAL && c == own) ? so : o; // SA.rs:169-170 (column lb)
                        const int qe = (MODE == SWB_GLOCAL && c == own) ? se : e;
                        int m, x, y; unsigned code;
                        kw_cell<MODE>(hd, ad, tM, tX, tY, Mu[c], Xu[c], Yu[c], sv[c], po, pe, qo, qe, m, x, y, code);
                        hd = Hu[c]; ad = Au[c];
                        Mu[c] = m; Xu[c] = x; Yu[c] = y;
                        Hu[c] = __vimax3_s32(m, x, y); Au[c] = kw_am(m, x, y);
                        tM = m; tX = x; tY = y;
                        if (DIRS) codes |= (codes_t)code << (8 * c);
                        // padded columns (j > lb) score PAD: their MATCH is clamped to 0 and can never be a strict maximum
                        if (MODE == SWB_LOCAL && m > sb) { sb = m; sj = j0 + c + 1; }
                    }
                    if (MODE == SWB_LOCAL && sb != sb_in) si = i;
                    if (MODE != SWB_LOCAL && i == la && own >= 0 && own < C) {    // cell (la, lb): SA.rs:302-308, first strict max M,X,Y
#pragma unroll
                        for (int c = 0; c < C; ++c)
                            if (c == own) {
                                fin_score = Mu[c]; fin_state = 0;
                                if (Xu[c] > fin_score) { fin_score = Xu[c]; fin_state = 1; }
                                if (Yu[c] > fin_score) { fin_score = Yu[c]; fin_state = 2; }
                            }
                        have_fin = true;
                    }
                    lastM = Mu[C - 1]; lastX = Xu[C - 1]; lastY = Yu[C - 1];
                    if (BAND && band.ckpt_out != nullptr && (row0 + i) % band.every == 0 && row0 + i < la_abs) {
                        int4 *crow = band.ckpt_out + (int64_t)((row0 + i) / band.every - 1) * band.stride;
#pragma unroll
                        for (int c = 0; c < C; ++c)
                            if (j0 + c + 1 <= lb) crow[j0 + c + 1] = make_int4(Mu[c], Xu[c], Yu[c], 0);
                    }
                    if (lane == 31 && sidx + 1 < n_strips) {
                        bnd_out[i] = make_int4(lastM, lastX, lastY, 0);
                        if (team_warps > 1 && ((i & 31) == 0 || i == la))         // publish rows 1..i of this strip's last column
                            kw_flag(prog[sidx]).store(i, cuda::memory_order_release);
                    }
                    if (DIRS) *reinterpret_cast<codes_t *>(dstrip + (int64_t)t * SW) = codes;
                }
            }
            __syncwarp();
            // fold this strip's candidate into the lane's best: score desc, row asc, column asc (SA.rs:286-294 scan order)
            if (MODE == SWB_LOCAL && (sb > best || (sb == best && sb > 0 && (si < bi || (si == bi && sj < bj))))) { best = sb; bi = si; bj = sj; }
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
                    arrived = kw_flag(*counter).fetch_add(1, cuda::memory_order_acq_rel);   // releases this warp's candidate, acquires the earlier ones
                }
                arrived = __shfl_sync(0xffffffffu, arrived, 0);
                __syncwarp();
                if (arrived == team_warps - 1) {                                  // last warp of the team: merge the candidates
                    int cb = 0, ci = -1, cj = -1;
                    for (int k = lane; k < team_warps; k += 32) {
                        const int ob = __ldcg(cand + k * 4), oi = __ldcg(cand + k * 4 + 1), oj = __ldcg(cand + k * 4 + 2);
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

// Host-side plan: columns per lane, team size, grid, scratch and direction-byte offsets.  Fills bnd_off / scr_off /
// dir_off of the pairs.  Teams spin on each other, so every block of a team launch must be resident at once: the
// resident blocks per SM are those of the EXACT instantiation that will be launched (mode, direction bytes, strip
// width, band form -- their register counts differ), queried from the runtime per device, and the launch is cooperative.
template <int MODE, bool DIRS, int C, bool BAND>
static const void *kw_fn() { return reinterpret_cast<const void *>(&kw_kernel<MODE, DIRS, C, BAND>); }

static const void *kw_pick(int mode, bool dirs, int cols, bool band)
{
#define KW_P(M) (dirs ? (cols == 2 ? (band ? kw_fn<M, true, 2, true>() : kw_fn<M, true, 2, false>()) : (band ? kw_fn<M, true, 8, true>() : kw_fn<M, true, 8, false>())) \
                      : (cols == 2 ? (band ? kw_fn<M, false, 2, true>() : kw_fn<M, false, 2, false>()) : (band ? kw_fn<M, false, 8, true>() : kw_fn<M, false, 8, false>())))
    switch (mode) {
    case SWB_GLOBAL: return KW_P(SWB_GLOBAL);
    case SWB_GLOCAL: return KW_P(SWB_GLOCAL);
    default: return KW_P(SWB_LOCAL);
    }
#undef KW_P
}

int kw_blocks_per_sm(const swb_engine *e, int cols, bool dirs, bool band)
{
    static std::mutex mu;
    static std::map<std::tuple<int, int, int, int, int, int>, int> cache;       // (device, mode, dirs, cols, band, nl)
    const int nl = e->n + 1;
    const auto key = std::make_tuple(e->device, e->mode, (int)dirs, cols == 2 ? 2 : 8, (int)band, nl);
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    int v = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, kw_pick(e->mode, dirs, cols == 2 ? 2 : 8, band), KW_WARPS * 32, (size_t)nl * nl * sizeof(int)) != cudaSuccess) { cudaGetLastError(); v = 1; }
    v = std::max(1, std::min(v, 8));
    cache[key] = v;
    return v;
}

void kw_plan(const swb_engine *e, KWPair *pairs, int n_pairs, bool with_dirs, KWPlan *plan, int force_cols, bool band)
{
    // Strip width.  Score only: 8 columns per lane -- with whole-round teams (cost model below) a step of 8 cells per lane
    // costs the same 0.44 us as a step of 2 (the step is bound by its shuffle / staging latency), so the wide strip wins:
    // 100 kb pair 49 ms vs 71 ms.  With direction bytes a step of 8 cells is ~340 instructions of one warp and no longer
    // latency-bound (0.71 us vs 0.51 us per step): a few long pairs then run better on 2-column strips and 4x the warps.
    int64_t strips8 = 0;
    for (int k = 0; k < n_pairs; ++k) strips8 += (pairs[k].lb + 255) / 256;
    int cols = (with_dirs && n_pairs > 0 && strips8 * 3 < e->sm_count * kw_blocks_per_sm(e, 8, with_dirs, band) * KW_WARPS) ? 2 : 8;
    if (force_cols == 2 || force_cols == 8) cols = force_cols;
    if (const char *f = getenv("SWB_KW_COLS")) { if (atoi(f) == 2) cols = 2; else if (atoi(f) == 8) cols = 8; }   // tests force both widths
    const int capacity = e->sm_count * kw_blocks_per_sm(e, cols, with_dirs, band) * KW_WARPS;   // warps that a cooperative launch keeps co-resident
    const int sw = 32 * cols;
    int T = 1;
    if (n_pairs > 0 && n_pairs < capacity / 4) {
        // A team of T warps sweeps a pair in ceil(strips / T) rounds; warp w trails warp w-1 by 32 rows, so the sweep lasts
        // about rounds * la + 32 * T steps.  Take the T (a whole number of rounds) that minimises it for the largest pair.
        int big = 0;
        for (int k = 1; k < n_pairs; ++k)
            if ((int64_t)pairs[k].la * pairs[k].lb > (int64_t)pairs[big].la * pairs[big].lb) big = k;
        const int strips = (pairs[big].lb + sw - 1) / sw;
        int cap = std::max(1, std::min(capacity / n_pairs, strips));
        // the boundary ring holds T + 1 columns of 16 B per row: keep it under 32 GB for very long pairs (la > 0.8 M rows)
        cap = (int)std::max<int64_t>(1, std::min<int64_t>(cap, (32ll << 30) / (16ll * (pairs[big].la + 2)) - 1));
        int64_t best_cost = -1;
        for (int rounds = 1; rounds <= strips; ++rounds) {
            const int t = (strips + rounds - 1) / rounds;
            if (t > cap) continue;
            const int64_t cost = (int64_t)rounds * pairs[big].la + 32ll * t;
            if (best_cost < 0 || cost < best_cost) { best_cost = cost; T = t; }
            if (t == 1) break;
        }
        if (const char *f = getenv("SWB_KW_TEAM")) { const int v = atoi(f); if (v >= 1) T = std::min(v, cap); }   // tuning aid
    }
    plan->cols = cols;
    plan->team_warps = T;
    plan->n_teams = (T == 1) ? n_pairs : std::max(1, std::min(n_pairs, capacity / T));
    const int64_t warps = (int64_t)plan->n_teams * T;
    plan->blocks = (int)((warps + KW_WARPS - 1) / KW_WARPS);
    int64_t bnd = 0, scr = 0, dirb = 0;
    for (int k = 0; k < n_pairs; ++k) {
        const int strips = (pairs[k].lb + sw - 1) / sw;
        pairs[k].bnd_off = bnd; pairs[k].scr_off = scr;
        pairs[k].dir_off = with_dirs ? dirb : -1;
        bnd += (int64_t)std::min(strips, T + 1) * (pairs[k].la + 2);
        scr += 1 + strips + 4 * (int64_t)T;
        if (with_dirs) dirb += kw_dir_bytes(pairs[k].la, pairs[k].lb, cols);
    }
    plan->bnd_elems = std::max<int64_t>(bnd, 1); plan->scratch_ints = std::max<int64_t>(scr, 1);
    plan->dir_bytes = dirb;
}

// Team launches (team_warps > 1) are cooperative: the runtime refuses a grid that cannot be resident at once (instead of
// letting its warps wait for blocks that are never scheduled), and they are chained through the engine's team event so
// that two of them never share the GPU, whatever streams they were issued on.
template <int MODE, bool DIRS, int C, bool BAND>
static cudaError_t kw_go(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pairs, int n_pairs,
                         const KWPlan &pl, const KWBand &band, uint8_t *d_dir, int4 *d_bnd, int *d_scratch, KWResult *d_res,
                         cudaStream_t st, const int32_t *d_pssm)
{
    const int nl = e->n + 1;
    const size_t smem = (size_t)nl * nl * sizeof(int);
    cudaError_t err = cudaMemsetAsync(d_scratch, 0, (size_t)pl.scratch_ints * sizeof(int), st);
    if (err != cudaSuccess) return err;
    if (pl.team_warps <= 1) {
        kw_kernel<MODE, DIRS, C, BAND><<<pl.blocks, KW_WARPS * 32, smem, st>>>(d_a, d_b, d_pairs, n_pairs, pl.team_warps, pl.n_teams, e->d_mat, nl,
                                                                             e->o, e->e, e->scale, d_dir, d_bnd, d_scratch, d_res, d_pssm, band);
        return cudaGetLastError();
    }
    if (pl.blocks > e->sm_count * kw_blocks_per_sm(e, C, DIRS, BAND)) return cudaErrorCooperativeLaunchTooLarge;
    swb_engine *me = const_cast<swb_engine *>(e);
    if (!me->ev_team) { if ((err = cudaEventCreateWithFlags(&me->ev_team, cudaEventDisableTiming)) != cudaSuccess) return err; }
    else if ((err = cudaStreamWaitEvent(st, me->ev_team, 0)) != cudaSuccess) return err;
    int team_warps = pl.team_warps, n_teams = pl.n_teams, o = e->o, ee = e->e, unit = e->scale, nl_ = nl;
    const int32_t *mat = e->d_mat;
    KWBand bnd_arg = band;
    void *args[] = {(void *)&d_a, (void *)&d_b, (void *)&d_pairs, (void *)&n_pairs, (void *)&team_warps, (void *)&n_teams, (void *)&mat, (void *)&nl_,
                    (void *)&o, (void *)&ee, (void *)&unit, (void *)&d_dir, (void *)&d_bnd, (void *)&d_scratch, (void *)&d_res, (void *)&d_pssm, (void *)&bnd_arg};
    err = cudaLaunchCooperativeKernel(reinterpret_cast<const void *>(&kw_kernel<MODE, DIRS, C, BAND>), dim3(pl.blocks), dim3(KW_WARPS * 32), args, smem, st);
    if (err != cudaSuccess) return err;
    return cudaEventRecord(me->ev_team, st);
}

template <int MODE, int C>
static cudaError_t kw_launch_mode(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pairs,
                                  int n_pairs, const KWPlan &pl, uint8_t *d_dir, int4 *d_bnd, int *d_scratch,
                                  KWResult *d_res, cudaStream_t st, const int32_t *d_pssm)
{
    if (d_dir) return kw_go<MODE, true, C, false>(e, d_a, d_b, d_pairs, n_pairs, pl, KWBand(), d_dir, d_bnd, d_scratch, d_res, st, d_pssm);
    return kw_go<MODE, false, C, false>(e, d_a, d_b, d_pairs, n_pairs, pl, KWBand(), d_dir, d_bnd, d_scratch, d_res, st, d_pssm);
}

template <int MODE, int C>
static cudaError_t kw_launch_band_mode(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pair,
                                       const KWPlan &pl, const KWBand &band, uint8_t *d_dir, int4 *d_bnd, int *d_scratch,
                                       KWResult *d_res, cudaStream_t st)
{
    if (d_dir) return kw_go<MODE, true, C, true>(e, d_a, d_b, d_pair, 1, pl, band, d_dir, d_bnd, d_scratch, d_res, st, nullptr);
    return kw_go<MODE, false, C, true>(e, d_a, d_b, d_pair, 1, pl, band, d_dir, d_bnd, d_scratch, d_res, st, nullptr);
}

cudaError_t kw_launch_band(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pair,
                           const KWPlan &plan, const KWBand &band, uint8_t *d_dir, int4 *d_bnd, int *d_scratch,
                           KWResult *d_res, cudaStream_t st)
{
#define KW_GO(M) (plan.cols == 2 ? kw_launch_band_mode<M, 2>(e, d_a, d_b, d_pair, plan, band, d_dir, d_bnd, d_scratch, d_res, st) \
                                 : kw_launch_band_mode<M, 8>(e, d_a, d_b, d_pair, plan, band, d_dir, d_bnd, d_scratch, d_res, st))
    switch (e->mode) {
    case SWB_GLOBAL: return KW_GO(SWB_GLOBAL);
    case SWB_GLOCAL: return KW_GO(SWB_GLOCAL);
    default: return KW_GO(SWB_LOCAL);
    }
#undef KW_GO
}

cudaError_t kw_launch(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pairs, int n_pairs,
                      const KWPlan &plan, uint8_t *d_dir, int4 *d_bnd, int *d_scratch, KWResult *d_res, cudaStream_t st,
                      const int32_t *d_pssm)
{
    if (n_pairs <= 0) return cudaSuccess;
#define KW_GO(M) (plan.cols == 2 ? kw_launch_mode<M, 2>(e, d_a, d_b, d_pairs, n_pairs, plan, d_dir, d_bnd, d_scratch, d_res, st, d_pssm) \
                                 : kw_launch_mode<M, 8>(e, d_a, d_b, d_pairs, n_pairs, plan, d_dir, d_bnd, d_scratch, d_res, st, d_pssm))
    switch (e->mode) {
    case SWB_GLOBAL: return KW_GO(SWB_GLOBAL);
    case SWB_GLOCAL: return KW_GO(SWB_GLOCAL);
    default: return KW_GO(SWB_LOCAL);
    }
#undef KW_GO
}

// ------------------------------------------------------------------------------------------------
// KT: backtrack (SA.rs:346-383).  One thread per pair follows the 2-bit predecessor codes from the
// start cell and emits the column lists in walk order (end -> start); the host reverses them
// (SA.rs:384-385).  idx value -1 = gap.  n_cols < 0 reports the reference's panic cases.
template <int COLS>
__device__ __forceinline__ int64_t kt_addr(int64_t dir_off, int la, int i, int j)
{
    constexpr int SW = 32 * COLS;                       // compile-time strip geometry: shifts, no divisions, in the walk
    const int jj = j - 1;
    const int sidx = jj / SW, within = jj % SW;
    const int t = (i - 1) + within / COLS;
    return dir_off + ((int64_t)sidx * (la + 31) + t) * SW + within;
}

// STATS: count instead of listing (KTStats; raw_a / raw_b = the sequences as given, for the equal-letter test).
template <int MODE, int COLS, bool STATS>
__global__ void kt_kernel(const KWPair *__restrict__ pairs, const KWResult *__restrict__ res, int n_pairs,
                          const uint8_t *__restrict__ dir, const int64_t *__restrict__ out_off,
                          int32_t *__restrict__ idx_a, int32_t *__restrict__ idx_b, int32_t *__restrict__ ncols, int sent,
                          const uint8_t *__restrict__ raw_a, const uint8_t *__restrict__ raw_b, KTStats *__restrict__ stats)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    const KWPair P = pairs[p];
    const KWResult R = res[p];
    int32_t *xa = STATS ? nullptr : idx_a + out_off[p], *xb = STATS ? nullptr : idx_b + out_off[p];
    const uint8_t *ra = STATS ? raw_a + P.a_off : nullptr, *rb = STATS ? raw_b + P.b_off : nullptr;
    KTStats S; S.ncols = 0; S.na = 0; S.nb = 0; S.match = 0; S.s1 = -1; S.e1 = -1; S.s2 = -1; S.e2 = -1;
    int n = 0;
    if (R.end_i < 0) { if (STATS) stats[p] = S; else ncols[p] = 0; return; }   // SA.rs:346-348
    int cx = R.end_i, cy = R.end_j, st = R.state;
    for (;;) {
        unsigned prev;
        if (cx >= 1 && cy >= 1) {
            // the walk is a chain of dependent loads; touch the line the path reaches KT_AHEAD diagonal steps from now
            if (cx > KT_AHEAD && cy > KT_AHEAD)
                asm volatile("prefetch.global.L1 [%0];" ::"l"(dir + kt_addr<COLS>(P.dir_off, P.la, cx - KT_AHEAD, cy - KT_AHEAD)));
            prev = ((unsigned)dir[kt_addr<COLS>(P.dir_off, P.la, cx, cy)] >> (2 * st)) & 3u;
        }
        else if (cx == 0 && cy == 0) prev = 0;                     // only reachable when la == lb == 0
        else if (cx == 0) {                                        // row 0: prev = GAPINX for all states (SA.rs:105-107)
            prev = 1;
            if (MODE == SWB_LOCAL && (st == 1 || sent == 0)) prev = 3;   // X(0,j) = 0 in local mode -> stop (SA.rs:354-357)
        } else {                                                   // column 0: prev = GAPINY (SA.rs:114-116)
            prev = 2;
            if (MODE == SWB_LOCAL && (st == 2 || sent == 0)) prev = 3;
        }
        if (MODE == SWB_LOCAL && prev == 3) break;
        // (the walk runs end -> start: the first index seen is the last of the alignment)
        if (st == 0) {
            if (cx == 0 || cy == 0) { n = -2; break; }             // reference underflows / panics
            if (STATS) {
                if (S.e1 < 0) S.e1 = cx - 1;
                if (S.e2 < 0) S.e2 = cy - 1;
                S.s1 = cx - 1; S.s2 = cy - 1; ++S.na; ++S.nb;
                S.match += (ra[cx - 1] == rb[cy - 1]) ? 1 : 0;
            } else { xa[n] = cx - 1; xb[n] = cy - 1; }
            ++n; --cx; --cy;
        } else if (st == 1) {
            if (cy == 0) { n = -2; break; }
            if (STATS) { if (S.e2 < 0) S.e2 = cy - 1; S.s2 = cy - 1; ++S.nb; }
            else { xa[n] = -1; xb[n] = cy - 1; }
            ++n; --cy;
        } else {
            if (cx == 0) { n = -2; break; }                        // SA.rs:369-371
            if (STATS) { if (S.e1 < 0) S.e1 = cx - 1; S.s1 = cx - 1; ++S.na; }
            else { xa[n] = cx - 1; xb[n] = -1; }
            ++n; --cx;
        }
        if (cx == 0 && cy == 0) break;
        st = (int)prev;
    }
    if (STATS) { S.ncols = n; stats[p] = S; }
    else ncols[p] = n;
}

// KT, warp form (a few long pairs): the walk is a chain of dependent byte loads, ~0.2 us each from L2.  Here a warp
// fetches the 32 x 32 block of direction bytes that ends at the cursor in one round trip (lane l loads row cx - l),
// parks it in shared memory and walks inside it -- at least 32 columns of the alignment per round trip.  The walk itself
// is the same code on every lane (uniform); lane 0 writes the lists.
#define KTW_WARPS 4
template <int MODE, int COLS>
__global__ void __launch_bounds__(KTW_WARPS * 32)
kt_warp_kernel(const KWPair *__restrict__ pairs, const KWResult *__restrict__ res, int n_pairs, const uint8_t *__restrict__ dir,
               const int64_t *__restrict__ out_off, int32_t *__restrict__ idx_a, int32_t *__restrict__ idx_b,
               int32_t *__restrict__ ncols, int sent)
{
    __shared__ uint8_t tile_all[KTW_WARPS][32][33];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int p = blockIdx.x * KTW_WARPS + wib;
    if (p >= n_pairs) return;
    uint8_t (*tile)[33] = tile_all[wib];
    const KWPair P = pairs[p];
    const KWResult R = res[p];
    int32_t *xa = idx_a + out_off[p], *xb = idx_b + out_off[p];
    int n = 0;
    if (R.end_i < 0) { if (lane == 0) ncols[p] = 0; return; }      // SA.rs:346-348
    int cx = R.end_i, cy = R.end_j, st = R.state;
    int tx = -1, ty = -1;                                          // the block in shared memory ends at (tx, ty): rows tx-31..tx, columns ty-31..ty
    for (;;) {
        unsigned prev;
        if (cx >= 1 && cy >= 1) {
            if (tx < 0 || cx <= tx - 32 || cy <= ty - 32) {
                tx = cx; ty = cy;
                __syncwarp();
                const int i = tx - lane;
                uint8_t v[32];                                     // all 32 loads in flight before the first use: one round trip
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const int j = ty - c;
                    v[c] = (i >= 1 && j >= 1) ? dir[kt_addr<COLS>(P.dir_off, P.la, i, j)] : (uint8_t)0;
                }
#pragma unroll
                for (int c = 0; c < 32; ++c) tile[lane][c] = v[c];
                __syncwarp();
            }
            prev = ((unsigned)tile[tx - cx][ty - cy] >> (2 * st)) & 3u;
        }
        else if (cx == 0 && cy == 0) prev = 0;                     // only reachable when la == lb == 0
        else if (cx == 0) {                                        // row 0: prev = GAPINX for all states (SA.rs:105-107)
            prev = 1;
            if (MODE == SWB_LOCAL && (st == 1 || sent == 0)) prev = 3;
        } else {                                                   // column 0: prev = GAPINY (SA.rs:114-116)
            prev = 2;
            if (MODE == SWB_LOCAL && (st == 2 || sent == 0)) prev = 3;
        }
        if (MODE == SWB_LOCAL && prev == 3) break;
        if (st == 0) {
            if (cx == 0 || cy == 0) { n = -2; break; }             // reference underflows / panics
            if (lane == 0) { xa[n] = cx - 1; xb[n] = cy - 1; }
            ++n; --cx; --cy;
        } else if (st == 1) {
            if (cy == 0) { n = -2; break; }
            if (lane == 0) { xa[n] = -1; xb[n] = cy - 1; }
            ++n; --cy;
        } else {
            if (cx == 0) { n = -2; break; }                        // SA.rs:369-371
            if (lane == 0) { xa[n] = cx - 1; xb[n] = -1; }
            ++n; --cx;
        }
        if (cx == 0 && cy == 0) break;
        st = (int)prev;
    }
    if (lane == 0) ncols[p] = n;
}

cudaError_t kt_launch(const swb_engine *e, const KWPair *d_pairs, const KWResult *d_res, int n_pairs,
                      const uint8_t *d_dir, int cols, const int64_t *d_out_off, int32_t *d_idx_a, int32_t *d_idx_b,
                      int32_t *d_ncols, cudaStream_t st)
{
    if (n_pairs <= 0) return cudaSuccess;
    const int threads = 32, blocks = (n_pairs + threads - 1) / threads;
    const int sent = 10 * e->o + 10 * e->e;
    const bool warp_form = n_pairs <= 2048 && !getenv("SWB_KT_THREAD");     // a few pairs: one warp each, block-cached walk
    const int wblocks = (n_pairs + KTW_WARPS - 1) / KTW_WARPS;
#define KT_GO(M) do { if (warp_form) { if (cols == 2) kt_warp_kernel<M, 2><<<wblocks, KTW_WARPS * 32, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, d_out_off, d_idx_a, d_idx_b, d_ncols, sent); \
                                       else kt_warp_kernel<M, 8><<<wblocks, KTW_WARPS * 32, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, d_out_off, d_idx_a, d_idx_b, d_ncols, sent); } \
                      else if (cols == 2) kt_kernel<M, 2, false><<<blocks, threads, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, d_out_off, d_idx_a, d_idx_b, d_ncols, sent, nullptr, nullptr, nullptr); \
                      else kt_kernel<M, 8, false><<<blocks, threads, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, d_out_off, d_idx_a, d_idx_b, d_ncols, sent, nullptr, nullptr, nullptr); } while (0)
    switch (e->mode) {
    case SWB_GLOBAL: KT_GO(SWB_GLOBAL); break;
    case SWB_GLOCAL: KT_GO(SWB_GLOCAL); break;
    default: KT_GO(SWB_LOCAL); break;
    }
#undef KT_GO
    return cudaGetLastError();
}

cudaError_t kt_stats_launch(const swb_engine *e, const KWPair *d_pairs, const KWResult *d_res, int n_pairs,
                            const uint8_t *d_dir, int cols, const uint8_t *d_raw_a, const uint8_t *d_raw_b,
                            KTStats *d_stats, cudaStream_t st)
{
    if (n_pairs <= 0) return cudaSuccess;
    const int threads = 32, blocks = (n_pairs + threads - 1) / threads;
    const int sent = 10 * e->o + 10 * e->e;
#define KT_GO(M) do { if (cols == 2) kt_kernel<M, 2, true><<<blocks, threads, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, nullptr, nullptr, nullptr, nullptr, sent, d_raw_a, d_raw_b, d_stats); \
                      else kt_kernel<M, 8, true><<<blocks, threads, 0, st>>>(d_pairs, d_res, n_pairs, d_dir, nullptr, nullptr, nullptr, nullptr, sent, d_raw_a, d_raw_b, d_stats); } while (0)
    switch (e->mode) {
    case SWB_GLOBAL: KT_GO(SWB_GLOBAL); break;
    case SWB_GLOCAL: KT_GO(SWB_GLOCAL); break;
    default: KT_GO(SWB_LOCAL); break;
    }
#undef KT_GO
    return cudaGetLastError();
}


// KT on a row band (checkpointed traceback): the walk of kt_kernel, resumed at a cursor and suspended at the band's
// top row.  dir holds the direction bytes of rows (row0, row0 + la_band] only.
template <int MODE, int COLS>
__global__ void kt_band_kernel(const uint8_t *__restrict__ dir, int la_band, int row0, KTCursor *__restrict__ cur,
                               int32_t *__restrict__ xa, int32_t *__restrict__ xb, int sent)
{
    // one warp, block-cached like kt_warp_kernel: the 32 x 32 direction bytes that end at the cursor per round trip
    __shared__ uint8_t tile[32][33];
    if (blockIdx.x != 0 || threadIdx.x >= 32) return;
    const int lane = threadIdx.x;
    KTCursor K = *cur;
    int cx = K.cx, cy = K.cy, st = K.st, n = K.n, status = 0;
    int tx = -1, ty = -1;
    for (;;) {
        if (row0 > 0 && cx == row0) { status = 1; break; }         // the predecessor codes of this row belong to the band above
        unsigned prev;
        if (cx >= 1 && cy >= 1) {
            if (tx < 0 || cx <= tx - 32 || cy <= ty - 32) {
                tx = cx; ty = cy;
                __syncwarp();
                const int i = tx - lane;
                uint8_t v[32];
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const int j = ty - c;
                    v[c] = (i > row0 && j >= 1) ? dir[kt_addr<COLS>(0, la_band, i - row0, j)] : (uint8_t)0;
                }
#pragma unroll
                for (int c = 0; c < 32; ++c) tile[lane][c] = v[c];
                __syncwarp();
            }
            prev = ((unsigned)tile[tx - cx][ty - cy] >> (2 * st)) & 3u;
        }
        else if (cx == 0 && cy == 0) prev = 0;
        else if (cx == 0) {                                        // row 0: prev = GAPINX for all states (SA.rs:105-107)
            prev = 1;
            if (MODE == SWB_LOCAL && (st == 1 || sent == 0)) prev = 3;
        } else {                                                   // column 0: prev = GAPINY (SA.rs:114-116)
            prev = 2;
            if (MODE == SWB_LOCAL && (st == 2 || sent == 0)) prev = 3;
        }
        if (MODE == SWB_LOCAL && prev == 3) break;
        if (st == 0) {
            if (cx == 0 || cy == 0) { status = -2; break; }
            if (lane == 0) { xa[n] = cx - 1; xb[n] = cy - 1; }
            ++n; --cx; --cy;
        } else if (st == 1) {
            if (cy == 0) { status = -2; break; }
            if (lane == 0) { xa[n] = -1; xb[n] = cy - 1; }
            ++n; --cy;
        } else {
            if (cx == 0) { status = -2; break; }
            if (lane == 0) { xa[n] = cx - 1; xb[n] = -1; }
            ++n; --cx;
        }
        if (cx == 0 && cy == 0) break;
        st = (int)prev;
    }
    if (lane == 0) { K.cx = cx; K.cy = cy; K.st = st; K.n = n; K.status = status; *cur = K; }
}

cudaError_t kt_band_launch(const swb_engine *e, const uint8_t *d_dir, int cols, int32_t la_band, int32_t row0,
                           KTCursor *d_cur, int32_t *d_idx_a, int32_t *d_idx_b, cudaStream_t st)
{
    const int sent = 10 * e->o + 10 * e->e;
#define KT_GO(M) do { if (cols == 2) kt_band_kernel<M, 2><<<1, 32, 0, st>>>(d_dir, la_band, row0, d_cur, d_idx_a, d_idx_b, sent); \
                      else kt_band_kernel<M, 8><<<1, 32, 0, st>>>(d_dir, la_band, row0, d_cur, d_idx_a, d_idx_b, sent); } while (0)
    switch (e->mode) {
    case SWB_GLOBAL: KT_GO(SWB_GLOBAL); break;
    case SWB_GLOCAL: KT_GO(SWB_GLOCAL); break;
    default: KT_GO(SWB_LOCAL); break;
    }
#undef KT_GO
    return cudaGetLastError();
}
