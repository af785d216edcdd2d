// Traceback path: swb_align / swb_align_batch.  Mirrors align_partial (SA.rs:393-551): encode ->
// fill (KW with direction bytes) -> backtrack (KT) -> strings with the retain_all padding rules.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <thread>
#include <cstdlib>
#include <cstring>

#include "swb_internal.h"

#define DIR_BUDGET_BYTES (24ll << 30)      // direction bytes resident per chunk of pairs (one pair may exceed it)

struct DevBuf {                            // stream-ordered allocation from the engine's pool, freed on scope exit
    void *p = nullptr; cudaStream_t st = nullptr;
    cudaError_t alloc(size_t bytes, cudaStream_t s) { st = s; return cudaMallocAsync(&p, bytes ? bytes : 16, s); }
    ~DevBuf() { if (p) cudaFreeAsync(p, st); }
    template <class T> T *as() { return reinterpret_cast<T *>(p); }
};

static char *dup_n(const std::string &s)
{
    char *r = (char *)malloc(s.size() + 1);
    if (r) { memcpy(r, s.data(), s.size()); r[s.size()] = 0; }
    return r;
}

// String re-inflation (SA.rs:470-550).  idx lists are forward (start -> end), -1 = gap.
static void inflate(const char *a, int64_t la, const char *b, int64_t lb, const int64_t *xa, const int64_t *xb,
                    int64_t n, int mode, bool retain_all, std::string &ra, std::string &rb)
{
    std::string c1, c2;
    int64_t s1 = -1, s2 = -1, e1 = -1, e2 = -1;
    c1.reserve(n); c2.reserve(n);
    for (int64_t k = 0; k < n; ++k) {
        if (xa[k] > -1) { if (s1 < 0) s1 = xa[k]; c1.push_back(a[xa[k]]); e1 = xa[k]; } else c1.push_back('-');
        if (xb[k] > -1) { if (s2 < 0) s2 = xb[k]; c2.push_back(b[xb[k]]); e2 = xb[k]; } else c2.push_back('-');
    }
    if (mode == SWB_LOCAL && !retain_all) { ra.swap(c1); rb.swap(c2); return; }      // SA.rs:502-504
    ra.clear(); rb.clear();
    ra.reserve(la + lb); rb.reserve(la + lb);
    if (s1 < 0 || s2 < 0) {                                                           // SA.rs:512-524
        ra.append(a, la); rb.append(la, '-');
        ra.append(lb, '-'); rb.append(b, lb);
        return;
    }
    ra.append(a, s1); rb.append(s1, '-');                                             // SA.rs:530-537
    ra.append(s2, '-'); rb.append(b, s2);
    ra += c1; rb += c2;
    ra.append(a + e1 + 1, la - e1 - 1); rb.append(la - e1 - 1, '-');                  // SA.rs:542-549
    ra.append(lb - e2 - 1, '-'); rb.append(b + e2 + 1, lb - e2 - 1);
}

static int fill_out(swb_alignment *out, float score, int64_t ei, int64_t ej, const std::vector<int64_t> &xa,
                    const std::vector<int64_t> &xb, const std::string &ra, const std::string &rb)
{
    out->score = score; out->end_a = ei; out->end_b = ej; out->n_cols = (int64_t)xa.size();
    out->idx_a = (int64_t *)malloc(sizeof(int64_t) * std::max<size_t>(1, xa.size()));
    out->idx_b = (int64_t *)malloc(sizeof(int64_t) * std::max<size_t>(1, xb.size()));
    out->aln_a = dup_n(ra); out->aln_b = dup_n(rb); out->aln_len = (int64_t)ra.size();
    if (!out->idx_a || !out->idx_b || !out->aln_a || !out->aln_b) return SWB_ERR_NOMEM;
    if (!xa.empty()) { memcpy(out->idx_a, xa.data(), xa.size() * sizeof(int64_t)); memcpy(out->idx_b, xb.data(), xb.size() * sizeof(int64_t)); }
    return SWB_OK;
}

extern "C" void swb_alignment_free(swb_alignment *a)
{
    if (!a) return;
    free(a->idx_a); free(a->idx_b); free(a->aln_a); free(a->aln_b);
    memset(a, 0, sizeof(*a));
}

// The strings' statistics from the core alignment's counts (inflate() above without building anything).
void swb_compose_stats(swb_aln_stats *o, int mode, bool retain_all, int64_t la, int64_t lb, const KTStats &S)
{
    if (mode == SWB_LOCAL && !retain_all) {                                               // SA.rs:502-504
        o->aln_len = S.ncols; o->a_letters = S.na; o->b_letters = S.nb; o->matches = S.match;
        return;
    }
    o->a_letters = la; o->b_letters = lb;
    if (S.s1 < 0 || S.s2 < 0) { o->aln_len = la + lb; o->matches = 0; return; }           // SA.rs:512-524
    o->aln_len = (int64_t)S.s1 + S.s2 + S.ncols + (la - S.e1 - 1) + (lb - S.e2 - 1);      // SA.rs:530-549
    o->matches = S.match;
}

// Direction-byte budget: a chunk of pairs keeps at most this many direction bytes resident, and a single pair that
// needs more takes the checkpointed path below.  SWB_CKPT_BYTES overrides it (the tests force tiny budgets).
static int64_t dir_budget()
{
    if (const char *f = getenv("SWB_CKPT_BYTES")) { const long long v = atoll(f); if (v >= 4096) return v; }
    return DIR_BUDGET_BYTES;
}

// int32 lanes of the wavefront kernel: every score of a pair must stay well inside +-2^29 (its PAD score).  Upwards a
// pair can reach min(la,lb) best substitutions, downwards (la+lb) extensions + two openings + the boundary sentinel +
// min(la,lb) worst substitutions.
bool swb_range_ok(const swb_engine *e, int64_t la, int64_t lb, int64_t smax_abs)
{
    const int64_t lim = 1ll << 28, mn = la < lb ? la : lb;
    const int64_t sp = smax_abs >= 0 ? smax_abs : std::max(e->smax, 0), sn = smax_abs >= 0 ? smax_abs : std::max(-e->smin, 0);
    const int64_t up = mn * sp;
    const int64_t down = (la + lb) * -(int64_t)e->e + 3 * -(int64_t)e->o + -(int64_t)(10 * e->o + 10 * e->e) + mn * sn;
    return up <= lim && down <= lim;
}

// Checkpointed traceback of ONE pair whose direction bytes exceed the budget (BASELINE configs[3]: "wavefront +
// checkpointed recompute"; the reference keeps the whole SWCell matrix, SA.rs:60-86, and cannot run such pairs at all).
//   1. forward sweep, score only (KW, no direction bytes): finds the start cell of the walk (SA.rs:282-308) and saves
//      the three states of every R-th row (16 B per column per checkpoint).
//   2. from the band that holds the start cell upwards: re-run KW with direction bytes on rows (bR, cx] x columns [1, cy]
//      only -- the walk never moves right or down, so everything beyond the cursor is skipped -- starting from the
//      checkpoint row bR, then let KT walk until it reaches row bR, and continue in the band above.
// Memory: R * lb direction bytes + (la / R) * lb * 16 checkpoint bytes instead of la * lb.
static int align_ckpt_pair(swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, int64_t a_off, int64_t b_off, int32_t la,
                           int32_t lb, int64_t budget, KWResult *hres, std::vector<int64_t> &xa, std::vector<int64_t> &xb)
{
    cudaStream_t st = e->stream;
    const int64_t row_bytes = ((int64_t)lb + 255) / 256 * 256;                    // direction bytes per row, widest strips
    int32_t R = (int32_t)std::max<int64_t>(32, std::min<int64_t>(la, budget / row_bytes - 31));
    const int32_t n_ck = (la - 1) / R;                                            // checkpoint rows R, 2R, .. (< la)
    const int64_t stride = (int64_t)lb + 2;
    DevBuf d_ck, d_pair, d_res, d_cur, d_xa, d_xb;
    CUDA_TRY(e, d_ck.alloc((size_t)std::max<int64_t>(1, (int64_t)n_ck * stride) * sizeof(int4), st));
    CUDA_TRY(e, d_pair.alloc(sizeof(KWPair), st)); CUDA_TRY(e, d_res.alloc(2 * sizeof(KWResult), st));
    CUDA_TRY(e, d_cur.alloc(sizeof(KTCursor), st));
    CUDA_TRY(e, d_xa.alloc(((size_t)la + lb + 1) * sizeof(int32_t), st)); CUDA_TRY(e, d_xb.alloc(((size_t)la + lb + 1) * sizeof(int32_t), st));
    if (!e->h_tb || e->tb_bytes < 256) {
        if (e->h_tb) cudaFreeHost(e->h_tb);
        e->h_tb = nullptr; e->tb_bytes = 0;
        CUDA_TRY(e, cudaMallocHost(&e->h_tb, 4096)); e->tb_bytes = 4096;
    }
    KWPair *hp = (KWPair *)e->h_tb;                                               // pinned staging: pair, result, cursor
    KWResult *hr = (KWResult *)((char *)e->h_tb + 64);
    KTCursor *hc = (KTCursor *)((char *)e->h_tb + 128);

    auto run_kw = [&](int32_t row0, int32_t rows, int32_t cols_used, bool dirs, const KWBand &band, KWPlan &plan) -> int {
        KWPair P; P.a_off = a_off + row0; P.b_off = b_off; P.la = rows; P.lb = cols_used; P.dir_off = dirs ? 0 : -1;
        P.bnd_off = 0; P.scr_off = 0; P.pssm_off = -1;
        // short, wide bands: the 32-row lag between the warps of a team dominates, fewer and wider strips win (measured)
        kw_plan(e, &P, 1, dirs, &plan, (dirs && rows < 16384) ? 8 : 0, true);
        if (dirs && (size_t)plan.dir_bytes + 16 > e->dir_bytes) {
            CUDA_TRY(e, cudaStreamSynchronize(st));
            if (e->d_dir) cudaFree(e->d_dir);
            e->d_dir = nullptr; e->dir_bytes = 0;
            CUDA_TRY(e, cudaMalloc(&e->d_dir, (size_t)plan.dir_bytes + 16));
            e->dir_bytes = (size_t)plan.dir_bytes + 16;
        }
        DevBuf d_bnd, d_scr;
        CUDA_TRY(e, d_bnd.alloc(plan.bnd_elems * sizeof(int4), st)); CUDA_TRY(e, d_scr.alloc(plan.scratch_ints * sizeof(int), st));
        CUDA_TRY(e, cudaStreamSynchronize(st));                                   // the staging slot is reused per launch
        *hp = P;
        CUDA_TRY(e, cudaMemcpyAsync(d_pair.p, hp, sizeof(KWPair), cudaMemcpyHostToDevice, st));
        CUDA_TRY(e, kw_launch_band(e, d_a, d_b, d_pair.as<KWPair>(), plan, band, dirs ? (uint8_t *)e->d_dir : nullptr,
                                   d_bnd.as<int4>(), d_scr.as<int>(), d_res.as<KWResult>() + (dirs ? 1 : 0), st));
        e->stats.launches += 1;
        return SWB_OK;
    };

    CUDA_TRY(e, cudaEventRecord(e->ev0, st));
    {   // 1. forward sweep with checkpoints
        KWBand band; band.row0 = 0; band.la_full = la; band.lb_full = lb; band.ckpt_out = d_ck.as<int4>(); band.every = R; band.stride = stride;
        KWPlan plan;
        int rc = run_kw(0, la, lb, false, band, plan);
        if (rc != SWB_OK) return rc;
        CUDA_TRY(e, cudaMemcpyAsync(hr, d_res.p, sizeof(KWResult), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(e, cudaStreamSynchronize(st));
        *hres = *hr;
    }
    xa.clear(); xb.clear();
    if (hres->end_i < 0) {                                                        // local score 0: no alignment (SA.rs:346-348)
        CUDA_TRY(e, cudaEventRecord(e->ev1, st)); CUDA_TRY(e, cudaStreamSynchronize(st));
        float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1); e->stats.kernel_ms += ms;
        return SWB_OK;
    }
    KTCursor C; C.cx = hres->end_i; C.cy = hres->end_j; C.st = hres->state; C.n = 0; C.status = 1;
    while (C.status == 1) {                                                       // 2. bands, bottom-up
        const int32_t b = (C.cx >= 1 && C.cy >= 1) ? (C.cx - 1) / R : 0;
        const int32_t row0 = b * R, rows = C.cx - row0;
        KWPlan plan;
        if (C.cx >= 1 && C.cy >= 1) {
            KWBand band; band.row0 = row0; band.la_full = la; band.lb_full = lb; band.every = R; band.stride = stride;
            band.ckpt_in = (b > 0) ? d_ck.as<int4>() + (int64_t)(b - 1) * stride : nullptr;
            int rc = run_kw(row0, rows, C.cy, true, band, plan);
            if (rc != SWB_OK) return rc;
        } else CUDA_TRY(e, cudaStreamSynchronize(st));                            // the rest runs along the matrix boundary: no bytes needed
        *hc = C;
        CUDA_TRY(e, cudaMemcpyAsync(d_cur.p, hc, sizeof(KTCursor), cudaMemcpyHostToDevice, st));
        CUDA_TRY(e, kt_band_launch(e, (const uint8_t *)e->d_dir, plan.cols, rows, row0, d_cur.as<KTCursor>(), d_xa.as<int32_t>(), d_xb.as<int32_t>(), st));
        CUDA_TRY(e, cudaMemcpyAsync(hc, d_cur.p, sizeof(KTCursor), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(e, cudaStreamSynchronize(st));
        e->stats.launches += 1;
        if (hc->status == 1 && hc->cx == C.cx && hc->cy == C.cy && hc->n == C.n) { e->err = "checkpointed traceback made no progress"; return SWB_ERR_CUDA; }
        C = *hc;
    }
    CUDA_TRY(e, cudaEventRecord(e->ev1, st));
    if (C.status < 0) { e->err = "traceback left the matrix (the reference panics here, SA.rs:360-371)"; return SWB_ERR_EMPTY; }
    std::vector<int32_t> ha((size_t)C.n), hb((size_t)C.n);                       // (a one-off pageable copy: this path is for a few huge pairs)
    if (C.n > 0) {
        CUDA_TRY(e, cudaMemcpyAsync(ha.data(), d_xa.p, (size_t)C.n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(e, cudaMemcpyAsync(hb.data(), d_xb.p, (size_t)C.n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(e, cudaStreamSynchronize(st));
    float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1); e->stats.kernel_ms += ms;
    e->stats.d2h_bytes += 8.0 * C.n;
    xa.resize(C.n); xb.resize(C.n);
    for (int64_t c = 0; c < C.n; ++c) { xa[c] = ha[C.n - 1 - c]; xb[c] = hb[C.n - 1 - c]; }   // reverse (SA.rs:384-385)
    return SWB_OK;
}

// AlnJob (swb_internal.h): the kernels see the DP extents (la, lb) at (a_off, b_off) of the device code arrays; the
// strings are spelled from the letters as given with the FULL lengths.  Local mode: a hit of a database search is traced
// inside the prefix box that ends at its end cell -- every cell of the box has the value it has in the full matrix and
// precedes the end cell in the reference's scan order (SA.rs:286-294), so the walk is the same.
// The packed traceback path (k3_tile.cu) for the jobs `sel`: K3 fill with tile borders + KT3 tile walk, chunked by the
// border-store budget; the strings are inflated on host threads as on the direction-byte path.
static int align_jobs_k3(swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const std::vector<AlnJob> &jobs, std::vector<size_t> sel,
                         bool retain_all, swb_alignment *outs)
{
    const float inv = 1.0f / (float)e->scale;
    const bool dyadic = (e->scale & (e->scale - 1)) == 0;
    auto to_f = [&](int64_t s) { return dyadic ? (float)s * inv : (float)((double)s / e->scale); };
    const int64_t budget = dir_budget();
    // transposition needs a symmetric score table (the kernels look up table[row residue][column residue])
    bool swap_ok = true, swap_all = false;
    for (int r = 0; r < e->n && swap_ok; ++r)
        for (int c = 0; c < r; ++c) if (e->mat[(size_t)r * e->n + c] != e->mat[(size_t)c * e->n + r]) { swap_ok = false; break; }
    if (const char *f = getenv("SWB_K3_SWAP")) { if (atoi(f) == 0) swap_ok = false; else swap_all = true; }   // tests force both forms
    // longest first: a warp takes one pair at a time from a counter
    std::sort(sel.begin(), sel.end(), [&](size_t x, size_t y) { return (int64_t)jobs[x].la * jobs[x].lb > (int64_t)jobs[y].la * jobs[y].lb; });
    size_t pos = 0;
    while (pos < sel.size()) {
        std::vector<K3Pair> hp;
        int64_t colb = 0, rowb = 0, out_total = 0;
        int max_sweeps = 1;
        size_t end = pos;
        while (end < sel.size()) {
            const AlnJob &J = jobs[sel[end]];
            // a tall box (many rows, few columns) is one long sweep of a single warp; transposed it is several short sweeps
            // that a team of warps pipelines (the hits of a search: 5000 query rows x a few hundred subject columns)
            const bool swp = swap_ok && (swap_all || J.la > 2 * (int64_t)J.lb);
            const int32_t la_k = swp ? J.lb : J.la, lb_k = swp ? J.la : J.lb;
            int64_t c4, rw; k3_sizes(la_k, lb_k, &c4, &rw);
            rw = (rw + 3) / 4 * 4;
            if (!hp.empty() && (colb + c4) * 16 + (rowb + rw) * 4 > budget) break;
            K3Pair P; P.a_off = J.a_off; P.b_off = J.b_off; P.la = la_k; P.lb = lb_k; P.colb_off = colb; P.rowb_off = rowb; P.out_off = out_total;
            P.start_state = 0; P.score = J.known_score == INT32_MIN ? 0 : J.known_score; P.swap = swp ? 1 : 0; P.find_col = 0; P.cx0 = 0; P.cy0 = 0;
            hp.push_back(P);
            max_sweeps = std::max(max_sweeps, (lb_k + 1023) / 1024);
            colb += c4; rowb += rw; out_total += (int64_t)J.la + J.lb + 1;
            ++end;
        }
        const int np = (int)hp.size();
        DevBuf d_pairs, d_res, d_colb, d_rowb, d_cnt, d_xa, d_xb, d_nc;
        CUDA_TRY(e, d_pairs.alloc(np * sizeof(K3Pair), e->stream)); CUDA_TRY(e, d_res.alloc(np * sizeof(KWResult), e->stream));
        CUDA_TRY(e, d_colb.alloc((size_t)colb * 16 + 64, e->stream)); CUDA_TRY(e, d_rowb.alloc((size_t)rowb * 4 + 64, e->stream));
        CUDA_TRY(e, d_cnt.alloc(sizeof(int), e->stream));
        CUDA_TRY(e, d_xa.alloc(out_total * sizeof(int32_t), e->stream)); CUDA_TRY(e, d_xb.alloc(out_total * sizeof(int32_t), e->stream));
        CUDA_TRY(e, d_nc.alloc(np * sizeof(int32_t), e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(d_pairs.p, hp.data(), np * sizeof(K3Pair), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, cudaEventRecord(e->ev0, e->stream));
        CUDA_TRY(e, k3_launch(e, d_a, d_b, d_pairs.as<K3Pair>(), np, d_colb.as<uint4>(), d_rowb.as<unsigned>(), d_cnt.as<int>(),
                              d_xa.as<int32_t>(), d_xb.as<int32_t>(), d_nc.as<int32_t>(), d_res.as<KWResult>(), e->stream, max_sweeps));
        CUDA_TRY(e, cudaEventRecord(e->ev1, e->stream));
        const size_t tb_need = (size_t)np * (sizeof(KWResult) + sizeof(int32_t)) + 2 * (size_t)out_total * sizeof(int32_t) + 64;
        if (tb_need > e->tb_bytes) {
            if (e->h_tb) cudaFreeHost(e->h_tb);
            e->h_tb = nullptr; e->tb_bytes = 0;
            CUDA_TRY(e, cudaMallocHost(&e->h_tb, tb_need));
            e->tb_bytes = tb_need;
        }
        KWResult *hr = (KWResult *)e->h_tb;
        int32_t *hnc = (int32_t *)(hr + np), *hxa = hnc + np, *hxb = hxa + out_total;
        CUDA_TRY(e, cudaMemcpyAsync(hr, d_res.p, np * sizeof(KWResult), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(hnc, d_nc.p, np * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(hxa, d_xa.p, out_total * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(hxb, d_xb.p, out_total * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1);
        e->stats.kernel_ms += ms; e->stats.launches += 2; e->stats.k3_pairs += np;
        e->stats.h2d_bytes += (double)np * sizeof(K3Pair);
        e->stats.d2h_bytes += (double)np * (sizeof(KWResult) + 4) + 8.0 * out_total;
        std::atomic<int> next(0), fail(SWB_OK);
        auto worker = [&]() {
            for (;;) {
                const int t = next.fetch_add(1);
                if (t >= np || fail.load() != SWB_OK) return;
                const AlnJob &J = jobs[sel[pos + t]];
                if (hnc[t] < 0) { fail.store(SWB_ERR_EMPTY); return; }
                const int64_t nc = hnc[t];
                std::vector<int64_t> xa(nc), xb(nc);
                const int32_t *la_list = hp[t].swap ? hxb : hxa, *lb_list = hp[t].swap ? hxa : hxb;     // transposed pairs: rows were b
                for (int64_t c = 0; c < nc; ++c) { xa[c] = la_list[hp[t].out_off + nc - 1 - c]; xb[c] = lb_list[hp[t].out_off + nc - 1 - c]; }   // reverse (SA.rs:384-385)
                std::string ra, rb;
                inflate(J.ra, J.la_full, J.rb, J.lb_full, xa.data(), xb.data(), nc, e->mode, retain_all, ra, rb);
                const int64_t ei = hp[t].swap ? hr[t].end_j : hr[t].end_i, ej = hp[t].swap ? hr[t].end_i : hr[t].end_j;
                if (fill_out(&outs[J.out], to_f(hr[t].score), ei, ej, xa, xb, ra, rb) != SWB_OK) { fail.store(SWB_ERR_NOMEM); return; }
            }
        };
        // (a thread costs ~20 us to start: worth it from a few dozen alignments on, not for the dozen hits a shard owns)
        const int n_thr = (np >= 48) ? (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), std::min(16u, (unsigned)np / 12u)) : 1;
        if (n_thr > 1) {
            std::vector<std::thread> pool;
            for (int w = 0; w < n_thr; ++w) pool.emplace_back(worker);
            for (auto &th : pool) th.join();
        } else worker();
        if (fail.load() == SWB_ERR_EMPTY) { e->err = "traceback left the matrix (the reference panics here, SA.rs:360-371)"; return SWB_ERR_EMPTY; }
        if (fail.load() != SWB_OK) { e->err = "out of memory"; return fail.load(); }
        pos = end;
    }
    return SWB_OK;
}

// One long pair through the int32 tile path (K3W fill over the whole GPU + KT3 walk): the border store (0.5 B per cell) and
// the row checkpoints live in the engine's direction-byte arena.
static int64_t k3w_min_cells()
{
    if (getenv("SWB_NO_K3W")) return INT64_MAX;
    if (const char *f = getenv("SWB_K3W_MIN")) { const long long v = atoll(f); if (v >= 1) return v; }    // the tests send small pairs here
    return 400000000ll;                        // below ~2e4 x 2e4 a batch of pairs keeps the GPU busier on the direction-byte path
}
static bool k3w_wanted(const swb_engine *e, const AlnJob &J, int64_t budget)
{
    if (e->mode == SWB_GLOCAL || e->o > e->e || J.la < 1 || J.lb < 1) return false;
    if ((int64_t)J.la * J.lb < k3w_min_cells()) return false;
    int64_t c4, rw, sc; k3w_sizes(J.la, J.lb, &c4, &rw, &sc);
    return c4 * 16 + rw * 4 + 256 <= budget;
}
static int align_pair_k3w(swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const AlnJob &J, bool retain_all, swb_alignment *out)
{
    const float inv = 1.0f / (float)e->scale;
    const bool dyadic = (e->scale & (e->scale - 1)) == 0;
    auto to_f = [&](int64_t s) { return dyadic ? (float)s * inv : (float)((double)s / e->scale); };
    // more columns than rows = more sweeps side by side and fewer row steps: a pair with la > lb runs transposed (see K3)
    // (global mode only: a local fill finds the end cell itself, and its first-maximum order -- rows, then columns,
    //  SA.rs:286-294 -- is not the transposed problem's)
    bool swp = e->mode == SWB_GLOBAL && J.la > J.lb && !getenv("SWB_K3W_NOSWAP");
    for (int r = 0; r < e->n && swp; ++r)
        for (int c = 0; c < r; ++c) if (e->mat[(size_t)r * e->n + c] != e->mat[(size_t)c * e->n + r]) { swp = false; break; }
    const int32_t la_k = swp ? J.lb : J.la, lb_k = swp ? J.la : J.lb;
    int64_t c4, rw, sc; k3w_sizes(la_k, lb_k, &c4, &rw, &sc);
    const size_t colb_bytes = ((size_t)c4 * 16 + 255) & ~(size_t)255, need = colb_bytes + (size_t)rw * 4 + 64;
    if (need > e->dir_bytes) {                                  // engine-owned arena, grown on demand and kept
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        if (e->d_dir) cudaFree(e->d_dir);
        e->d_dir = nullptr; e->dir_bytes = 0;
        CUDA_TRY(e, cudaMalloc(&e->d_dir, need));
        e->dir_bytes = need;
    }
    uint4 *d_colb = reinterpret_cast<uint4 *>(e->d_dir);
    unsigned *d_rowb = reinterpret_cast<unsigned *>((char *)e->d_dir + colb_bytes);
    const int64_t out_total = (int64_t)J.la + J.lb + 1;
    K3Pair P; P.a_off = J.a_off; P.b_off = J.b_off; P.la = la_k; P.lb = lb_k; P.colb_off = 0; P.rowb_off = 0; P.out_off = 0;
    P.start_state = 0; P.score = 0; P.swap = swp ? 1 : 0; P.find_col = 0; P.cx0 = 0; P.cy0 = 0;
    DevBuf d_pair, d_res, d_scr, d_xa, d_xb, d_nc;
    CUDA_TRY(e, d_pair.alloc(sizeof(K3Pair), e->stream)); CUDA_TRY(e, d_res.alloc(sizeof(KWResult), e->stream));
    CUDA_TRY(e, d_scr.alloc((size_t)sc * sizeof(int), e->stream));
    CUDA_TRY(e, d_xa.alloc(out_total * sizeof(int32_t), e->stream)); CUDA_TRY(e, d_xb.alloc(out_total * sizeof(int32_t), e->stream));
    CUDA_TRY(e, d_nc.alloc(sizeof(int32_t), e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(d_pair.p, &P, sizeof(K3Pair), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaEventRecord(e->ev0, e->stream));
    CUDA_TRY(e, k3w_launch(e, d_a, d_b, d_pair.as<K3Pair>(), P, d_colb, d_rowb, d_scr.as<int>(), d_xa.as<int32_t>(), d_xb.as<int32_t>(),
                           d_nc.as<int32_t>(), d_res.as<KWResult>(), e->stream));
    CUDA_TRY(e, cudaEventRecord(e->ev1, e->stream));
    const size_t tb_need = sizeof(KWResult) + sizeof(int32_t) + 2 * (size_t)out_total * sizeof(int32_t) + 64;
    if (tb_need > e->tb_bytes) {
        if (e->h_tb) cudaFreeHost(e->h_tb);
        e->h_tb = nullptr; e->tb_bytes = 0;
        CUDA_TRY(e, cudaMallocHost(&e->h_tb, tb_need));
        e->tb_bytes = tb_need;
    }
    KWResult *hr = (KWResult *)e->h_tb;
    int32_t *hnc = (int32_t *)(hr + 1), *hxa = hnc + 1, *hxb = hxa + out_total;
    CUDA_TRY(e, cudaMemcpyAsync(hr, d_res.p, sizeof(KWResult), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(hnc, d_nc.p, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    const int64_t nc = hnc[0];
    if (nc < 0) { e->err = "traceback left the matrix (the reference panics here, SA.rs:360-371)"; return SWB_ERR_EMPTY; }
    if (nc > 0) {                                               // the column lists: only the part that was written
        CUDA_TRY(e, cudaMemcpyAsync(hxa, d_xa.p, nc * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(hxb, d_xb.p, nc * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    }
    float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1);
    e->stats.kernel_ms += ms; e->stats.launches += 3; e->stats.k3_pairs += 1;
    e->stats.d2h_bytes += 8.0 * nc + sizeof(KWResult) + 4;
    std::vector<int64_t> xa(nc), xb(nc);
    const int32_t *la_list = swp ? hxb : hxa, *lb_list = swp ? hxa : hxb;                          // transposed: rows were b
    for (int64_t c = 0; c < nc; ++c) { xa[c] = la_list[nc - 1 - c]; xb[c] = lb_list[nc - 1 - c]; }   // reverse (SA.rs:384-385)
    std::string ra, rb;
    inflate(J.ra, J.la_full, J.rb, J.lb_full, xa.data(), xb.data(), nc, e->mode, retain_all, ra, rb);
    const int64_t ei = swp ? hr[0].end_j : hr[0].end_i, ej = swp ? hr[0].end_i : hr[0].end_j;
    if (fill_out(out, to_f(hr[0].score), ei, ej, xa, xb, ra, rb) != SWB_OK) { e->err = "out of memory"; return SWB_ERR_NOMEM; }
    return SWB_OK;
}

// Device core of the traceback path: KW with direction bytes + KT for the jobs whose bytes fit the budget (chunked),
// checkpointed recompute for the others.  outs: full alignments; souts: the counting form.  Exactly one is non-null.
int align_jobs_device(swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const std::vector<AlnJob> &jobs, bool retain_all,
                      swb_alignment *outs, swb_aln_stats *souts, const int32_t *d_pssm, const uint8_t *d_ra, const uint8_t *d_rb)
{
    const float inv = 1.0f / (float)e->scale;
    const bool dyadic = (e->scale & (e->scale - 1)) == 0;
    auto to_f = [&](int64_t s) { return dyadic ? (float)s * inv : (float)((double)s / e->scale); };
    const int64_t budget = dir_budget();
    std::vector<size_t> work, packed;
    // the packed path (K3 + KT3): global-mode pairs inside the static s16 bound; local-mode pairs whose end cell and score
    // are known from the scan (hits of a database search) and below the s16 limit
    const int64_t k3_limit = 32767 - 2 * (int64_t)std::max(e->smax, 0) + e->o - 64;
    for (size_t w = 0; w < jobs.size() && outs && !d_pssm; ++w) {
        const AlnJob &J = jobs[w];
        if (!k3_eligible(e, J.la, J.lb)) continue;
        if (e->mode == SWB_LOCAL && (J.known_score == INT32_MIN || J.known_score <= 0 || J.known_score >= k3_limit)) continue;
        packed.push_back(w);
    }
    if (!packed.empty()) {
        int rc = align_jobs_k3(e, d_a, d_b, jobs, packed, retain_all, outs);
        if (rc != SWB_OK) return rc;
    }
    std::vector<char> is_packed(jobs.size(), 0);
    for (size_t w : packed) is_packed[w] = 1;
    for (size_t w = 0; w < jobs.size() && outs && !d_pssm; ++w) {       // long pairs outside the s16 range: the int32 tile path, one at a time
        if (is_packed[w] || !k3w_wanted(e, jobs[w], budget)) continue;
        int rc = align_pair_k3w(e, d_a, d_b, jobs[w], retain_all, &outs[jobs[w].out]);
        if (rc != SWB_OK) return rc;
        is_packed[w] = 1;
    }
    for (size_t w = 0; w < jobs.size(); ++w) {
        const AlnJob &J = jobs[w];
        if (is_packed[w]) continue;
        if (d_pssm || kw_dir_bytes(J.la, J.lb, 8) <= budget) { work.push_back(w); continue; }
        // direction bytes do not fit: checkpointed traceback, one pair at a time
        KWResult R; std::vector<int64_t> xa, xb; std::string ra, rb;
        int rc = align_ckpt_pair(e, d_a, d_b, J.a_off, J.b_off, J.la, J.lb, budget, &R, xa, xb);
        if (rc != SWB_OK) return rc;
        e->stats.ckpt_pairs += 1;
        if (souts) {                              // counting form: the same counts KT takes on the device, from the lists
            KTStats S; S.ncols = (int32_t)xa.size(); S.na = 0; S.nb = 0; S.match = 0; S.s1 = S.e1 = S.s2 = S.e2 = -1;
            for (size_t c = 0; c < xa.size(); ++c) {
                if (xa[c] >= 0) { if (S.s1 < 0) S.s1 = (int32_t)xa[c]; S.e1 = (int32_t)xa[c]; ++S.na; }
                if (xb[c] >= 0) { if (S.s2 < 0) S.s2 = (int32_t)xb[c]; S.e2 = (int32_t)xb[c]; ++S.nb; }
                if (xa[c] >= 0 && xb[c] >= 0 && J.ra[xa[c]] == J.rb[xb[c]]) ++S.match;
            }
            souts[J.out].score = to_f(R.score); souts[J.out].end_a = R.end_i; souts[J.out].end_b = R.end_j;
            swb_compose_stats(&souts[J.out], e->mode, retain_all, J.la_full, J.lb_full, S);
            continue;
        }
        inflate(J.ra, J.la_full, J.rb, J.lb_full, xa.data(), xb.data(), (int64_t)xa.size(), e->mode, retain_all, ra, rb);
        rc = fill_out(&outs[J.out], to_f(R.score), R.end_i, R.end_j, xa, xb, ra, rb);
        if (rc != SWB_OK) { e->err = "out of memory"; return rc; }
    }
    size_t pos = 0;
    while (pos < work.size()) {
        // chunk of pairs whose direction bytes fit the budget
        std::vector<KWPair> hp; std::vector<int64_t> out_off;
        int64_t dir_total = 0, out_total = 0;
        size_t end = pos;
        while (end < work.size()) {
            const AlnJob &J = jobs[work[end]];
            const int64_t db = kw_dir_bytes(J.la, J.lb, 8);              // upper bound over both strip widths
            if (!hp.empty() && dir_total + db > budget) break;
            KWPair P; P.a_off = J.a_off; P.b_off = J.b_off; P.la = J.la; P.lb = J.lb; P.dir_off = dir_total; P.bnd_off = 0; P.scr_off = 0; P.pssm_off = d_pssm ? 0 : -1;
            hp.push_back(P); out_off.push_back(out_total);
            dir_total += db; out_total += (int64_t)J.la + J.lb + 1;
            ++end;
        }
        const int np = (int)hp.size();
        KWPlan plan; kw_plan(e, hp.data(), np, true, &plan);       // assigns the direction-byte offsets
        dir_total = plan.dir_bytes;
        DevBuf d_pairs, d_res, d_bnd, d_scr, d_off, d_xa, d_xb, d_nc;
        if ((size_t)dir_total + 16 > e->dir_bytes) {           // engine-owned direction store, grown on demand and kept
            CUDA_TRY(e, cudaStreamSynchronize(e->stream));
            if (e->d_dir) cudaFree(e->d_dir);
            e->d_dir = nullptr; e->dir_bytes = 0;
            CUDA_TRY(e, cudaMalloc(&e->d_dir, (size_t)dir_total + 16));
            e->dir_bytes = (size_t)dir_total + 16;
        }
        CUDA_TRY(e, d_pairs.alloc(np * sizeof(KWPair), e->stream)); CUDA_TRY(e, d_res.alloc(np * sizeof(KWResult), e->stream));
        CUDA_TRY(e, d_bnd.alloc(plan.bnd_elems * sizeof(int4), e->stream));
        CUDA_TRY(e, d_scr.alloc(plan.scratch_ints * sizeof(int), e->stream));
        DevBuf d_st;
        if (souts) out_total = 0;                                  // no column lists in the counting form
        if (souts) CUDA_TRY(e, d_st.alloc(np * sizeof(KTStats), e->stream));
        else {
            CUDA_TRY(e, d_off.alloc(np * sizeof(int64_t), e->stream));
            CUDA_TRY(e, d_xa.alloc(out_total * sizeof(int32_t), e->stream)); CUDA_TRY(e, d_xb.alloc(out_total * sizeof(int32_t), e->stream));
            CUDA_TRY(e, d_nc.alloc(np * sizeof(int32_t), e->stream));
            CUDA_TRY(e, cudaMemcpyAsync(d_off.p, out_off.data(), np * sizeof(int64_t), cudaMemcpyHostToDevice, e->stream));
        }
        CUDA_TRY(e, cudaMemcpyAsync(d_pairs.p, hp.data(), np * sizeof(KWPair), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, cudaEventRecord(e->ev0, e->stream));
        CUDA_TRY(e, kw_launch(e, d_a, d_b, d_pairs.as<KWPair>(), np, plan, (uint8_t *)e->d_dir,
                              d_bnd.as<int4>(), d_scr.as<int>(), d_res.as<KWResult>(), e->stream, d_pssm));
        if (souts)
            CUDA_TRY(e, kt_stats_launch(e, d_pairs.as<KWPair>(), d_res.as<KWResult>(), np, (const uint8_t *)e->d_dir, plan.cols,
                                        d_ra, d_rb, d_st.as<KTStats>(), e->stream));
        else
            CUDA_TRY(e, kt_launch(e, d_pairs.as<KWPair>(), d_res.as<KWResult>(), np, (const uint8_t *)e->d_dir, plan.cols, d_off.as<int64_t>(),
                                  d_xa.as<int32_t>(), d_xb.as<int32_t>(), d_nc.as<int32_t>(), e->stream));
        CUDA_TRY(e, cudaEventRecord(e->ev1, e->stream));
        // results land in the engine's pinned traceback arena (a pageable destination would stage and block)
        const size_t tb_need = (size_t)np * (sizeof(KWResult) + sizeof(int32_t) + sizeof(KTStats)) + 2 * (size_t)out_total * sizeof(int32_t) + 64;
        if (tb_need > e->tb_bytes) {
            if (e->h_tb) cudaFreeHost(e->h_tb);
            e->h_tb = nullptr; e->tb_bytes = 0;
            CUDA_TRY(e, cudaMallocHost(&e->h_tb, tb_need));
            e->tb_bytes = tb_need;
        }
        KWResult *hr = (KWResult *)e->h_tb;
        int32_t *hnc = (int32_t *)(hr + np), *hxa = hnc + np, *hxb = hxa + out_total;
        CUDA_TRY(e, cudaMemcpyAsync(hr, d_res.p, np * sizeof(KWResult), cudaMemcpyDeviceToHost, e->stream));
        KTStats *hst = (KTStats *)(hxb + out_total);
        if (souts) CUDA_TRY(e, cudaMemcpyAsync(hst, d_st.p, np * sizeof(KTStats), cudaMemcpyDeviceToHost, e->stream));
        else {
            CUDA_TRY(e, cudaMemcpyAsync(hnc, d_nc.p, np * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(e, cudaMemcpyAsync(hxa, d_xa.p, out_total * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(e, cudaMemcpyAsync(hxb, d_xb.p, out_total * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
        }
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1);
        e->stats.kernel_ms += ms; e->stats.launches += 2;
        e->stats.h2d_bytes += (double)np * (sizeof(KWPair) + sizeof(int64_t));
        e->stats.d2h_bytes += (double)np * (sizeof(KWResult) + (souts ? sizeof(KTStats) : 4)) + 8.0 * out_total;
        if (souts) {
            for (int t = 0; t < np; ++t) {
                const AlnJob &J = jobs[work[pos + t]];
                swb_aln_stats *o = &souts[J.out];
                o->score = to_f(hr[t].score); o->end_a = hr[t].end_i; o->end_b = hr[t].end_j;
                if (hst[t].ncols < 0) { o->status = SWB_ERR_EMPTY; continue; }      // the reference panics on this pair (SA.rs:360-371)
                swb_compose_stats(o, e->mode, retain_all, J.la_full, J.lb_full, hst[t]);
            }
            pos = end;
            continue;
        }
        // reverse + string inflation per pair: independent, spread over host threads for large batches
        std::atomic<int> next(0), fail(SWB_OK);
        auto worker = [&]() {
            for (;;) {
                const int t = next.fetch_add(1);
                if (t >= np || fail.load() != SWB_OK) return;
                const AlnJob &J = jobs[work[pos + t]];
                if (hnc[t] < 0) { fail.store(SWB_ERR_EMPTY); return; }
                const int64_t nc = hnc[t];
                std::vector<int64_t> xa(nc), xb(nc);
                for (int64_t c = 0; c < nc; ++c) {              // reverse (SA.rs:384-385)
                    xa[c] = hxa[out_off[t] + nc - 1 - c]; xb[c] = hxb[out_off[t] + nc - 1 - c];
                }
                std::string ra, rb;
                inflate(J.ra, J.la_full, J.rb, J.lb_full, xa.data(), xb.data(), nc, e->mode, retain_all, ra, rb);
                if (fill_out(&outs[J.out], to_f(hr[t].score), hr[t].end_i, hr[t].end_j, xa, xb, ra, rb) != SWB_OK) { fail.store(SWB_ERR_NOMEM); return; }
            }
        };
        const int n_thr = (np >= 8) ? (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u) : 1;
        if (n_thr > 1) {
            std::vector<std::thread> pool;
            for (int w = 0; w < n_thr; ++w) pool.emplace_back(worker);
            for (auto &th : pool) th.join();
        } else worker();
        if (fail.load() == SWB_ERR_EMPTY) { e->err = "traceback left the matrix (the reference panics here, SA.rs:360-371)"; return SWB_ERR_EMPTY; }
        if (fail.load() != SWB_OK) { e->err = "out of memory"; return fail.load(); }
        pos = end;
    }
    return SWB_OK;
}

// Releases what a failed batch call has already filled in (the caller's array goes back to all-zero records).
static void free_outs(swb_alignment *outs, int64_t n)
{
    if (!outs) return;
    for (int64_t k = 0; k < n; ++k) swb_alignment_free(&outs[k]);
}

// outs: full alignments (lists + strings); souts: the counting form (swb_align_stats_batch).  Exactly one is non-null.
// full_a / full_b (may be null): lengths of the whole sequences when a_all / b_all hold only the prefixes to align
// (align_partial's partial_region, SA.rs:396-405: the DP runs on the prefixes, the padding loops over the whole sequences).
static int align_batch_core(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                            const int64_t *off_b, int64_t n, unsigned flags, swb_alignment *outs, const float *pssm,
                            swb_aln_stats *souts, const int64_t *part_a, const int64_t *part_b)
{
    const bool retain_all = (flags & SWB_RETAIN_ALL) != 0;
    const float inv = 1.0f / (float)e->scale;
    const bool dyadic = (e->scale & (e->scale - 1)) == 0;
    auto to_f = [&](int64_t s) { return dyadic ? (float)s * inv : (float)((double)s / e->scale); };
    // Local mode, pairs the score kernels take (both lengths <= 8192): scan first (K1, ~4 TCUPS).  The end cell bounds the
    // part of the matrix the traceback can touch (the prefix box, see AlnJob) and the score proves the s16 range, so the
    // pair can take the packed traceback path; pairs that score 0 need no device work at all.
    std::vector<float> pre_sc; std::vector<int64_t> pre_ea, pre_eb;
    double pre_ms = 0; int64_t pre_launches = 0;
    bool prescan = outs && !pssm && !part_a && !part_b && e->mode == SWB_LOCAL && e->o <= e->e && n > 0 && !getenv("SWB_NO_K3");
    for (int64_t k = 0; k < n && prescan; ++k)
        if (off_a[k + 1] - off_a[k] > 8192 || off_b[k + 1] - off_b[k] > 8192) prescan = false;
    if (prescan) {
        pre_sc.resize(n); pre_ea.resize(n); pre_eb.resize(n);
        const int rc = swb_score_batch(e, a_all, off_a, b_all, off_b, n, pre_sc.data(), pre_ea.data(), pre_eb.data());
        if (rc != SWB_OK) return rc;
        pre_ms = e->stats.kernel_ms; pre_launches = e->stats.launches;
    }
    e->stats = swb_stats{}; e->stats.scale = e->scale; e->stats.exact_scale = e->exact_scale;
    e->stats.kernel_ms = pre_ms; e->stats.launches = pre_launches;

    // encode on the host (small inputs on this path); unknown letters as SA.rs:669-679
    const int64_t na = n ? off_a[n] : 0, nb = n ? off_b[n] : 0;
    std::vector<uint8_t> ca(na + 16), cb(nb + 16);
    for (int64_t k = 0; k < na && !pssm; ++k) { int c = e->code_of[(unsigned char)a_all[k]]; if (c < 0) { e->err = std::string("unknown letter ") + a_all[k] + ". please set X to allow scoring for undefined letter pair."; return SWB_ERR_LETTER; } ca[k] = (uint8_t)c; }
    for (int64_t k = 0; k < nb && !pssm; ++k) { int c = e->code_of[(unsigned char)b_all[k]]; if (c < 0) { e->err = std::string("unknown letter ") + b_all[k] + ". please set X to allow scoring for undefined letter pair."; return SWB_ERR_LETTER; } cb[k] = (uint8_t)c; }

    // int32 lanes: the scores of a pair must stay well inside +-2^29 (the PAD score of the wavefront kernel)
    std::vector<AlnJob> jobs;                        // pairs that need the device
    for (int64_t k = 0; k < n; ++k) {
        const int64_t la_full = off_a[k + 1] - off_a[k], lb_full = off_b[k + 1] - off_b[k];
        if (la_full < 0 || lb_full < 0 || la_full > 0x3FFFFFF0 || lb_full > 0x3FFFFFF0) { e->err = "bad offsets"; return SWB_ERR_ARG; }
        // partial_region (SA.rs:396-405): the DP sees the prefixes, the padding the whole sequences
        const int64_t la = part_a ? std::min(std::max<int64_t>(part_a[k], 0), la_full) : la_full;
        const int64_t lb = part_b ? std::min(std::max<int64_t>(part_b[k], 0), lb_full) : lb_full;
        if (!swb_range_ok(e, la, lb, pssm ? 8000 : -1)) { e->err = "scores of a pair this long do not fit the int32 lanes"; return SWB_ERR_RANGE; }
        e->stats.cells += (double)la * (double)lb;
        if (la > 0 && lb > 0 && prescan && pre_ea[k] < 1) {          // local score 0: no alignment (SA.rs:346-348)
            std::vector<int64_t> xa, xb; std::string ra, rb;
            inflate(a_all + off_a[k], la_full, b_all + off_b[k], lb_full, nullptr, nullptr, 0, e->mode, retain_all, ra, rb);
            int rc = fill_out(&outs[k], 0.0f, -1, -1, xa, xb, ra, rb);
            if (rc != SWB_OK) { e->err = "out of memory"; return rc; }
            continue;
        }
        if (la > 0 && lb > 0) {
            AlnJob J; J.a_off = off_a[k]; J.b_off = off_b[k]; J.la = (int32_t)la; J.lb = (int32_t)lb;
            J.ra = a_all + off_a[k]; J.rb = b_all + off_b[k]; J.la_full = la_full; J.lb_full = lb_full; J.out = k;
            if (prescan) {                                             // trace inside the prefix box that ends at the end cell
                J.la = (int32_t)pre_ea[k]; J.lb = (int32_t)pre_eb[k];
                J.known_score = (int32_t)std::llround((double)pre_sc[k] * e->scale);
            }
            jobs.push_back(J);
            continue;
        }
        // an empty sequence: the walk runs along the boundary row/column only (SA.rs:100-118, 296-383)
        std::vector<int64_t> xa, xb; std::string ra, rb; float score = 0.0f; int64_t ei = -1, ej = -1;
        if (e->mode != SWB_LOCAL) {
            if (la == 0 && lb == 0) { e->err = "both sequences are empty (the reference panics here, SA.rs:360-363)"; return SWB_ERR_EMPTY; }
            const int32_t so = (e->mode == SWB_GLOBAL) ? e->o : 0, se = (e->mode == SWB_GLOBAL) ? e->e : 0;
            score = to_f((la + lb) * (int64_t)se + (so - se)); ei = la; ej = lb;
            if (10 * e->o + 10 * e->e == 0) { e->err = "zero gap penalties with an empty sequence (the reference panics here)"; return SWB_ERR_EMPTY; }
            for (int64_t t = 0; t < la; ++t) { xa.push_back(t); xb.push_back(-1); }
            for (int64_t t = 0; t < lb; ++t) { xa.push_back(-1); xb.push_back(t); }
        }
        if (souts) {
            KTStats S; S.ncols = (int32_t)xa.size(); S.na = (int32_t)la; S.nb = (int32_t)lb; S.match = 0;
            S.s1 = la ? 0 : -1; S.e1 = (int32_t)la - 1; S.s2 = lb ? 0 : -1; S.e2 = (int32_t)lb - 1;
            if (e->mode == SWB_LOCAL) { S.ncols = 0; S.na = 0; S.nb = 0; S.s1 = S.s2 = S.e1 = S.e2 = -1; }
            souts[k].score = score; souts[k].end_a = ei; souts[k].end_b = ej;
            swb_compose_stats(&souts[k], e->mode, retain_all, la_full, lb_full, S);
            continue;
        }
        inflate(a_all + off_a[k], la_full, b_all + off_b[k], lb_full, xa.data(), xb.data(), (int64_t)xa.size(), e->mode, retain_all, ra, rb);
        int rc = fill_out(&outs[k], score, ei, ej, xa, xb, ra, rb);
        if (rc != SWB_OK) { e->err = "out of memory"; return rc; }
    }
    if (jobs.empty()) return SWB_OK;

    DevBuf d_a, d_b, d_pssm;
    if (pssm) {                                      // PositionSpecificMatrix (SA.rs:583-623): scores[a + b*a_length] on the engine's grid
        const size_t np_ = (size_t)na * (size_t)nb;
        std::vector<int32_t> ps(np_ ? np_ : 1);
        for (size_t k = 0; k < np_; ++k) {
            const double v = (double)pssm[k] * e->scale, r = std::nearbyint(v);
            if (std::fabs(v - r) > 1e-4 * e->scale || std::fabs(r) > 8000) { e->err = "position-specific scores are not on the engine's integer grid (or exceed 8000)"; return SWB_ERR_SCALE; }
            ps[k] = (int32_t)r;
        }
        CUDA_TRY(e, d_pssm.alloc(ps.size() * sizeof(int32_t), e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(d_pssm.p, ps.data(), ps.size() * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));      // ps is a local pageable vector
        e->stats.h2d_bytes += (double)ps.size() * sizeof(int32_t);
    }
    CUDA_TRY(e, d_a.alloc(ca.size(), e->stream)); CUDA_TRY(e, d_b.alloc(cb.size(), e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(d_a.p, ca.data(), ca.size(), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(d_b.p, cb.data(), cb.size(), cudaMemcpyHostToDevice, e->stream));
    e->stats.h2d_bytes += (double)ca.size() + (double)cb.size();
    DevBuf d_ra, d_rb;                                // the letters as given: the equal-letter test of the counting form
    if (souts) {
        CUDA_TRY(e, d_ra.alloc((size_t)na + 16, e->stream)); CUDA_TRY(e, d_rb.alloc((size_t)nb + 16, e->stream));
        if (na) CUDA_TRY(e, cudaMemcpyAsync(d_ra.p, a_all, na, cudaMemcpyHostToDevice, e->stream));
        if (nb) CUDA_TRY(e, cudaMemcpyAsync(d_rb.p, b_all, nb, cudaMemcpyHostToDevice, e->stream));
        e->stats.h2d_bytes += (double)na + (double)nb;
    }
    return align_jobs_device(e, d_a.as<uint8_t>(), d_b.as<uint8_t>(), jobs, retain_all, outs, souts,
                             pssm ? d_pssm.as<int32_t>() : nullptr, d_ra.as<uint8_t>(), d_rb.as<uint8_t>());
}

static int align_batch_impl(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                            const int64_t *off_b, int64_t n, unsigned flags, swb_alignment *outs, const float *pssm,
                            swb_aln_stats *souts = nullptr, const int64_t *part_a = nullptr, const int64_t *part_b = nullptr)
{
    if (!e) return SWB_ERR_ARG;
    if (n < 0 || (n > 0 && (!off_a || !off_b || (!outs && !souts)))) { e->err = "bad argument"; return SWB_ERR_ARG; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    if (outs) memset(outs, 0, sizeof(swb_alignment) * (size_t)n);
    if (souts) memset(souts, 0, sizeof(swb_aln_stats) * (size_t)n);
    const int rc = align_batch_core(e, a_all, off_a, b_all, off_b, n, flags, outs, pssm, souts, part_a, part_b);
    if (rc != SWB_OK) free_outs(outs, n);            // nothing half-filled is handed back (the records are all-zero again)
    return rc;
}

extern "C" int swb_align_batch(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                               const int64_t *off_b, int64_t n, unsigned flags, swb_alignment *outs)
{
    if (e && n > 0 && !outs) { e->err = "bad argument"; return SWB_ERR_ARG; }
    return align_batch_impl(e, a_all, off_a, b_all, off_b, n, flags, outs, nullptr);
}

// align(.., retain_all) of n pairs, reduced on the device to what the clustering loop reads off the strings
// (OCL-main.rs:205-233): no column lists cross PCIe and no strings are built.
extern "C" int swb_align_stats_batch(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                                     const int64_t *off_b, int64_t n, unsigned flags, swb_aln_stats *outs)
{
    if (e && n > 0 && !outs) { e->err = "bad argument"; return SWB_ERR_ARG; }
    return align_batch_impl(e, a_all, off_a, b_all, off_b, n, flags, nullptr, nullptr, outs);
}

// align() of an engine built over a PositionSpecificMatrix (SA.rs:583-623, KAT psm_test SA.rs:1057-1086): the score of
// (position i of a, position j of b) is scores[i + j*la]; the letters are only used for the output strings.
extern "C" int swb_align_pssm(swb_engine *e, const float *scores, const char *a, int64_t la, const char *b, int64_t lb,
                              int retain_all, swb_alignment *out)
{
    if (!e) return SWB_ERR_ARG;
    if (!out || !scores || la < 1 || lb < 1 || !a || !b) { e->err = "bad argument (position-specific alignment needs two non-empty sequences)"; return SWB_ERR_ARG; }
    const int64_t oa[2] = {0, la}, ob[2] = {0, lb};
    return align_batch_impl(e, a, oa, b, ob, 1, retain_all ? SWB_RETAIN_ALL : 0u, out, scores);
}

// align_partial (SA.rs:393-551): DP on the prefixes a[0..pa), b[0..pb); the retain_all padding runs over the whole
// sequences (SA.rs:512-549); score_only returns after the score scan (SA.rs:438-467) with empty strings.
extern "C" int swb_align_partial(swb_engine *e, const char *a, int64_t la, const char *b, int64_t lb, int64_t pa, int64_t pb,
                                 int retain_all, int score_only, swb_alignment *out)
{
    if (!e) return SWB_ERR_ARG;
    if (!out || la < 0 || lb < 0 || pa < 0 || pb < 0) { e->err = "bad argument"; return SWB_ERR_ARG; }
    static const char empty[1] = {0};
    pa = std::min(pa, la); pb = std::min(pb, lb);
    const int64_t oa[2] = {0, la}, ob[2] = {0, lb};
    if (score_only) {
        memset(out, 0, sizeof(*out));
        if (e->mode != SWB_LOCAL && pa == 0 && pb == 0) { e->err = "both sequences are empty (the reference panics here, SA.rs:360-363)"; return SWB_ERR_EMPTY; }
        const int64_t pa2[2] = {0, pa}, pb2[2] = {0, pb};
        float sc = 0; int64_t ea = -1, eb = -1;
        int rc = swb_score_batch(e, a ? a : empty, pa2, b ? b : empty, pb2, 1, &sc, &ea, &eb);
        if (rc != SWB_OK) return rc;
        out->score = sc; out->end_a = ea; out->end_b = eb;
        out->idx_a = (int64_t *)malloc(sizeof(int64_t)); out->idx_b = (int64_t *)malloc(sizeof(int64_t));
        out->aln_a = dup_n(std::string()); out->aln_b = dup_n(std::string());
        if (!out->idx_a || !out->idx_b || !out->aln_a || !out->aln_b) { swb_alignment_free(out); e->err = "out of memory"; return SWB_ERR_NOMEM; }
        return SWB_OK;
    }
    return align_batch_impl(e, a ? a : empty, oa, b ? b : empty, ob, 1, retain_all ? SWB_RETAIN_ALL : 0u, out, nullptr, nullptr, &pa, &pb);
}

extern "C" int swb_align(swb_engine *e, const char *a, int64_t la, const char *b, int64_t lb, int retain_all,
                         swb_alignment *out)
{
    if (!e) return SWB_ERR_ARG;
    if (!out || la < 0 || lb < 0) { e->err = "bad argument"; return SWB_ERR_ARG; }
    const int64_t oa[2] = {0, la}, ob[2] = {0, lb};
    static const char empty[1] = {0};
    return swb_align_batch(e, a ? a : empty, oa, b ? b : empty, ob, 1, retain_all ? SWB_RETAIN_ALL : 0u, out);
}
