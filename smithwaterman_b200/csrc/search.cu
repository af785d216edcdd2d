// Database search driver behind the C ABI: a database (shard) resident on the device, one query per call, top-K on the
// device, traceback of the hits from the resident residues.
//
// The reference has no such driver; its shape is the all-vs-all loop of the CLIs with one fixed query
// (SA-main.rs:35-45, OCL-main.rs:295-316: every `sw.align(ss1, ss2, true)` of one record against all records of the
// other file).  Per (query, subject) pair the scan has the semantics of align_partial(.., score_only) (SA.rs:438-467)
// and the hits those of align(.., retain_all) (SA.rs:389).
//
// Sharding: subjects are dealt round-robin over the stable longest-first order, every shard is resident
// on its own GPU, and the only exchange between shards is the per-shard top-K list (K x 32 bytes), merged by
// swb_search_merge with the key (score desc, database index asc) -- which makes the result independent of the number of
// shards.  Each hit is traced back by the shard that owns it (swb_search_align), so the traceback work divides by the
// number of GPUs as well.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <thread>

#include "swb_internal.h"

#define SEL_THREADS 256
#define SWB_MAX_TOPK 4096

// ---- selection: exact top-K of n (score, index) records by radix select on 64-bit keys ------------------------------
// key = (score + 2^31) << 32 | (2^32 - 1 - local index): larger key = better hit (score desc, index asc); keys are unique.
struct SelState {
    unsigned long long prefix, mask;    // the digits of the K-th largest key fixed so far
    unsigned k_rem;                     // rank still to resolve inside the current prefix class
    unsigned n_out;                     // gather cursor
    unsigned n_ovf;                     // subjects whose s16 lanes overflowed
    unsigned pad;
    unsigned hist[256];
};

__device__ __forceinline__ unsigned long long sel_key(const K1Result *res, int64_t i)
{
    const int score = res[i].score;
    return ((unsigned long long)((unsigned)score ^ 0x80000000u) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)i);
}

__global__ void __launch_bounds__(SEL_THREADS) sel_init_kernel(SelState *S, unsigned k)
{
    if (threadIdx.x == 0) { S->prefix = 0; S->mask = 0; S->k_rem = k; S->n_out = 0; S->n_ovf = 0; }
    S->hist[threadIdx.x] = 0;
}

// Subjects flagged by K1 / K2 (s16 range exceeded): the list for the int32 re-run.
__global__ void __launch_bounds__(SEL_THREADS) sel_ovf_kernel(const K1Result *__restrict__ res, int64_t n, SelState *S, int32_t *__restrict__ list)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && (res[i].flags & 1)) list[atomicAdd(&S->n_ovf, 1u)] = (int32_t)i;
}

__global__ void __launch_bounds__(SEL_THREADS) sel_hist_kernel(const K1Result *__restrict__ res, int64_t n, SelState *S, int shift)
{
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const unsigned long long prefix = S->prefix, mask = S->mask;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = sel_key(res, i);
        if ((key & mask) == prefix) atomicAdd(&h[(unsigned)(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (h[threadIdx.x]) atomicAdd(&S->hist[threadIdx.x], h[threadIdx.x]);
}

// One block: the digit of the K-th largest key at this position; clears the histogram for the next round.
__global__ void __launch_bounds__(SEL_THREADS) sel_pick_kernel(SelState *S, int shift)
{
    __shared__ unsigned h[256];
    h[threadIdx.x] = S->hist[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned k = S->k_rem, above = 0;
        int d = 255;
        for (; d > 0; --d) {
            if (above + h[d] >= k) break;
            above += h[d];
        }
        S->k_rem = k - above;
        S->prefix |= (unsigned long long)d << shift;
        S->mask |= 255ull << shift;
    }
    S->hist[threadIdx.x] = 0;
}

__global__ void __launch_bounds__(SEL_THREADS) sel_gather_kernel(const K1Result *__restrict__ res, int64_t n, SelState *S, unsigned long long *__restrict__ keys, unsigned cap)
{
    const unsigned long long kth = S->prefix;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = sel_key(res, i);
        if (key >= kth) { const unsigned p = atomicAdd(&S->n_out, 1u); if (p < cap) keys[p] = key; }
    }
}

// One block: sorts the selected keys (bitonic, descending) and writes the hit records.
__global__ void __launch_bounds__(1024) sel_emit_kernel(const unsigned long long *__restrict__ keys, unsigned k, unsigned pow2, const K1Result *__restrict__ res,
                                                        const int64_t *__restrict__ gidx, int scale, float inv, int dyadic, swb_hit *__restrict__ hits)
{
    extern __shared__ unsigned long long sk[];
    for (unsigned t = threadIdx.x; t < pow2; t += blockDim.x) sk[t] = t < k ? keys[t] : 0ull;
    __syncthreads();
    for (unsigned size = 2; size <= pow2; size <<= 1)
        for (unsigned stride = size >> 1; stride > 0; stride >>= 1) {
            for (unsigned t = threadIdx.x; t < pow2; t += blockDim.x) {
                const unsigned u = t ^ stride;
                if (u > t) {
                    const bool desc = (t & size) == 0;
                    const unsigned long long a = sk[t], b = sk[u];
                    if (desc ? a < b : a > b) { sk[t] = b; sk[u] = a; }
                }
            }
            __syncthreads();
        }
    for (unsigned t = threadIdx.x; t < k; t += blockDim.x) {
        const unsigned long long key = sk[t];
        const int64_t i = (int64_t)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
        const K1Result r = res[i];
        swb_hit h;
        h.score = dyadic ? (float)r.score * inv : (float)((double)r.score / (double)scale);
        h.shard = 0;
        h.index = gidx[i]; h.end_q = r.end_i; h.end_t = r.end_j;
        hits[t] = h;
    }
}

__global__ void __launch_bounds__(SEL_THREADS) sel_patch_kernel(K1Result *__restrict__ res, const int32_t *__restrict__ idx, const KWResult *__restrict__ kw, int n)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    K1Result o; o.score = kw[t].score; o.end_i = kw[t].end_i; o.end_j = kw[t].end_j; o.flags = 0;
    res[idx[t]] = o;
}

// ---- shard selection (host only; the same deal on every rank) ---------------------------------------------------------
extern "C" int swb_db_shard(const int64_t *off, int64_t n, int rank, int world, int64_t *idx, int64_t *n_idx)
{
    if (!off || n < 0 || world < 1 || rank < 0 || rank >= world || !idx || !n_idx) return SWB_ERR_ARG;
    std::vector<int64_t> order((size_t)n);
    std::iota(order.begin(), order.end(), (int64_t)0);
    std::stable_sort(order.begin(), order.end(), [&](int64_t x, int64_t y) { return off[x + 1] - off[x] > off[y + 1] - off[y]; });
    int64_t m = 0;
    for (int64_t p = rank; p < n; p += world) idx[m++] = order[(size_t)p];
    std::sort(idx, idx + m);
    *n_idx = m;
    return SWB_OK;
}

extern "C" void swb_db_free(swb_engine *e, swb_db *db)
{
    if (!db) return;
    cudaSetDevice(db->device);
    if (e) cudaStreamSynchronize(e->stream);
    cudaFree(db->d_codes); cudaFree(db->d_off); cudaFree(db->d_gidx); cudaFree(db->d_items); cudaFree(db->d_sweeps);
    cudaFree(db->d_res); cudaFree(db->d_query); cudaFree(db->d_sel);
    if (db->h_hits) cudaFreeHost(db->h_hits);
    delete db;
}

static size_t sel_bytes(int64_t n) { return sizeof(SelState) + SWB_MAX_TOPK * sizeof(unsigned long long) + ((size_t)n + 1) * sizeof(int32_t); }

// letters / off: the subjects of this shard, already gathered (off[0] == 0); gidx ascending.
static int db_build(swb_engine *e, std::vector<char> &&letters, std::vector<int64_t> &&off, std::vector<int64_t> &&gidx, swb_db **out)
{
    swb_db *db = new swb_db();
    struct Guard { swb_engine *e; swb_db *db; bool ok = false; ~Guard() { if (!ok) swb_db_free(e, db); } } guard{e, db};
    db->device = e->device;
    db->n = (int64_t)off.size() - 1; db->n_res = off.back();
    db->h_letters = std::move(letters); db->h_off = std::move(off); db->h_gidx = std::move(gidx);
    const int64_t n = db->n, nr = db->n_res;
    db->lb.resize((size_t)n);
    std::vector<int32_t> live;
    for (int64_t k = 0; k < n; ++k) {
        const int64_t l = db->h_off[k + 1] - db->h_off[k];
        if (l < 0 || l > 0x3FFFFFF0) { e->err = "bad offsets"; return SWB_ERR_ARG; }
        db->lb[k] = (int32_t)l;
        if (l == 0) db->degenerate.push_back((int32_t)k); else live.push_back((int32_t)k);
        if (k > 0 && db->h_gidx[k] <= db->h_gidx[k - 1]) { e->err = "database indices must be ascending"; return SWB_ERR_ARG; }
    }
    const int64_t nr_p = (nr + 15) / 16 * 16 + 32;
    CUDA_TRY(e, cudaMalloc(&db->d_codes, (size_t)nr_p));
    CUDA_TRY(e, cudaMalloc(&db->d_off, (size_t)(n + 1) * sizeof(int64_t)));
    CUDA_TRY(e, cudaMalloc(&db->d_gidx, (size_t)std::max<int64_t>(n, 1) * sizeof(int64_t)));
    CUDA_TRY(e, cudaMalloc(&db->d_res, (size_t)std::max<int64_t>(n, 1) * sizeof(K1Result)));
    CUDA_TRY(e, cudaMalloc(&db->d_sel, sel_bytes(n)));
    CUDA_TRY(e, cudaMallocHost(&db->h_hits, SWB_MAX_TOPK * sizeof(swb_hit) + 256));
    CUDA_TRY(e, cudaMemsetAsync(db->d_codes, 0, (size_t)nr_p, e->stream));
    if (nr) CUDA_TRY(e, cudaMemcpyAsync(db->d_codes, db->h_letters.data(), (size_t)nr, cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(db->d_off, db->h_off.data(), (size_t)(n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, e->stream));
    if (n) CUDA_TRY(e, cudaMemcpyAsync(db->d_gidx, db->h_gidx.data(), (size_t)n * sizeof(int64_t), cudaMemcpyHostToDevice, e->stream));
    int32_t *d_err = reinterpret_cast<int32_t *>(db->d_sel);           // scratch is free until the first search
    CUDA_TRY(e, cudaMemsetAsync(d_err, 0, sizeof(int32_t), e->stream));
    if (nr) CUDA_TRY(e, batch_encode(e, db->d_codes, nr, d_err, e->stream));
    int32_t *flag = reinterpret_cast<int32_t *>(db->h_hits);
    CUDA_TRY(e, cudaMemcpyAsync(flag, d_err, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
    // K2 work items of the whole shard (the packing does not depend on the query)
    if (!live.empty()) {
        const size_t cap = k2_items_cap(db->lb, live);
        std::vector<int32_t> items(cap * 33);
        const int64_t ni = k2_build_items(db->lb, live, items.data(), items.data() + cap * 32, cap);
        if (ni < 0) { e->err = "K2 item table overflow"; return SWB_ERR_NOMEM; }
        CUDA_TRY(e, cudaMalloc(&db->d_items, (size_t)ni * 32 * sizeof(int32_t)));
        CUDA_TRY(e, cudaMalloc(&db->d_sweeps, (size_t)ni * sizeof(int32_t)));
        CUDA_TRY(e, cudaMemcpy(db->d_items, items.data(), (size_t)ni * 32 * sizeof(int32_t), cudaMemcpyHostToDevice));
        CUDA_TRY(e, cudaMemcpy(db->d_sweeps, items.data() + cap * 32, (size_t)ni * sizeof(int32_t), cudaMemcpyHostToDevice));
        db->n_items = (int)ni; db->n_k2 = (int)live.size();
    }
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    if (*flag) { e->err = "unknown letter. please set X to allow scoring for undefined letter pair."; return SWB_ERR_LETTER; }   // SA.rs:676
    e->stats.h2d_bytes = (double)nr + (double)(2 * n + 1) * sizeof(int64_t) + (double)db->n_items * 33 * sizeof(int32_t);
    guard.ok = true;
    *out = db;
    return SWB_OK;
}

extern "C" int swb_db_upload(swb_engine *e, const char *letters, const int64_t *off, int64_t n, const int64_t *global_index, swb_db **out)
{
    if (!e) return SWB_ERR_ARG;
    if (!out || n < 0 || !off || n > 0x7FFFFFF0 || (n > 0 && off[n] > off[0] && !letters)) { e->err = "bad argument"; return SWB_ERR_ARG; }
    *out = nullptr;
    CUDA_TRY(e, cudaSetDevice(e->device));
    const int64_t base = off[0];
    std::vector<int64_t> o((size_t)n + 1), g((size_t)n);
    for (int64_t k = 0; k <= n; ++k) o[k] = off[k] - base;
    for (int64_t k = 0; k < n; ++k) g[k] = global_index ? global_index[k] : k;
    std::vector<char> l(letters ? letters + base : nullptr, letters ? letters + off[n] : nullptr);
    return db_build(e, std::move(l), std::move(o), std::move(g), out);
}

extern "C" int swb_db_upload_shard(swb_engine *e, const char *letters, const int64_t *off, int64_t n_total, int rank, int world, swb_db **out)
{
    if (!e) return SWB_ERR_ARG;
    if (!out || n_total < 0 || !off || n_total > 0x7FFFFFF0 || world < 1 || rank < 0 || rank >= world) { e->err = "bad argument"; return SWB_ERR_ARG; }
    *out = nullptr;
    CUDA_TRY(e, cudaSetDevice(e->device));
    std::vector<int64_t> g((size_t)n_total);
    int64_t m = 0;
    if (swb_db_shard(off, n_total, rank, world, g.data(), &m) != SWB_OK) { e->err = "bad argument"; return SWB_ERR_ARG; }
    g.resize((size_t)m);
    std::vector<int64_t> o((size_t)m + 1, 0);
    for (int64_t k = 0; k < m; ++k) o[k + 1] = o[k] + (off[g[k] + 1] - off[g[k]]);
    std::vector<char> l((size_t)o[m]);
    for (int64_t k = 0; k < m; ++k) if (o[k + 1] > o[k]) memcpy(l.data() + o[k], letters + off[g[k]], (size_t)(o[k + 1] - o[k]));
    return db_build(e, std::move(l), std::move(o), std::move(g), out);
}

extern "C" int64_t swb_db_count(const swb_db *db) { return db ? db->n : 0; }
extern "C" int64_t swb_db_residues(const swb_db *db) { return db ? db->n_res : 0; }

// ---- the search -----------------------------------------------------------------------------------------------------
static int run_select(swb_engine *e, swb_db *db, unsigned k, cudaStream_t st)
{
    SelState *S = reinterpret_cast<SelState *>(db->d_sel);
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(S + 1);
    const int64_t n = db->n;
    const int blocks = (int)std::min<int64_t>((n + SEL_THREADS - 1) / SEL_THREADS, 4 * e->sm_count);
    sel_init_kernel<<<1, SEL_THREADS, 0, st>>>(S, k);
    for (int shift = 56; shift >= 0; shift -= 8) {
        sel_hist_kernel<<<blocks, SEL_THREADS, 0, st>>>(db->d_res, n, S, shift);
        sel_pick_kernel<<<1, SEL_THREADS, 0, st>>>(S, shift);
    }
    sel_gather_kernel<<<blocks, SEL_THREADS, 0, st>>>(db->d_res, n, S, keys, SWB_MAX_TOPK);
    unsigned pow2 = 1;
    while (pow2 < k) pow2 <<= 1;
    const float inv = 1.0f / (float)e->scale;
    const int dyadic = (e->scale & (e->scale - 1)) == 0;
    sel_emit_kernel<<<1, 1024, pow2 * sizeof(unsigned long long), st>>>(keys, k, pow2, db->d_res, db->d_gidx, e->scale, inv, dyadic,
                                                                         reinterpret_cast<swb_hit *>(db->h_hits));
    CUDA_TRY(e, cudaGetLastError());
    e->stats.launches += 19;
    return SWB_OK;
}

// Score of every subject into db->d_res (device), s16 overflows re-run in int32.  Leaves the stream synchronised.
static int scan_db(swb_engine *e, const char *query, int64_t lq, swb_db *db, unsigned topk_eff)
{
    cudaStream_t st = e->stream;
    const int64_t n = db->n;
    SelState *S = reinterpret_cast<SelState *>(db->d_sel);
    int32_t *ovf_list = reinterpret_cast<int32_t *>(reinterpret_cast<char *>(db->d_sel) + sizeof(SelState) + SWB_MAX_TOPK * sizeof(unsigned long long));
    unsigned *h_cnt = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(db->h_hits) + SWB_MAX_TOPK * sizeof(swb_hit));
    if (e->stage_bytes < 4096) {
        if (e->h_stage) cudaFreeHost(e->h_stage);
        e->h_stage = nullptr; e->stage_bytes = 0;
        CUDA_TRY(e, cudaMallocHost(&e->h_stage, 65536)); e->stage_bytes = 65536;
    }
    e->stage_used = 0;
    swb_batch *b = nullptr;
    CUDA_TRY(e, cudaEventRecord(e->ev0, st));
    int rc = batch_upload_impl(e, query, nullptr, nullptr, nullptr, n, lq, true, st, st, &b, db);
    if (rc != SWB_OK) return rc;
    struct Rel { swb_engine *e; swb_batch *b; ~Rel() { batch_release(e, b); } } rel{e, b};
    int64_t launches = 0;
    rc = batch_launch(e, b, &launches);
    if (rc != SWB_OK) return rc;
    CUDA_TRY(e, batch_collect_kw(e, b, st));
    // subjects of length 0 (and every subject when the query is empty): closed-form boundary values (SA.rs:88-118, 438-467)
    std::vector<K1Result> deg;
    std::vector<int32_t> deg_idx;
    if (lq == 0) { deg_idx.resize((size_t)n); std::iota(deg_idx.begin(), deg_idx.end(), 0); } else deg_idx = db->degenerate;
    if (!deg_idx.empty()) {
        deg.resize(deg_idx.size());
        for (size_t t = 0; t < deg_idx.size(); ++t) {
            const int64_t lb = db->lb[deg_idx[t]], len = lq + lb;
            K1Result r; r.flags = 0;
            if (e->mode == SWB_LOCAL) { r.score = 0; r.end_i = -1; r.end_j = -1; }
            else {
                const int32_t so = (e->mode == SWB_GLOBAL) ? e->o : 0, se = (e->mode == SWB_GLOBAL) ? e->e : 0;
                r.score = (int32_t)(len ? len * se + (so - se) : 0); r.end_i = (int32_t)lq; r.end_j = (int32_t)lb;
            }
            deg[t] = r;
            CUDA_TRY(e, cudaMemcpyAsync(db->d_res + deg_idx[t], &deg[t], sizeof(K1Result), cudaMemcpyHostToDevice, st));
        }
    }
    // the overflow list and (speculatively) the selection; one synchronisation brings both counts back
    sel_init_kernel<<<1, SEL_THREADS, 0, st>>>(S, topk_eff);
    if (n) sel_ovf_kernel<<<(int)((n + SEL_THREADS - 1) / SEL_THREADS), SEL_THREADS, 0, st>>>(db->d_res, n, S, ovf_list);
    CUDA_TRY(e, cudaMemcpyAsync(h_cnt, &S->n_ovf, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    if (topk_eff) { rc = run_select(e, db, topk_eff, st); if (rc != SWB_OK) return rc; }
    CUDA_TRY(e, cudaStreamSynchronize(st));
    e->stats.launches += launches + 2;
    const unsigned nr = *h_cnt;
    if (nr > 0) {                                     // int32 re-run of the overflowed subjects (KW, no direction bytes), then select again
        std::vector<int32_t> redo(nr);
        CUDA_TRY(e, cudaMemcpy(redo.data(), ovf_list, nr * sizeof(int32_t), cudaMemcpyDeviceToHost));
        std::sort(redo.begin(), redo.end());
        std::vector<KWPair> hp(nr);
        for (unsigned t = 0; t < nr; ++t) {
            hp[t].a_off = 0; hp[t].b_off = db->h_off[redo[t]]; hp[t].la = (int32_t)lq; hp[t].lb = db->lb[redo[t]];
            hp[t].dir_off = -1; hp[t].pssm_off = -1;
        }
        KWPlan pl; kw_plan(e, hp.data(), (int)nr, false, &pl);
        KWPair *dp = nullptr; KWResult *dr = nullptr; int4 *dbn = nullptr; int *ds = nullptr; int32_t *di = nullptr;
        CUDA_TRY(e, cudaMallocAsync((void **)&dp, nr * sizeof(KWPair), st)); CUDA_TRY(e, cudaMallocAsync((void **)&dr, nr * sizeof(KWResult), st));
        CUDA_TRY(e, cudaMallocAsync((void **)&dbn, pl.bnd_elems * sizeof(int4), st)); CUDA_TRY(e, cudaMallocAsync((void **)&ds, pl.scratch_ints * sizeof(int), st));
        CUDA_TRY(e, cudaMallocAsync((void **)&di, nr * sizeof(int32_t), st));
        CUDA_TRY(e, cudaMemcpyAsync(dp, hp.data(), nr * sizeof(KWPair), cudaMemcpyHostToDevice, st));
        CUDA_TRY(e, cudaMemcpyAsync(di, redo.data(), nr * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        const uint8_t *ca = nullptr, *cb = nullptr;
        batch_device_views(b, &ca, &cb);
        CUDA_TRY(e, kw_launch(e, ca, cb, dp, (int)nr, pl, nullptr, dbn, ds, dr, st));
        sel_patch_kernel<<<(nr + SEL_THREADS - 1) / SEL_THREADS, SEL_THREADS, 0, st>>>(db->d_res, di, dr, (int)nr);
        if (topk_eff) { rc = run_select(e, db, topk_eff, st); if (rc != SWB_OK) return rc; }
        CUDA_TRY(e, cudaStreamSynchronize(st));
        cudaFreeAsync(dp, st); cudaFreeAsync(dr, st); cudaFreeAsync(dbn, st); cudaFreeAsync(ds, st); cudaFreeAsync(di, st);
        e->stats.i32_reruns += nr; e->stats.launches += 2;
    }
    CUDA_TRY(e, cudaEventRecord(e->ev1, st));
    CUDA_TRY(e, cudaEventSynchronize(e->ev1));
    float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1);
    e->stats.kernel_ms = ms;
    db->last_lq = lq;
    return SWB_OK;
}

extern "C" int swb_search(swb_engine *e, const char *query, int64_t lq, swb_db *db, int topk, swb_hit *hits, int64_t *n_hits)
{
    if (!e) return SWB_ERR_ARG;
    if (!db || lq < 0 || (lq > 0 && !query) || topk < 0 || topk > SWB_MAX_TOPK || (topk > 0 && !hits) || !n_hits || lq > 0x3FFFFFF0) { e->err = "bad argument"; return SWB_ERR_ARG; }
    if (db->device != e->device) { e->err = "the database lives on another device"; return SWB_ERR_ARG; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    e->stats = swb_stats{}; e->stats.scale = e->scale; e->stats.exact_scale = e->exact_scale;
    *n_hits = 0;
    const unsigned k = (unsigned)std::min<int64_t>(topk, db->n);
    int rc = scan_db(e, query, lq, db, k);
    if (rc != SWB_OK) return rc;
    if (k) memcpy(hits, db->h_hits, (size_t)k * sizeof(swb_hit));
    *n_hits = k;
    e->stats.d2h_bytes += (double)k * sizeof(swb_hit) + 8;
    return SWB_OK;
}

// The whole score vector of the last search (tests, full-vector cross-checks between shardings): local subject order.
extern "C" int swb_db_scores(swb_engine *e, swb_db *db, float *scores, int64_t *end_q, int64_t *end_t)
{
    if (!e) return SWB_ERR_ARG;
    if (!db || db->last_lq < 0 || (db->n > 0 && (!scores || !end_q || !end_t))) { e->err = "no search has run on this database"; return SWB_ERR_ARG; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    std::vector<K1Result> r((size_t)db->n);
    if (db->n) CUDA_TRY(e, cudaMemcpy(r.data(), db->d_res, (size_t)db->n * sizeof(K1Result), cudaMemcpyDeviceToHost));
    const float inv = 1.0f / (float)e->scale;
    const bool dyadic = (e->scale & (e->scale - 1)) == 0;
    for (int64_t k = 0; k < db->n; ++k) {
        scores[k] = dyadic ? (float)r[k].score * inv : (float)((double)r[k].score / e->scale);
        end_q[k] = r[k].end_i; end_t[k] = r[k].end_j;
    }
    return SWB_OK;
}

// Host only: the global top-K of the concatenated per-shard lists; key (score desc, database index asc).
extern "C" int swb_search_merge(const swb_hit *parts, int64_t n_parts, int topk, swb_hit *out, int64_t *n_out)
{
    if (n_parts < 0 || topk < 0 || (n_parts > 0 && !parts) || !n_out || (topk > 0 && n_parts > 0 && !out)) return SWB_ERR_ARG;
    std::vector<swb_hit> v(parts, parts + n_parts);
    std::sort(v.begin(), v.end(), [](const swb_hit &x, const swb_hit &y) { return x.score > y.score || (x.score == y.score && x.index < y.index); });
    const int64_t m = std::min<int64_t>(topk, n_parts);
    for (int64_t k = 0; k < m; ++k) out[k] = v[(size_t)k];
    *n_out = m;
    return SWB_OK;
}

// Traceback of the hits this shard owns (align(query, subject, retain_all), SA.rs:389).  owned[k] = 1 and alns[k] filled
// for those; the others are left zeroed (another shard traces them).  Local mode traces inside the prefix box that ends
// at the hit's end cell (see AlnJob).
extern "C" int swb_search_align(swb_engine *e, const char *query, int64_t lq, swb_db *db, const swb_hit *hits, int64_t n_hits,
                                unsigned flags, swb_alignment *alns, int32_t *owned)
{
    if (!e) return SWB_ERR_ARG;
    if (!db || lq < 0 || (lq > 0 && !query) || n_hits < 0 || (n_hits > 0 && (!hits || !alns))) { e->err = "bad argument"; return SWB_ERR_ARG; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    memset(alns, 0, sizeof(swb_alignment) * (size_t)n_hits);
    e->stats = swb_stats{}; e->stats.scale = e->scale; e->stats.exact_scale = e->exact_scale;
    const bool retain_all = (flags & SWB_RETAIN_ALL) != 0;
    std::vector<AlnJob> jobs;
    for (int64_t k = 0; k < n_hits; ++k) {
        if (owned) owned[k] = 0;
        auto it = std::lower_bound(db->h_gidx.begin(), db->h_gidx.end(), hits[k].index);
        if (it == db->h_gidx.end() || *it != hits[k].index) continue;
        const int64_t loc = it - db->h_gidx.begin();
        if (owned) owned[k] = 1;
        const int64_t lb = db->lb[loc];
        AlnJob J; J.a_off = 0; J.b_off = db->h_off[loc]; J.la = (int32_t)lq; J.lb = (int32_t)lb;
        J.ra = query; J.rb = db->h_letters.data() + db->h_off[loc]; J.la_full = lq; J.lb_full = lb; J.out = k;
        if (e->mode == SWB_LOCAL && hits[k].end_q >= 1 && hits[k].end_t >= 1 && hits[k].end_q <= lq && hits[k].end_t <= lb) {
            J.la = (int32_t)hits[k].end_q; J.lb = (int32_t)hits[k].end_t;
            J.known_score = (int32_t)std::llround((double)hits[k].score * e->scale);      // the scan's score: the hit may take the packed path
        }
        if (lq == 0 || lb == 0 || (e->mode == SWB_LOCAL && hits[k].end_q < 1)) {     // nothing to trace on the device: closed forms
            const int64_t oa[2] = {0, lq}, ob[2] = {0, lb};
            swb_stats keep = e->stats;
            int rc = swb_align_batch(e, query ? query : "", oa, J.rb, ob, 1, flags, &alns[k]);
            keep.cells += e->stats.cells; e->stats = keep;
            if (rc != SWB_OK) { for (int64_t t = 0; t < n_hits; ++t) swb_alignment_free(&alns[t]); return rc; }
            continue;
        }
        e->stats.cells += (double)J.la * (double)J.lb;
        jobs.push_back(J);
    }
    if (jobs.empty()) return SWB_OK;
    // the query on the device (encoded on the host: a few kB)
    if (lq + 32 > db->query_cap) {
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        cudaFree(db->d_query); db->d_query = nullptr; db->query_cap = 0; db->h_query.clear();
        CUDA_TRY(e, cudaMalloc(&db->d_query, (size_t)lq + 64)); db->query_cap = lq + 64;
    }
    std::vector<uint8_t> cq((size_t)lq + 32, (uint8_t)e->n);
    for (int64_t t = 0; t < lq; ++t) {
        const int c = e->code_of[(unsigned char)query[t]];
        if (c < 0) { e->err = std::string("unknown letter ") + query[t] + ". please set X to allow scoring for undefined letter pair."; return SWB_ERR_LETTER; }
        cq[t] = (uint8_t)c;
    }
    if (db->h_query != cq) {                                    // the same query as in the last call: its codes are on the device already
        CUDA_TRY(e, cudaMemcpyAsync(db->d_query, cq.data(), cq.size(), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        db->h_query = cq;
    }
    int rc = align_jobs_device(e, db->d_query, db->d_codes, jobs, retain_all, alns, nullptr, nullptr, nullptr, nullptr);
    if (rc == SWB_OK)
        for (const AlnJob &J : jobs)
            if (e->mode == SWB_LOCAL && (alns[J.out].score != hits[J.out].score || alns[J.out].end_a != hits[J.out].end_q || alns[J.out].end_b != hits[J.out].end_t)) {
                e->err = "scan and traceback disagree"; rc = SWB_ERR_CUDA; break;
            }
    if (rc != SWB_OK) for (int64_t t = 0; t < n_hits; ++t) swb_alignment_free(&alns[t]);
    return rc;
}

// One process, several GPUs: engines[g] / dbs[g] hold shard g (swb_db_upload_shard(.., g, n_shards, ..)).  One host thread
// per shard scans and selects, the lists are merged on the host, every shard traces its own hits.  alns may be NULL.
extern "C" int swb_search_sharded(swb_engine **engines, swb_db **dbs, int n_shards, const char *query, int64_t lq, int topk,
                                  unsigned flags, swb_hit *hits, int64_t *n_hits, swb_alignment *alns)
{
    if (!engines || !dbs || n_shards < 1 || !n_hits || topk < 0 || topk > SWB_MAX_TOPK || (topk > 0 && !hits)) return SWB_ERR_ARG;
    std::vector<swb_hit> parts((size_t)n_shards * (size_t)std::max(topk, 1));
    std::vector<int64_t> cnt((size_t)n_shards, 0);
    std::vector<int> rcs((size_t)n_shards, SWB_OK);
    auto each = [&](auto fn) {
        std::vector<std::thread> th;
        for (int g = 1; g < n_shards; ++g) th.emplace_back(fn, g);
        fn(0);
        for (auto &t : th) t.join();
        for (int g = 0; g < n_shards; ++g) if (rcs[g] != SWB_OK) return rcs[g];
        return (int)SWB_OK;
    };
    int rc = each([&](int g) { rcs[g] = swb_search(engines[g], query, lq, dbs[g], topk, parts.data() + (size_t)g * std::max(topk, 1), &cnt[g]); });
    if (rc != SWB_OK) return rc;
    std::vector<swb_hit> flat;
    for (int g = 0; g < n_shards; ++g) {
        for (int64_t k = 0; k < cnt[g]; ++k) { swb_hit h = parts[(size_t)g * std::max(topk, 1) + k]; h.shard = g; flat.push_back(h); }
    }
    rc = swb_search_merge(flat.data(), (int64_t)flat.size(), topk, hits, n_hits);
    if (rc != SWB_OK || !alns || *n_hits == 0) return rc;
    const int64_t nh = *n_hits;
    memset(alns, 0, sizeof(swb_alignment) * (size_t)nh);
    std::vector<std::vector<swb_alignment>> tmp((size_t)n_shards, std::vector<swb_alignment>((size_t)nh));
    std::vector<std::vector<int32_t>> own((size_t)n_shards, std::vector<int32_t>((size_t)nh, 0));
    rc = each([&](int g) { rcs[g] = swb_search_align(engines[g], query, lq, dbs[g], hits, nh, flags, tmp[g].data(), own[g].data()); });
    for (int g = 0; g < n_shards; ++g)
        for (int64_t k = 0; k < nh; ++k) {
            if (rcs[g] == SWB_OK && own[g][k] && hits[k].shard == g && !alns[k].aln_a) alns[k] = tmp[g][k];
            else swb_alignment_free(&tmp[g][k]);
        }
    if (rc != SWB_OK) for (int64_t k = 0; k < nh; ++k) swb_alignment_free(&alns[k]);
    return rc;
}
