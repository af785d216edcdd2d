// Batch-list path: score + end cell of n independent pairs (the `sa_opencl -list` loop of the
// reference, OCL-main.rs:295-316, with align_partial(.., score_only) semantics SA.rs:438-467).
// Host side: shape sort + work-item table.  Device side: residue encoding, packing into the
// K1 layout, K1 (packed s16x2) and KW (int32) launches.  No CPU compute path exists: the host only
// handles pairs with an empty sequence, whose result is a closed-form boundary value (SA.rs:100-118).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>

#include "swb_internal.h"

#define K1_MAX_LEN 8192            // longer pairs go to the wavefront kernel
#define K2_MIN_QUERY 1024           // shared queries at least this long use K2 (its 124-step skew is < 12 % there)

struct BatchDev {                  // device-only members that do not fit the plain struct in the header
    uint8_t *codes_a = nullptr, *codes_b = nullptr;   // encoded residues, caller layout
    int64_t *off_a = nullptr, *off_b = nullptr;
    int32_t *order = nullptr;                          // sorted position -> caller index (K1 pairs)
    int32_t *err = nullptr;
    KWPair *kw_pairs = nullptr; KWResult *kw_res = nullptr; int4 *kw_bnd = nullptr; int *kw_scr = nullptr; KWPlan kw_plan_;
    int n_kw = 0;
    std::vector<KWPair> h_kw; std::vector<int32_t> kw_idx;   // caller index of each KW pair
    std::vector<int32_t> degenerate;                         // caller indices with la==0 or lb==0
    int n_k1 = 0;
    int n_k2 = 0, n_items = 0;                               // K2 (shared long query): pairs taken, warp work items
    int32_t *items = nullptr, *sweeps = nullptr;             // device: [n_items][32] caller index per lane (-1 idle), sweeps per item
    int64_t base_a = 0, base_b = 0;                          // caller offset of the first residue held on the device
    int64_t shared_la = -1;                                  // >= 0: all pairs share one `a` (database search)
    K1Result *h_res = nullptr; KWResult *h_kwres = nullptr; int32_t *h_err = nullptr;   // pinned staging (engine arena or own)
    bool own_stage = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_done = nullptr, ev_ready = nullptr;
};
static BatchDev *dev_of(swb_batch *b) { return b->dev; }

// ---- device helpers ---------------------------------------------------------------------------
__global__ void encode_kernel(uint8_t *buf, int64_t nbytes, const int16_t *__restrict__ code_of, int32_t *err, int pad)
{
    __shared__ int16_t tab[256];
    for (int k = threadIdx.x; k < 256; k += blockDim.x) tab[k] = code_of[k];
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 16;
    for (int64_t base = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16; base < nbytes; base += stride) {
        uint4 v = *reinterpret_cast<const uint4 *>(buf + base);           // buffers are padded to 16 bytes
        unsigned w[4] = {v.x, v.y, v.z, v.w};
        bool bad = false;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            unsigned r = 0;
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                int c = tab[(w[q] >> (8 * s)) & 0xFFu];
                if (c < 0) { bad = bad || (base + q * 4 + s < nbytes); c = pad; }   // flagged; PAD keeps later kernels in range
                r |= ((unsigned)(c & 0xFF)) << (8 * s);
            }
            w[q] = r;
        }
        if (bad) atomicExch(err, 1);
        *reinterpret_cast<uint4 *>(buf + base) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// One block per K1 work item: gathers the 32 pairs' residues into the [4-row block][lane] row layout
// and the [pass][lane][32] column layout, PAD-filled, and writes the per-pair lengths.
__global__ void k1_pack_kernel(const uint8_t *__restrict__ codes_a, const uint8_t *__restrict__ codes_b,
                               const int64_t *__restrict__ off_a, const int64_t *__restrict__ off_b,
                               const int32_t *__restrict__ order, const K1Group *__restrict__ groups,
                               uint8_t *__restrict__ a_pk, uint8_t *__restrict__ b_pk, int32_t *__restrict__ la_arr,
                               int32_t *__restrict__ lb_arr, int pad, int64_t base_a, int64_t base_b, int shared_la)
{
    const K1Group G = groups[blockIdx.x];
    const int lane = threadIdx.x & 31, sub = threadIdx.x >> 5, nsub = blockDim.x >> 5;
    const bool valid = lane < G.n_pairs;
    int64_t oa = 0, ob = 0; int la = 0, lb = 0;
    if (valid) {
        const int src = order[G.first_pair + lane];
        if (shared_la < 0) { oa = off_a[src] - base_a; la = (int)(off_a[src + 1] - off_a[src]); }
        ob = off_b[src] - base_b; lb = (int)(off_b[src + 1] - off_b[src]);
    }
    if (shared_la >= 0) { oa = 0; la = shared_la; }                // database search: every lane reads the same `a` (the query)
    if (valid && sub == 0) { la_arr[G.first_pair + lane] = la; lb_arr[G.first_pair + lane] = lb; }
    const int n_blk = (G.la_max + 1 + 3) / 4 + 1;              // 4 rows per word, one word of slack (all PAD)
    for (int c = sub; c < n_blk && (shared_la < 0 || blockIdx.x == 0); c += nsub) {   // shared query: one row block for all items
        unsigned r = 0;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const int row = c * 4 + s;
            const unsigned code = (row < la) ? codes_a[oa + row] : (unsigned)pad;
            r |= code << (8 * s);
        }
        reinterpret_cast<unsigned *>(a_pk + G.a_off)[c * 32 + lane] = r;
    }
    for (int h = sub; h < G.n_pass * 2; h += nsub) {       // 16-byte halves of the 32-column passes
        unsigned w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            unsigned r = 0;
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int col = h * 16 + q * 4 + s;
                const unsigned code = (col < lb) ? codes_b[ob + col] : (unsigned)pad;
                r |= code << (8 * s);
            }
            w[q] = r;
        }
        reinterpret_cast<uint4 *>(b_pk + G.b_off)[((h >> 1) * 32 + lane) * 2 + (h & 1)] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

static int64_t pad16(int64_t x) { return (x + 15) / 16 * 16; }

// Device memory comes from the stream-ordered pool of the engine's device (cudaMallocAsync): after the first call the
// allocations of a batch cost microseconds and never synchronise the device.
#define DEV_ALLOC_ON(e_, st, ptr, bytes) CUDA_TRY(e_, cudaMallocAsync((void **)&(ptr), (size_t)(bytes) ? (size_t)(bytes) : 16, st))
#define DEV_ALLOC(e_, ptr, bytes) DEV_ALLOC_ON(e_, (e_)->stream, ptr, bytes)
static void dev_free(swb_engine *e, void *p) { if (p) cudaFreeAsync(p, e->stream); }

// Pinned host staging for results: a grow-only arena owned by the engine (reset by swb_score_batch), or a private
// allocation for stand-alone staged batches.
static void *stage_alloc(swb_engine *e, size_t bytes)
{
    bytes = (bytes + 255) / 256 * 256;
    if (e->stage_used + bytes > e->stage_bytes) return nullptr;
    void *p = (char *)e->h_stage + e->stage_used;
    e->stage_used += bytes;
    return p;
}

static void batch_release(swb_engine *e, swb_batch *b)
{
    if (!b) return;
    cudaSetDevice(e->device);
    BatchDev *d = dev_of(b);
    if (d) {
        dev_free(e, d->codes_a); dev_free(e, d->codes_b); dev_free(e, d->off_a); dev_free(e, d->off_b); dev_free(e, d->order);
        dev_free(e, d->err); dev_free(e, d->items); dev_free(e, d->sweeps); dev_free(e, d->kw_pairs); dev_free(e, d->kw_res); dev_free(e, d->kw_bnd); dev_free(e, d->kw_scr);
        if (d->own_stage) { cudaFreeHost(d->h_res); cudaFreeHost(d->h_kwres); cudaFreeHost(d->h_err); }
        if (d->ev0) cudaEventDestroy(d->ev0);
        if (d->ev1) cudaEventDestroy(d->ev1);
        if (d->ev_done) cudaEventDestroy(d->ev_done);
        if (d->ev_ready) cudaEventDestroy(d->ev_ready);
        delete d;
        b->dev = nullptr;
    }
    dev_free(e, b->d_a); dev_free(e, b->d_b); dev_free(e, b->d_groups); dev_free(e, b->d_la); dev_free(e, b->d_lb);
    dev_free(e, b->d_res); dev_free(e, b->d_counter);
    delete b;
}

extern "C" void swb_batch_free(swb_engine *e, swb_batch *b) { if (e) batch_release(e, b); }

// Upload = shape sort + work-item table on the host, then (all asynchronous on the engine's stream) H2D of the raw
// residues and offsets, in-place encoding, packing into the K1 layout.  `off_*` are the caller's absolute offsets;
// residues are taken from a_all[off_a[0] .. off_a[n]).  use_arena: results staging comes from the engine's arena.
static int batch_upload_impl(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                             const int64_t *off_b, int64_t n, int64_t shared_la, bool use_arena, cudaStream_t up, swb_batch **out)
{
    *out = nullptr;
    swb_batch *b = new swb_batch();
    BatchDev *d = new BatchDev();
    b->dev = d;
    b->n = n;
    struct Guard { swb_engine *e; swb_batch *b; bool ok = false; ~Guard() { if (!ok) batch_release(e, b); } } guard{e, b};
    const bool shared = shared_la >= 0;
    d->shared_la = shared_la;
    d->base_a = (n && !shared) ? off_a[0] : 0; d->base_b = n ? off_b[0] : 0;
    const int64_t na = shared ? shared_la : (n ? off_a[n] - d->base_a : 0), nb = n ? off_b[n] - d->base_b : 0;
    if (na < 0 || nb < 0 || (na > 0 && !a_all) || (nb > 0 && !b_all)) { e->err = "bad sequence buffer / offsets"; return SWB_ERR_ARG; }
    const bool use_k1 = (e->mode == SWB_LOCAL && e->o <= e->e);

    // ---- device buffers + H2D of the raw residues first (stream-ordered, no host synchronisation): the copy engine works
    //      while the host sorts
    const int64_t na_p = pad16(na) + 16, nb_p = pad16(nb) + 16;
    DEV_ALLOC_ON(e, up, d->codes_a, na_p); DEV_ALLOC_ON(e, up, d->codes_b, nb_p);
    if (!shared) DEV_ALLOC_ON(e, up, d->off_a, (n + 1) * sizeof(int64_t));
    DEV_ALLOC_ON(e, up, d->off_b, (n + 1) * sizeof(int64_t));
    DEV_ALLOC_ON(e, up, d->err, sizeof(int32_t));
    CUDA_TRY(e, cudaMemsetAsync(d->err, 0, sizeof(int32_t), up));
    CUDA_TRY(e, cudaMemsetAsync(d->codes_a + pad16(na) - (na ? 16 : 0), 0, na ? 32 : 16, up));
    CUDA_TRY(e, cudaMemsetAsync(d->codes_b + pad16(nb) - (nb ? 16 : 0), 0, nb ? 32 : 16, up));
    if (na) CUDA_TRY(e, cudaMemcpyAsync(d->codes_a, a_all + d->base_a, na, cudaMemcpyHostToDevice, up));
    if (nb) CUDA_TRY(e, cudaMemcpyAsync(d->codes_b, b_all + d->base_b, nb, cudaMemcpyHostToDevice, up));
    if (n) {
        if (!shared) CUDA_TRY(e, cudaMemcpyAsync(d->off_a, off_a, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, up));
        CUDA_TRY(e, cudaMemcpyAsync(d->off_b, off_b, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, up));
    }
    e->stats.h2d_bytes += (double)na + (double)nb + 2.0 * (n + 1) * sizeof(int64_t);
    if (na) encode_kernel<<<(int)std::min<int64_t>((na_p / 16 + 255) / 256, e->sm_count * 8), 256, 0, up>>>(d->codes_a, na, e->d_code_tab, d->err, e->n);
    if (nb) encode_kernel<<<(int)std::min<int64_t>((nb_p / 16 + 255) / 256, e->sm_count * 8), 256, 0, up>>>(d->codes_b, nb, e->d_code_tab, d->err, e->n);
    CUDA_TRY(e, cudaGetLastError());

    // ---- shape sort (counting sort on (passes, la), longest first) and routing
    b->la.resize(n); b->lb.resize(n);
    // a long shared query (database search) goes to K2: lanes split the columns, every work item lasts la + 124 row steps
    const bool use_k2 = use_k1 && shared && shared_la >= K2_MIN_QUERY && shared_la <= 0x0FFFFFFF;
    std::vector<int32_t> k1_idx, k2_idx;
    k1_idx.reserve(n);
    int la_cap = 0, pass_cap = 0;
    b->cells = 0;
    for (int64_t k = 0; k < n; ++k) {
        const int64_t la = shared ? shared_la : off_a[k + 1] - off_a[k], lb = off_b[k + 1] - off_b[k];
        if (la < 0 || lb < 0 || la > 0x3FFFFFFF || lb > 0x3FFFFFFF) { e->err = "bad offsets"; return SWB_ERR_ARG; }
        b->la[k] = (int32_t)la; b->lb[k] = (int32_t)lb;
        b->cells += (double)la * (double)lb;
        if (la == 0 || lb == 0) d->degenerate.push_back((int32_t)k);
        else if (use_k2) k2_idx.push_back((int32_t)k);
        else if (use_k1 && la <= K1_MAX_LEN && lb <= K1_MAX_LEN) {
            k1_idx.push_back((int32_t)k);
            la_cap = std::max(la_cap, (int)la);
            pass_cap = std::max(pass_cap, (int)((lb + K1_STRIP - 1) / K1_STRIP));
        } else d->kw_idx.push_back((int32_t)k);
    }
    d->n_k1 = (int)k1_idx.size();
    b->order.resize(d->n_k1);
    if (d->n_k1) {
        const size_t nb_ = (size_t)(pass_cap + 1) * (la_cap + 1);
        std::vector<int32_t> cnt(nb_ + 1, 0);
        auto bucket = [&](int32_t k) {
            const int np = (b->lb[k] + K1_STRIP - 1) / K1_STRIP;
            return (size_t)(pass_cap - np) * (la_cap + 1) + (size_t)(la_cap - b->la[k]);   // descending
        };
        for (int32_t k : k1_idx) cnt[bucket(k) + 1]++;
        for (size_t q = 0; q < nb_; ++q) cnt[q + 1] += cnt[q];
        for (int32_t k : k1_idx) b->order[cnt[bucket(k)]++] = k;
        const int ng = (d->n_k1 + 31) / 32;
        b->groups.resize(ng);
        int64_t a_off = 0, b_off = 0;
        for (int g = 0; g < ng; ++g) {
            K1Group &G = b->groups[g];
            G.first_pair = g * 32; G.n_pairs = std::min(32, d->n_k1 - g * 32);
            int lam = 0, npm = 0;
            for (int t = 0; t < G.n_pairs; ++t) {
                const int32_t k = b->order[G.first_pair + t];
                lam = std::max(lam, b->la[k]);
                npm = std::max(npm, (b->lb[k] + K1_STRIP - 1) / K1_STRIP);
            }
            G.la_max = lam; G.n_pass = npm; G.a_off = shared ? 0 : a_off; G.b_off = b_off;
            if (!shared || g == 0) a_off += (int64_t)((lam + 1 + 3) / 4 + 1) * 32 * 4;
            b_off += (int64_t)npm * 32 * 32;
        }
        b->a_bytes = (size_t)a_off; b->b_bytes = (size_t)b_off; b->la_max = la_cap;
    }

    if (d->n_k1) {
        const int ng = (int)b->groups.size();
        DEV_ALLOC_ON(e, up, b->d_a, b->a_bytes + 32 * 16);        // one block of slack: K1 prefetches a row block ahead
        DEV_ALLOC_ON(e, up, b->d_b, b->b_bytes);
        DEV_ALLOC_ON(e, up, b->d_groups, ng * sizeof(K1Group));
        DEV_ALLOC_ON(e, up, b->d_la, d->n_k1 * sizeof(int32_t)); DEV_ALLOC_ON(e, up, b->d_lb, d->n_k1 * sizeof(int32_t));
        DEV_ALLOC_ON(e, up, b->d_res, n * sizeof(K1Result));
        CUDA_TRY(e, cudaMemsetAsync(b->d_res, 0, n * sizeof(K1Result), up));      // pairs K1 does not take keep flags == 0
        DEV_ALLOC_ON(e, up, b->d_counter, sizeof(int32_t));
        DEV_ALLOC_ON(e, up, d->order, d->n_k1 * sizeof(int32_t));
        b->d_order = d->order;
        // the work-item table and the order go through pinned staging when the arena has room: an asynchronous copy from
        // pageable memory makes the host wait for the residue copies queued before it
        const void *src_g = b->groups.data(), *src_o = b->order.data();
        if (use_arena) {
            void *pg = stage_alloc(e, ng * sizeof(K1Group)), *po = stage_alloc(e, d->n_k1 * sizeof(int32_t));
            if (pg && po) {
                memcpy(pg, src_g, ng * sizeof(K1Group)); memcpy(po, src_o, d->n_k1 * sizeof(int32_t));
                src_g = pg; src_o = po;
            }
        }
        CUDA_TRY(e, cudaMemcpyAsync(b->d_groups, src_g, ng * sizeof(K1Group), cudaMemcpyHostToDevice, up));
        CUDA_TRY(e, cudaMemcpyAsync(d->order, src_o, d->n_k1 * sizeof(int32_t), cudaMemcpyHostToDevice, up));
        e->stats.h2d_bytes += (double)ng * sizeof(K1Group) + (double)d->n_k1 * sizeof(int32_t);
        k1_pack_kernel<<<ng, 128, 0, up>>>(d->codes_a, d->codes_b, d->off_a, d->off_b, d->order, b->d_groups,
                                                  b->d_a, b->d_b, b->d_la, b->d_lb, e->n, d->base_a, d->base_b, (int)shared_la);
        CUDA_TRY(e, cudaGetLastError());
        const size_t need = k1_bnd_bytes(e->sm_count, b->la_max);
        if (need > e->bnd_bytes) {                          // engine-owned scratch, grown on demand (rare)
            CUDA_TRY(e, cudaDeviceSynchronize());
            if (e->d_bnd) cudaFree(e->d_bnd);
            e->d_bnd = nullptr; e->bnd_bytes = 0;
            CUDA_TRY(e, cudaMalloc(&e->d_bnd, need));
            e->bnd_bytes = need;
            // keep the pass-boundary scratch resident in L2: it is rewritten and re-read every pass by the resident warps
            int max_persist = 0, max_win = 0;
            cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, e->device);
            cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, e->device);
            if (max_persist > 0 && max_win > 0) {
                cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)max_persist);
                cudaStreamAttrValue av = {};
                av.accessPolicyWindow.base_ptr = e->d_bnd;
                av.accessPolicyWindow.num_bytes = std::min<size_t>(need, (size_t)std::min(max_win, max_persist));   // the low row blocks
                av.accessPolicyWindow.hitRatio = 1.0f;
                av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                cudaStreamSetAttribute(e->stream, cudaStreamAttributeAccessPolicyWindow, &av);
                if (getenv("SWB_TRACE")) fprintf(stderr, "[swb] L2 persisting window: %zu bytes, max persist %d, hit ratio %.2f\n", av.accessPolicyWindow.num_bytes, max_persist, av.accessPolicyWindow.hitRatio);
                cudaGetLastError();
            }
        }
    }
    // ---- K2 work items: subjects of >= 32 passes take a warp each (several sweeps), the others are bin-packed
    //      (first fit, longest first) into warps of 32 lanes, one lane per 32-column pass
    d->n_k2 = (int)k2_idx.size();
    if (d->n_k2) {
        std::vector<std::vector<int32_t>> bucket(32);
        std::vector<int32_t> longs;
        for (int32_t k : k2_idx) {
            const int np = (b->lb[k] + K1_STRIP - 1) / K1_STRIP;
            if (np >= 32) longs.push_back(k); else bucket[np].push_back(k);
        }
        std::sort(longs.begin(), longs.end(), [&](int32_t x, int32_t y) { return b->lb[x] > b->lb[y]; });
        size_t small_passes = 0; unsigned mask = 0;
        for (int p = 1; p < 32; ++p) { small_passes += bucket[p].size() * (size_t)p; if (!bucket[p].empty()) mask |= 1u << p; }
        const size_t cap_items = longs.size() + small_passes / 32 + small_passes / 256 + 64;    // first fit wastes < 12 %
        int32_t *h_items = nullptr, *h_sweeps = nullptr;
        const size_t items_bytes = (cap_items * 33) * sizeof(int32_t);
        if (items_bytes > e->items_bytes) {
            CUDA_TRY(e, cudaStreamSynchronize(up));
            if (e->h_items) cudaFreeHost(e->h_items);
            e->h_items = nullptr; e->items_bytes = 0;
            CUDA_TRY(e, cudaMallocHost(&e->h_items, items_bytes));
            e->items_bytes = items_bytes;
        }
        else if (e->ev_items) CUDA_TRY(e, cudaEventSynchronize(e->ev_items));   // an earlier chunk's copy out of this buffer
        h_items = (int32_t *)e->h_items; h_sweeps = h_items + cap_items * 32;
        size_t ni = 0;
        for (int32_t k : longs) {
            for (int l = 0; l < 32; ++l) h_items[ni * 32 + l] = k;
            h_sweeps[ni++] = ((b->lb[k] + K1_STRIP - 1) / K1_STRIP + 31) / 32;
        }
        while (mask) {
            int cap = 32, l = 0;
            if (ni >= cap_items) { e->err = "K2 item table overflow"; return SWB_ERR_NOMEM; }
            while (cap > 0) {
                const unsigned fit = mask & ((cap >= 31) ? 0xFFFFFFFFu : ((2u << cap) - 1u));
                if (!fit) break;
                const int p = 31 - __builtin_clz(fit);
                const int32_t k = bucket[p].back(); bucket[p].pop_back();
                if (bucket[p].empty()) mask &= ~(1u << p);
                for (int t = 0; t < p; ++t) h_items[ni * 32 + l++] = k;
                cap -= p;
            }
            for (; l < 32; ++l) h_items[ni * 32 + l] = -1;
            h_sweeps[ni++] = 1;
        }
        d->n_items = (int)ni;
        DEV_ALLOC_ON(e, up, d->items, ni * 32 * sizeof(int32_t)); DEV_ALLOC_ON(e, up, d->sweeps, ni * sizeof(int32_t));
        DEV_ALLOC_ON(e, up, b->d_res, n * sizeof(K1Result));
        CUDA_TRY(e, cudaMemsetAsync(b->d_res, 0, n * sizeof(K1Result), up));
        DEV_ALLOC_ON(e, up, b->d_counter, sizeof(int32_t));
        CUDA_TRY(e, cudaMemcpyAsync(d->items, h_items, ni * 32 * sizeof(int32_t), cudaMemcpyHostToDevice, up));
        CUDA_TRY(e, cudaMemcpyAsync(d->sweeps, h_sweeps, ni * sizeof(int32_t), cudaMemcpyHostToDevice, up));
        e->stats.h2d_bytes += (double)ni * 33 * sizeof(int32_t);
        if (!e->ev_items) CUDA_TRY(e, cudaEventCreateWithFlags(&e->ev_items, cudaEventDisableTiming));
        CUDA_TRY(e, cudaEventRecord(e->ev_items, up));
        const size_t need = k2_bnd_bytes(e->sm_count, (int)shared_la);
        if (need > e->bnd_bytes) {
            CUDA_TRY(e, cudaDeviceSynchronize());
            if (e->d_bnd) cudaFree(e->d_bnd);
            e->d_bnd = nullptr; e->bnd_bytes = 0;
            CUDA_TRY(e, cudaMalloc(&e->d_bnd, need));
            e->bnd_bytes = need;
        }
    }
    // ---- pairs for the int32 wavefront kernel
    d->n_kw = (int)d->kw_idx.size();
    if (d->n_kw) {
        d->h_kw.resize(d->n_kw);
        for (int t = 0; t < d->n_kw; ++t) {
            const int32_t k = d->kw_idx[t];
            KWPair &P = d->h_kw[t];
            P.a_off = shared ? 0 : off_a[k] - d->base_a; P.b_off = off_b[k] - d->base_b; P.la = b->la[k]; P.lb = b->lb[k];
            P.dir_off = -1; P.pssm_off = -1;
        }
        kw_plan(e, d->h_kw.data(), d->n_kw, false, &d->kw_plan_);
        DEV_ALLOC_ON(e, up, d->kw_pairs, d->n_kw * sizeof(KWPair));
        DEV_ALLOC_ON(e, up, d->kw_res, d->n_kw * sizeof(KWResult));
        DEV_ALLOC_ON(e, up, d->kw_bnd, d->kw_plan_.bnd_elems * sizeof(int4));
        DEV_ALLOC_ON(e, up, d->kw_scr, d->kw_plan_.scratch_ints * sizeof(int));
        CUDA_TRY(e, cudaMemcpyAsync(d->kw_pairs, d->h_kw.data(), d->n_kw * sizeof(KWPair), cudaMemcpyHostToDevice, up));
        e->stats.h2d_bytes += (double)d->n_kw * sizeof(KWPair);
    }
    CUDA_TRY(e, cudaEventCreate(&d->ev0)); CUDA_TRY(e, cudaEventCreate(&d->ev1));
    CUDA_TRY(e, cudaEventCreateWithFlags(&d->ev_done, cudaEventDisableTiming));
    // ---- pinned staging for the results
    if (use_arena) {
        d->h_res = (K1Result *)stage_alloc(e, (size_t)std::max<int64_t>(1, n) * sizeof(K1Result));
        d->h_kwres = (KWResult *)stage_alloc(e, (size_t)std::max(1, d->n_kw) * sizeof(KWResult));
        d->h_err = (int32_t *)stage_alloc(e, sizeof(int32_t));
    }
    if (!d->h_res || !d->h_kwres || !d->h_err) {
        d->h_res = nullptr; d->h_kwres = nullptr; d->h_err = nullptr;
        d->own_stage = true;
        CUDA_TRY(e, cudaMallocHost((void **)&d->h_res, (size_t)std::max<int64_t>(1, n) * sizeof(K1Result)));
        CUDA_TRY(e, cudaMallocHost((void **)&d->h_kwres, (size_t)std::max(1, d->n_kw) * sizeof(KWResult)));
        CUDA_TRY(e, cudaMallocHost((void **)&d->h_err, sizeof(int32_t)));
    }
    *d->h_err = 0;
    CUDA_TRY(e, cudaMemcpyAsync(d->h_err, d->err, sizeof(int32_t), cudaMemcpyDeviceToHost, up));
    CUDA_TRY(e, cudaEventCreateWithFlags(&d->ev_ready, cudaEventDisableTiming));
    CUDA_TRY(e, cudaEventRecord(d->ev_ready, up));
    e->stats.cells += b->cells;
    guard.ok = true;
    *out = b;
    return SWB_OK;
}

extern "C" int swb_batch_upload(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                                const int64_t *off_b, int64_t n, swb_batch **out)
{
    if (!e) return SWB_ERR_ARG;
    if (!out || n < 0 || (n > 0 && (!off_a || !off_b))) { e->err = "bad argument"; return SWB_ERR_ARG; }
    if (n > 0x7FFFFFF0) { e->err = "too many pairs"; return SWB_ERR_ARG; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    e->stats.cells = 0; e->stats.h2d_bytes = 0; e->stats.d2h_bytes = 0; e->stats.i32_reruns = 0;
    swb_batch *b = nullptr;
    int rc = batch_upload_impl(e, a_all, off_a, b_all, off_b, n, -1, false, e->stream, &b);
    if (rc != SWB_OK) return rc;
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));           // the staged API reports letter errors here
    if (*b->dev->h_err) {
        e->err = "unknown letter. please set X to allow scoring for undefined letter pair.";   // SA.rs:676
        batch_release(e, b);
        return SWB_ERR_LETTER;
    }
    *out = b;
    return SWB_OK;
}

static int batch_launch(swb_engine *e, swb_batch *b, int64_t *launches)
{
    BatchDev *d = b->dev;
    CUDA_TRY(e, cudaStreamWaitEvent(e->stream, d->ev_ready, 0));     // upload may have run on the copy stream
    if (d->n_k1) { CUDA_TRY(e, k1_launch(e, b, e->sm_count, e->stream)); ++*launches; }
    if (d->n_k2) {
        CUDA_TRY(e, k2_launch(e, d->codes_a, (int)d->shared_la, d->codes_b, d->off_b, d->base_b, d->items, d->sweeps, d->n_items, e->d_bnd,
                              b->d_counter, b->d_res, e->sm_count, e->stream));
        ++*launches;
    }
    if (d->n_kw) {
        CUDA_TRY(e, kw_launch(e, d->codes_a, d->codes_b, d->kw_pairs, d->n_kw, d->kw_plan_, nullptr, d->kw_bnd, d->kw_scr, d->kw_res, e->stream));
        ++*launches;
    }
    return SWB_OK;
}

extern "C" int swb_batch_run(swb_engine *e, swb_batch *b)
{
    if (!e || !b || !b->dev) return SWB_ERR_ARG;
    BatchDev *d = b->dev;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int64_t launches = 0;
    CUDA_TRY(e, cudaEventRecord(d->ev0, e->stream));
    int rc = batch_launch(e, b, &launches);
    if (rc != SWB_OK) return rc;
    CUDA_TRY(e, cudaEventRecord(d->ev1, e->stream));
    b->last_launches = launches;
    e->stats.launches = launches;
    return SWB_OK;
}

extern "C" int swb_batch_last_run_ms(swb_engine *e, swb_batch *b, float *ms, int64_t *launches)
{
    if (!e || !b || !b->dev || !ms) return SWB_ERR_ARG;
    CUDA_TRY(e, cudaSetDevice(e->device));
    CUDA_TRY(e, cudaEventSynchronize(b->dev->ev1));
    CUDA_TRY(e, cudaEventElapsedTime(ms, b->dev->ev0, b->dev->ev1));
    b->last_ms = *ms;
    e->stats.kernel_ms = *ms;
    if (launches) *launches = b->last_launches;
    return SWB_OK;
}

extern "C" int swb_batch_run_n(swb_engine *e, swb_batch *b, int steps, float *total_ms, int64_t *launches)
{
    if (!e || !b || !b->dev || !total_ms || steps < 1) return SWB_ERR_ARG;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int64_t n_launch = 0;
    CUDA_TRY(e, cudaEventRecord(e->ev0, e->stream));
    for (int s = 0; s < steps; ++s) {
        int rc = batch_launch(e, b, &n_launch);
        if (rc != SWB_OK) return rc;
    }
    CUDA_TRY(e, cudaEventRecord(e->ev1, e->stream));
    CUDA_TRY(e, cudaEventSynchronize(e->ev1));
    CUDA_TRY(e, cudaEventElapsedTime(total_ms, e->ev0, e->ev1));
    if (launches) *launches = n_launch;
    e->stats.kernel_ms = *total_ms; e->stats.launches = n_launch;
    return SWB_OK;
}

// D2H of the results into pinned staging, asynchronous; ev_done fires when they have landed.
static int batch_fetch_enqueue(swb_engine *e, swb_batch *b)
{
    BatchDev *d = b->dev;
    if (d->n_k1 || d->n_k2) CUDA_TRY(e, cudaMemcpyAsync(d->h_res, b->d_res, b->n * sizeof(K1Result), cudaMemcpyDeviceToHost, e->stream));
    if (d->n_kw) CUDA_TRY(e, cudaMemcpyAsync(d->h_kwres, d->kw_res, d->n_kw * sizeof(KWResult), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaEventRecord(d->ev_done, e->stream));
    e->stats.d2h_bytes += ((d->n_k1 || d->n_k2) ? (double)b->n * sizeof(K1Result) : 0.0) + (double)d->n_kw * sizeof(KWResult);
    return SWB_OK;
}

// Waits for ev_done, scatters the results to the caller's order (outputs are indexed relative to the batch's first
// pair), handles empty sequences and re-runs s16 overflows in int32.
static int batch_fetch_finish(swb_engine *e, swb_batch *b, float *scores, int64_t *end_a, int64_t *end_b)
{
    BatchDev *d = b->dev;
    CUDA_TRY(e, cudaEventSynchronize(d->ev_done));
    if (*d->h_err) {
        e->err = "unknown letter. please set X to allow scoring for undefined letter pair.";   // SA.rs:676
        return SWB_ERR_LETTER;
    }
    const float inv = 1.0f / (float)e->scale;       // exact for dyadic scales
    const bool dyadic = (e->scale & (e->scale - 1)) == 0;
    auto to_f = [&](int32_t s) { return dyadic ? (float)s * inv : (float)((double)s / e->scale); };
    std::vector<int32_t> redo;                       // pairs whose s16 lanes overflowed
    if (d->n_k1 || d->n_k2) {
        const K1Result *hr = d->h_res;                 // caller order (K1 scatters); non-K1 entries are zero and overwritten below
        for (int64_t k = 0; k < b->n; ++k) {
            const K1Result r = hr[k];
            if (r.flags & 1) { redo.push_back((int32_t)k); continue; }
            scores[k] = to_f(r.score); end_a[k] = r.end_i; end_b[k] = r.end_j;
        }
    }
    for (int t = 0; t < d->n_kw; ++t) {
        const int32_t k = d->kw_idx[t];
        scores[k] = to_f(d->h_kwres[t].score); end_a[k] = d->h_kwres[t].end_i; end_b[k] = d->h_kwres[t].end_j;
    }
    // pairs with an empty sequence: closed-form boundary cells (SA.rs:88-118, 438-467)
    for (int32_t k : d->degenerate) {
        const int64_t la = b->la[k], lb = b->lb[k];
        if (e->mode == SWB_LOCAL) { scores[k] = 0.0f; end_a[k] = -1; end_b[k] = -1; continue; }
        const int32_t so = (e->mode == SWB_GLOBAL) ? e->o : 0, se = (e->mode == SWB_GLOBAL) ? e->e : 0;
        const int64_t len = la + lb;
        const int64_t ls = len ? len * se + (so - se) : 0;           // (0,0) holds max(0,-1,-1) = 0
        scores[k] = to_f((int32_t)ls); end_a[k] = la; end_b[k] = lb;
    }
    if (!redo.empty()) {                              // int32 re-run of the overflowed pairs (KW, no direction bytes)
        const int nr = (int)redo.size();
        std::vector<KWPair> hp(nr);
        std::vector<int64_t> oa(b->n + 1), ob(b->n + 1);
        if (d->shared_la < 0) CUDA_TRY(e, cudaMemcpyAsync(oa.data(), d->off_a, (b->n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(ob.data(), d->off_b, (b->n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        for (int t = 0; t < nr; ++t) {
            const int32_t k = redo[t];
            hp[t].a_off = d->shared_la >= 0 ? 0 : oa[k] - d->base_a; hp[t].b_off = ob[k] - d->base_b; hp[t].la = b->la[k]; hp[t].lb = b->lb[k];
            hp[t].dir_off = -1; hp[t].pssm_off = -1;
        }
        KWPlan pl; kw_plan(e, hp.data(), nr, false, &pl);
        KWPair *dp = nullptr; KWResult *dr = nullptr; int4 *db = nullptr; int *ds = nullptr;
        DEV_ALLOC(e, dp, nr * sizeof(KWPair)); DEV_ALLOC(e, dr, nr * sizeof(KWResult)); DEV_ALLOC(e, db, pl.bnd_elems * sizeof(int4));
        DEV_ALLOC(e, ds, pl.scratch_ints * sizeof(int));
        CUDA_TRY(e, cudaMemcpyAsync(dp, hp.data(), nr * sizeof(KWPair), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, kw_launch(e, d->codes_a, d->codes_b, dp, nr, pl, nullptr, db, ds, dr, e->stream));
        std::vector<KWResult> rr(nr);
        CUDA_TRY(e, cudaMemcpyAsync(rr.data(), dr, nr * sizeof(KWResult), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        dev_free(e, dp); dev_free(e, dr); dev_free(e, db); dev_free(e, ds);
        for (int t = 0; t < nr; ++t) {
            const int32_t k = redo[t];
            scores[k] = to_f(rr[t].score); end_a[k] = rr[t].end_i; end_b[k] = rr[t].end_j;
        }
        e->stats.i32_reruns += nr; e->stats.launches += 1;
    }
    return SWB_OK;
}

extern "C" int swb_batch_fetch(swb_engine *e, swb_batch *b, float *scores, int64_t *end_a, int64_t *end_b)
{
    if (!e || !b || !b->dev) return SWB_ERR_ARG;
    if (b->n > 0 && (!scores || !end_a || !end_b)) { e->err = "null output buffer"; return SWB_ERR_ARG; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    e->stats.d2h_bytes = 0; e->stats.i32_reruns = 0;
    int rc = batch_fetch_enqueue(e, b);
    if (rc != SWB_OK) return rc;
    return batch_fetch_finish(e, b, scores, end_a, end_b);
}

// Host buffers in, host buffers out.  The pair list is cut into chunks that flow through a two-stream software
// pipeline: the copy stream uploads, encodes and packs chunk c+1 while the compute stream aligns chunk c and the host
// sorts chunk c+2 / scatters the results of chunk c-1 (with pinned caller memory the copies are truly asynchronous).
#define SCORE_CHUNK_PAIRS 393216
#define SCORE_CHUNK_FIRST 131072

static int score_pipeline(swb_engine *e, const char *a_all, const int64_t *off_a, int64_t shared_la, const char *b_all,
                          const int64_t *off_b, int64_t n, float *scores, int64_t *end_a, int64_t *end_b)
{
    if (n > 0x7FFFFFF0) { e->err = "too many pairs"; return SWB_ERR_ARG; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    e->stats.cells = 0; e->stats.h2d_bytes = 0; e->stats.d2h_bytes = 0; e->stats.i32_reruns = 0; e->stats.launches = 0;
    // results arena: 16 B per pair + slack per chunk
    // chunking pays only while the copy is a visible fraction of the work: a long shared query (database search) makes the
    // upload negligible next to the DP, and every extra launch adds the tail of its own longest work item
    int64_t chunk_pairs = SCORE_CHUNK_PAIRS;
    if (n > 0) {
        const double nb_tot = (double)(off_b[n] - off_b[0]), na_tot = shared_la >= 0 ? (double)shared_la : (double)(off_a[n] - off_a[0]);
        const double est_cells = shared_la >= 0 ? na_tot * nb_tot : (na_tot / n) * (nb_tot / n) * n;
        const double copy_ms = (na_tot + nb_tot) / 50e6, dp_ms = est_cells / 4e9;
        if (copy_ms < 0.05 * dp_ms) chunk_pairs = n;
    }
    // graded chunks: a small first chunk gets K1 going after ~1 ms of copy; each following chunk is as much larger as the
    // copy engine is faster than K1 (x1.8 on configs[1]), so its upload hides behind the previous chunk's DP
    std::vector<int64_t> cuts(1, 0);
    for (int64_t sz = (chunk_pairs == n) ? n : SCORE_CHUNK_FIRST; cuts.back() < n; sz = std::min<int64_t>(chunk_pairs, sz * 9 / 5)) {
        int64_t next = std::min<int64_t>(n, cuts.back() + sz);
        if (n - next < SCORE_CHUNK_FIRST) next = n;          // no crumbs at the end
        cuts.push_back(next);
    }
    if (cuts.size() == 1) cuts.push_back(0);
    const int64_t n_chunks = (int64_t)cuts.size() - 1;
    const size_t need = (size_t)n * (sizeof(K1Result) + sizeof(KWResult) + sizeof(int32_t) + 2) + (size_t)n_chunks * 8192;
    if (need > e->stage_bytes) {
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        if (e->h_stage) cudaFreeHost(e->h_stage);
        e->h_stage = nullptr; e->stage_bytes = 0;
        CUDA_TRY(e, cudaMallocHost(&e->h_stage, need));
        e->stage_bytes = need;
    }
    e->stage_used = 0;
    int rc = SWB_OK;
    std::vector<swb_batch *> live((size_t)n_chunks, nullptr);   // chunks in flight; finished in order as their results land
    int64_t next_finish = 0;
    const bool trace = getenv("SWB_TRACE") != nullptr;       // host-side phase times of the pipeline, to stderr
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_up = 0, t_launch = 0, t_fin = 0, t_rel = 0; const double t_begin = now();
    CUDA_TRY(e, cudaEventRecord(e->ev0, e->stream));
    std::vector<cudaEvent_t> tev;                             // trace only: per chunk {upload begin, packed, K1 done}
    auto mark = [&](cudaStream_t st) { if (trace) { cudaEvent_t ev; cudaEventCreate(&ev); cudaEventRecord(ev, st); tev.push_back(ev); } };
    if (trace) CUDA_TRY(e, cudaStreamWaitEvent(e->stream_up, e->ev0, 0));
    for (int64_t c = 0; c < n_chunks && rc == SWB_OK; ++c) {
        const int64_t start = cuts[c], cnt = cuts[c + 1] - cuts[c];
        swb_batch *cur = nullptr;
        double t0 = now();
        mark(e->stream_up);
        rc = batch_upload_impl(e, a_all, shared_la >= 0 ? nullptr : off_a + start, b_all, off_b + start, cnt, shared_la, true,
                               e->stream_up, &cur);
        double t1 = now(); t_up += t1 - t0;
        mark(e->stream_up);
        int64_t launches = 0;
        if (rc == SWB_OK) rc = batch_launch(e, cur, &launches);
        mark(e->stream);
        if (rc == SWB_OK) rc = batch_fetch_enqueue(e, cur);
        e->stats.launches += launches;
        double t2 = now(); t_launch += t2 - t1;
        live[c] = cur;
        // scatter the chunks whose results have already landed; never block here: the next chunk's sort and upload come first
        while (rc == SWB_OK && next_finish < c && cudaEventQuery(live[next_finish]->dev->ev_done) == cudaSuccess) {
            double t3 = now();
            rc = batch_fetch_finish(e, live[next_finish], scores + cuts[next_finish], end_a + cuts[next_finish], end_b + cuts[next_finish]);
            double t4 = now(); t_fin += t4 - t3;
            batch_release(e, live[next_finish]); live[next_finish] = nullptr;
            t_rel += now() - t4;
            ++next_finish;
        }
    }
    for (int64_t c = next_finish; c < n_chunks; ++c) {
        if (!live[c]) continue;
        double t3 = now();
        if (rc == SWB_OK) rc = batch_fetch_finish(e, live[c], scores + cuts[c], end_a + cuts[c], end_b + cuts[c]);
        t_fin += now() - t3;
        batch_release(e, live[c]); live[c] = nullptr;
    }
    if (rc == SWB_OK) {
        CUDA_TRY(e, cudaEventRecord(e->ev1, e->stream));
        CUDA_TRY(e, cudaEventSynchronize(e->ev1));
        float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1);
        e->stats.kernel_ms = ms;                      // device time of the whole pipelined call
        if (trace) fprintf(stderr, "[swb] score_batch n=%lld chunks=%lld total %.2f ms: upload(host) %.2f launch %.2f finish(wait+scatter) %.2f release %.2f device-span %.2f\n",
                           (long long)n, (long long)n_chunks, now() - t_begin, t_up, t_launch, t_fin, t_rel, ms);
        for (size_t k = 0; k + 2 < tev.size() + 0 && trace; k += 3) {
            float a = 0, b2 = 0, c2 = 0;
            cudaEventElapsedTime(&a, e->ev0, tev[k]); cudaEventElapsedTime(&b2, e->ev0, tev[k + 1]); cudaEventElapsedTime(&c2, e->ev0, tev[k + 2]);
            fprintf(stderr, "[swb]   chunk %zu: upload begins %.2f  packed %.2f  K1 done %.2f ms\n", k / 3, a, b2, c2);
        }
        for (cudaEvent_t ev : tev) cudaEventDestroy(ev);
    }
    return rc;
}

extern "C" int swb_score_batch(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                               const int64_t *off_b, int64_t n, float *scores, int64_t *end_a, int64_t *end_b)
{
    if (!e) return SWB_ERR_ARG;
    if (n < 0 || (n > 0 && (!off_a || !off_b || !scores || !end_a || !end_b))) { e->err = "bad argument"; return SWB_ERR_ARG; }
    return score_pipeline(e, a_all, off_a, -1, b_all, off_b, n, scores, end_a, end_b);
}

// Database search, scoring stage: one query against n database sequences (score_only semantics per pair,
// SA.rs:438-467).  The query is uploaded and packed once and shared by all K1 work items.
extern "C" int swb_search_scores(swb_engine *e, const char *query, int64_t lq, const char *db_all, const int64_t *off_db,
                                 int64_t n, float *scores, int64_t *end_q, int64_t *end_t)
{
    if (!e) return SWB_ERR_ARG;
    if (lq < 0 || n < 0 || (lq > 0 && !query) || (n > 0 && (!off_db || !scores || !end_q || !end_t))) { e->err = "bad argument"; return SWB_ERR_ARG; }
    if (lq > 0x3FFFFFFF) { e->err = "query too long"; return SWB_ERR_ARG; }
    return score_pipeline(e, query, nullptr, lq, db_all, off_db, n, scores, end_q, end_t);
}
