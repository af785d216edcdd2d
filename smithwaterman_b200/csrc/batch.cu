// Batch-list path: score + end cell of n independent pairs (the `sa_opencl -list` loop of the
// reference, OCL-main.rs:295-316, with align_partial(.., score_only) semantics SA.rs:438-467).
// Host side: shape sort + work-item table.  Device side: residue encoding, packing into the
// K1 layout, K1 (packed s16x2) and KW (int32) launches.  No CPU compute path exists: the host only
// handles pairs with an empty sequence, whose result is a closed-form boundary value (SA.rs:100-118).
#include <algorithm>
#include <cstring>
#include <numeric>

#include "swb_internal.h"

#define K1_MAX_LEN 4096            // longer pairs go to the wavefront kernel

struct BatchDev {                  // device-only members that do not fit the plain struct in the header
    uint8_t *codes_a = nullptr, *codes_b = nullptr;   // encoded residues, caller layout
    int64_t *off_a = nullptr, *off_b = nullptr;
    int32_t *order = nullptr;                          // sorted position -> caller index (K1 pairs)
    int32_t *err = nullptr;
    KWPair *kw_pairs = nullptr; KWResult *kw_res = nullptr; int4 *kw_bnd = nullptr;
    int n_kw = 0;
    std::vector<KWPair> h_kw; std::vector<int32_t> kw_idx;   // caller index of each KW pair
    std::vector<int32_t> degenerate;                         // caller indices with la==0 or lb==0
    int n_k1 = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};
static BatchDev *dev_of(swb_batch *b) { return b->dev; }

// ---- device helpers ---------------------------------------------------------------------------
__global__ void encode_kernel(uint8_t *buf, int64_t nbytes, const int16_t *__restrict__ code_of, int32_t *err)
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
                const int c = tab[(w[q] >> (8 * s)) & 0xFFu];
                if (c < 0 && base + q * 4 + s < nbytes) bad = true;
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
                               int32_t *__restrict__ lb_arr, int pad)
{
    const K1Group G = groups[blockIdx.x];
    const int lane = threadIdx.x & 31, sub = threadIdx.x >> 5, nsub = blockDim.x >> 5;
    const bool valid = lane < G.n_pairs;
    int64_t oa = 0, ob = 0; int la = 0, lb = 0;
    if (valid) {
        const int src = order[G.first_pair + lane];
        oa = off_a[src]; la = (int)(off_a[src + 1] - oa);
        ob = off_b[src]; lb = (int)(off_b[src + 1] - ob);
        if (sub == 0) { la_arr[G.first_pair + lane] = la; lb_arr[G.first_pair + lane] = lb; }
    }
    const int n_blk = (G.la_max + 1 + 3) / 4 + 1;              // 4 rows per word, one word of slack (all PAD)
    for (int c = sub; c < n_blk; c += nsub) {
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

static void batch_release(swb_engine *e, swb_batch *b)
{
    if (!b) return;
    cudaSetDevice(e->device);
    BatchDev *d = dev_of(b);
    if (d) {
        cudaFree(d->codes_a); cudaFree(d->codes_b); cudaFree(d->off_a); cudaFree(d->off_b); cudaFree(d->order);
        cudaFree(d->err); cudaFree(d->kw_pairs); cudaFree(d->kw_res); cudaFree(d->kw_bnd);
        if (d->ev0) cudaEventDestroy(d->ev0);
        if (d->ev1) cudaEventDestroy(d->ev1);
        delete d;
        b->dev = nullptr;
    }
    cudaFree(b->d_a); cudaFree(b->d_b); cudaFree(b->d_groups); cudaFree(b->d_la); cudaFree(b->d_lb);
    cudaFree(b->d_res); cudaFree(b->d_counter);
    delete b;
}

extern "C" void swb_batch_free(swb_engine *e, swb_batch *b) { if (e) batch_release(e, b); }

extern "C" int swb_batch_upload(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                                const int64_t *off_b, int64_t n, swb_batch **out)
{
    if (!e) return SWB_ERR_ARG;
    if (!out || n < 0 || (n > 0 && (!off_a || !off_b))) { e->err = "bad argument"; return SWB_ERR_ARG; }
    if (n > 0x7FFFFFF0) { e->err = "too many pairs"; return SWB_ERR_ARG; }
    *out = nullptr;
    CUDA_TRY(e, cudaSetDevice(e->device));
    swb_batch *b = new swb_batch();
    BatchDev *d = new BatchDev();
    b->dev = d;
    b->n = n;
    const int64_t na = n ? off_a[n] : 0, nb = n ? off_b[n] : 0;
    if ((na > 0 && !a_all) || (nb > 0 && !b_all)) { e->err = "null sequence buffer"; batch_release(e, b); return SWB_ERR_ARG; }
    const bool use_k1 = (e->mode == SWB_LOCAL && e->o <= e->e);

    // ---- shape sort (counting sort on (passes, la), longest first) and routing
    b->la.resize(n); b->lb.resize(n);
    std::vector<int32_t> k1_idx;
    k1_idx.reserve(n);
    int la_cap = 0, pass_cap = 0;
    b->cells = 0;
    for (int64_t k = 0; k < n; ++k) {
        const int64_t la = off_a[k + 1] - off_a[k], lb = off_b[k + 1] - off_b[k];
        if (la < 0 || lb < 0 || la > 0x3FFFFFFF || lb > 0x3FFFFFFF) { e->err = "bad offsets"; batch_release(e, b); return SWB_ERR_ARG; }
        b->la[k] = (int32_t)la; b->lb[k] = (int32_t)lb;
        b->cells += (double)la * (double)lb;
        if (la == 0 || lb == 0) d->degenerate.push_back((int32_t)k);
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
        // work items
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
            G.la_max = lam; G.n_pass = npm; G.a_off = a_off; G.b_off = b_off;
            a_off += (int64_t)((lam + 1 + 3) / 4 + 1) * 32 * 4;
            b_off += (int64_t)npm * 32 * 32;
        }
        b->a_bytes = (size_t)a_off; b->b_bytes = (size_t)b_off; b->la_max = la_cap;
    }

    // ---- device buffers + H2D
    const int64_t na_p = pad16(na) + 16, nb_p = pad16(nb) + 16;
    CUDA_TRY(e, cudaMalloc(&d->codes_a, na_p)); CUDA_TRY(e, cudaMalloc(&d->codes_b, nb_p));
    CUDA_TRY(e, cudaMalloc(&d->off_a, (n + 1) * sizeof(int64_t))); CUDA_TRY(e, cudaMalloc(&d->off_b, (n + 1) * sizeof(int64_t)));
    CUDA_TRY(e, cudaMalloc(&d->err, sizeof(int32_t)));
    CUDA_TRY(e, cudaMemsetAsync(d->err, 0, sizeof(int32_t), e->stream));
    CUDA_TRY(e, cudaMemsetAsync(d->codes_a + pad16(na) - (na ? 16 : 0), 0, na ? 32 : 16, e->stream));
    CUDA_TRY(e, cudaMemsetAsync(d->codes_b + pad16(nb) - (nb ? 16 : 0), 0, nb ? 32 : 16, e->stream));
    if (na) CUDA_TRY(e, cudaMemcpyAsync(d->codes_a, a_all, na, cudaMemcpyHostToDevice, e->stream));
    if (nb) CUDA_TRY(e, cudaMemcpyAsync(d->codes_b, b_all, nb, cudaMemcpyHostToDevice, e->stream));
    if (n) {
        CUDA_TRY(e, cudaMemcpyAsync(d->off_a, off_a, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(d->off_b, off_b, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, e->stream));
    }
    e->stats.h2d_bytes = (double)na + (double)nb + 2.0 * (n + 1) * sizeof(int64_t);
    int16_t *d_tab = nullptr;
    CUDA_TRY(e, cudaMalloc(&d_tab, 256 * sizeof(int16_t)));
    CUDA_TRY(e, cudaMemcpyAsync(d_tab, e->code_of, 256 * sizeof(int16_t), cudaMemcpyHostToDevice, e->stream));
    if (na) encode_kernel<<<std::min<int64_t>((na_p / 16 + 255) / 256, e->sm_count * 8), 256, 0, e->stream>>>(d->codes_a, na, d_tab, d->err);
    if (nb) encode_kernel<<<std::min<int64_t>((nb_p / 16 + 255) / 256, e->sm_count * 8), 256, 0, e->stream>>>(d->codes_b, nb, d_tab, d->err);
    CUDA_TRY(e, cudaGetLastError());

    if (d->n_k1) {
        const int ng = (int)b->groups.size();
        CUDA_TRY(e, cudaMalloc(&b->d_a, b->a_bytes + 32 * 16));   // one chunk of slack: K1 prefetches a chunk ahead
        CUDA_TRY(e, cudaMalloc(&b->d_b, b->b_bytes));
        CUDA_TRY(e, cudaMalloc(&b->d_groups, ng * sizeof(K1Group)));
        CUDA_TRY(e, cudaMalloc(&b->d_la, d->n_k1 * sizeof(int32_t))); CUDA_TRY(e, cudaMalloc(&b->d_lb, d->n_k1 * sizeof(int32_t)));
        CUDA_TRY(e, cudaMalloc(&b->d_res, d->n_k1 * sizeof(K1Result)));
        CUDA_TRY(e, cudaMalloc(&b->d_counter, sizeof(int32_t)));
        CUDA_TRY(e, cudaMalloc(&d->order, d->n_k1 * sizeof(int32_t)));
        CUDA_TRY(e, cudaMemcpyAsync(b->d_groups, b->groups.data(), ng * sizeof(K1Group), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(d->order, b->order.data(), d->n_k1 * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
        e->stats.h2d_bytes += (double)ng * sizeof(K1Group) + (double)d->n_k1 * sizeof(int32_t);
        k1_pack_kernel<<<ng, 128, 0, e->stream>>>(d->codes_a, d->codes_b, d->off_a, d->off_b, d->order, b->d_groups,
                                                  b->d_a, b->d_b, b->d_la, b->d_lb, e->n);
        CUDA_TRY(e, cudaGetLastError());
        const size_t need = k1_bnd_bytes(e->sm_count, b->la_max);
        if (need > e->bnd_bytes) {
            CUDA_TRY(e, cudaStreamSynchronize(e->stream));
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
        int64_t bnd_off = 0;
        for (int t = 0; t < d->n_kw; ++t) {
            const int32_t k = d->kw_idx[t];
            KWPair &P = d->h_kw[t];
            P.a_off = off_a[k]; P.b_off = off_b[k]; P.la = b->la[k]; P.lb = b->lb[k]; P.dir_off = -1; P.bnd_off = bnd_off;
            bnd_off += kw_bnd_elems(P.la, P.lb);
        }
        CUDA_TRY(e, cudaMalloc(&d->kw_pairs, d->n_kw * sizeof(KWPair)));
        CUDA_TRY(e, cudaMalloc(&d->kw_res, d->n_kw * sizeof(KWResult)));
        CUDA_TRY(e, cudaMalloc(&d->kw_bnd, bnd_off * sizeof(int4)));
        CUDA_TRY(e, cudaMemcpyAsync(d->kw_pairs, d->h_kw.data(), d->n_kw * sizeof(KWPair), cudaMemcpyHostToDevice, e->stream));
        e->stats.h2d_bytes += (double)d->n_kw * sizeof(KWPair);
    }
    CUDA_TRY(e, cudaEventCreate(&d->ev0)); CUDA_TRY(e, cudaEventCreate(&d->ev1));
    int32_t h_err = 0;
    CUDA_TRY(e, cudaMemcpyAsync(&h_err, d->err, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    cudaFree(d_tab);
    if (h_err) {
        e->err = "unknown letter. please set X to allow scoring for undefined letter pair.";   // SA.rs:676
        batch_release(e, b);
        return SWB_ERR_LETTER;
    }
    e->stats.cells = b->cells;
    *out = b;
    return SWB_OK;
}

extern "C" int swb_batch_run(swb_engine *e, swb_batch *b)
{
    if (!e || !b || !b->dev) return SWB_ERR_ARG;
    BatchDev *d = b->dev;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int64_t launches = 0;
    CUDA_TRY(e, cudaEventRecord(d->ev0, e->stream));
    if (d->n_k1) { CUDA_TRY(e, k1_launch(e, b, e->sm_count, e->stream)); ++launches; }
    if (d->n_kw) {
        CUDA_TRY(e, kw_launch(e, d->codes_a, d->codes_b, d->kw_pairs, d->n_kw, nullptr, d->kw_bnd, d->kw_res, e->stream));
        ++launches;
    }
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
    BatchDev *d = b->dev;
    CUDA_TRY(e, cudaSetDevice(e->device));
    int64_t n_launch = 0;
    CUDA_TRY(e, cudaEventRecord(e->ev0, e->stream));
    for (int s = 0; s < steps; ++s) {
        if (d->n_k1) { CUDA_TRY(e, k1_launch(e, b, e->sm_count, e->stream)); ++n_launch; }
        if (d->n_kw) {
            CUDA_TRY(e, kw_launch(e, d->codes_a, d->codes_b, d->kw_pairs, d->n_kw, nullptr, d->kw_bnd, d->kw_res, e->stream));
            ++n_launch;
        }
    }
    CUDA_TRY(e, cudaEventRecord(e->ev1, e->stream));
    CUDA_TRY(e, cudaEventSynchronize(e->ev1));
    CUDA_TRY(e, cudaEventElapsedTime(total_ms, e->ev0, e->ev1));
    if (launches) *launches = n_launch;
    e->stats.kernel_ms = *total_ms; e->stats.launches = n_launch;
    return SWB_OK;
}

extern "C" int swb_batch_fetch(swb_engine *e, swb_batch *b, float *scores, int64_t *end_a, int64_t *end_b)
{
    if (!e || !b || !b->dev) return SWB_ERR_ARG;
    if (b->n > 0 && (!scores || !end_a || !end_b)) { e->err = "null output buffer"; return SWB_ERR_ARG; }
    BatchDev *d = b->dev;
    CUDA_TRY(e, cudaSetDevice(e->device));
    const float inv = 1.0f / (float)e->scale;       // exact for dyadic scales
    const bool dyadic = (e->scale & (e->scale - 1)) == 0;
    auto to_f = [&](int32_t s) { return dyadic ? (float)s * inv : (float)((double)s / e->scale); };
    e->stats.d2h_bytes = 0; e->stats.i32_reruns = 0;
    std::vector<KWResult> kwres(d->n_kw);
    b->h_res.resize(d->n_k1);
    if (d->n_k1) CUDA_TRY(e, cudaMemcpyAsync(b->h_res.data(), b->d_res, d->n_k1 * sizeof(K1Result), cudaMemcpyDeviceToHost, e->stream));
    if (d->n_kw) CUDA_TRY(e, cudaMemcpyAsync(kwres.data(), d->kw_res, d->n_kw * sizeof(KWResult), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    e->stats.d2h_bytes += (double)d->n_k1 * sizeof(K1Result) + (double)d->n_kw * sizeof(KWResult);

    std::vector<int32_t> redo;                       // caller indices whose s16 lanes overflowed
    for (int pos = 0; pos < d->n_k1; ++pos) {
        const K1Result &r = b->h_res[pos];
        const int32_t k = b->order[pos];
        if (r.flags & 1) { redo.push_back(k); continue; }
        scores[k] = to_f(r.score); end_a[k] = r.end_i; end_b[k] = r.end_j;
    }
    for (int t = 0; t < d->n_kw; ++t) {
        const int32_t k = d->kw_idx[t];
        scores[k] = to_f(kwres[t].score); end_a[k] = kwres[t].end_i; end_b[k] = kwres[t].end_j;
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
        CUDA_TRY(e, cudaMemcpy(oa.data(), d->off_a, (b->n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost));
        CUDA_TRY(e, cudaMemcpy(ob.data(), d->off_b, (b->n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost));
        int64_t bnd_off = 0;
        for (int t = 0; t < nr; ++t) {
            const int32_t k = redo[t];
            hp[t].a_off = oa[k]; hp[t].b_off = ob[k]; hp[t].la = b->la[k]; hp[t].lb = b->lb[k];
            hp[t].dir_off = -1; hp[t].bnd_off = bnd_off;
            bnd_off += kw_bnd_elems(hp[t].la, hp[t].lb);
        }
        KWPair *dp = nullptr; KWResult *dr = nullptr; int4 *db = nullptr;
        CUDA_TRY(e, cudaMalloc(&dp, nr * sizeof(KWPair))); CUDA_TRY(e, cudaMalloc(&dr, nr * sizeof(KWResult)));
        CUDA_TRY(e, cudaMalloc(&db, bnd_off * sizeof(int4)));
        CUDA_TRY(e, cudaMemcpyAsync(dp, hp.data(), nr * sizeof(KWPair), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, kw_launch(e, d->codes_a, d->codes_b, dp, nr, nullptr, db, dr, e->stream));
        std::vector<KWResult> rr(nr);
        CUDA_TRY(e, cudaMemcpyAsync(rr.data(), dr, nr * sizeof(KWResult), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        cudaFree(dp); cudaFree(dr); cudaFree(db);
        for (int t = 0; t < nr; ++t) {
            const int32_t k = redo[t];
            scores[k] = to_f(rr[t].score); end_a[k] = rr[t].end_i; end_b[k] = rr[t].end_j;
        }
        e->stats.i32_reruns = nr; e->stats.launches += 1;
    }
    return SWB_OK;
}

extern "C" int swb_score_batch(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                               const int64_t *off_b, int64_t n, float *scores, int64_t *end_a, int64_t *end_b)
{
    swb_batch *b = nullptr;
    int rc = swb_batch_upload(e, a_all, off_a, b_all, off_b, n, &b);
    if (rc != SWB_OK) return rc;
    rc = swb_batch_run(e, b);
    if (rc == SWB_OK) rc = swb_batch_fetch(e, b, scores, end_a, end_b);
    if (rc == SWB_OK) { float ms; swb_batch_last_run_ms(e, b, &ms, nullptr); }
    swb_batch_free(e, b);
    return rc;
}
