// Greedy length-sorted clustering, the `-cluster` mode of sa_opencl (OCL-main.rs:149-243), driven from the host with
// the sequences resident on the device: per representative ONE batch (KW fill with direction bytes + the counting form
// of the KT walk) against all candidates that are still unclustered; only 48-byte pair descriptors go up and 32-byte
// counts come down.  The decisions are the reference's (f64 ratios of the counts read off align(rep, cand, false)).
#include <algorithm>
#include <cstring>
#include <numeric>
#include <string_view>

#include "swb_internal.h"

#define CL_DIR_BUDGET_BYTES (24ll << 30)

namespace {
struct Pool {                               // stream-ordered scratch, freed on scope exit
    void *p = nullptr; cudaStream_t st = nullptr;
    cudaError_t alloc(size_t bytes, cudaStream_t s) { st = s; return cudaMallocAsync(&p, bytes ? bytes : 16, s); }
    ~Pool() { if (p) cudaFreeAsync(p, st); }
    template <class T> T *as() { return reinterpret_cast<T *>(p); }
};
struct Pinned {
    void *p = nullptr;
    cudaError_t alloc(size_t bytes) { return cudaMallocHost(&p, bytes ? bytes : 16); }
    ~Pinned() { if (p) cudaFreeHost(p); }
    template <class T> T *as() { return reinterpret_cast<T *>(p); }
};
}  // namespace

// KW (direction bytes) + counting KT for `np` pairs whose residues are on the device; chunks by the direction-byte budget.
static int run_counts(swb_engine *e, const uint8_t *d_codes, const uint8_t *d_raw, KWPair *hp, int np, KWResult *h_res,
                      KTStats *h_st)
{
    int pos = 0;
    while (pos < np) {
        int end = pos; int64_t dir_total = 0;
        while (end < np) {
            const int64_t db = kw_dir_bytes(hp[end].la, hp[end].lb, 8);
            if (end > pos && dir_total + db > CL_DIR_BUDGET_BYTES) break;
            dir_total += db; ++end;
        }
        const int nc = end - pos;
        KWPlan plan; kw_plan(e, hp + pos, nc, true, &plan);
        if ((size_t)plan.dir_bytes + 16 > e->dir_bytes) {
            CUDA_TRY(e, cudaStreamSynchronize(e->stream));
            if (e->d_dir) cudaFree(e->d_dir);
            e->d_dir = nullptr; e->dir_bytes = 0;
            CUDA_TRY(e, cudaMalloc(&e->d_dir, (size_t)plan.dir_bytes + 16));
            e->dir_bytes = (size_t)plan.dir_bytes + 16;
        }
        Pool d_pairs, d_res, d_bnd, d_scr, d_st;
        CUDA_TRY(e, d_pairs.alloc(nc * sizeof(KWPair), e->stream)); CUDA_TRY(e, d_res.alloc(nc * sizeof(KWResult), e->stream));
        CUDA_TRY(e, d_bnd.alloc(plan.bnd_elems * sizeof(int4), e->stream)); CUDA_TRY(e, d_scr.alloc(plan.scratch_ints * sizeof(int), e->stream));
        CUDA_TRY(e, d_st.alloc(nc * sizeof(KTStats), e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(d_pairs.p, hp + pos, nc * sizeof(KWPair), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, kw_launch(e, d_codes, d_codes, d_pairs.as<KWPair>(), nc, plan, (uint8_t *)e->d_dir, d_bnd.as<int4>(), d_scr.as<int>(),
                              d_res.as<KWResult>(), e->stream));
        CUDA_TRY(e, kt_stats_launch(e, d_pairs.as<KWPair>(), d_res.as<KWResult>(), nc, (const uint8_t *)e->d_dir, plan.cols, d_raw, d_raw,
                                    d_st.as<KTStats>(), e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(h_res + pos, d_res.p, nc * sizeof(KWResult), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(h_st + pos, d_st.p, nc * sizeof(KTStats), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        e->stats.launches += 2;
        e->stats.h2d_bytes += (double)nc * sizeof(KWPair); e->stats.d2h_bytes += (double)nc * (sizeof(KWResult) + sizeof(KTStats));
        pos = end;
    }
    return SWB_OK;
}

extern "C" int swb_cluster(swb_engine *e, const char *seq_all, const int64_t *off, int64_t n, double identity,
                           double coverage_short, double coverage_long, int64_t *order, int64_t *cluster_of,
                           int64_t *identical, int64_t *n_alignments)
{
    if (!e) return SWB_ERR_ARG;
    if (n < 0 || (n > 0 && (!off || !order || !cluster_of || !identical)) || n > 0x7FFFFFF0) { e->err = "bad argument"; return SWB_ERR_ARG; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    e->stats = swb_stats{}; e->stats.scale = e->scale; e->stats.exact_scale = e->exact_scale;
    if (n_alignments) *n_alignments = 0;
    if (n == 0) return SWB_OK;
    // ---- longest first, stable (OCL-main.rs:151)
    std::vector<int64_t> len(n);
    for (int64_t k = 0; k < n; ++k) {
        len[k] = off[k + 1] - off[k];
        if (len[k] < 0 || len[k] > 0x3FFFFFF0) { e->err = "bad offsets"; return SWB_ERR_ARG; }
    }
    std::iota(order, order + n, (int64_t)0);
    std::stable_sort(order, order + n, [&](int64_t x, int64_t y) { return len[x] > len[y]; });
    if (!swb_range_ok(e, len[order[0]], n > 1 ? len[order[1]] : len[order[0]])) {     // the two longest: the largest pair the loop can form
        e->err = "scores of a pair this long do not fit the int32 lanes"; return SWB_ERR_RANGE;
    }
    // ---- the sequences in sorted order: letters as given + encoded, resident on the device for the whole run
    std::vector<int64_t> so(n + 1, 0);
    for (int64_t k = 0; k < n; ++k) so[k + 1] = so[k] + len[order[k]];
    const int64_t total = so[n];
    if (total > 0 && !seq_all) { e->err = "bad sequence buffer"; return SWB_ERR_ARG; }
    std::vector<char> raw((size_t)total + 16, 0);
    std::vector<uint8_t> codes((size_t)total + 16, 0);
    for (int64_t k = 0; k < n; ++k) memcpy(raw.data() + so[k], seq_all + off[order[k]], (size_t)len[order[k]]);
    for (int64_t k = 0; k < total; ++k) {
        const int c = e->code_of[(unsigned char)raw[k]];
        if (c < 0) { e->err = std::string("unknown letter ") + raw[k] + ". please set X to allow scoring for undefined letter pair."; return SWB_ERR_LETTER; }
        codes[k] = (uint8_t)c;
    }
    Pool d_codes, d_raw;
    CUDA_TRY(e, d_codes.alloc(codes.size(), e->stream)); CUDA_TRY(e, d_raw.alloc(raw.size(), e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(d_codes.p, codes.data(), codes.size(), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(d_raw.p, raw.data(), raw.size(), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    e->stats.h2d_bytes += 2.0 * (double)raw.size();
    auto sv = [&](int64_t k) { return std::string_view(raw.data() + so[k], (size_t)(so[k + 1] - so[k])); };
    auto L = [&](int64_t k) { return so[k + 1] - so[k]; };

    // ---- exact duplicates (OCL-main.rs:160-180)
    for (int64_t k = 0; k < n; ++k) { cluster_of[k] = k; identical[k] = k; }
    for (int64_t ii = 0; ii < n; ++ii) {
        if (identical[ii] != ii) continue;
        for (int64_t jj = ii + 1; jj < n; ++jj) {
            if (identical[jj] != jj) continue;
            if (L(ii) != L(jj)) break;
            if (sv(ii) == sv(jj)) identical[jj] = ii;
        }
    }
    // ---- representatives in order; one device batch per representative (OCL-main.rs:181-243)
    Pinned h_pairs, h_res, h_st;
    CUDA_TRY(e, h_pairs.alloc((size_t)n * sizeof(KWPair))); CUDA_TRY(e, h_res.alloc((size_t)n * sizeof(KWResult)));
    CUDA_TRY(e, h_st.alloc((size_t)n * sizeof(KTStats)));
    KWPair *hp = h_pairs.as<KWPair>(); KWResult *hr = h_res.as<KWResult>(); KTStats *hs = h_st.as<KTStats>();
    std::vector<int64_t> need;
    int64_t n_aln = 0;
    cudaEvent_t ev0 = e->ev0, ev1 = e->ev1;
    CUDA_TRY(e, cudaEventRecord(ev0, e->stream));
    for (int64_t ii = 0; ii < n; ++ii) {
        if (cluster_of[ii] != ii || identical[ii] != ii) continue;
        const std::string_view si = sv(ii);
        need.clear();
        for (int64_t jj = ii + 1; jj < n; ++jj) {
            if (cluster_of[jj] != jj || identical[jj] != jj) continue;
            const std::string_view sj = sv(jj);
            if (si.find(sj) != std::string_view::npos) {                                  // si.contains(&sj)
                const double lcov = (double)sj.size() / (double)si.size();               // 0/0 -> NaN -> false
                if (lcov >= coverage_long) cluster_of[jj] = ii;
                continue;
            }
            KWPair &P = hp[need.size()];
            P.a_off = so[ii]; P.b_off = so[jj]; P.la = (int32_t)L(ii); P.lb = (int32_t)L(jj);
            P.dir_off = 0; P.bnd_off = 0; P.scr_off = 0; P.pssm_off = -1;
            need.push_back(jj);
        }
        if (need.empty()) continue;
        const int np = (int)need.size();
        int rc = run_counts(e, d_codes.as<uint8_t>(), d_raw.as<uint8_t>(), hp, np, hr, hs);
        if (rc != SWB_OK) return rc;
        n_aln += np;
        for (int t = 0; t < np; ++t) {
            const int64_t jj = need[t];
            e->stats.cells += (double)L(ii) * (double)L(jj);
            if (hs[t].ncols < 0) { e->err = "traceback left the matrix (the reference panics here, SA.rs:360-371)"; return SWB_ERR_EMPTY; }
            swb_aln_stats st; memset(&st, 0, sizeof(st));
            swb_compose_stats(&st, e->mode, false, L(ii), L(jj), hs[t]);
            const double lcov = (double)st.a_letters / (double)L(ii), scov = (double)st.b_letters / (double)L(jj);
            const double ident = (double)st.matches / (double)st.aln_len;               // 0/0 -> NaN -> false, as in the reference
            if (lcov >= coverage_long && scov >= coverage_short && ident >= identity) cluster_of[jj] = ii;
        }
    }
    CUDA_TRY(e, cudaEventRecord(ev1, e->stream));
    CUDA_TRY(e, cudaEventSynchronize(ev1));
    float ms = 0; cudaEventElapsedTime(&ms, ev0, ev1);
    e->stats.kernel_ms = ms;                          // device span of the whole loop (host decisions included)
    if (n_alignments) *n_alignments = n_aln;
    return SWB_OK;
}
