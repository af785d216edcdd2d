// Host layer of the engine behind include/swb200.h: scoring model, integer rescaling, engine
// life-cycle.  Mirrors SequenceAlignment::new (SA.rs:26-45), SubstitutionMatrix (SA.rs:626-733)
// and seq_to_index / get_string_index (SA.rs:649-679) of the reference.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "swb_internal.h"

static thread_local std::string g_create_err;      // per thread: distinct engines may be created concurrently

// BLOSUM62 as published by NCBI (the table the reference embeds at SA.rs:706-730), letter order
// A R N D C Q E G H I L K M F P S T W Y V B Z X *.
static const char BL62_LETTERS[25] = "ARNDCQEGHILKMFPSTWYVBZX*";
static const signed char BL62[24][24] = {
    {4, -1, -2, -2, 0, -1, -1, 0, -2, -1, -1, -1, -1, -2, -1, 1, 0, -3, -2, 0, -2, -1, 0, -4},
    {-1, 5, 0, -2, -3, 1, 0, -2, 0, -3, -2, 2, -1, -3, -2, -1, -1, -3, -2, -3, -1, 0, -1, -4},
    {-2, 0, 6, 1, -3, 0, 0, 0, 1, -3, -3, 0, -2, -3, -2, 1, 0, -4, -2, -3, 3, 0, -1, -4},
    {-2, -2, 1, 6, -3, 0, 2, -1, -1, -3, -4, -1, -3, -3, -1, 0, -1, -4, -3, -3, 4, 1, -1, -4},
    {0, -3, -3, -3, 9, -3, -4, -3, -3, -1, -1, -3, -1, -2, -3, -1, -1, -2, -2, -1, -3, -3, -2, -4},
    {-1, 1, 0, 0, -3, 5, 2, -2, 0, -3, -2, 1, 0, -3, -1, 0, -1, -2, -1, -2, 0, 3, -1, -4},
    {-1, 0, 0, 2, -4, 2, 5, -2, 0, -3, -3, 1, -2, -3, -1, 0, -1, -3, -2, -2, 1, 4, -1, -4},
    {0, -2, 0, -1, -3, -2, -2, 6, -2, -4, -4, -2, -3, -3, -2, 0, -2, -2, -3, -3, -1, -2, -1, -4},
    {-2, 0, 1, -1, -3, 0, 0, -2, 8, -3, -3, -1, -2, -1, -2, -1, -2, -2, 2, -3, 0, 0, -1, -4},
    {-1, -3, -3, -3, -1, -3, -3, -4, -3, 4, 2, -3, 1, 0, -3, -2, -1, -3, -1, 3, -3, -3, -1, -4},
    {-1, -2, -3, -4, -1, -2, -3, -4, -3, 2, 4, -2, 2, 0, -3, -2, -1, -2, -1, 1, -4, -3, -1, -4},
    {-1, 2, 0, -1, -3, 1, 1, -2, -1, -3, -2, 5, -1, -3, -1, 0, -1, -3, -2, -2, 0, 1, -1, -4},
    {-1, -1, -2, -3, -1, 0, -2, -3, -2, 1, 2, -1, 5, 0, -2, -1, -1, -1, -1, 1, -3, -1, -1, -4},
    {-2, -3, -3, -3, -2, -3, -3, -3, -1, 0, 0, -3, 0, 6, -4, -2, -2, 1, 3, -1, -3, -3, -1, -4},
    {-1, -2, -2, -1, -3, -1, -1, -2, -2, -3, -3, -1, -2, -4, 7, -1, -1, -4, -3, -2, -2, -1, -2, -4},
    {1, -1, 1, 0, -1, 0, 0, 0, -1, -2, -2, 0, -1, -2, -1, 4, 1, -3, -2, -2, 0, 0, 0, -4},
    {0, -1, 0, -1, -1, -1, -1, -2, -2, -1, -1, -1, -1, -2, -1, 1, 5, -2, -2, 0, -1, -1, 0, -4},
    {-3, -3, -4, -4, -2, -2, -3, -2, -2, -3, -2, -3, -1, 1, -4, -3, -2, 11, 2, -3, -4, -3, -2, -4},
    {-2, -2, -2, -3, -2, -1, -2, -3, 2, -1, -1, -2, -1, 3, -3, -2, -2, 2, 7, -1, -3, -2, -1, -4},
    {0, -3, -3, -3, -1, -2, -2, -3, -3, 3, 1, -2, 1, -1, -2, -2, 0, -3, -1, 4, -3, -2, -1, -4},
    {-2, -1, 3, 4, -3, 0, 1, -1, 0, -3, -4, 0, -3, -3, -2, 0, -1, -4, -3, -3, 4, 1, -1, -4},
    {-1, 0, 0, 1, -3, 3, 4, -2, 0, -3, -3, 1, -1, -3, -1, 0, -1, -3, -2, -2, 1, 4, -1, -4},
    {0, -1, -1, -1, -2, -1, -1, -1, -1, -1, -1, -1, -1, -1, -2, 0, 0, -2, -1, -1, -1, -1, -1, -4},
    {-4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, -4, 1}};

extern "C" int32_t swb_matrix_blosum62(char *letters, float *scores)
{
    for (int i = 0; i < 24; ++i) {
        letters[i] = BL62_LETTERS[i];
        for (int j = 0; j < 24; ++j) scores[i * 24 + j] = (float)BL62[i][j];
    }
    return 24;
}

extern "C" int32_t swb_matrix_match_mismatch(float match, float mismatch, char *letters, float *scores)
{
    for (int i = 0; i < 26; ++i) {                      // SA.rs:681-695: letters A..Z
        letters[i] = (char)('A' + i);
        for (int j = 0; j < 26; ++j) scores[i * 26 + j] = (i == j) ? match : mismatch;
    }
    return 26;
}

extern "C" int swb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

// Smallest factor that puts every value on an integer grid.  Dyadic factors are exact (the f32
// reference is then exact as well and integer arithmetic is an identity transformation,
// SURVEY.md 7.3-3); decimal factors (0.1 steps) are accepted within 1e-4 and flagged inexact.
static bool find_scale(const std::vector<double> &vals, int32_t *scale, int32_t *exact)
{
    for (int s = 1; s <= 1024; s *= 2) {
        bool ok = true;
        for (double v : vals) if (v * s != std::nearbyint(v * s)) { ok = false; break; }
        if (ok) { *scale = s; *exact = 1; return true; }
    }
    const int dec[] = {5, 10, 20, 25, 50, 100, 200, 250, 500, 1000};
    for (int s : dec) {
        bool ok = true;
        for (double v : vals) if (std::fabs(v * s - std::nearbyint(v * s)) > 1e-4 * s) { ok = false; break; }
        if (ok) { *scale = s; *exact = 0; return true; }
    }
    return false;
}

extern "C" int swb_engine_create(const swb_matrix *m, float gap_open, float gap_extend, int mode, int device,
                                 swb_engine **out)
{
    if (!out) { g_create_err = "out is null"; return SWB_ERR_ARG; }
    *out = nullptr;
    if (!m || !m->letters || !m->scores || m->n_letters < 1 || m->n_letters > SWB_MAX_LETTERS) {
        g_create_err = "matrix must have 1..31 single-byte letters";
        return SWB_ERR_ARG;
    }
    if (mode != SWB_GLOBAL && mode != SWB_GLOCAL && mode != SWB_LOCAL) { g_create_err = "bad mode"; return SWB_ERR_ARG; }
    swb_engine *e = new swb_engine();
    e->device = device; e->mode = mode; e->n = m->n_letters;
    for (int i = 0; i < 256; ++i) e->code_of[i] = -1;
    for (int i = 0; i < e->n; ++i) {
        e->letters[i] = m->letters[i];
        if (e->code_of[(unsigned char)m->letters[i]] >= 0) {        // SA.rs:751-753
            g_create_err = std::string(1, m->letters[i]) + " was already found."; delete e; return SWB_ERR_ARG;
        }
        e->code_of[(unsigned char)m->letters[i]] = (int16_t)i;
    }
    const int xcode = e->code_of[(unsigned char)'X'];                // unknown -> X (SA.rs:669-679)
    if (xcode >= 0) for (int i = 0; i < 256; ++i) if (e->code_of[i] < 0) e->code_of[i] = (int16_t)xcode;
    e->go = gap_open > 0.0f ? -gap_open : gap_open;                  // SA.rs:28-35
    e->ge = gap_extend > 0.0f ? -gap_extend : gap_extend;
    e->matf.assign(m->scores, m->scores + (size_t)e->n * e->n);
    std::vector<double> vals(e->matf.begin(), e->matf.end());
    vals.push_back(e->go); vals.push_back(e->ge); vals.push_back(1.0);   // 1.0: the (0,0) cell holds -1.0 (SA.rs:90-91)
    for (double v : vals) if (!std::isfinite(v)) { g_create_err = "non-finite score"; delete e; return SWB_ERR_ARG; }
    if (!find_scale(vals, &e->scale, &e->exact_scale)) {
        g_create_err = "matrix / penalties are not on a common integer grid"; delete e; return SWB_ERR_SCALE;
    }
    e->mat.resize((size_t)e->n * e->n);
    double amax = 0;
    for (size_t k = 0; k < e->mat.size(); ++k) {
        const double v = std::nearbyint((double)e->matf[k] * e->scale);
        amax = std::max(amax, std::fabs(v));
        e->mat[k] = (int32_t)v;
    }
    e->o = (int32_t)std::nearbyint((double)e->go * e->scale);
    e->e = (int32_t)std::nearbyint((double)e->ge * e->scale);
    if (amax > 8000 || e->o < -8000 || e->e < -8000) {
        g_create_err = "scaled scores exceed the supported magnitude (8000)"; delete e; return SWB_ERR_RANGE;
    }
    e->smax = *std::max_element(e->mat.begin(), e->mat.end());
    e->smin = *std::min_element(e->mat.begin(), e->mat.end());

    cudaError_t ce = cudaSetDevice(device);
    if (ce != cudaSuccess) { g_create_err = std::string("cudaSetDevice: ") + cudaGetErrorString(ce); delete e; return SWB_ERR_CUDA; }
    cudaDeviceProp prop;
    ce = cudaGetDeviceProperties(&prop, device);
    if (ce != cudaSuccess) { g_create_err = cudaGetErrorString(ce); delete e; return SWB_ERR_CUDA; }
    e->sm_count = prop.multiProcessorCount;
    // device table with a PAD row/column (code n): strongly negative, never part of an optimum
    const int nl = e->n + 1;
    const int pad = -(std::max(e->smax, 0) + std::abs(e->smin) + 64);
    std::vector<int32_t> tab((size_t)nl * nl, pad);
    for (int i = 0; i < e->n; ++i) for (int j = 0; j < e->n; ++j) tab[(size_t)i * nl + j] = e->mat[(size_t)i * e->n + j];
    if ((ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (ce = cudaStreamCreateWithFlags(&e->stream_up, cudaStreamNonBlocking)) != cudaSuccess ||
        (ce = cudaStreamCreateWithFlags(&e->stream2, cudaStreamNonBlocking)) != cudaSuccess ||
        (ce = cudaEventCreate(&e->ev0)) != cudaSuccess || (ce = cudaEventCreate(&e->ev1)) != cudaSuccess ||
        (ce = cudaMalloc(&e->d_mat, tab.size() * sizeof(int32_t))) != cudaSuccess ||
        (ce = cudaMemcpy(e->d_mat, tab.data(), tab.size() * sizeof(int32_t), cudaMemcpyHostToDevice)) != cudaSuccess ||
        (ce = cudaMalloc(&e->d_code_tab, 256 * sizeof(int16_t))) != cudaSuccess ||
        (ce = cudaMemcpy(e->d_code_tab, e->code_of, 256 * sizeof(int16_t), cudaMemcpyHostToDevice)) != cudaSuccess) {
        g_create_err = std::string("CUDA init: ") + cudaGetErrorString(ce);
        swb_engine_destroy(e);
        return SWB_ERR_CUDA;
    }
    {   // keep freed blocks in the stream-ordered pool (cudaMallocAsync) instead of returning them to the driver
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    e->stats.scale = e->scale; e->stats.exact_scale = e->exact_scale;
    *out = e;
    return SWB_OK;
}

extern "C" void swb_engine_destroy(swb_engine *e)
{
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->d_mat) cudaFree(e->d_mat);
    if (e->d_code_tab) cudaFree(e->d_code_tab);
    if (e->d_bnd) cudaFree(e->d_bnd);
    if (e->d_dir) cudaFree(e->d_dir);
    if (e->h_stage) cudaFreeHost(e->h_stage);
    if (e->h_items) cudaFreeHost(e->h_items);
    if (e->h_tb) cudaFreeHost(e->h_tb);
    if (e->ev_items) cudaEventDestroy(e->ev_items);
    if (e->ev_team) cudaEventDestroy(e->ev_team);
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->stream_up) cudaStreamDestroy(e->stream_up);
    if (e->stream2) cudaStreamDestroy(e->stream2);
    delete e;
}

extern "C" const char *swb_last_error(const swb_engine *e) { return e ? e->err.c_str() : g_create_err.c_str(); }

extern "C" int swb_get_stats(const swb_engine *e, swb_stats *out)
{
    if (!e || !out) return SWB_ERR_ARG;
    *out = e->stats;
    return SWB_OK;
}

extern "C" int swb_dpx_peak(int device, int which, double *thread_instr_per_s)
{
    if (!thread_instr_per_s) return SWB_ERR_ARG;
    cudaDeviceProp prop;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) return SWB_ERR_CUDA;
    return k0_run(which, prop.multiProcessorCount, thread_instr_per_s) == cudaSuccess ? SWB_OK : SWB_ERR_CUDA;
}
