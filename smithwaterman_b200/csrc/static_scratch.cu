// swb_score_batch_static -- score + end cell of a small batch of pairs WITHOUT any device allocation.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scores are those of
// swb_score_batch; the kernel is the KW wavefront kernel (kw_trace.cu: fill_matrix SA.rs:55-272, score only), whose
// scratch -- the pair table, the strip-boundary rings, the per-pair progress words and the result records -- together
// with the encoded residues lives in ONE file-scope __device__ array of fixed size instead of cudaMallocAsync'ed buffers.
//
//   * every GPU that loads the library holds SWB_STATIC_SCRATCH_BYTES (about 9.5 MiB) until the process exits;
//   * all calls on a device share the array and it keeps the last call's contents: what must start at zero (the status
//     word here, the progress words inside kw_launch) is cleared on the caller's stream at the start of every call;
//   * two calls that overlap on the GPU (different streams, no dependency between them) RACE on the array.  Calls on
//     one stream are ordered by the stream and safe.
//
// Inputs and outputs are the caller's DEVICE buffers; the call enqueues two helper kernels, one memset of 256 bytes, the
// memset + launch of kw_launch, and returns without synchronising.
#include <mutex>

#include "swb_internal.h"

#define SS_TEAMS 256                                    // warps of the launch; pair k uses ring k % SS_TEAMS
#define SS_SCR_INTS 12                                  // per pair: arrivals + strips (<= 4) + 4 candidate words, rounded up
#define SS_RING_ELEMS (2 * (SWB_STATIC_MAX_LEN + 2))    // int4 per ring: two strip columns of la + 2 rows

static constexpr size_t ss_align(size_t x) { return (x + 255) / 256 * 256; }
static constexpr size_t SS_OFF_PAIRS = 256;             // [0, 256): header, word 0 = status
static constexpr size_t SS_OFF_RES = SS_OFF_PAIRS + ss_align(sizeof(KWPair) * SWB_STATIC_MAX_PAIRS);
static constexpr size_t SS_OFF_SCR = SS_OFF_RES + ss_align(sizeof(KWResult) * SWB_STATIC_MAX_PAIRS);
static constexpr size_t SS_OFF_CODES = SS_OFF_SCR + ss_align(sizeof(int) * SS_SCR_INTS * SWB_STATIC_MAX_PAIRS);
static constexpr size_t SS_OFF_BND = SS_OFF_CODES + ss_align(SWB_STATIC_MAX_RESIDUES + 256);   // codes of a, then (256-aligned) of b
static constexpr size_t SS_BYTES = SS_OFF_BND + ss_align(sizeof(int4) * SS_RING_ELEMS * SS_TEAMS);
static_assert(SS_BYTES == SWB_STATIC_SCRATCH_BYTES, "include/swb200.h states the size of the static scratch array");

// The scratch.  In host code this name is a stand-in, not a device address: ss_scratch_ptr() resolves it per device.
static __device__ __align__(256) unsigned char g_ss_scratch[SS_BYTES];

#define SS_MAX_DEVICES 64
static std::mutex g_ss_mu;
static unsigned char *g_ss_ptr[SS_MAX_DEVICES];          // device address of g_ss_scratch on device d (resolved once, with d current)

static cudaError_t ss_scratch_ptr(int device, unsigned char **out)
{
    if (device < 0 || device >= SS_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_ss_mu);
    if (!g_ss_ptr[device]) {
        int cur = -1;
        cudaError_t err = cudaGetDevice(&cur);
        if (err != cudaSuccess) return err;
        if (cur != device) return cudaErrorInvalidDevice;          // the caller made the engine's device current
        void *p = nullptr;
        err = cudaGetSymbolAddress(&p, g_ss_scratch);
        if (err != cudaSuccess) return err;
        g_ss_ptr[device] = (unsigned char *)p;
    }
    *out = g_ss_ptr[device];
    return cudaSuccess;
}

// Letters -> residue codes into the scratch (the caller's buffers are not touched), and the pair table.
__global__ void __launch_bounds__(256)
ss_prepare_kernel(const uint8_t *__restrict__ a_all, const int64_t *__restrict__ off_a, int64_t total_a,
                  const uint8_t *__restrict__ b_all, const int64_t *__restrict__ off_b, int64_t total_b, int n,
                  const int16_t *__restrict__ code_of, int pad, uint8_t *__restrict__ codes_a, uint8_t *__restrict__ codes_b,
                  KWPair *__restrict__ pairs, int *__restrict__ status)
{
    __shared__ int16_t tab[256];
    for (int k = threadIdx.x; k < 256; k += blockDim.x) tab[k] = code_of[k];
    __syncthreads();
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
    bool bad = false;
    for (int64_t i = tid; i < total_a; i += stride) { int c = tab[a_all[i]]; if (c < 0) { bad = true; c = pad; } codes_a[i] = (uint8_t)c; }
    for (int64_t i = tid; i < total_b; i += stride) { int c = tab[b_all[i]]; if (c < 0) { bad = true; c = pad; } codes_b[i] = (uint8_t)c; }
    if (bad) atomicCAS(status, 0, SWB_ERR_LETTER);                 // SA.rs:676: unknown letter and no X in the matrix
    for (int64_t k = tid; k < n; k += stride) {
        const int64_t a0 = off_a[k], a1 = off_a[k + 1], b0 = off_b[k], b1 = off_b[k + 1];
        const bool ok = a0 >= 0 && a0 <= a1 && a1 <= total_a && a1 - a0 <= SWB_STATIC_MAX_LEN &&
                        b0 >= 0 && b0 <= b1 && b1 <= total_b && b1 - b0 <= SWB_STATIC_MAX_LEN;
        if (!ok) atomicCAS(status, 0, SWB_ERR_ARG);
        KWPair P;
        P.a_off = ok ? a0 : 0; P.b_off = ok ? b0 : 0;
        P.la = ok ? (int32_t)(a1 - a0) : 0; P.lb = ok ? (int32_t)(b1 - b0) : 0;   // a rejected pair is scored as two empty sequences
        P.dir_off = -1; P.pssm_off = -1;
        P.bnd_off = (k % SS_TEAMS) * (int64_t)SS_RING_ELEMS;       // pairs k, k + SS_TEAMS, ... are swept by the same warp, one after the other
        P.scr_off = k * SS_SCR_INTS;
        pairs[k] = P;
    }
}

// KW records -> the caller's layout (f32 score, i64 end cell); pairs with an empty sequence get the closed-form boundary
// value as in swb_score_batch (SA.rs:88-118, 438-467).
__global__ void __launch_bounds__(256)
ss_finalize_kernel(const KWPair *__restrict__ pairs, const KWResult *__restrict__ res, int n, int mode, int o, int e, int scale,
                   float *__restrict__ scores, int64_t *__restrict__ end_a, int64_t *__restrict__ end_b,
                   const int *__restrict__ status, int32_t *__restrict__ status_out)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k == 0 && status_out) *status_out = *status;
    if (k >= n) return;
    const int la = pairs[k].la, lb = pairs[k].lb;
    int32_t s; int64_t ei, ej;
    if (la > 0 && lb > 0) { const KWResult r = res[k]; s = r.score; ei = r.end_i; ej = r.end_j; }
    else if (mode == SWB_LOCAL) { s = 0; ei = -1; ej = -1; }
    else {
        const int so = (mode == SWB_GLOBAL) ? o : 0, se = (mode == SWB_GLOBAL) ? e : 0;
        const int len = la + lb;
        s = len ? len * se + (so - se) : 0; ei = la; ej = lb;
    }
    const bool dyadic = (scale & (scale - 1)) == 0;
    scores[k] = dyadic ? (float)s * (1.0f / (float)scale) : (float)((double)s / (double)scale);
    if (end_a) end_a[k] = ei;
    if (end_b) end_b[k] = ej;
}

extern "C" int swb_score_batch_static(swb_engine *e, const char *d_a_all, const int64_t *d_off_a, int64_t total_a,
                                      const char *d_b_all, const int64_t *d_off_b, int64_t total_b, int64_t n,
                                      float *d_scores, int64_t *d_end_a, int64_t *d_end_b, int32_t *d_status, void *stream)
{
    if (!e) return SWB_ERR_ARG;
    if (n < 0 || total_a < 0 || total_b < 0) { e->err = "swb_score_batch_static: negative size"; return SWB_ERR_ARG; }
    if (n > SWB_STATIC_MAX_PAIRS) {
        e->err = "swb_score_batch_static: " + std::to_string(n) + " pairs, the static scratch holds at most " + std::to_string(SWB_STATIC_MAX_PAIRS);
        return SWB_ERR_ARG;
    }
    if (ss_align((size_t)total_a) + (size_t)total_b > SWB_STATIC_MAX_RESIDUES) {
        e->err = "swb_score_batch_static: " + std::to_string(total_a) + " + " + std::to_string(total_b) +
                 " letters, the static scratch holds at most " + std::to_string(SWB_STATIC_MAX_RESIDUES) + " (a rounded up to 256)";
        return SWB_ERR_ARG;
    }
    if (n > 0 && (!d_off_a || !d_off_b || !d_scores || (total_a > 0 && !d_a_all) || (total_b > 0 && !d_b_all))) {
        e->err = "swb_score_batch_static: null buffer"; return SWB_ERR_ARG;
    }
    if (!swb_range_ok(e, SWB_STATIC_MAX_LEN, SWB_STATIC_MAX_LEN)) { e->err = "scores of a pair this long do not fit the int32 lanes"; return SWB_ERR_RANGE; }
    CUDA_TRY(e, cudaSetDevice(e->device));
    unsigned char *base = nullptr;
    CUDA_TRY(e, ss_scratch_ptr(e->device, &base));
    cudaStream_t st = (cudaStream_t)stream;
    int *status = reinterpret_cast<int *>(base);
    KWPair *pairs = reinterpret_cast<KWPair *>(base + SS_OFF_PAIRS);
    KWResult *res = reinterpret_cast<KWResult *>(base + SS_OFF_RES);
    int *scr = reinterpret_cast<int *>(base + SS_OFF_SCR);
    uint8_t *codes_a = base + SS_OFF_CODES, *codes_b = codes_a + ss_align((size_t)total_a);
    int4 *bnd = reinterpret_cast<int4 *>(base + SS_OFF_BND);

    CUDA_TRY(e, cudaMemsetAsync(base, 0, 256, st));                // the header (status word) starts at zero in every call
    if (n == 0) {
        if (d_status) CUDA_TRY(e, cudaMemcpyAsync(d_status, status, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
        return SWB_OK;
    }
    const int64_t work = std::max<int64_t>(std::max(total_a, total_b), n);
    const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((work + 255) / 256, (int64_t)e->sm_count * 8));
    ss_prepare_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const uint8_t *>(d_a_all), d_off_a, total_a,
                                             reinterpret_cast<const uint8_t *>(d_b_all), d_off_b, total_b, (int)n, e->d_code_tab, e->n,
                                             codes_a, codes_b, pairs, status);
    CUDA_TRY(e, cudaGetLastError());
    KWPlan pl;                                                     // one warp per pair, 256-column strips, no teams: a plain launch
    pl.cols = 8; pl.team_warps = 1;
    pl.n_teams = (int)std::min<int64_t>(n, SS_TEAMS);
    pl.blocks = (pl.n_teams + KW_WARPS - 1) / KW_WARPS;
    pl.bnd_elems = (int64_t)SS_RING_ELEMS * SS_TEAMS;
    pl.scratch_ints = n * SS_SCR_INTS;                             // cleared by kw_launch (cudaMemsetAsync on st)
    pl.dir_bytes = 0;
    CUDA_TRY(e, kw_launch(e, codes_a, codes_b, pairs, (int)n, pl, nullptr, bnd, scr, res, st));
    ss_finalize_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(pairs, res, (int)n, e->mode, e->o, e->e, e->scale, d_scores, d_end_a, d_end_b,
                                                              status, d_status);
    CUDA_TRY(e, cudaGetLastError());
    return SWB_OK;
}
