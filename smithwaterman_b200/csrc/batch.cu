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

#define K2_MIN_QUERY 1024           // shared queries at least this long use K2 (its 124-step skew is < 12 % there)

struct BatchDev {                  // device-only members that do not fit the plain struct in the header
    uint8_t *codes_a = nullptr, *codes_b = nullptr;   // encoded residues, caller layout
    int64_t *off_a = nullptr, *off_b = nullptr;
    int32_t *order = nullptr;                          // sorted position -> caller index (K1 pairs)
    int32_t *sort_cnt = nullptr; int64_t *sort_sz = nullptr;   // device shape sort scratch
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
    // results in the caller's layout, produced on the device (finalize kernels) and copied out as three plain arrays
    float *d_sc = nullptr; int64_t *d_ea = nullptr, *d_eb = nullptr;
    int32_t *d_kw_idx = nullptr;                             // caller index of each KW pair (device copy)
    int32_t *d_ovf = nullptr;                                // [0] = number of s16 overflows, [1..] their caller indices
    int32_t *h_flags = nullptr;                              // pinned: [0] unknown-letter flag, [1] overflow count
    float *h_sc = nullptr; int64_t *h_ea = nullptr, *h_eb = nullptr;   // pinned staging, used when the caller's arrays are pageable
    bool own_stage = false, own_res = false, direct = false, use_arena = false;
    float *dst_sc = nullptr; int64_t *dst_ea = nullptr, *dst_eb = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_done = nullptr, ev_ready = nullptr;
    cudaEvent_t ev_h2d = nullptr;                            // joins the copy stream into the helper stream (upload)
    cudaEvent_t ev_trace = nullptr;                          // SWB_TRACE only: residues and offsets have landed
    cudaStream_t alloc_stream = nullptr;                     // stream of the batch's allocations (and of their release)
    bool launched = false, done_recorded = false;
    bool borrow_b = false, borrow_items = false, borrow_res = false;   // resident database: codes_b / off_b, the K2 items, d_res belong to the swb_db
};
static BatchDev *dev_of(swb_batch *b) { return b->dev; }

// ---- device helpers ---------------------------------------------------------------------------
// All helper kernels of the upload / result path are built as 128-thread blocks of at most 32 registers: K1 is capped at
// 120 registers (k1_score.cu), so one helper block fits beside the resident K1 block of every SM.  The next chunk's
// encode / sort / pack kernels and the previous chunk's result kernels therefore run WHILE K1 computes instead of
// waiting for it to drain, and never stall their streams behind a kernel that holds all SMs.
#define SWB_HELPER __launch_bounds__(128, 16)
#define HELPER_THREADS 128

__global__ void SWB_HELPER zero_kernel(uint32_t *p, int64_t words)
{
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < words; k += (int64_t)gridDim.x * blockDim.x) p[k] = 0u;
}
static cudaError_t zero_async(void *p, size_t bytes, cudaStream_t st)      // bytes and p multiples of 4
{
    const int64_t words = (int64_t)(bytes / 4);
    if (!words) return cudaSuccess;
    zero_kernel<<<(int)std::min<int64_t>((words + HELPER_THREADS - 1) / HELPER_THREADS, 1184), HELPER_THREADS, 0, st>>>((uint32_t *)p, words);
    return cudaGetLastError();
}

// Exclusive prefix sum by ONE block (the inputs are the sort's bucket counts and the work items' sizes: 10^4 elements).
template <typename T>
__global__ void SWB_HELPER scan_kernel(const T *__restrict__ in, T *__restrict__ out, int64_t n)
{
    __shared__ T wsum[HELPER_THREADS / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    T carry = 0;
    for (int64_t base = 0; base < n; base += HELPER_THREADS) {
        const int64_t i = base + threadIdx.x;
        const T v = i < n ? in[i] : (T)0;
        T x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const T y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) wsum[w] = x;
        __syncthreads();
        T woff = 0, total = 0;
#pragma unroll
        for (int q = 0; q < HELPER_THREADS / 32; ++q) { if (q < w) woff += wsum[q]; total += wsum[q]; }
        if (i < n) out[i] = carry + woff + x - v;
        carry += total;
        __syncthreads();
    }
}
static int64_t pad16(int64_t x) { return (x + 15) / 16 * 16; }

__global__ void SWB_HELPER encode_kernel(uint8_t *buf, int64_t nbytes, const int16_t *__restrict__ code_of, int32_t *err, int pad)
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
                if (base + q * 4 + s >= nbytes) c = pad;                             // behind the last residue: PAD, whatever was there
                else if (c < 0) { bad = true; c = pad; }                            // flagged; PAD keeps later kernels in range
                r |= ((unsigned)(c & 0xFF)) << (8 * s);
            }
            w[q] = r;
        }
        if (bad) atomicExch(err, 1);
        *reinterpret_cast<uint4 *>(buf + base) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

cudaError_t batch_encode(swb_engine *e, uint8_t *d_buf, int64_t nbytes, int32_t *d_err, cudaStream_t st)
{
    const int64_t np = pad16(nbytes) + 16;
    encode_kernel<<<(int)std::min<int64_t>((np / 16 + HELPER_THREADS - 1) / HELPER_THREADS, e->sm_count * 16), HELPER_THREADS, 0, st>>>(d_buf, nbytes, e->d_code_tab, d_err, e->n);
    return cudaGetLastError();
}

// One block per K1 work item: gathers the 32 pairs' residues into the [4-row block][lane] row layout
// and the [pass][lane][32] column layout, PAD-filled, and writes the per-pair lengths.
__global__ void SWB_HELPER k1_pack_kernel(const uint8_t *__restrict__ codes_a, const uint8_t *__restrict__ codes_b,
                               const int64_t *__restrict__ off_a, const int64_t *__restrict__ off_b,
                               const int32_t *__restrict__ order, K1Group *groups,
                               uint8_t *__restrict__ a_pk, uint8_t *__restrict__ b_pk, int32_t *__restrict__ la_arr,
                               int32_t *__restrict__ lb_arr, int pad, int64_t base_a, int64_t base_b, int shared_la,
                               const int64_t *__restrict__ aoff, const int64_t *__restrict__ boff, int right_align)
{
    K1Group G = groups[blockIdx.x];
    G.a_off = shared_la >= 0 ? 0 : aoff[blockIdx.x]; G.b_off = boff[blockIdx.x];     // exclusive sums of the items' block sizes
    __syncthreads();
    if (threadIdx.x == 0) { groups[blockIdx.x].a_off = G.a_off; groups[blockIdx.x].b_off = G.b_off; }   // K1 reads them from here
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
    // glocal (right_align): b ends at the last column of the item's last pass, PAD columns lead (k1_global.cu)
    const int shift = right_align ? G.n_pass * K1_STRIP - lb : 0;
    for (int h = sub; h < G.n_pass * 2; h += nsub) {       // 16-byte halves of the 32-column passes
        unsigned w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            unsigned r = 0;
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int col = h * 16 + q * 4 + s - shift;
                const unsigned code = (col >= 0 && col < lb) ? codes_b[ob + col] : (unsigned)pad;
                r |= code << (8 * s);
            }
            w[q] = r;
        }
        reinterpret_cast<uint4 *>(b_pk + G.b_off)[((h >> 1) * 32 + lane) * 2 + (h & 1)] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---- shape sort on the device -----------------------------------------------------------------------------------
// K1 wants its work items (32 pairs each) homogeneous in shape and the long ones first.  The sort is a counting sort on
// (passes descending, la descending) done by four small kernels on the upload stream, straight from the offsets that
// are on the device anyway: the host never touches per-pair sort data (it only sizes the buffers from an upper bound),
// so a chunk's preparation costs the host a streaming pass over the offsets.
#define K1_MAX_LEN 8192            // longer pairs go to the wavefront kernel

// Which pairs K1 takes, the same rule on the host (routing, buffer bounds) and on the device (sort).  Local mode: both
// lengths in 1..8192 (an s16 overflow is flagged by the kernel and re-run in int32).  Global mode has no clamp and no
// flag: the pair must be unable to leave the s16 lanes -- at most min(la,lb) matches of the best score upwards, at
// most two gap openings plus la+lb extensions, the sentinel offset and one worst substitution downwards.
struct K1Take {
    int global_mode;
    int64_t up_per_match, up_base;     // min(la,lb) * up_per_match + up_base <= 32000
    int64_t down_per_res, down_base;   // (la+lb) * down_per_res + down_base <= 32000
};
__host__ __device__ __forceinline__ bool k1_take(int64_t la, int64_t lb, const K1Take &T)
{
    if (la <= 0 || lb <= 0 || la > K1_MAX_LEN || lb > K1_MAX_LEN) return false;
    if (!T.global_mode) return true;
    const int64_t mn = la < lb ? la : lb;
    return mn * T.up_per_match + T.up_base <= 32000 && (la + lb) * T.down_per_res + T.down_base <= 32000;
}
static K1Take k1_take_rule(const swb_engine *e)
{
    K1Take T;
    T.global_mode = (e->mode != SWB_LOCAL) ? 1 : 0;     // global and glocal: no clamp, no overflow flag -> static bound
    const int64_t ao = -(int64_t)e->o, ae = -(int64_t)e->e, sp = std::max(e->smax, 0), sn = std::max(-e->smin, 0);
    T.up_per_match = sp; T.up_base = ao + sp + sn + 64;
    T.down_per_res = ae; T.down_base = 3 * ao + (10 * ao + 10 * ae) + sp + sn + 64;
    return T;
}

__device__ __forceinline__ bool sort_take(const int64_t *__restrict__ off_a, const int64_t *__restrict__ off_b, int64_t k,
                                          int shared_la, int la_cap, int la_gran, int pass_cap, const K1Take &T, int &la, int &np, int &bucket)
{
    const int64_t la64 = shared_la >= 0 ? (int64_t)shared_la : off_a[k + 1] - off_a[k], lb64 = off_b[k + 1] - off_b[k];
    if (!k1_take(la64, lb64, T)) return false;
    la = (int)la64; np = ((int)lb64 + K1_STRIP - 1) / K1_STRIP;
    bucket = (pass_cap - np) * (la_cap / la_gran + 1) + (la_cap - la) / la_gran;     // la in steps of la_gran: at most 65536 buckets
    return true;
}

// SCATTER == false: histogram of the buckets; true: order[cursor[bucket]++] = pair.  One atomic per distinct bucket of a warp.
template <bool SCATTER>
__global__ void SWB_HELPER sort_count_kernel(const int64_t *__restrict__ off_a, const int64_t *__restrict__ off_b, int64_t n, int shared_la,
                                  int la_cap, int la_gran, int pass_cap, K1Take T, int32_t *__restrict__ cnt, int32_t *__restrict__ order)
{
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    int la = 0, np = 0, bucket = -1;
    const bool take = k < n && sort_take(off_a, off_b, k, shared_la, la_cap, la_gran, pass_cap, T, la, np, bucket);
    const unsigned takers = __ballot_sync(0xffffffffu, take);
    if (!take) return;
    const unsigned peers = __match_any_sync(takers, bucket);
    const int leader = __ffs(peers) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(&cnt[bucket], __popc(peers));
    if (SCATTER) {
        base = __shfl_sync(peers, base, leader);
        order[base + __popc(peers & ((1u << lane) - 1u))] = (int32_t)k;
    }
}

// One warp per work item: shape of the item (longest a, most passes) and the bytes of its packed residue blocks.
__global__ void SWB_HELPER sort_groups_kernel(const int64_t *__restrict__ off_a, const int64_t *__restrict__ off_b, int shared_la,
                                   const int32_t *__restrict__ order, int n_k1, int n_groups, K1Group *__restrict__ groups,
                                   int64_t *__restrict__ asz, int64_t *__restrict__ bsz)
{
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (g >= n_groups) return;
    const int np_ = min(32, n_k1 - g * 32);
    int la = 0, np = 0;
    if (lane < np_) {
        const int k = order[g * 32 + lane];
        la = shared_la >= 0 ? shared_la : (int)(off_a[k + 1] - off_a[k]);
        np = ((int)(off_b[k + 1] - off_b[k]) + K1_STRIP - 1) / K1_STRIP;
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        la = max(la, __shfl_xor_sync(0xffffffffu, la, d));
        np = max(np, __shfl_xor_sync(0xffffffffu, np, d));
    }
    if (lane == 0) {
        K1Group G;
        G.a_off = 0; G.b_off = 0; G.la_max = la; G.n_pass = np; G.first_pair = g * 32; G.n_pairs = np_;
        groups[g] = G;
        asz[g] = (shared_la >= 0 && g > 0) ? 0 : (int64_t)((la + 1 + 3) / 4 + 1) * 32 * 4;   // shared query: one row block for all
        bsz[g] = (int64_t)np * 32 * 32;
    }
}

// Results in the caller's layout (f32 score, i64 end cells): the D2H copy lands in the caller's arrays and the host has
// no per-pair work left.  Pairs whose s16 lanes overflowed are appended to a short list for the int32 re-run.
__device__ __forceinline__ float score_to_f(int32_t s, int scale, float inv, bool dyadic)
{
    return dyadic ? (float)s * inv : (float)((double)s / (double)scale);
}
__global__ void SWB_HELPER finalize_k1_kernel(const K1Result *__restrict__ res, int64_t n, int scale, float inv, int dyadic,
                                   float *__restrict__ sc, int64_t *__restrict__ ea, int64_t *__restrict__ eb, int32_t *__restrict__ ovf)
{
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int4 r = reinterpret_cast<const int4 *>(res)[k];       // score, end_i, end_j, flags
    if (r.w & 1) ovf[1 + atomicAdd(&ovf[0], 1)] = (int32_t)k;
    sc[k] = score_to_f(r.x, scale, inv, dyadic != 0); ea[k] = r.y; eb[k] = r.z;
}
__global__ void SWB_HELPER finalize_kw_kernel(const KWResult *__restrict__ res, const int32_t *__restrict__ idx, int n_kw, int scale, float inv,
                                   int dyadic, float *__restrict__ sc, int64_t *__restrict__ ea, int64_t *__restrict__ eb)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_kw) return;
    const KWResult r = res[t];
    const int k = idx[t];
    sc[k] = score_to_f(r.score, scale, inv, dyadic != 0); ea[k] = r.end_i; eb[k] = r.end_j;
}

// KW results -> the unified per-pair record array (database search: the top-K selection reads one array)
__global__ void SWB_HELPER collect_kw_kernel(const KWResult *__restrict__ res, const int32_t *__restrict__ idx, int n_kw, K1Result *__restrict__ out)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_kw) return;
    const KWResult r = res[t];
    K1Result o; o.score = r.score; o.end_i = r.end_i; o.end_j = r.end_j; o.flags = 0;
    out[idx[t]] = o;
}
cudaError_t batch_collect_kw(swb_engine *e, swb_batch *b, cudaStream_t st)
{
    BatchDev *d = b->dev;
    if (!d->n_kw) return cudaSuccess;
    collect_kw_kernel<<<(d->n_kw + HELPER_THREADS - 1) / HELPER_THREADS, HELPER_THREADS, 0, st>>>(d->kw_res, d->d_kw_idx, d->n_kw, b->d_res);
    return cudaGetLastError();
}
int batch_device_views(swb_batch *b, const uint8_t **codes_a, const uint8_t **codes_b)
{
    *codes_a = b->dev->codes_a; *codes_b = b->dev->codes_b;
    return SWB_OK;
}

// cudaHostAlloc'ed / registered host memory: an asynchronous copy to it does not stall the host
static bool host_pinned(const void *p, size_t bytes)
{
    if (!p || !bytes) return false;
    cudaPointerAttributes at0, at1;
    if (cudaPointerGetAttributes(&at0, p) != cudaSuccess || cudaPointerGetAttributes(&at1, (const char *)p + bytes - 1) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at0.type == cudaMemoryTypeHost && at1.type == cudaMemoryTypeHost;
}

// Device memory comes from the stream-ordered pool of the engine's device (cudaMallocAsync): after the first call the
// allocations of a batch cost microseconds and never synchronise the device.
#define DEV_ALLOC_ON(e_, st, ptr, bytes) CUDA_TRY(e_, cudaMallocAsync((void **)&(ptr), (size_t)(bytes) ? (size_t)(bytes) : 16, st))
#define DEV_ALLOC(e_, ptr, bytes) DEV_ALLOC_ON(e_, (e_)->stream, ptr, bytes)

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

static cudaStream_t kstream(const swb_engine *e, const swb_batch *b);
// K2 work items: subjects of >= 32 passes take a warp each (several sweeps), the others are bin-packed (first fit,
// longest first) into warps of 32 lanes, one lane per 32-column pass.
size_t k2_items_cap(const std::vector<int32_t> &lb, const std::vector<int32_t> &idx)
{
    size_t longs = 0, small_passes = 0;
    for (int32_t k : idx) {
        const int np = (lb[k] + K1_STRIP - 1) / K1_STRIP;
        if (np >= 32) ++longs; else small_passes += (size_t)np;
    }
    return longs + small_passes / 32 + small_passes / 256 + 64;    // first fit wastes < 12 %
}
int64_t k2_build_items(const std::vector<int32_t> &lb, const std::vector<int32_t> &idx, int32_t *h_items, int32_t *h_sweeps, size_t cap_items)
{
    std::vector<std::vector<int32_t>> bucket(32);
    std::vector<int32_t> longs;
    for (int32_t k : idx) {
        const int np = (lb[k] + K1_STRIP - 1) / K1_STRIP;
        if (np >= 32) longs.push_back(k); else bucket[np].push_back(k);
    }
    std::sort(longs.begin(), longs.end(), [&](int32_t x, int32_t y) { return lb[x] > lb[y]; });
    unsigned mask = 0;
    for (int p = 1; p < 32; ++p) if (!bucket[p].empty()) mask |= 1u << p;
    size_t ni = 0;
    for (int32_t k : longs) {
        if (ni >= cap_items) return -1;
        for (int l = 0; l < 32; ++l) h_items[ni * 32 + l] = k;
        h_sweeps[ni++] = ((lb[k] + K1_STRIP - 1) / K1_STRIP + 31) / 32;
    }
    while (mask) {
        int cap = 32, l = 0;
        if (ni >= cap_items) return -1;
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
    return (int64_t)ni;
}

// The blocks go back to the pool on the stream they were allocated on: a block freed on a compute stream would be ordered
// behind whatever kernels are queued there (the K1 of a later chunk), and the next chunk's allocations on the upload
// stream, which reuse it, would wait for those kernels.  The batch's own work is ordered before the free by its events.
void batch_release(swb_engine *e, swb_batch *b)
{
    if (!b) return;
    if (b->dev) {                                     // what belongs to a resident database is not the batch's to free
        if (b->dev->borrow_b) { b->dev->codes_b = nullptr; b->dev->off_b = nullptr; }
        if (b->dev->borrow_items) { b->dev->items = nullptr; b->dev->sweeps = nullptr; }
        if (b->dev->borrow_res) b->d_res = nullptr;
    }
    cudaSetDevice(e->device);
    BatchDev *d = dev_of(b);
    cudaStream_t fs = (d && d->alloc_stream) ? d->alloc_stream : e->stream;
    if (d && d->launched && fs != kstream(e, b)) {
        if (!d->done_recorded && d->ev_done) cudaEventRecord(d->ev_done, kstream(e, b));     // work in flight, results never fetched
        if (d->ev_done) cudaStreamWaitEvent(fs, d->ev_done, 0);
    }
    auto dev_free = [fs](void *p) { if (p) cudaFreeAsync(p, fs); };
    if (d) {
        dev_free(d->codes_a); dev_free(d->codes_b); dev_free(d->off_a); dev_free(d->off_b); dev_free(d->order);
        dev_free(d->sort_cnt); dev_free(d->sort_sz);
        dev_free(d->err); dev_free(d->items); dev_free(d->sweeps); dev_free(d->kw_pairs); dev_free(d->kw_res); dev_free(d->kw_bnd); dev_free(d->kw_scr);
        dev_free(d->d_sc); dev_free(d->d_ea); dev_free(d->d_eb); dev_free(d->d_kw_idx); dev_free(d->d_ovf);
        if (d->own_stage) cudaFreeHost(d->h_flags);
        if (d->own_res) { cudaFreeHost(d->h_sc); cudaFreeHost(d->h_ea); cudaFreeHost(d->h_eb); }
        if (d->ev0) cudaEventDestroy(d->ev0);
        if (d->ev1) cudaEventDestroy(d->ev1);
        if (d->ev_done) cudaEventDestroy(d->ev_done);
        if (d->ev_ready) cudaEventDestroy(d->ev_ready);
        if (d->ev_h2d) cudaEventDestroy(d->ev_h2d);
        if (d->ev_trace) cudaEventDestroy(d->ev_trace);
        delete d;
        b->dev = nullptr;
    }
    dev_free(b->d_a); dev_free(b->d_b); dev_free(b->d_groups); dev_free(b->d_la); dev_free(b->d_lb);
    dev_free(b->d_res); dev_free(b->d_counter);
    delete b;
}

extern "C" void swb_batch_free(swb_engine *e, swb_batch *b) { if (e) batch_release(e, b); }

// Pass-boundary scratch of K1 / K2, owned by the engine, grown on demand (rare): two slots, one per compute stream, so
// that the kernels of consecutive chunks can be resident together (the next chunk's blocks fill the SMs the previous
// chunk's tail leaves idle).  Each slot's low row blocks are pinned in L2: they are rewritten and re-read every pass.
static int ensure_bnd(swb_engine *e, size_t need)
{
    if (need <= e->bnd_bytes) return SWB_OK;
    need = (need + 255) / 256 * 256;
    CUDA_TRY(e, cudaDeviceSynchronize());
    if (e->d_bnd) cudaFree(e->d_bnd);
    e->d_bnd = nullptr; e->bnd_bytes = 0;
    CUDA_TRY(e, cudaMalloc(&e->d_bnd, 2 * need));
    e->bnd_bytes = need;
    int max_persist = 0, max_win = 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, e->device);
    cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, e->device);
    if (max_persist > 0 && max_win > 0) {
        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)max_persist);
        for (int slot = 0; slot < 2; ++slot) {
            cudaStreamAttrValue av = {};
            av.accessPolicyWindow.base_ptr = (char *)e->d_bnd + slot * need;
            av.accessPolicyWindow.num_bytes = std::min<size_t>(need, (size_t)std::min(max_win, max_persist));   // the low row blocks
            av.accessPolicyWindow.hitRatio = 1.0f;
            av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            cudaStreamSetAttribute(slot ? e->stream2 : e->stream, cudaStreamAttributeAccessPolicyWindow, &av);
        }
        if (getenv("SWB_TRACE")) fprintf(stderr, "[swb] boundary scratch 2 x %zu bytes, L2 persisting limit %d\n", need, max_persist);
        cudaGetLastError();
    }
    return SWB_OK;
}
static void *bnd_slot(const swb_engine *e, const swb_batch *b) { return (char *)e->d_bnd + (size_t)b->slot * e->bnd_bytes; }
static cudaStream_t kstream(const swb_engine *e, const swb_batch *b) { return b->slot ? e->stream2 : e->stream; }

// Upload = shape sort + work-item table on the host, then (all asynchronous on the engine's stream) H2D of the raw
// residues and offsets, in-place encoding, packing into the K1 layout.  `off_*` are the caller's absolute offsets;
// residues are taken from a_all[off_a[0] .. off_a[n]).  use_arena: results staging comes from the engine's arena.
int batch_upload_impl(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                      const int64_t *off_b, int64_t n, int64_t shared_la, bool use_arena, cudaStream_t up, cudaStream_t hs, swb_batch **out,
                      const swb_db *rdb)
{
    if (rdb) { off_b = rdb->h_off.data(); b_all = nullptr; n = rdb->n; }
    *out = nullptr;
    swb_batch *b = new swb_batch();
    BatchDev *d = new BatchDev();
    b->dev = d;
    b->n = n;
    struct Guard { swb_engine *e; swb_batch *b; bool ok = false; ~Guard() { if (!ok) batch_release(e, b); } } guard{e, b};
    const bool shared = shared_la >= 0;
    d->shared_la = shared_la;
    d->alloc_stream = up;
    // `up` carries the allocations and the H2D copies, `hs` the helper kernels; join() orders everything issued on `up` so
    // far before what follows on `hs`.  (The pipelined path passes the same stream for both: with a separate helper
    // stream the copies run back to back, but the helper kernels then always run beside a resident K1 instead of in the
    // gaps between two K1 launches, where they are 5x faster -- measured 25.6 vs 24.6 ms per 10^6 pairs.)
    CUDA_TRY(e, cudaEventCreateWithFlags(&d->ev_h2d, cudaEventDisableTiming));
    auto join = [&]() -> cudaError_t {
        if (up == hs) return cudaSuccess;
        cudaError_t ce = cudaEventRecord(d->ev_h2d, up);
        return ce != cudaSuccess ? ce : cudaStreamWaitEvent(hs, d->ev_h2d, 0);
    };
    d->base_a = (n && !shared) ? off_a[0] : 0; d->base_b = n ? off_b[0] : 0;
    const int64_t na = shared ? shared_la : (n ? off_a[n] - d->base_a : 0), nb = n ? off_b[n] - d->base_b : 0;
    if (na < 0 || nb < 0 || (na > 0 && !a_all) || (nb > 0 && !b_all && !rdb)) { e->err = "bad sequence buffer / offsets"; return SWB_ERR_ARG; }
    const bool use_k1 = e->o <= e->e;                // all three modes; the packed Gotoh cell needs |open| >= |ext|
    const K1Take take = k1_take_rule(e);

    // ---- device buffers + H2D of the raw residues first (stream-ordered, no host synchronisation): the copy engine works
    //      while the host sorts
    const int64_t na_p = pad16(na) + 16, nb_p = pad16(nb) + 16;
    DEV_ALLOC_ON(e, up, d->codes_a, na_p);
    if (rdb) { d->codes_b = rdb->d_codes; d->off_b = rdb->d_off; d->borrow_b = true; }       // resident: nothing to copy or encode
    else { DEV_ALLOC_ON(e, up, d->codes_b, nb_p); DEV_ALLOC_ON(e, up, d->off_b, (n + 1) * sizeof(int64_t)); }
    if (!shared) DEV_ALLOC_ON(e, up, d->off_a, (n + 1) * sizeof(int64_t));
    DEV_ALLOC_ON(e, up, d->err, sizeof(int32_t));
    if (na) CUDA_TRY(e, cudaMemcpyAsync(d->codes_a, a_all + d->base_a, na, cudaMemcpyHostToDevice, up));
    if (nb && !rdb) CUDA_TRY(e, cudaMemcpyAsync(d->codes_b, b_all + d->base_b, nb, cudaMemcpyHostToDevice, up));
    if (n) {
        if (!shared) CUDA_TRY(e, cudaMemcpyAsync(d->off_a, off_a, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, up));
        if (!rdb) CUDA_TRY(e, cudaMemcpyAsync(d->off_b, off_b, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, up));
    }
    e->stats.h2d_bytes += (double)na + (rdb ? 0.0 : (double)nb + (double)(n + 1) * sizeof(int64_t)) + (shared ? 0.0 : (double)(n + 1) * sizeof(int64_t));
    if (getenv("SWB_TRACE")) { cudaEventCreate(&d->ev_trace); cudaEventRecord(d->ev_trace, up); }
    CUDA_TRY(e, join());
    // (own kernels instead of cudaMemsetAsync: see SWB_HELPER)
    CUDA_TRY(e, zero_async(d->err, sizeof(int32_t), hs));
    CUDA_TRY(e, zero_async(d->codes_a + pad16(na), 16, hs));       // the slack behind the residues
    if (!rdb) CUDA_TRY(e, zero_async(d->codes_b + pad16(nb), 16, hs));
    if (na) encode_kernel<<<(int)std::min<int64_t>((na_p / 16 + HELPER_THREADS - 1) / HELPER_THREADS, e->sm_count * 16), HELPER_THREADS, 0, hs>>>(d->codes_a, na, e->d_code_tab, d->err, e->n);
    if (nb && !rdb) encode_kernel<<<(int)std::min<int64_t>((nb_p / 16 + HELPER_THREADS - 1) / HELPER_THREADS, e->sm_count * 16), HELPER_THREADS, 0, hs>>>(d->codes_b, nb, e->d_code_tab, d->err, e->n);
    CUDA_TRY(e, cudaGetLastError());

    // ---- routing (one streaming pass over the offsets); the shape sort itself runs on the device
    b->la.resize(n); b->lb.resize(n);
    // a long shared query (database search) goes to K2: lanes split the columns, every work item lasts la + 124 row steps
    const bool use_k2 = use_k1 && e->mode == SWB_LOCAL && shared && shared_la >= K2_MIN_QUERY && shared_la <= 0x0FFFFFFF;
    std::vector<int32_t> k2_idx;
    int la_cap = 0, pass_cap = 0;
    int64_t n_k1 = 0, sum_blk = 0, sum_np = 0;                 // K1 pairs, their row-block words and passes (buffer bounds)
    b->cells = 0;
    const bool cached_k2 = rdb && use_k2 && rdb->d_items;      // the routing of a resident database under K2 does not depend on the query
    if (cached_k2) {
        d->degenerate = rdb->degenerate;
        b->cells = (double)shared_la * (double)rdb->n_res;
    }
    for (int64_t k = 0; k < n && !cached_k2; ++k) {
        const int64_t la = shared ? shared_la : off_a[k + 1] - off_a[k], lb = off_b[k + 1] - off_b[k];
        if (la < 0 || lb < 0 || la > 0x3FFFFFFF || lb > 0x3FFFFFFF) { e->err = "bad offsets"; return SWB_ERR_ARG; }
        b->la[k] = (int32_t)la; b->lb[k] = (int32_t)lb;
        b->cells += (double)la * (double)lb;
        if (la == 0 || lb == 0) d->degenerate.push_back((int32_t)k);
        else if (use_k2) k2_idx.push_back((int32_t)k);
        else if (use_k1 && k1_take(la, lb, take)) {                               // the device sort applies the same rule
            const int np = (int)((lb + K1_STRIP - 1) / K1_STRIP);
            ++n_k1; sum_blk += (la + 1 + 3) / 4 + 1; sum_np += np;
            la_cap = std::max(la_cap, (int)la); pass_cap = std::max(pass_cap, np);
        } else {
            if (!swb_range_ok(e, la, lb)) { e->err = "scores of a pair this long do not fit the int32 lanes"; return SWB_ERR_RANGE; }
            d->kw_idx.push_back((int32_t)k);
        }
    }
    d->n_k1 = (int)n_k1;
    if (d->n_k1) {
        const int ng = (d->n_k1 + 31) / 32;
        b->n_groups = ng; b->la_max = la_cap;
        // Upper bounds of the packed blocks: an item is as tall as its longest a and as wide as its widest b.  In the sorted
        // order the excess over the pairs' own sizes telescopes inside a bucket of equal passes; items that straddle two
        // buckets and the last partial item add at most one full item each.
        // The sort key holds la in steps of la_gran so that the one-block scan over the buckets stays short (long pairs:
        // 8192 x 256 passes would be 2.1 M buckets); inside a step the order is arbitrary, which costs an item at most
        // la_gran - 1 extra rows per pair.
        const int la_gran = (int)(((int64_t)(pass_cap + 1) * (la_cap + 1) + 65535) / 65536);
        const int64_t blk_cap = (la_cap + 1 + 3) / 4 + 1;
        b->a_bytes = (size_t)(shared ? blk_cap * 128 : 4 * (sum_blk + n_k1 * ((la_gran + 2) / 4 + 1) + 32 * blk_cap * (2 * (int64_t)pass_cap + 1)));
        b->b_bytes = (size_t)(32 * (sum_np + 64 * (int64_t)pass_cap));
        const int64_t n_buckets = (int64_t)(pass_cap + 1) * (la_cap / la_gran + 1);
        DEV_ALLOC_ON(e, up, d->sort_cnt, 2 * n_buckets * sizeof(int32_t));       // histogram | cursors
        DEV_ALLOC_ON(e, up, d->sort_sz, 4 * (size_t)ng * sizeof(int64_t));      // asz | bsz | aoff | boff
        DEV_ALLOC_ON(e, up, b->d_a, b->a_bytes + 32 * 16);        // one block of slack: K1 prefetches a row block ahead
        DEV_ALLOC_ON(e, up, b->d_b, b->b_bytes);
        DEV_ALLOC_ON(e, up, b->d_groups, ng * sizeof(K1Group));
        DEV_ALLOC_ON(e, up, b->d_la, d->n_k1 * sizeof(int32_t)); DEV_ALLOC_ON(e, up, b->d_lb, d->n_k1 * sizeof(int32_t));
        if (rdb) { b->d_res = rdb->d_res; d->borrow_res = true; } else DEV_ALLOC_ON(e, up, b->d_res, n * sizeof(K1Result));
        CUDA_TRY(e, join());
        if (n_k1 < n) CUDA_TRY(e, zero_async(b->d_res, n * sizeof(K1Result), hs));      // pairs K1 does not take keep flags == 0
        DEV_ALLOC_ON(e, up, b->d_counter, sizeof(int32_t));
        DEV_ALLOC_ON(e, up, d->order, d->n_k1 * sizeof(int32_t));
        b->d_order = d->order;
        int32_t *cnt = d->sort_cnt, *cur = d->sort_cnt + n_buckets;
        int64_t *asz = d->sort_sz, *bsz = asz + ng, *aoff = bsz + ng, *boff = aoff + ng;
        const int sla = shared ? (int)shared_la : -1;
        const int nblk = (int)((n + HELPER_THREADS - 1) / HELPER_THREADS);
        CUDA_TRY(e, join());
        CUDA_TRY(e, zero_async(cnt, n_buckets * sizeof(int32_t), hs));
        sort_count_kernel<false><<<nblk, HELPER_THREADS, 0, hs>>>(d->off_a, d->off_b, n, sla, la_cap, la_gran, pass_cap, take, cnt, nullptr);
        scan_kernel<int32_t><<<1, HELPER_THREADS, 0, hs>>>(cnt, cur, n_buckets);
        sort_count_kernel<true><<<nblk, HELPER_THREADS, 0, hs>>>(d->off_a, d->off_b, n, sla, la_cap, la_gran, pass_cap, take, cur, d->order);
        sort_groups_kernel<<<(ng + HELPER_THREADS / 32 - 1) / (HELPER_THREADS / 32), HELPER_THREADS, 0, hs>>>(d->off_a, d->off_b, sla, d->order, d->n_k1, ng, b->d_groups, asz, bsz);
        scan_kernel<int64_t><<<1, HELPER_THREADS, 0, hs>>>(asz, aoff, ng);
        scan_kernel<int64_t><<<1, HELPER_THREADS, 0, hs>>>(bsz, boff, ng);
        k1_pack_kernel<<<ng, HELPER_THREADS, 0, hs>>>(d->codes_a, d->codes_b, d->off_a, d->off_b, d->order, b->d_groups,
                                                  b->d_a, b->d_b, b->d_la, b->d_lb, e->n, d->base_a, d->base_b, sla, aoff, boff, e->mode == SWB_GLOCAL ? 1 : 0);
        CUDA_TRY(e, cudaGetLastError());
        { int rc_ = ensure_bnd(e, k1_bnd_bytes(e->sm_count, b->la_max)); if (rc_ != SWB_OK) return rc_; }
    }
    // ---- K2 work items: subjects of >= 32 passes take a warp each (several sweeps), the others are bin-packed
    //      (first fit, longest first) into warps of 32 lanes, one lane per 32-column pass
    d->n_k2 = cached_k2 ? rdb->n_k2 : (int)k2_idx.size();
    if (cached_k2 && d->n_k2) {
        d->items = rdb->d_items; d->sweeps = rdb->d_sweeps; d->n_items = rdb->n_items; d->borrow_items = true;
        b->d_res = rdb->d_res; d->borrow_res = true;
        CUDA_TRY(e, join());
        if (d->n_k2 < n) CUDA_TRY(e, zero_async(b->d_res, n * sizeof(K1Result), hs));
        DEV_ALLOC_ON(e, up, b->d_counter, sizeof(int32_t));
        { int rc_ = ensure_bnd(e, k2_bnd_bytes(e->sm_count, (int)shared_la)); if (rc_ != SWB_OK) return rc_; }
    } else if (d->n_k2) {
        const size_t cap_items = k2_items_cap(b->lb, k2_idx);
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
        const int64_t ni_ = k2_build_items(b->lb, k2_idx, h_items, h_sweeps, cap_items);
        if (ni_ < 0) { e->err = "K2 item table overflow"; return SWB_ERR_NOMEM; }
        const size_t ni = (size_t)ni_;
        d->n_items = (int)ni;
        DEV_ALLOC_ON(e, up, d->items, ni * 32 * sizeof(int32_t)); DEV_ALLOC_ON(e, up, d->sweeps, ni * sizeof(int32_t));
        if (rdb) { b->d_res = rdb->d_res; d->borrow_res = true; } else DEV_ALLOC_ON(e, up, b->d_res, n * sizeof(K1Result));
        CUDA_TRY(e, join());
        if (d->n_k2 < n) CUDA_TRY(e, zero_async(b->d_res, n * sizeof(K1Result), hs));
        DEV_ALLOC_ON(e, up, b->d_counter, sizeof(int32_t));
        CUDA_TRY(e, cudaMemcpyAsync(d->items, h_items, ni * 32 * sizeof(int32_t), cudaMemcpyHostToDevice, up));
        CUDA_TRY(e, cudaMemcpyAsync(d->sweeps, h_sweeps, ni * sizeof(int32_t), cudaMemcpyHostToDevice, up));
        e->stats.h2d_bytes += (double)ni * 33 * sizeof(int32_t);
        if (!e->ev_items) CUDA_TRY(e, cudaEventCreateWithFlags(&e->ev_items, cudaEventDisableTiming));
        CUDA_TRY(e, cudaEventRecord(e->ev_items, up));
        { int rc_ = ensure_bnd(e, k2_bnd_bytes(e->sm_count, (int)shared_la)); if (rc_ != SWB_OK) return rc_; }
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
    if (rdb && !b->d_res) {                                    // KW-only routing (non-local modes, open < extend): results are collected into the database's array
        b->d_res = rdb->d_res; d->borrow_res = true;
        CUDA_TRY(e, join());
        CUDA_TRY(e, zero_async(b->d_res, n * sizeof(K1Result), hs));
    }
    CUDA_TRY(e, cudaEventCreate(&d->ev0)); CUDA_TRY(e, cudaEventCreate(&d->ev1));
    CUDA_TRY(e, cudaEventCreateWithFlags(&d->ev_done, cudaEventDisableTiming));
    // ---- results: device arrays in the caller's layout, the overflow list, pinned flags
    DEV_ALLOC_ON(e, up, d->d_sc, n * sizeof(float)); DEV_ALLOC_ON(e, up, d->d_ea, n * sizeof(int64_t)); DEV_ALLOC_ON(e, up, d->d_eb, n * sizeof(int64_t));
    DEV_ALLOC_ON(e, up, d->d_ovf, (n + 1) * sizeof(int32_t));
    if (d->n_kw) {
        DEV_ALLOC_ON(e, up, d->d_kw_idx, d->n_kw * sizeof(int32_t));
        CUDA_TRY(e, cudaMemcpyAsync(d->d_kw_idx, d->kw_idx.data(), d->n_kw * sizeof(int32_t), cudaMemcpyHostToDevice, up));
        e->stats.h2d_bytes += (double)d->n_kw * sizeof(int32_t);
    }
    CUDA_TRY(e, join());                                       // everything copied and allocated above is visible to `hs` from here
    CUDA_TRY(e, zero_async(d->d_ovf, sizeof(int32_t), hs));
    if (use_arena) d->h_flags = (int32_t *)stage_alloc(e, 2 * sizeof(int32_t));
    if (!d->h_flags) {
        d->own_stage = true;
        CUDA_TRY(e, cudaMallocHost((void **)&d->h_flags, 2 * sizeof(int32_t)));
    }
    d->use_arena = use_arena;
    d->h_flags[0] = 0; d->h_flags[1] = 0;
    CUDA_TRY(e, cudaMemcpyAsync(d->h_flags, d->err, sizeof(int32_t), cudaMemcpyDeviceToHost, hs));
    CUDA_TRY(e, cudaEventCreateWithFlags(&d->ev_ready, cudaEventDisableTiming));
    CUDA_TRY(e, cudaEventRecord(d->ev_ready, hs));
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
    int rc = batch_upload_impl(e, a_all, off_a, b_all, off_b, n, -1, false, e->stream, e->stream, &b);
    if (rc != SWB_OK) return rc;
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));           // the staged API reports letter errors here
    if (b->dev->h_flags[0]) {
        e->err = "unknown letter. please set X to allow scoring for undefined letter pair.";   // SA.rs:676
        batch_release(e, b);
        return SWB_ERR_LETTER;
    }
    *out = b;
    return SWB_OK;
}

int batch_launch(swb_engine *e, swb_batch *b, int64_t *launches)
{
    BatchDev *d = b->dev;
    cudaStream_t st = kstream(e, b);
    CUDA_TRY(e, cudaStreamWaitEvent(st, d->ev_ready, 0));     // upload may have run on the copy stream
    d->launched = true; d->done_recorded = false;
    if (d->n_k1) {
        CUDA_TRY(e, e->mode != SWB_LOCAL ? k1_global_launch(e, b, e->sm_count, bnd_slot(e, b), st) : k1_launch(e, b, e->sm_count, bnd_slot(e, b), st));
        ++*launches;
    }
    if (d->n_k2) {
        CUDA_TRY(e, k2_launch(e, d->codes_a, (int)d->shared_la, d->codes_b, d->off_b, d->base_b, d->items, d->sweeps, d->n_items,
                              bnd_slot(e, b), b->d_counter, b->d_res, e->sm_count, st));
        ++*launches;
    }
    if (d->n_kw) {
        CUDA_TRY(e, kw_launch(e, d->codes_a, d->codes_b, d->kw_pairs, d->n_kw, d->kw_plan_, nullptr, d->kw_bnd, d->kw_scr, d->kw_res, st));
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

// Results -> caller layout on the device, then D2H: straight into the caller's arrays when those are pinned, through
// pinned staging otherwise (an asynchronous copy to pageable memory would stall the host until the kernels are done).
// ev_done fires when everything has landed.
static int batch_fetch_enqueue(swb_engine *e, swb_batch *b, float *scores, int64_t *end_a, int64_t *end_b)
{
    BatchDev *d = b->dev;
    const int64_t n = b->n;
    cudaStream_t st = kstream(e, b);
    d->dst_sc = scores; d->dst_ea = end_a; d->dst_eb = end_b;
    const float inv = 1.0f / (float)e->scale;       // exact for dyadic scales
    const int dyadic = (e->scale & (e->scale - 1)) == 0;
    if (n == 0) { CUDA_TRY(e, cudaEventRecord(d->ev_done, st)); d->done_recorded = true; return SWB_OK; }
    CUDA_TRY(e, zero_async(d->d_ovf, sizeof(int32_t), st));
    if (d->n_k1 || d->n_k2)
        finalize_k1_kernel<<<(int)((n + HELPER_THREADS - 1) / HELPER_THREADS), HELPER_THREADS, 0, st>>>(b->d_res, n, e->scale, inv, dyadic, d->d_sc, d->d_ea, d->d_eb, d->d_ovf);
    if (d->n_kw)
        finalize_kw_kernel<<<(d->n_kw + HELPER_THREADS - 1) / HELPER_THREADS, HELPER_THREADS, 0, st>>>(d->kw_res, d->d_kw_idx, d->n_kw, e->scale, inv, dyadic, d->d_sc, d->d_ea, d->d_eb);
    CUDA_TRY(e, cudaGetLastError());
    d->direct = host_pinned(scores, n * sizeof(float)) && host_pinned(end_a, n * sizeof(int64_t)) && host_pinned(end_b, n * sizeof(int64_t));
    if (!d->direct && !d->h_sc) {
        if (d->use_arena) {
            d->h_sc = (float *)stage_alloc(e, n * sizeof(float)); d->h_ea = (int64_t *)stage_alloc(e, n * sizeof(int64_t));
            d->h_eb = (int64_t *)stage_alloc(e, n * sizeof(int64_t));
        }
        if (!d->h_sc || !d->h_ea || !d->h_eb) {       // no arena (stand-alone staged batch) or arena exhausted
            d->h_sc = nullptr; d->h_ea = nullptr; d->h_eb = nullptr;
            d->own_res = true;
            CUDA_TRY(e, cudaMallocHost((void **)&d->h_sc, n * sizeof(float)));
            CUDA_TRY(e, cudaMallocHost((void **)&d->h_ea, n * sizeof(int64_t)));
            CUDA_TRY(e, cudaMallocHost((void **)&d->h_eb, n * sizeof(int64_t)));
        }
    }
    CUDA_TRY(e, cudaMemcpyAsync(d->direct ? scores : d->h_sc, d->d_sc, n * sizeof(float), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaMemcpyAsync(d->direct ? end_a : d->h_ea, d->d_ea, n * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaMemcpyAsync(d->direct ? end_b : d->h_eb, d->d_eb, n * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaMemcpyAsync(d->h_flags + 1, d->d_ovf, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaEventRecord(d->ev_done, st));
    d->done_recorded = true;
    e->stats.d2h_bytes += (double)n * (sizeof(float) + 2 * sizeof(int64_t)) + 2 * sizeof(int32_t);
    return SWB_OK;
}

// Waits for ev_done; what is left for the host: the staging copy for pageable callers, pairs with an empty sequence
// (closed form) and the int32 re-run of s16 overflows.
static int batch_fetch_finish(swb_engine *e, swb_batch *b)
{
    BatchDev *d = b->dev;
    float *scores = d->dst_sc; int64_t *end_a = d->dst_ea, *end_b = d->dst_eb;
    CUDA_TRY(e, cudaEventSynchronize(d->ev_done));
    if (d->h_flags[0]) {
        e->err = "unknown letter. please set X to allow scoring for undefined letter pair.";   // SA.rs:676
        return SWB_ERR_LETTER;
    }
    const float inv = 1.0f / (float)e->scale;       // exact for dyadic scales
    const bool dyadic = (e->scale & (e->scale - 1)) == 0;
    auto to_f = [&](int32_t s) { return dyadic ? (float)s * inv : (float)((double)s / e->scale); };
    if (!d->direct && b->n) {
        memcpy(scores, d->h_sc, b->n * sizeof(float)); memcpy(end_a, d->h_ea, b->n * sizeof(int64_t)); memcpy(end_b, d->h_eb, b->n * sizeof(int64_t));
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
    const int nr = d->h_flags[1];
    if (nr > 0) {                                     // int32 re-run of the overflowed pairs (KW, no direction bytes)
        std::vector<int32_t> redo(nr);
        std::vector<KWPair> hp(nr);
        std::vector<int64_t> oa(b->n + 1), ob(b->n + 1);
        CUDA_TRY(e, cudaMemcpyAsync(redo.data(), d->d_ovf + 1, nr * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
        if (d->shared_la < 0) CUDA_TRY(e, cudaMemcpyAsync(oa.data(), d->off_a, (b->n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaMemcpyAsync(ob.data(), d->off_b, (b->n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));
        std::sort(redo.begin(), redo.end());
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
        cudaFreeAsync(dp, e->stream); cudaFreeAsync(dr, e->stream); cudaFreeAsync(db, e->stream); cudaFreeAsync(ds, e->stream);
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
    int rc = batch_fetch_enqueue(e, b, scores, end_a, end_b);
    if (rc != SWB_OK) return rc;
    return batch_fetch_finish(e, b);
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
    int64_t chunk_pairs = SCORE_CHUNK_PAIRS, chunk_first = SCORE_CHUNK_FIRST;
    if (const char *v = getenv("SWB_CHUNK_PAIRS")) chunk_pairs = std::max<int64_t>(1024, atoll(v));      // tuning knobs
    if (const char *v = getenv("SWB_CHUNK_FIRST")) chunk_first = std::max<int64_t>(1024, atoll(v));
    if (n > 0) {
        const double nb_tot = (double)(off_b[n] - off_b[0]), na_tot = shared_la >= 0 ? (double)shared_la : (double)(off_a[n] - off_a[0]);
        const double est_cells = shared_la >= 0 ? na_tot * nb_tot : (na_tot / n) * (nb_tot / n) * n;
        const double copy_ms = (na_tot + nb_tot) / 50e6, dp_ms = est_cells / 4e9;
        if (copy_ms < 0.05 * dp_ms) chunk_pairs = n;
    }
    // graded chunks: a small first chunk gets K1 going after ~1 ms of copy; each following chunk is as much larger as the
    // copy engine is faster than K1 (x1.8 on configs[1]), so its upload hides behind the previous chunk's DP
    std::vector<int64_t> cuts(1, 0);
    for (int64_t sz = (chunk_pairs == n) ? n : chunk_first; cuts.back() < n; sz = std::min<int64_t>(chunk_pairs, sz * 9 / 5)) {
        int64_t next = std::min<int64_t>(n, cuts.back() + sz);
        if (n - next < chunk_first) next = n;                // no crumbs at the end
        cuts.push_back(next);
    }
    // a tapered tail (tuning knob, off by default): when the copy engine is the bottleneck (eight ranks behind one host) the DP
    // of the last chunk runs after everything else has finished, so SWB_CHUNK_LAST cuts the last chunk down.  On one GPU,
    // where the DP is the bottleneck, the extra launch costs its own floor: measured 24.7 -> 26.4 ms (65536) / 27.0 ms (32768)
    int64_t chunk_last = 0;
    if (const char *v = getenv("SWB_CHUNK_LAST")) chunk_last = std::max<int64_t>(0, atoll(v));
    if (chunk_last >= 1024 && cuts.size() >= 3) {
        const int64_t beg = cuts[cuts.size() - 2], len = n - beg;
        if (len > 2 * chunk_last) cuts.insert(cuts.end() - 1, n - chunk_last);
    }
    if (cuts.size() == 1) cuts.push_back(0);
    const int64_t n_chunks = (int64_t)cuts.size() - 1;
    const size_t need = (size_t)n * (sizeof(float) + 2 * sizeof(int64_t)) + (size_t)n_chunks * 8192;    // staging for pageable callers
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
                               e->stream_up, e->stream_up, &cur);
        double t1 = now(); t_up += t1 - t0;
        if (trace && cur && cur->dev->ev_trace) { tev.push_back(cur->dev->ev_trace); cur->dev->ev_trace = nullptr; } else if (trace) mark(e->stream_up);
        mark(e->stream_up);
        int64_t launches = 0;
        if (rc == SWB_OK) { cur->slot = (int)(c & 1); rc = batch_launch(e, cur, &launches); }   // alternate the compute streams
        mark(rc == SWB_OK ? kstream(e, cur) : e->stream);
        if (rc == SWB_OK) rc = batch_fetch_enqueue(e, cur, scores + start, end_a + start, end_b + start);
        e->stats.launches += launches;
        double t2 = now(); t_launch += t2 - t1;
        live[c] = cur;
        // scatter the chunks whose results have already landed; never block here: the next chunk's sort and upload come first
        while (rc == SWB_OK && next_finish < c && cudaEventQuery(live[next_finish]->dev->ev_done) == cudaSuccess) {
            double t3 = now();
            rc = batch_fetch_finish(e, live[next_finish]);
            double t4 = now(); t_fin += t4 - t3;
            batch_release(e, live[next_finish]); live[next_finish] = nullptr;
            t_rel += now() - t4;
            ++next_finish;
        }
    }
    for (int64_t c = next_finish; c < n_chunks; ++c) {
        if (!live[c]) continue;
        double t3 = now();
        if (rc == SWB_OK) rc = batch_fetch_finish(e, live[c]);
        t_fin += now() - t3;
        batch_release(e, live[c]); live[c] = nullptr;
    }
    if (rc != SWB_OK) { cudaStreamSynchronize(e->stream_up); cudaStreamSynchronize(e->stream); cudaStreamSynchronize(e->stream2); }   // nothing of this call stays in flight
    if (rc == SWB_OK) {
        CUDA_TRY(e, cudaEventRecord(e->ev1, e->stream));
        CUDA_TRY(e, cudaEventSynchronize(e->ev1));
        float ms = 0; cudaEventElapsedTime(&ms, e->ev0, e->ev1);
        e->stats.kernel_ms = ms;                      // device time of the whole pipelined call
        if (trace) fprintf(stderr, "[swb] score_batch n=%lld chunks=%lld total %.2f ms: upload(host) %.2f launch %.2f finish(wait+scatter) %.2f release %.2f device-span %.2f\n",
                           (long long)n, (long long)n_chunks, now() - t_begin, t_up, t_launch, t_fin, t_rel, ms);
        for (size_t k = 0; k + 3 < tev.size() + 0 && trace; k += 4) {
            float a = 0, h = 0, b2 = 0, c2 = 0;
            cudaEventElapsedTime(&a, e->ev0, tev[k]); cudaEventElapsedTime(&h, e->ev0, tev[k + 1]); cudaEventElapsedTime(&b2, e->ev0, tev[k + 2]);
            cudaEventElapsedTime(&c2, e->ev0, tev[k + 3]);
            fprintf(stderr, "[swb]   chunk %zu: upload begins %.2f  H2D done %.2f  packed %.2f  K1 done %.2f ms\n", k / 4, a, h, b2, c2);
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
