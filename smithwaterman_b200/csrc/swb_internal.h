// Internal declarations shared by the host layer and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <climits>
#include <string>
#include <vector>

#include "../../include/swb200.h"

#define SWB_MAX_LETTERS 31          // residue codes 0..n-1, PAD code = n  (table stride n+1 <= 32)

// ---- K1 (inter-sequence batch, packed s16x2) geometry ---------------------------------------
#define K1_STRIP 32                 // DP columns per pass of one thread: 16 in the low s16 lane, 16 in the high
#define K1_HALF 16
#define K1_ROWCHUNK 16              // row residues fetched per 128-bit load
#define K1_THREADS 256
#define K1_BND_SLACK 16            // row steps run in blocks of 4: up to 3 steps past la_max+1 touch the boundary buffer

struct K1Group {                    // one warp-sized work item: 32 pairs of similar shape
    int64_t a_off;                  // byte offset of the group's row-residue block   [4-row block][lane] u32
    int64_t b_off;                  // byte offset of the group's column-residue block [pass][lane][32]
    int32_t la_max;                 // longest a in the group (row steps = la_max + 1)
    int32_t n_pass;                 // ceil(max lb / 32)
    int32_t first_pair;             // index of lane 0's pair in the sorted order
    int32_t n_pairs;                // 1..32
};

struct K1Result {                   // per pair (caller order)
    int32_t score;                  // scaled integer score
    int32_t end_i, end_j;           // 1-based first-maximum cell, -1,-1 if score == 0
    int32_t flags;                  // bit0: s16 range exceeded -> re-run in int32
};

struct swb_engine {
    int device = 0;
    int mode = SWB_LOCAL;
    int n = 0;                      // letters
    char letters[SWB_MAX_LETTERS + 1] = {0};
    int16_t code_of[256];           // byte -> residue code, -1 = unknown and no X
    float go = 0, ge = 0;           // as given (sign-normalised negative, SA.rs:28-35)
    int32_t scale = 1, exact_scale = 1;
    int32_t o = 0, e = 0;           // scaled, <= 0
    int32_t smax = 0, smin = 0;     // scaled matrix extremes
    std::vector<int32_t> mat;       // n*n scaled
    std::vector<float> matf;        // n*n as given
    cudaStream_t stream = nullptr;      // compute stream (kernels, result D2H)
    cudaStream_t stream2 = nullptr;     // second compute stream: odd chunks of the pipelined host-buffer path
    cudaStream_t stream_up = nullptr;   // upload stream of the pipelined host-buffer path (allocations, H2D, encode, sort, pack)
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev_team = nullptr;      // completion of the last team (cooperative) wavefront launch: team launches are chained through it
    int sm_count = 0;
    int32_t *d_mat = nullptr;       // (n+1)*(n+1) int32 scaled, PAD row/col included
    int16_t *d_code_tab = nullptr;  // byte -> residue code (device copy of code_of)
    // scratch owned by the engine, grown on demand
    void *d_bnd = nullptr; size_t bnd_bytes = 0;                        // two slots of bnd_bytes each (batch.cu ensure_bnd)
    void *d_dir = nullptr; size_t dir_bytes = 0;                        // direction bytes of the traceback path (align.cu)
    void *h_tb = nullptr; size_t tb_bytes = 0;                          // pinned arena of the traceback results (align.cu)
    void *h_items = nullptr; size_t items_bytes = 0; cudaEvent_t ev_items = nullptr;   // pinned K2 work-item table
    void *h_stage = nullptr; size_t stage_bytes = 0, stage_used = 0;   // pinned results arena of swb_score_batch
    swb_stats stats = {};
    std::string err;
};

struct BatchDev;
struct swb_batch {
    int64_t n = 0;
    int slot = 0;                   // compute stream / boundary-scratch slot this batch runs on
    BatchDev *dev = nullptr;        // device-side part owned by batch.cu
    std::vector<int32_t> la, lb;    // per caller index
    // device (K1 work items are sorted and built on the device, batch.cu)
    uint8_t *d_a = nullptr, *d_b = nullptr;
    K1Group *d_groups = nullptr;
    int32_t n_groups = 0;
    int32_t *d_la = nullptr, *d_lb = nullptr;   // sorted order
    K1Result *d_res = nullptr;                  // indexed by the caller's pair index within the batch (K1 scatters)
    const int32_t *d_order = nullptr;           // sorted position -> caller index
    int32_t *d_counter = nullptr;
    size_t a_bytes = 0, b_bytes = 0;            // upper bounds of the packed residue blocks
    int32_t la_max = 0;
    float last_ms = 0; int64_t last_launches = 0;
    double cells = 0;
};

// A database (shard) resident on the device: encoded once, searched many times (search.cu; swb_db_upload / swb_search).
struct swb_db {
    int device = 0;
    int64_t n = 0, n_res = 0;
    std::vector<char> h_letters;            // the letters as given (host copy: the strings of the hits are spelled from it)
    std::vector<int64_t> h_off;             // n + 1 offsets into h_letters / d_codes
    std::vector<int64_t> h_gidx;            // caller (global) index of every subject, ascending
    std::vector<int32_t> lb;                // subject lengths
    std::vector<int32_t> degenerate;        // subjects of length 0
    uint8_t *d_codes = nullptr;             // encoded residues (PAD-filled slack of 32 bytes)
    int64_t *d_off = nullptr;
    int64_t *d_gidx = nullptr;
    // K2 work items (first-fit packing of the subjects' 32-column passes into warps), built once per database
    int32_t *d_items = nullptr, *d_sweeps = nullptr;
    int n_items = 0, n_k2 = 0;
    // per-search scratch, kept between searches
    K1Result *d_res = nullptr;              // score / end cell / flags per subject (local index)
    uint8_t *d_query = nullptr; int64_t query_cap = 0;
    std::vector<uint8_t> h_query;           // the codes d_query holds (swb_search_align uploads a query once)
    void *d_sel = nullptr;                  // top-K selection state + keys
    void *h_hits = nullptr;                 // pinned: the hits of the last search + counters
    int64_t last_lq = -1;                   // query length of the last search (its scores are in d_res)
};

#define CUDA_TRY(e_, call)                                                                      \
    do {                                                                                        \
        cudaError_t err__ = (call);                                                             \
        if (err__ != cudaSuccess) {                                                             \
            (e_)->err = std::string(#call) + ": " + cudaGetErrorString(err__);                  \
            return SWB_ERR_CUDA;                                                                \
        }                                                                                       \
    } while (0)

// ---- kernel launchers (defined in the .cu files) ----------------------------------------------
// K1: local, score + first-maximum cell, packed s16x2, one thread per pair.
cudaError_t k1_launch(const swb_engine *e, const swb_batch *b, int n_blocks, void *d_bnd, cudaStream_t st);
cudaError_t k1_global_launch(const swb_engine *e, const swb_batch *b, int n_blocks, void *d_bnd, cudaStream_t st);   // k1_global.cu
size_t k1_bnd_bytes(int n_blocks, int la_max);
int k1_blocks_per_sm(const swb_engine *e);
// K2: shared long query, lanes split the columns (database search).
size_t k2_bnd_bytes(int n_blocks, int la);
cudaError_t k2_launch(const swb_engine *e, const uint8_t *d_query, int la, const uint8_t *d_codes_b, const int64_t *d_off_b,
                      int64_t base_b, const int32_t *d_item_pairs, const int32_t *d_item_sweeps, int n_items, void *d_gbnd,
                      int *d_counter, K1Result *d_res, int n_blocks, cudaStream_t st);

// KW: one warp per pair, exact 3-state int32 recurrence (all modes), optional direction bytes.
#define KW_WARPS 8                  // warps per block of kw_kernel (a plan's `blocks` = ceil(team warps / KW_WARPS))
struct KWPair {
    int64_t a_off, b_off;           // offsets into the code arrays
    int32_t la, lb;
    int64_t dir_off;                // offset of this pair's direction bytes (skewed layout), -1 = none
    int64_t bnd_off;                // offset (in int4 units) of the strip-boundary ring   (filled by kw_plan)
    int64_t scr_off;                // offset (in ints) of the pair's team scratch          (filled by kw_plan)
    int64_t pssm_off;               // offset (in ints) of the pair's position-specific scores [b][a], -1 = use the matrix
};
struct KWPlan {
    int cols = 8;                   // DP columns per lane (strip width 32*cols): 8, or 2 for a few long pairs
    int team_warps = 1;             // warps that sweep one pair together
    int n_teams = 0, blocks = 0;
    int64_t bnd_elems = 0;          // int4 elements of the boundary rings of all pairs
    int64_t scratch_ints = 0;       // progress counters + candidates of all pairs
    int64_t dir_bytes = 0;          // direction bytes of all pairs (with_dirs)
};
void kw_plan(const swb_engine *e, KWPair *pairs, int n_pairs, bool with_dirs, KWPlan *plan, int force_cols = 0, bool band = false);   // also assigns dir_off
int kw_blocks_per_sm(const swb_engine *e, int cols, bool dirs, bool band);   // resident blocks per SM of the exact instantiation (runtime query, per device)
struct KWResult {
    int32_t score;                  // scaled
    int32_t end_i, end_j;           // start cell of the traceback (1-based), -1 = empty
    int32_t state;                  // state at the start cell (0 M, 1 X, 2 Y)
};
cudaError_t kw_launch(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pairs,
                      int n_pairs, const KWPlan &plan, uint8_t *d_dir, int4 *d_bnd, int *d_scratch, KWResult *d_res,
                      cudaStream_t st, const int32_t *d_pssm = nullptr);
// Checkpointed traceback (pairs whose direction bytes do not fit): the same kernel run on ONE pair as a ROW BAND.
// The pair handed to the kernel is the sub-matrix rows (row0, row0+la] x columns [1, lb] of a la_full x lb_full matrix;
// its top row (row0) comes from a checkpoint row instead of the boundary, and/or it writes checkpoint rows itself.
struct KWBand {
    int32_t row0 = 0;               // rows above the band (a_off of the pair already points at row row0+1)
    int32_t la_full = 0, lb_full = 0;
    const int4 *ckpt_in = nullptr;  // (M, X, Y) of row row0 at [j], j = 1..lb; nullptr = the matrix boundary (row0 == 0)
    int4 *ckpt_out = nullptr;       // rows every, 2*every, ... (< la_full): row k*every at [(k-1)*stride + j]
    int32_t every = 0;
    int64_t stride = 0;
};
cudaError_t kw_launch_band(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const KWPair *d_pair,
                           const KWPlan &plan, const KWBand &band, uint8_t *d_dir, int4 *d_bnd, int *d_scratch,
                           KWResult *d_res, cudaStream_t st);
// KT on a band: resumes the walk at the cursor, stops at the band's top row (status 1), at the alignment's start
// (status 0) or at one of the reference's panic cases (status < 0); appends to the reversed column lists at [n].
struct KTCursor { int32_t cx, cy, st, n, status; };
cudaError_t kt_band_launch(const swb_engine *e, const uint8_t *d_dir, int cols, int32_t la_band, int32_t row0,
                           KTCursor *d_cur, int32_t *d_idx_a, int32_t *d_idx_b, cudaStream_t st);
int64_t kw_dir_bytes(int32_t la, int32_t lb, int cols);   // bytes of the skewed direction store of one pair
// KT: pointer walk over the direction bytes; one thread per pair.  Writes reversed column lists.
cudaError_t kt_launch(const swb_engine *e, const KWPair *d_pairs, const KWResult *d_res, int n_pairs,
                      const uint8_t *d_dir, int cols, const int64_t *d_out_off, int32_t *d_idx_a, int32_t *d_idx_b,
                      int32_t *d_ncols, cudaStream_t st);

// KT, counting form: the same walk, but instead of the column lists it returns what the clustering loop needs.
struct KTStats {
    int32_t ncols;                  // columns of the core alignment; < 0 = the reference's panic cases
    int32_t na, nb;                 // letters of a / b in the core
    int32_t match;                  // columns with two equal letters (raw bytes)
    int32_t s1, e1, s2, e2;         // first / last residue index of a and of b in the core, -1 = none
};
cudaError_t kt_stats_launch(const swb_engine *e, const KWPair *d_pairs, const KWResult *d_res, int n_pairs,
                            const uint8_t *d_dir, int cols, const uint8_t *d_raw_a, const uint8_t *d_raw_b,
                            KTStats *d_stats, cudaStream_t st);

// The statistics of the strings align(.., retain_all) would return, from the core alignment's counts (align.cu).
void swb_compose_stats(swb_aln_stats *o, int mode, bool retain_all, int64_t la, int64_t lb, const KTStats &S);

// ---- batch path internals shared with the search driver (batch.cu) ------------------------------------------------
// Upload / prepare one batch.  rdb != nullptr: the `b` side is the resident database (nothing is copied or encoded, the
// cached routing and K2 work items are used); a_all then is the query (shared_la >= 0).
int batch_upload_impl(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all, const int64_t *off_b,
                      int64_t n, int64_t shared_la, bool use_arena, cudaStream_t up, cudaStream_t hs, swb_batch **out,
                      const swb_db *rdb = nullptr);
int batch_launch(swb_engine *e, swb_batch *b, int64_t *launches);
void batch_release(swb_engine *e, swb_batch *b);
// K2 work items of a set of subjects (host, first fit, longest first); returns the number of items, fills items[32*cap], sweeps[cap]
size_t k2_items_cap(const std::vector<int32_t> &lb, const std::vector<int32_t> &idx);
int64_t k2_build_items(const std::vector<int32_t> &lb, const std::vector<int32_t> &idx, int32_t *h_items, int32_t *h_sweeps, size_t cap_items);
// unified per-subject results of a launched batch (K1 / K2 write b->d_res directly; this adds the KW pairs)
cudaError_t batch_collect_kw(swb_engine *e, swb_batch *b, cudaStream_t st);
int batch_device_views(swb_batch *b, const uint8_t **codes_a, const uint8_t **codes_b);

cudaError_t batch_encode(swb_engine *e, uint8_t *d_buf, int64_t nbytes, int32_t *d_err, cudaStream_t st);   // letters -> codes in place (buffer padded to 16 B)

// ---- traceback core shared by swb_align_batch and the search driver (align.cu) -----------------------------------------
struct AlnJob {
    int64_t a_off, b_off;           // offsets into the device code arrays (and the position-specific table when given)
    int32_t la, lb;                 // DP extents handed to the kernels
    const char *ra, *rb;            // letters as given
    int64_t la_full, lb_full;       // lengths of the whole sequences (retain_all padding, SA.rs:512-549)
    int64_t out;                    // index into outs / souts
    int32_t known_score = INT32_MIN;// local mode: the scan's scaled score of the pair whose end cell is (la, lb) -- allows the packed path
};
int align_jobs_device(swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const std::vector<AlnJob> &jobs, bool retain_all,
                      swb_alignment *outs, swb_aln_stats *souts, const int32_t *d_pssm, const uint8_t *d_ra, const uint8_t *d_rb);
bool swb_range_ok(const swb_engine *e, int64_t la, int64_t lb, int64_t smax_abs = -1);   // do the scores of a pair fit the int32 lanes of KW?

// ---- K3 / KT3: packed fill with tile borders + tile-recompute traceback (k3_tile.cu) ------------------------------------
struct K3Pair {
    int64_t a_off, b_off;           // offsets into the device code arrays
    int32_t la, lb;                 // DP extents (local mode: the prefix box that ends at the end cell)
    int64_t colb_off;               // uint4 units: last column of every pass, [pass][row block + 1 slack][2]
    int64_t rowb_off;               // 32-bit words: row checkpoints, [checkpoint][pass][80]
    int64_t out_off;                // offset of the pair's reversed column lists
    int32_t start_state;            // local: 0 (MATCH); global: taken from the recomputed cell (la, lb)
    int32_t score;                  // local: the scan's score (scaled); global: filled in by the walker
    int32_t swap;                   // 1: the kernels run the TRANSPOSED problem (rows = b, columns = a; la / lb above are the transposed
                                    //    extents): a tall box becomes a wide one, whose sweeps pipeline over a team of warps
    int32_t find_col;               // wide local form: 1 = cy0 is the LAST column of the pass that holds the start cell, the walker finds the column
    int32_t cx0, cy0;               // wide (int32) form: the cell the walk starts at, inside a border store laid out for la x lb
};                                  //    (0, 0 = (la, lb); -1 = local score 0, nothing to walk); written by k3w_pick_kernel in local mode
// K3W: the int32 form of K3 for ONE long pair whose scores leave the s16 range (100 kb DNA pairs): sweeps spread over the
// whole GPU (cooperative launch, progress words in global memory), local mode finds its end cell itself.
void k3w_sizes(int32_t la, int32_t lb, int64_t *colb_uint4, int64_t *rowb_words, int64_t *scratch_ints);
cudaError_t k3w_launch(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, K3Pair *d_pair, const K3Pair &h_pair,
                       uint4 *d_colb, unsigned *d_rowb, int *d_scratch, int32_t *d_idx_a, int32_t *d_idx_b, int32_t *d_ncols,
                       KWResult *d_res, cudaStream_t st);
bool k3_eligible(const swb_engine *e, int64_t la, int64_t lb);       // mode, |open| >= |extend| and the static s16 bound (global)
void k3_sizes(int32_t la, int32_t lb, int64_t *colb_uint4, int64_t *rowb_words);
cudaError_t k3_launch(const swb_engine *e, const uint8_t *d_a, const uint8_t *d_b, const K3Pair *d_pairs, int n_pairs,
                      uint4 *d_colb, unsigned *d_rowb, int *d_counter, int32_t *d_idx_a, int32_t *d_idx_b, int32_t *d_ncols,
                      KWResult *d_res, cudaStream_t st, int max_sweeps = 1);

// K0
cudaError_t k0_run(int which, int sm_count, double *thread_instr_per_s);
