/*
 * swb200.h -- C ABI of the B200-native Smith-Waterman / Needleman-Wunsch engine.
 *
 * This is the drop-in boundary for the alignment path of yamule/smithwaterman.  The reference has
 * no FFI layer; the entry points below are what a Rust `-sys` crate (or cgo/ctypes) would bind to
 * replace the following reference interfaces ("SA.rs" = rust/sequence_alignment/src/
 * sequence_alignment.rs, "OCL.rs" = rust/sa_opencl/src/opencl_sequence_alignment.rs,
 * "OCL-main.rs" = rust/sa_opencl/src/main.rs):
 *
 *   swb_engine_create      SequenceAlignment::new            SA.rs:26-45   (OCL.rs:103 new)
 *   swb_engine_destroy     Drop of SequenceAlignment / OpenCLSequenceAlignment
 *   swb_align              SequenceAlignment::align          SA.rs:389-392 (OCL.rs:796 align)
 *   swb_align_pssm         align over a PositionSpecificMatrix  SA.rs:583-623 (ScoringMatrix trait SA.rs:574-580)
 *   swb_align_batch        the serial loop of `sa_opencl -list`  OCL-main.rs:295-316 (one call, n pairs)
 *   swb_align_stats_batch  align + the identity / coverage counts of `sa_opencl -cluster`  OCL-main.rs:187-243
 *   swb_cluster            the clustering loop of `sa_opencl -cluster`  OCL-main.rs:149-243
 *   swb_score_batch        align_partial(.., score_only=true) per pair  SA.rs:393-467
 *   swb_search_scores      the all-vs-all loops of the CLIs with one fixed query (SA-main.rs:35-45), n subjects per call
 *   swb_db_* / swb_search* the same loop as a database search: the subjects stay resident on the device (encoded once),
 *                          a call scans one query, selects the top-K on the device and traces the hits back
 *                          (SA-main.rs:35-45 / OCL-main.rs:295-316 with one fixed `ss1`; align() SA.rs:389 per hit)
 *   swb_align_partial      SequenceAlignment::align_partial  SA.rs:393-551 (partial_region, score_only)
 *   swb_batch_*            the same as swb_score_batch, split into upload / run / fetch so that a
 *                          caller can keep inputs resident in HBM (bench.py `value` vs `e2e`)
 *   swb_alignment_free     Drop of the (Vec<String>,Vec<String>,f32) result
 *   swb_last_error         replaces the reference's panics (SA.rs:370,636,676; OCL.rs:38)
 *
 * Conventions: plain pointers and sizes, no C++/torch types.  Sequences are byte strings, one
 * residue per byte (the reference holds one `String` per residue, SA.rs:799-803; every letter its
 * matrices know is one ASCII byte).  Unknown letters map to the matrix's `X` row, or fail with
 * SWB_ERR_LETTER when the matrix has no `X` (SA.rs:669-679).  Penalties are sign-insensitive
 * (SA.rs:28-35).  An engine is NOT thread-safe (the reference takes `&mut self`); distinct
 * engines may be used concurrently.  There is no CPU fallback: every entry point that computes
 * needs a CUDA device and fails with SWB_ERR_CUDA otherwise.
 */
#ifndef SWB200_H
#define SWB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* alignment_type, SA.rs:11-13 */
#define SWB_GLOBAL 0
#define SWB_GLOCAL 1
#define SWB_LOCAL 2

/* status codes (0 = ok) */
#define SWB_OK 0
#define SWB_ERR_ARG 1      /* bad argument (null pointer, bad mode, too many letters, ...) */
#define SWB_ERR_SCALE 2    /* matrix / penalties are not on a common integer grid (<= 1/1024 or 1/1000) */
#define SWB_ERR_LETTER 3   /* unknown residue and the matrix has no X row (SA.rs:676) */
#define SWB_ERR_CUDA 4     /* CUDA runtime error or no device */
#define SWB_ERR_EMPTY 5    /* both sequences empty in a non-local mode (reference panics, SA.rs:360-363) */
#define SWB_ERR_NOMEM 6
#define SWB_ERR_RANGE 7    /* scores do not fit the int32 lanes */

/* flags of the batch calls */
#define SWB_RETAIN_ALL 1u  /* retain_all of SA.rs:389 (ignored = forced for non-local modes, SA.rs:506) */

typedef struct swb_engine swb_engine;
typedef struct swb_batch swb_batch;

/* Substitution matrix, SubstitutionMatrix of SA.rs:626-631: n letters, scores[a*n+b] (f32). */
typedef struct swb_matrix {
    int32_t n_letters;
    const char *letters;     /* n_letters bytes, index = position (SA.rs:627-628) */
    const float *scores;     /* n_letters*n_letters, row-major [a][b] (SA.rs:629) */
} swb_matrix;

/* Result of one alignment, the (Vec<i64>,Vec<i64>,f32) of backtrack SA.rs:273 plus the strings of
 * align_partial SA.rs:470-550.  Owned by the library; release with swb_alignment_free. */
typedef struct swb_alignment {
    float score;             /* SA.rs:386 */
    int64_t end_a, end_b;    /* 1-based DP cell the traceback starts from; -1,-1 = empty local alignment */
    int64_t n_cols;          /* columns of the core alignment (length of idx_a / idx_b) */
    int64_t *idx_a;          /* residue index per column, -1 = gap (SA.rs:359-373) */
    int64_t *idx_b;
    int64_t aln_len;         /* strlen(aln_a) == strlen(aln_b) */
    char *aln_a;             /* padded strings per retain_all (SA.rs:502-550), NUL-terminated */
    char *aln_b;
} swb_alignment;

/* Counters of the last compute call on an engine. */
typedef struct swb_stats {
    double cells;            /* sum la*lb of the requested work */
    double kernel_ms;        /* device time of the DP kernels (CUDA events on the engine's stream) */
    double h2d_bytes, d2h_bytes;
    int64_t launches;        /* kernels launched */
    int64_t i32_reruns;      /* pairs re-run in int32 lanes after s16 overflow */
    int32_t scale;           /* integer rescaling factor in use (10/0.5 -> 2) */
    int32_t exact_scale;     /* 1 = every input was exactly representable on the integer grid */
    int64_t ckpt_pairs;      /* pairs traced back through checkpointed recompute (direction bytes over the budget) */
    int64_t k3_pairs;        /* pairs traced back on the packed path (s16x2 fill with tile borders + tile-recompute walk) */
} swb_stats;

/* Built-in matrices (SA.rs:697-733 BLOSUM62, 24 letters; SA.rs:681-695 match/mismatch, A-Z).
 * `scores` must hold n*n floats and `letters` n bytes; returns n (24 / 26). */
int32_t swb_matrix_blosum62(char *letters, float *scores);
int32_t swb_matrix_match_mismatch(float match, float mismatch, char *letters, float *scores);

int swb_engine_create(const swb_matrix *m, float gap_open, float gap_extend, int mode, int device,
                      swb_engine **out);
void swb_engine_destroy(swb_engine *e);
const char *swb_last_error(const swb_engine *e);   /* e may be NULL: last create error */
int swb_get_stats(const swb_engine *e, swb_stats *out);
int swb_device_count(void);

/* One pair with traceback (SA.rs:389).  out is filled on SWB_OK. */
int swb_align(swb_engine *e, const char *a, int64_t la, const char *b, int64_t lb, int retain_all,
              swb_alignment *out);
/* align_partial (SA.rs:393-551): the DP runs on the prefixes a[0..pa), b[0..pb) (partial_region; pa = la, pb = lb for
 * None) while the retain_all padding spells the WHOLE sequences (SA.rs:512-549 loop over s1.seq.len()); score_only != 0
 * returns the score and end cell with empty strings (SA.rs:438-467). */
int swb_align_partial(swb_engine *e, const char *a, int64_t la, const char *b, int64_t lb, int64_t pa, int64_t pb,
                      int retain_all, int score_only, swb_alignment *out);
/* One pair scored by a PositionSpecificMatrix (SA.rs:583-623; psm_test SA.rs:1057-1086): the score of (position i of a,
 * position j of b) is scores[i + j*la] (f32, la*lb values on the engine's penalty grid); the engine's matrix is not used,
 * the letters only spell the output strings.  Gap penalties and mode are the engine's. */
int swb_align_pssm(swb_engine *e, const float *scores, const char *a, int64_t la, const char *b, int64_t lb,
                   int retain_all, swb_alignment *out);
/* n pairs with traceback; sequences k = a_all[off_a[k]..off_a[k+1]), same for b; outs[n]. */
int swb_align_batch(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                    const int64_t *off_b, int64_t n, unsigned flags, swb_alignment *outs);
void swb_alignment_free(swb_alignment *a);

/* What the clustering loop of `sa_opencl -cluster` reads off an alignment (OCL-main.rs:205-233: the non-gap letters of
 * each aligned string, the columns with two equal letters, the string length), without building the strings: the
 * traceback walk counts on the device and 48 bytes per pair come back.  Values are those of the strings
 * swb_align_batch would return for the same `flags`. */
typedef struct swb_aln_stats {
    float score;
    int32_t status;          /* 0, or SWB_ERR_EMPTY where the reference panics on this pair */
    int64_t end_a, end_b;    /* as swb_alignment */
    int64_t aln_len;         /* strlen(aln_a) */
    int64_t a_letters;       /* characters of aln_a that are not '-' */
    int64_t b_letters;
    int64_t matches;         /* columns k with aln_a[k] == aln_b[k], neither a gap (compared as given, not as encoded) */
} swb_aln_stats;
int swb_align_stats_batch(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                          const int64_t *off_b, int64_t n, unsigned flags, swb_aln_stats *outs);

/* The clustering loop of `sa_opencl -cluster` itself (OCL-main.rs:149-243): sequences sorted longest first (stable),
 * exact duplicates folded, then every still-unclustered sequence becomes a representative and takes the later ones that
 * it contains (length ratio >= coverage_long) or that align to it with coverage of the long / short sequence and
 * identity at or above the thresholds (f64 ratios of the counts of align(rep, cand, false), as the reference).  The
 * sequences stay on the device for the whole run; one batch per representative.  Outputs, all of length n and indexed
 * by SORTED position: order[p] = input index of the sequence at position p; cluster_of[p] = position of its
 * representative (p itself for representatives and, as in the reference, for folded duplicates); identical[p] =
 * position of the first exact copy (p if none before it).  n_alignments (may be NULL): alignments computed. */
int swb_cluster(swb_engine *e, const char *seq_all, const int64_t *off, int64_t n, double identity,
                double coverage_short, double coverage_long, int64_t *order, int64_t *cluster_of, int64_t *identical,
                int64_t *n_alignments);

/* Score + end cell of n pairs, HOST buffers in and out (H2D, kernels, D2H inside the call).
 * scores[k] as the reference's f32; end_a/end_b: first-maximum cell (local, SA.rs:286-294;
 * -1,-1 when no positive score) or (la,lb) (global/glocal, SA.rs:296-309). */
int swb_score_batch(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                    const int64_t *off_b, int64_t n, float *scores, int64_t *end_a, int64_t *end_b);

/* Score + end cell of n pairs WITHOUT any device allocation: DEVICE buffers in and out, the kernel's scratch in a
 * file-scope __device__ array of the library (csrc/static_scratch.cu).
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  Results are those of swb_score_batch.
 *   d_a_all / d_b_all    letters on the device, total_a / total_b bytes;   d_off_a / d_off_b: n + 1 offsets, on the device
 *   d_scores[n], d_end_a[n], d_end_b[n] (the end arrays may be NULL), d_status[1] (may be NULL): on the device
 *   stream               a cudaStream_t (NULL = the default stream); the work is enqueued there and the call returns
 *                        without synchronising; *d_status = 0, SWB_ERR_LETTER (unknown letter, no X) or SWB_ERR_ARG (an
 *                        offset out of order / out of range or a sequence longer than SWB_STATIC_MAX_LEN: such a pair is
 *                        scored as two empty sequences).
 * Limits (SWB_ERR_ARG beyond them, nothing is enqueued): n <= SWB_STATIC_MAX_PAIRS, every sequence <=
 * SWB_STATIC_MAX_LEN letters, total_a (rounded up to 256) + total_b <= SWB_STATIC_MAX_RESIDUES.
 * Every GPU that loads the library holds SWB_STATIC_SCRATCH_BYTES until exit.  All calls share that array: calls that
 * overlap on the GPU (different streams without a dependency) race; calls on one stream are safe.  The first call on a
 * device resolves the array's address (cudaGetSymbolAddress); no call allocates or frees device memory. */
#define SWB_STATIC_MAX_PAIRS 4096
#define SWB_STATIC_MAX_LEN 1024
#define SWB_STATIC_MAX_RESIDUES 1048576
#define SWB_STATIC_SCRATCH_BYTES 9945600
int swb_score_batch_static(swb_engine *e, const char *d_a_all, const int64_t *d_off_a, int64_t total_a,
                           const char *d_b_all, const int64_t *d_off_b, int64_t total_b, int64_t n, float *d_scores,
                           int64_t *d_end_a, int64_t *d_end_b, int32_t *d_status, void *stream);

/* Database search, scoring stage: one query (the `a` sequence) against n database sequences (`b`); results as
 * swb_score_batch.  The caller selects its top-K from `scores` (ties: lower index first keeps multi-GPU runs
 * deterministic) and obtains the alignments of those hits with swb_align_batch. */
int swb_search_scores(swb_engine *e, const char *query, int64_t lq, const char *db_all, const int64_t *off_db,
                      int64_t n, float *scores, int64_t *end_q, int64_t *end_t);

/* ---- Database search with a resident database -----------------------------------------------------------------------
 * swb_db_upload keeps n subjects (letters[off[k]..off[k+1]), global index global_index[k], ascending; NULL = 0..n-1)
 * encoded on the engine's device; swb_db_upload_shard takes the WHOLE database and keeps shard `rank` of `world`: the
 * subjects are dealt round-robin over the stable longest-first order (swb_db_shard returns that index list; host only),
 * so every shard holds the same length mix.  The database keeps a host copy of its letters (the hits' strings are
 * spelled from it); the caller's buffers may be released after the call.  A database belongs to one engine's device. */
typedef struct swb_db swb_db;
typedef struct swb_hit {
    float score;             /* as the reference's f32 */
    int32_t shard;           /* set by swb_search_sharded: the shard that owns the subject; 0 otherwise */
    int64_t index;           /* global index of the subject */
    int64_t end_q, end_t;    /* first-maximum cell (local) or (lq, lt) (global / glocal), as swb_score_batch */
} swb_hit;
#define SWB_MAX_TOPK 4096
int swb_db_shard(const int64_t *off, int64_t n, int rank, int world, int64_t *idx, int64_t *n_idx);
int swb_db_upload(swb_engine *e, const char *letters, const int64_t *off, int64_t n, const int64_t *global_index,
                  swb_db **out);
int swb_db_upload_shard(swb_engine *e, const char *letters, const int64_t *off, int64_t n_total, int rank, int world,
                        swb_db **out);
void swb_db_free(swb_engine *e, swb_db *db);
int64_t swb_db_count(const swb_db *db);
int64_t swb_db_residues(const swb_db *db);
/* One query against the resident database: score + end cell of every subject (align_partial(.., score_only) per pair,
 * SA.rs:438-467; s16 overflows re-run in int32), then the top-K by (score desc, global index asc) selected ON THE
 * DEVICE -- topk x 32 bytes cross PCIe, not the score vector.  hits[topk]; *n_hits = min(topk, n). */
int swb_search(swb_engine *e, const char *query, int64_t lq, swb_db *db, int topk, swb_hit *hits, int64_t *n_hits);
/* The full score vector of the last swb_search on this database, in the database's local subject order. */
int swb_db_scores(swb_engine *e, swb_db *db, float *scores, int64_t *end_q, int64_t *end_t);
/* Host only: global top-K of the concatenated per-shard lists (the one exchange between shards: gather n_shards x K
 * records with NCCL / MPI / plain memory, then merge).  Same key, so the list does not depend on the sharding. */
int swb_search_merge(const swb_hit *parts, int64_t n_parts, int topk, swb_hit *out, int64_t *n_out);
/* Traceback (align(query, subject, retain_all), SA.rs:389) of the hits whose subject this database holds: owned[k] = 1
 * and alns[k] filled for those, the others stay zeroed -- every shard traces its own hits, so the traceback divides by
 * the number of GPUs like the scan.  owned may be NULL. */
int swb_search_align(swb_engine *e, const char *query, int64_t lq, swb_db *db, const swb_hit *hits, int64_t n_hits,
                     unsigned flags, swb_alignment *alns, int32_t *owned);
/* One process, n_shards GPUs: engines[g] and dbs[g] = shard g of n_shards.  Scan + select on one host thread per shard,
 * host merge, hits traced by their owners.  alns may be NULL (no traceback). */
int swb_search_sharded(swb_engine **engines, swb_db **dbs, int n_shards, const char *query, int64_t lq, int topk,
                       unsigned flags, swb_hit *hits, int64_t *n_hits, swb_alignment *alns);

/* The same, staged: upload (encode + pack + H2D) once, run the kernels any number of times on the
 * resident batch, fetch (D2H + unpermute).  swb_batch_run is asynchronous on the engine's stream;
 * swb_batch_fetch synchronises. */
int swb_batch_upload(swb_engine *e, const char *a_all, const int64_t *off_a, const char *b_all,
                     const int64_t *off_b, int64_t n, swb_batch **out);
int swb_batch_run(swb_engine *e, swb_batch *b);
int swb_batch_fetch(swb_engine *e, swb_batch *b, float *scores, int64_t *end_a, int64_t *end_b);
void swb_batch_free(swb_engine *e, swb_batch *b);
/* device time (ms) of the last swb_batch_run, measured with CUDA events on the engine's stream;
 * synchronises.  launches = kernels in that run. */
int swb_batch_last_run_ms(swb_engine *e, swb_batch *b, float *ms, int64_t *launches);
/* steps back-to-back runs bracketed by two CUDA events on the engine's stream; synchronises and
 * returns the total device time (ms) and the number of kernels launched. */
int swb_batch_run_n(swb_engine *e, swb_batch *b, int steps, float *total_ms, int64_t *launches);

/* DPX / integer-pipe micro-benchmark (kernel K0): thread-instructions per second of the given
 * instruction mix on the whole chip; which: 0 VIADDMNMX.S16x2, 1 VIMNMX3.S16x2, 2 VIADD.16x2,
 * 3 the 4-op Gotoh cell mix, 4 VIADDMNMX (s32), 5 IMAD, 6 mix + IMAD interleaved, 7 HMNMX2, 8 HFMA2,
 * 9 VIADDMNMX + HMNMX2, 10 VIADDMNMX + HFMA2, 11 VIADDMNMX + IMAD (pipe-sharing probes). */
int swb_dpx_peak(int device, int which, double *thread_instr_per_s);

#ifdef __cplusplus
}
#endif
#endif /* SWB200_H */
