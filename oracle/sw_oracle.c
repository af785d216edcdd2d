/*
 * sw_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * A plain-C restatement of the reference's CPU alignment path,
 *   rust/sequence_alignment/src/sequence_alignment.rs   ("SA.rs" below)
 * used ONLY by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs as the checker / timed CPU baseline.  The engine under
 * smithwaterman_b200/ never links, imports or calls anything in this file.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this file against
 * every Rust #[test] known answer of SA.rs:921-1221 / OCL.rs:1429-1448 and all
 * 300 EMBOSS results of test/emboss_results (strings and scores), committed as
 * fixtures under tests/golden/ by tests/golden/make_golden.py.
 *
 * The reference is Rust; no rustc/cargo exists in this image, so oracle/_ref
 * (a build of the reference's own sources) is not possible: kind = "port".
 *
 * Arithmetic is f32 exactly as in the reference (same operand order, same
 * comparison operators); build with -O2 -ffp-contract=off (no fast-math).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* SA.rs:7-9 */
enum { CELL_MATCH = 0, CELL_GAPINX = 1, CELL_GAPINY = 2 };
/* SA.rs:11-13 */
enum { ALIGN_GLOBAL = 0, ALIGN_GLOCAL = 1, ALIGN_LOCAL = 2 };

typedef struct {
    float *sc;      /* 3 floats per cell, [ (i*(lb+1)+j)*3 + state ]  (SA.rs:557-560, flattened) */
    uint8_t *pv;    /* 3 predecessor codes per cell */
    size_t cap;     /* cells allocated */
} swo_buf;

static int swo_reserve(swo_buf *b, size_t cells)
{
    if (cells <= b->cap) return 0;
    free(b->sc); free(b->pv);
    b->sc = (float *)malloc(cells * 3 * sizeof(float));
    b->pv = (uint8_t *)malloc(cells * 3);
    b->cap = (b->sc && b->pv) ? cells : 0;
    return b->cap ? 0 : -1;
}

/* SA.rs:26-35: penalties are forced negative without touching magnitude. */
static void swo_penalties(float go, float ge, float *o, float *e)
{
    *o = go; *e = ge;
    if (*o > 0.0f) *o *= -1.0f;
    if (*e > 0.0f) *e *= -1.0f;
}

#define SWO_FILL_NAME swo_fill
#define SWO_STORE swo_buf
#define SWO_PV_T uint8_t
#define SWO_SC(B, i, j) ((B)->sc + ((i) * W + (j)) * 3)
#define SWO_PV(B, i, j) ((B)->pv + ((i) * W + (j)) * 3)
#define SWO_SCORE(mat, n, pssm, ai, bj) ((pssm) ? (mat)[(size_t)(ai) + (size_t)(bj) * (n)] : (mat)[(size_t)(ai) * (n) + (size_t)(bj)])
#include "sw_oracle_fill.inc"
#undef SWO_FILL_NAME
#undef SWO_STORE
#undef SWO_PV_T
#undef SWO_SC
#undef SWO_PV
#undef SWO_SCORE

/*
 * backtrack, SA.rs:273-387.  Writes reversed-then-reversed (i.e. forward) index
 * lists into idx_a/idx_b (capacity >= la+lb), -1 = gap; returns the column count,
 * or -2 for the reference's panic / usize underflow cases (SA.rs:360-363, 369-371).
 */
static int64_t swo_backtrack(const swo_buf *B, size_t la, size_t lb, int mode,
                             int64_t *idx_a, int64_t *idx_b, float *score,
                             int64_t *end_i, int64_t *end_j)
{
    const size_t W = lb + 1;
    int64_t sx = -1, sy = -1; float maxscore; int place;
    if (mode == ALIGN_LOCAL) {                    /* SA.rs:282-295 */
        maxscore = 0.0f;
        for (size_t ii = 0; ii <= la; ii++)
            for (size_t jj = 0; jj <= lb; jj++) {
                float v = B->sc[(ii * W + jj) * 3 + CELL_MATCH];
                if (v > maxscore) { maxscore = v; sx = (int64_t)ii; sy = (int64_t)jj; }
            }
        place = CELL_MATCH;
    } else {                                      /* SA.rs:296-309 */
        sx = (int64_t)la; sy = (int64_t)lb;
        const float *c = B->sc + ((size_t)sx * W + (size_t)sy) * 3;
        maxscore = c[0]; place = 0;
        for (int k = 1; k < 3; k++) if (maxscore < c[k]) { maxscore = c[k]; place = k; }
    }
    *end_i = sx; *end_j = sy;
    if (sx < 0) { *score = 0.0f; return 0; }      /* SA.rs:346-348 */
    size_t cx = (size_t)sx, cy = (size_t)sy; int dir = place; int64_t n = 0;
    for (;;) {                                    /* SA.rs:352-383 */
        int prev = B->pv[(cx * W + cy) * 3 + dir];
        if (mode == ALIGN_LOCAL && B->sc[(cx * W + cy) * 3 + dir] == 0.0f) break;
        if (dir == CELL_MATCH) {
            if (cx == 0 || cy == 0) return -2;
            idx_a[n] = (int64_t)cx - 1; idx_b[n] = (int64_t)cy - 1; n++; cx--; cy--;
        } else if (dir == CELL_GAPINX) {
            if (cy == 0) return -2;
            idx_a[n] = -1; idx_b[n] = (int64_t)cy - 1; n++; cy--;
        } else {
            if (cx == 0) return -2;
            idx_a[n] = (int64_t)cx - 1; idx_b[n] = -1; n++; cx--;
        }
        if (cx == 0 && cy == 0) break;
        dir = prev;
    }
    for (int64_t k = 0; k < n / 2; k++) {         /* SA.rs:384-385 */
        int64_t t = idx_a[k]; idx_a[k] = idx_a[n - 1 - k]; idx_a[n - 1 - k] = t;
        t = idx_b[k]; idx_b[k] = idx_b[n - 1 - k]; idx_b[n - 1 - k] = t;
    }
    *score = maxscore;
    return n;
}

/* score_only early-out of align_partial, SA.rs:438-467 (also reports the first-max cell). */
static float swo_score_only(const swo_buf *B, size_t la, size_t lb, int mode,
                            int64_t *end_i, int64_t *end_j)
{
    const size_t W = lb + 1; float maxscore;
    if (mode == ALIGN_LOCAL) {
        maxscore = 0.0f; *end_i = -1; *end_j = -1;
        for (size_t ii = 0; ii <= la; ii++)
            for (size_t jj = 0; jj <= lb; jj++) {
                float v = B->sc[(ii * W + jj) * 3 + CELL_MATCH];
                if (v > maxscore) { maxscore = v; *end_i = (int64_t)ii; *end_j = (int64_t)jj; }
            }
    } else {
        const float *c = B->sc + (la * W + lb) * 3;
        maxscore = c[0];
        for (int k = 1; k < 3; k++) if (maxscore < c[k]) maxscore = c[k];
        *end_i = (int64_t)la; *end_j = (int64_t)lb;
    }
    return maxscore;
}

/* ---------------------------------------------------------------- public entry points */

typedef struct swo_engine { swo_buf buf; } swo_engine;

swo_engine *swo_create(void) { return (swo_engine *)calloc(1, sizeof(swo_engine)); }
void swo_destroy(swo_engine *e) { if (e) { free(e->buf.sc); free(e->buf.pv); free(e); } }

/*
 * One alignment (fill + backtrack), the numeric part of align_partial SA.rs:393-469.
 * Returns the number of alignment columns (>=0), -1 on allocation failure, -2 where
 * the reference panics.  end_i/end_j: 1-based DP cell the traceback starts from
 * (-1,-1 for an empty local alignment).
 */
int64_t swo_align(swo_engine *E, const int32_t *a, int64_t la, const int32_t *b, int64_t lb,
                  const float *mat, int64_t n, int pssm, float go, float ge, int mode,
                  int score_only, int64_t *idx_a, int64_t *idx_b, float *score,
                  int64_t *end_i, int64_t *end_j)
{
    float o, e; swo_penalties(go, ge, &o, &e);
    if (swo_reserve(&E->buf, (size_t)(la + 1) * (size_t)(lb + 1))) return -1;
    swo_fill(&E->buf, a, (size_t)la, b, (size_t)lb, mat, (size_t)n, pssm, o, e, mode);
    if (score_only) { *score = swo_score_only(&E->buf, (size_t)la, (size_t)lb, mode, end_i, end_j); return 0; }
    return swo_backtrack(&E->buf, (size_t)la, (size_t)lb, mode, idx_a, idx_b, score, end_i, end_j);
}

/* Raw cell access for cell-level tests: copies 3 scores + 3 prev codes of cell (i,j) of the last fill. */
void swo_cell(const swo_engine *E, int64_t lb, int64_t i, int64_t j, float *sc3, uint8_t *pv3)
{
    size_t k = ((size_t)i * (size_t)(lb + 1) + (size_t)j) * 3;
    memcpy(sc3, E->buf.sc + k, 3 * sizeof(float)); memcpy(pv3, E->buf.pv + k, 3);
}

/*
 * Batch of score-only alignments (dense f32 fill exactly as above, then the SA.rs:438-467
 * scan), OpenMP over pairs.  This is the timed "reference CPU path" of bench.py
 * (kind = "port"): the reference itself runs the same dense fill serially
 * (SA-main.rs:35-45).  seqs: concatenated residue indices; off_x[k]..off_x[k+1].
 */
int swo_score_batch(const int32_t *a_all, const int64_t *off_a, const int32_t *b_all, const int64_t *off_b,
                    int64_t npairs, const float *mat, int64_t n, float go, float ge, int mode,
                    float *scores, int64_t *end_i, int64_t *end_j, int nthreads)
{
    int rc = 0;
#ifdef _OPENMP
#pragma omp parallel num_threads(nthreads > 0 ? nthreads : 1)
#endif
    {
        swo_engine *E = swo_create();
        float o, e; swo_penalties(go, ge, &o, &e);
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 4)
#endif
        for (int64_t k = 0; k < npairs; k++) {
            size_t la = (size_t)(off_a[k + 1] - off_a[k]), lb = (size_t)(off_b[k + 1] - off_b[k]);
            if (swo_reserve(&E->buf, (la + 1) * (lb + 1))) { rc = -1; continue; }
            swo_fill(&E->buf, a_all + off_a[k], la, b_all + off_b[k], lb, mat, (size_t)n, 0, o, e, mode);
            scores[k] = swo_score_only(&E->buf, la, lb, mode, &end_i[k], &end_j[k]);
        }
        swo_destroy(E);
    }
    return rc;
}
