/*
 * sw_oracle_faithful.c -- CPU ORACLE, "ref-faithful" layout (test infrastructure, NOT product code).
 *
 * The same fill_matrix body as sw_oracle.c (sw_oracle_fill.inc, SA.rs:55-272) over the DATA STRUCTURE of the reference:
 * a vector of rows of cells, every cell owning two small heap arrays (scores: Vec<f32> of 3, prev: Vec<usize> of 3,
 * SA.rs:556-572), the substitution score fetched through a function pointer (Box<dyn ScoringMatrix>, SA.rs:17, 127), the
 * cell matrix kept and only grown between calls (SA.rs:67-71).  Used only by bench.py's cpu_baseline leg to time how the
 * reference itself lays its matrix out, next to the flat-array port; results are identical to
 * sw_oracle.c by construction and tested to be (tests/test_oracle_golden.py).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { CELL_MATCH = 0, CELL_GAPINX = 1, CELL_GAPINY = 2 };
enum { ALIGN_GLOBAL = 0, ALIGN_GLOCAL = 1, ALIGN_LOCAL = 2 };

typedef struct { float *scores; size_t *prev; } swf_cell;            /* SWCell, SA.rs:557-560 */
typedef struct { swf_cell **rows; size_t n_rows, n_cols; } swf_store; /* Vec<Vec<SWCell>> */
typedef float (*swf_get_score)(const float *mat, size_t n, int pssm, int32_t a, int32_t b);

static float swf_matrix_score(const float *mat, size_t n, int pssm, int32_t a, int32_t b)
{
    return pssm ? mat[(size_t)a + (size_t)b * n] : mat[(size_t)a * n + (size_t)b];
}
static swf_get_score volatile swf_dyn = swf_matrix_score;            /* an indirect call per cell, as dyn ScoringMatrix */

static int swf_reserve(swf_store *S, size_t rows, size_t cols)       /* grow only (SA.rs:67-71) */
{
    if (rows <= S->n_rows && cols <= S->n_cols) return 0;
    for (size_t i = 0; i < S->n_rows; i++) {
        for (size_t j = 0; j < S->n_cols; j++) { free(S->rows[i][j].scores); free(S->rows[i][j].prev); }
        free(S->rows[i]);
    }
    free(S->rows);
    if (rows < S->n_rows) rows = S->n_rows;
    if (cols < S->n_cols) cols = S->n_cols;
    S->rows = (swf_cell **)malloc(rows * sizeof(swf_cell *));
    if (!S->rows) { S->n_rows = S->n_cols = 0; return -1; }
    for (size_t i = 0; i < rows; i++) {
        S->rows[i] = (swf_cell *)malloc(cols * sizeof(swf_cell));
        for (size_t j = 0; j < cols; j++) {
            S->rows[i][j].scores = (float *)calloc(3, sizeof(float));
            S->rows[i][j].prev = (size_t *)calloc(3, sizeof(size_t));
        }
    }
    S->n_rows = rows; S->n_cols = cols;
    return 0;
}

static void swf_free(swf_store *S)
{
    for (size_t i = 0; i < S->n_rows; i++) {
        for (size_t j = 0; j < S->n_cols; j++) { free(S->rows[i][j].scores); free(S->rows[i][j].prev); }
        free(S->rows[i]);
    }
    free(S->rows);
    S->rows = NULL; S->n_rows = S->n_cols = 0;
}

#define SWO_FILL_NAME swf_fill
#define SWO_STORE swf_store
#define SWO_PV_T size_t
#define SWO_SC(B, i, j) ((B)->rows[i][j].scores)
#define SWO_PV(B, i, j) ((B)->rows[i][j].prev)
#define SWO_SCORE(mat, n, pssm, ai, bj) swf_dyn(mat, n, pssm, ai, bj)
#include "sw_oracle_fill.inc"

/* Scores of a batch (fill + the score_only scan of SA.rs:438-467), OpenMP over pairs; one cell matrix per thread. */
int swf_score_batch(const int32_t *a_all, const int64_t *off_a, const int32_t *b_all, const int64_t *off_b,
                    int64_t npairs, const float *mat, int64_t n, float go, float ge, int mode,
                    float *scores, int64_t *end_i, int64_t *end_j, int nthreads)
{
    int rc = 0;
    float o = go > 0.0f ? -go : go, e = ge > 0.0f ? -ge : ge;        /* SA.rs:28-35 */
#ifdef _OPENMP
#pragma omp parallel num_threads(nthreads > 0 ? nthreads : 1)
#endif
    {
        swf_store S; memset(&S, 0, sizeof(S));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 4)
#endif
        for (int64_t k = 0; k < npairs; k++) {
            const size_t la = (size_t)(off_a[k + 1] - off_a[k]), lb = (size_t)(off_b[k + 1] - off_b[k]);
            if (swf_reserve(&S, la + 1, lb + 1)) { rc = -1; continue; }
            swf_fill(&S, a_all + off_a[k], la, b_all + off_b[k], lb, mat, (size_t)n, 0, o, e, mode);
            float best; int64_t bi = -1, bj = -1;
            if (mode == ALIGN_LOCAL) {
                best = 0.0f;
                for (size_t ii = 0; ii <= la; ii++)
                    for (size_t jj = 0; jj <= lb; jj++)
                        if (S.rows[ii][jj].scores[CELL_MATCH] > best) { best = S.rows[ii][jj].scores[CELL_MATCH]; bi = (int64_t)ii; bj = (int64_t)jj; }
            } else {
                const float *c = S.rows[la][lb].scores;
                best = c[0];
                for (int s = 1; s < 3; s++) if (best < c[s]) best = c[s];
                bi = (int64_t)la; bj = (int64_t)lb;
            }
            scores[k] = best; end_i[k] = bi; end_j[k] = bj;
        }
        swf_free(&S);
    }
    return rc;
}
