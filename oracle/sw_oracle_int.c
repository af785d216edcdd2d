/*
 * sw_oracle_int.c -- CPU ORACLE, integer twin (test infrastructure, NOT product code).
 *
 * Linear-memory, score-only restatement of fill_matrix (SA.rs:55-272) + the
 * score_only scan (SA.rs:438-467) on integer-rescaled scores: every f32 value of
 * the reference is a multiple of 1/scale, so (value*scale) is an exact int32 and
 * comparisons are unchanged.  tests/test_oracle_golden.py proves it equal to the
 * dense f32 oracle (sw_oracle.c) on the goldens and on fuzzed pairs; it exists so
 * that pairs whose dense matrix does not fit (10^10 cells) still have a checker.
 * Three rolling rows of (M, X, Y); O(lb) memory.
 */
#include <stdint.h>
#include <stdlib.h>

enum { ALIGN_GLOBAL = 0, ALIGN_GLOCAL = 1, ALIGN_LOCAL = 2 };

static inline int32_t max3i(int32_t a, int32_t b, int32_t c)
{
    int32_t m = a > b ? a : b; return m > c ? m : c;
}

/*
 * o,e: interior penalties (<=0, already scaled).  unit: the integer that stands
 * for 1.0 (the (0,0) cell holds X=Y=-1.0, SA.rs:90-91).  mat: n*n int32, scaled.
 * Returns 0, or -1 on allocation failure.  Local: score = first strict maximum of
 * the MATCH state in (i outer, j inner) order, end = that cell (1-based; -1,-1 if
 * none).  Non-local: score = max of the 3 states at (la,lb).
 */
int swo_int_score(const int32_t *a, int64_t la, const int32_t *b, int64_t lb,
                  const int32_t *mat, int64_t n, int32_t o, int32_t e, int32_t unit, int mode,
                  int64_t *score, int64_t *end_i, int64_t *end_j)
{
    const int32_t so = (mode == ALIGN_GLOBAL) ? o : 0, se = (mode == ALIGN_GLOBAL) ? e : 0;
    const int32_t sent = 10 * o + 10 * e;
    int32_t *row = (int32_t *)malloc((size_t)(lb + 1) * 3 * sizeof(int32_t));
    if (!row) return -1;
    int32_t *M = row, *X = row + (lb + 1), *Y = row + 2 * (lb + 1);
    int64_t best = 0, bi = -1, bj = -1;

    M[0] = 0; X[0] = -unit; Y[0] = -unit;                       /* SA.rs:88-93 */
    for (int64_t j = 1; j <= lb; j++) {                          /* SA.rs:100-108 */
        int32_t ls = (int32_t)j * se + (so - se);
        M[j] = ls + sent; X[j] = ls; Y[j] = ls + sent;
    }
    for (int64_t i = 1; i <= la; i++) {
        int32_t dM = M[0], dX = X[0], dY = Y[0];                 /* cell (i-1, j-1) */
        int32_t ls = (int32_t)i * se + (so - se);                /* SA.rs:109-118 */
        M[0] = ls + sent; X[0] = ls + sent; Y[0] = ls;
        const int32_t *mrow = mat + (size_t)a[i - 1] * n;
        const int32_t po = (mode != ALIGN_LOCAL && i == la) ? so : o;   /* SA.rs:166-167 */
        const int32_t pe = (mode != ALIGN_LOCAL && i == la) ? se : e;
        for (int64_t j = 1; j <= lb; j++) {
            const int32_t qo = (mode != ALIGN_LOCAL && j == lb) ? so : o;   /* SA.rs:169-170 */
            const int32_t qe = (mode != ALIGN_LOCAL && j == lb) ? se : e;
            int32_t lM = M[j], lX = X[j], lY = Y[j];             /* cell (i-1, j) */
            int32_t m = max3i(dM, dX, dY) + mrow[b[j - 1]];
            int32_t x = max3i(M[j - 1] + po, X[j - 1] + pe, Y[j - 1] + po);   /* cell (i, j-1) */
            int32_t y = max3i(lM + qo, lY + qe, lX + qo);
            if (mode == ALIGN_LOCAL) {                           /* SA.rs:256-260 */
                if (m < 0) m = 0;
                if (x < 0) x = 0;
                if (y < 0) y = 0;
                if (m > best) { best = m; bi = i; bj = j; }      /* SA.rs:444-449 */
            }
            dM = lM; dX = lX; dY = lY;
            M[j] = m; X[j] = x; Y[j] = y;
        }
    }
    if (mode == ALIGN_LOCAL) { *score = best; *end_i = bi; *end_j = bj; }
    else { *score = max3i(M[lb], X[lb], Y[lb]); *end_i = la; *end_j = lb; }
    free(row);
    return 0;
}

/* OpenMP batch of the above over concatenated sequences (offset arrays of npairs+1). */
int swo_int_score_batch(const int32_t *a_all, const int64_t *off_a, const int32_t *b_all, const int64_t *off_b,
                        int64_t npairs, const int32_t *mat, int64_t n, int32_t o, int32_t e, int32_t unit, int mode,
                        int64_t *scores, int64_t *end_i, int64_t *end_j, int nthreads)
{
    int rc = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (int64_t k = 0; k < npairs; k++) {
        if (swo_int_score(a_all + off_a[k], off_a[k + 1] - off_a[k], b_all + off_b[k], off_b[k + 1] - off_b[k],
                          mat, n, o, e, unit, mode, &scores[k], &end_i[k], &end_j[k])) rc = -1;
    }
    return rc;
}
