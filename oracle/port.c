/* TEST INFRASTRUCTURE ONLY — the CPU oracle ("port") for the marker-scoring hot path.
 *
 * A plain-C restatement of the arithmetic in the reference's two native files, written
 * from their behaviour (operation order included) and citing the lines it follows:
 *   - blocked mean / variance:      /root/reference/src/compute_blocked_stats.cpp:15-123
 *   - transforms none/lognorm/norm: /root/reference/src/compute_blocked_stats.cpp:129-202
 *   - Wilcoxon U and tie term:      /root/reference/src/overlap_exprs.cpp:34-128
 *   - drivers:                      /root/reference/src/overlap_exprs.cpp:140-220
 * Parity status: PINNED. tests/test_oracle.py checks every function here against
 * oracle/_ref/liboracle_ref.so (the reference's own object code compiled unmodified, when
 * present), against tests/golden/ vectors produced by that object code, against
 * scipy.stats.mannwhitneyu / numpy, and against the hand-worked vector of SURVEY.md
 * Appendix D.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this
 * library, and only as the checker / the timed CPU baseline. The product (libscranb200.so)
 * never calls it and has no CPU fallback.
 *
 * Build: gcc -std=gnu11 -O2 -ffp-contract=off -fPIC -shared (oracle/Makefile).
 * Matrices are CSC (p: int64[ncol+1], i: int32 rows ascending in a column, x: double) or
 * dense column-major; outputs are caller-allocated.
 */
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_NA_INT INT32_MIN

static const char* g_error = "";
const char* oracle_port_last_error(void) { return g_error; }

/* ------------------------------------------------------------------------------------ */
/* Blocked statistics                                                                   */
/* ------------------------------------------------------------------------------------ */

/* One running-moment update; compute_blocked_stats.cpp:61-66 (sparse) and :75-78 (dense). */
static inline void running_update(double v, double* M, double* S, double* NZ) {
    *NZ += 1.0;
    const double delta = v - *M;
    *M += delta / *NZ;
    *S += delta * (v - *M);
}

/* compute_blocked_stats.cpp:136 (lognorm), :168 (norm), :196-198 (none). */
static inline double apply_transform(int transform, double v, double sf, double pseudo) {
    if (transform == 1) return log(v / sf + pseudo) / M_LN2;
    if (transform == 2) return v / sf;
    return v;
}

/* transform: 0 none, 1 lognorm, 2 norm. block[c] in [0, nblocks) or ORACLE_NA_INT (skipped,
 * compute_blocked_stats.cpp:39-42). mean_out / var_out / nz_out: nrow x nblocks column-major;
 * nz_out (nullable) returns the per-block count of values that entered the running update
 * (the reference's internal 'nnzero' matrix, :22). */
int oracle_port_blocked_stats(int64_t nrow, int64_t ncol, const int64_t* p, const int32_t* i, const double* x,
                              int sparse, const int32_t* block, int nblocks, int transform, const double* sf,
                              double pseudo, double* mean_out, double* var_out, double* nz_out) {
    const size_t cells = (size_t)nrow * (size_t)nblocks;
    double* nz = nz_out ? nz_out : (double*)calloc(cells ? cells : 1, sizeof(double));
    int64_t* count = (int64_t*)calloc(nblocks > 0 ? nblocks : 1, sizeof(int64_t));
    if (!nz || !count) {
        g_error = "out of memory";
        return 1;
    }
    memset(mean_out, 0, cells * sizeof(double));
    memset(var_out, 0, cells * sizeof(double));
    if (nz_out) memset(nz_out, 0, cells * sizeof(double));

    /* Zeros may be skipped only when the transform maps 0 to 0 (:31, :142, :175, :201). */
    const int keeps_zero = (transform == 1) ? (pseudo == 1.0) : 1;

    for (int64_t c = 0; c < ncol; ++c) {
        const int32_t b = block[c];
        if (b == ORACLE_NA_INT) continue;
        if (b < 0 || b >= nblocks) {
            g_error = "block id out of range";
            free(count);
            if (!nz_out) free(nz);
            return 1;
        }
        ++count[b];
        double* M = mean_out + (size_t)b * nrow;
        double* S = var_out + (size_t)b * nrow;
        double* NZ = nz + (size_t)b * nrow;
        const double s = sf ? sf[c] : 1.0;

        if (sparse && keeps_zero) {
            /* Stored entries are NOT filtered for zero on this branch (:57-67). */
            for (int64_t k = p[c]; k < p[c + 1]; ++k) {
                const int32_t g = i[k];
                running_update(apply_transform(transform, x[k], s, pseudo), M + g, S + g, NZ + g);
            }
        } else if (sparse) {
            /* Sparse input but zeros matter: walk the densified column (:69-80). */
            int64_t k = p[c];
            for (int64_t g = 0; g < nrow; ++g) {
                double raw = 0.0;
                if (k < p[c + 1] && i[k] == g) raw = x[k++];
                running_update(apply_transform(transform, raw, s, pseudo), M + g, S + g, NZ + g);
            }
        } else {
            const double* col = x + (size_t)c * nrow;
            for (int64_t g = 0; g < nrow; ++g) {
                const double v = apply_transform(transform, col[g], s, pseudo);
                if (!keeps_zero || v != 0.0) running_update(v, M + g, S + g, NZ + g); /* :74 */
            }
        }
    }

    for (int b = 0; b < nblocks; ++b) {
        double* M = mean_out + (size_t)b * nrow;
        double* S = var_out + (size_t)b * nrow;
        const double* NZ = nz + (size_t)b * nrow;
        if (count[b] <= 0) { /* :86-88 */
            for (int64_t g = 0; g < nrow; ++g) M[g] = NAN;
        }
        if (count[b] <= 1) { /* :91-94: no zero fill either */
            for (int64_t g = 0; g < nrow; ++g) S[g] = NAN;
            continue;
        }
        if (keeps_zero) { /* :96-114 */
            const double total = (double)count[b];
            for (int64_t g = 0; g < nrow; ++g) {
                const double curnz = NZ[g];
                const double ratio = curnz / total;
                S[g] += M[g] * M[g] * ratio * (total - curnz);
                M[g] *= ratio;
            }
        }
        const double rdf = (double)(count[b] - 1); /* :116-119 */
        for (int64_t g = 0; g < nrow; ++g) S[g] /= rdf;
    }

    free(count);
    if (!nz_out) free(nz);
    return 0;
}

/* ------------------------------------------------------------------------------------ */
/* Wilcoxon U / tie term                                                                */
/* ------------------------------------------------------------------------------------ */

typedef struct {
    int ngroups;
    const int64_t* gptr;   /* ngroups + 1 offsets into gcells */
    const int32_t* gcells; /* 0-based cell indices */
    double* sorted;        /* same layout as gcells: per-group sorted non-zero values */
    int64_t* nfilled;      /* entries of 'sorted' in use, including the zero placeholder */
    int64_t* nzero;        /* number of zero-valued cells in the group */
} ranker;

static int cmp_double(const void* a, const void* b) {
    const double x = *(const double*)a, y = *(const double*)b;
    return (x > y) - (x < y);
}

static int ranker_init(ranker* r, int ngroups, const int64_t* gptr, const int32_t* gcells, int64_t ncells) {
    r->ngroups = ngroups;
    r->gptr = gptr;
    r->gcells = gcells;
    const int64_t total = gptr[ngroups];
    for (int64_t k = 0; k < total; ++k) { /* overlap_exprs.cpp:22-27 */
        if (gcells[k] < 0 || gcells[k] >= ncells) {
            g_error = "indices in 'groups' out of range";
            return 1;
        }
    }
    r->sorted = (double*)malloc((size_t)(total > 0 ? total : 1) * sizeof(double));
    r->nfilled = (int64_t*)calloc(ngroups > 0 ? ngroups : 1, sizeof(int64_t));
    r->nzero = (int64_t*)calloc(ngroups > 0 ? ngroups : 1, sizeof(int64_t));
    if (!r->sorted || !r->nfilled || !r->nzero) {
        g_error = "out of memory";
        return 1;
    }
    return 0;
}

static void ranker_free(ranker* r) {
    free(r->sorted);
    free(r->nfilled);
    free(r->nzero);
}

/* Per gene: keep the non-zero values of each group, one placeholder 0 standing for all the
 * zeros of the group, sorted ascending (overlap_exprs.cpp:34-60). */
static void ranker_load(ranker* r, const double* row) {
    for (int g = 0; g < r->ngroups; ++g) {
        double* dst = r->sorted + r->gptr[g];
        const int64_t len = r->gptr[g + 1] - r->gptr[g];
        int64_t kept = 0;
        for (int64_t k = 0; k < len; ++k) {
            const double v = row[r->gcells[r->gptr[g] + k]];
            if (v != 0.0) dst[kept++] = v; /* 'if (val)': drops +0.0 and -0.0, keeps NaN */
        }
        r->nzero[g] = len - kept;
        if (r->nzero[g]) dst[kept++] = 0.0;
        r->nfilled[g] = kept;
        qsort(dst, (size_t)kept, sizeof(double), cmp_double);
    }
}

/* Sweep the tie runs of {a - shift : a in left} U {b in right} in ascending order
 * (overlap_exprs.cpp:62-128). *U gets #(a-shift > b) + 0.5 #(a-shift == b); *T gets the sum of
 * t(t^2-1) over tie runs, both accumulated in double in the reference's order. */
static void ranker_contrast(const ranker* r, int left, int right, double shift, double* U, double* T) {
    const double* A = r->sorted + r->gptr[left];
    const double* B = r->sorted + r->gptr[right];
    const int64_t na = r->nfilled[left], nb = r->nfilled[right];
    int64_t ia = 0, ib = 0;
    int below = 0; /* the reference keeps this in 'int' (:73) and widens in the product (:119) */
    double score = 0.0, tieval = 0.0;

    while (ia < na || ib < nb) {
        double cur;
        if (ia < na && ib < nb) {
            const double sa = A[ia] - shift;
            cur = sa < B[ib] ? sa : B[ib];
        } else if (ia < na) {
            cur = A[ia] - shift;
        } else {
            cur = B[ib];
        }

        int64_t ta = 0;
        if (ia < na) {
            if (A[ia] != 0.0) { /* a real value (:94-99) */
                const int64_t start = ia;
                while (ia < na && A[ia] - shift == cur) ++ia;
                ta = ia - start;
            } else if (-shift == cur) { /* the placeholder, worth nzero[left] ties (:99-102) */
                ta = r->nzero[left];
                ++ia;
            }
        }
        int64_t tb = 0;
        if (ib < nb) {
            if (B[ib] != 0.0) { /* :107-112 */
                const int64_t start = ib;
                while (ib < nb && B[ib] == cur) ++ib;
                tb = ib - start;
            } else if (0.0 == cur) { /* :113-116 */
                tb = r->nzero[right];
                ++ib;
            }
        }

        score += ((double)below + (double)tb * 0.5) * (double)(int)ta; /* :119 */
        below += (int)tb;                                              /* :120 */
        const double nties = (double)(int)(ta + tb);                   /* :123 */
        tieval += nties * (nties * nties - 1.0);                       /* :124 */
    }
    *U = score;
    *T = tieval;
}

static void densify_row(int64_t ncol, const int64_t* rptr, const int32_t* rcol, const double* rval, int64_t g,
                        double* row) {
    memset(row, 0, (size_t)ncol * sizeof(double));
    for (int64_t k = rptr[g]; k < rptr[g + 1]; ++k) row[rcol[k]] = rval[k];
}

/* CSC -> row-major copy so that rows can be densified quickly (stands in for beachmat's
 * get_row, overlap_exprs.cpp:171,211). */
static int build_rows(int64_t nrow, int64_t ncol, const int64_t* p, const int32_t* i, const double* x,
                      int64_t** rptr_out, int32_t** rcol_out, double** rval_out) {
    const int64_t nnz = p[ncol];
    int64_t* rptr = (int64_t*)calloc((size_t)nrow + 1, sizeof(int64_t));
    int32_t* rcol = (int32_t*)malloc((size_t)(nnz > 0 ? nnz : 1) * sizeof(int32_t));
    double* rval = (double*)malloc((size_t)(nnz > 0 ? nnz : 1) * sizeof(double));
    int64_t* fill = (int64_t*)malloc(((size_t)nrow + 1) * sizeof(int64_t));
    if (!rptr || !rcol || !rval || !fill) {
        g_error = "out of memory";
        return 1;
    }
    for (int64_t k = 0; k < nnz; ++k) ++rptr[i[k] + 1];
    for (int64_t g = 0; g < nrow; ++g) rptr[g + 1] += rptr[g];
    memcpy(fill, rptr, ((size_t)nrow + 1) * sizeof(int64_t));
    for (int64_t c = 0; c < ncol; ++c) {
        for (int64_t k = p[c]; k < p[c + 1]; ++k) {
            const int64_t at = fill[i[k]]++;
            rcol[at] = (int32_t)c;
            rval[at] = x[k];
        }
    }
    free(fill);
    *rptr_out = rptr;
    *rcol_out = rcol;
    *rval_out = rval;
    return 0;
}

typedef struct {
    int sparse;
    int64_t nrow, ncol;
    const double* dense;
    int64_t* rptr;
    int32_t* rcol;
    double* rval;
} row_source;

static int row_source_init(row_source* rs, int64_t nrow, int64_t ncol, const int64_t* p, const int32_t* i,
                           const double* x, int sparse) {
    rs->sparse = sparse;
    rs->nrow = nrow;
    rs->ncol = ncol;
    rs->dense = x;
    rs->rptr = NULL;
    rs->rcol = NULL;
    rs->rval = NULL;
    if (sparse) return build_rows(nrow, ncol, p, i, x, &rs->rptr, &rs->rcol, &rs->rval);
    return 0;
}

static void row_source_get(const row_source* rs, int64_t g, double* row) {
    if (rs->sparse) {
        densify_row(rs->ncol, rs->rptr, rs->rcol, rs->rval, g, row);
    } else {
        for (int64_t c = 0; c < rs->ncol; ++c) row[c] = rs->dense[(size_t)c * rs->nrow + g];
    }
}

static void row_source_free(row_source* rs) {
    free(rs->rptr);
    free(rs->rcol);
    free(rs->rval);
}

/* overlap_exprs_paired (overlap_exprs.cpp:196-220). left/right are 1-based (:215).
 * out: npairs x nrow column-major; ties_out (nullable, same shape) is an extra the reference
 * computes but does not return from this entry point. Genes [g0, g1) only (g1<=0: all). */
int oracle_port_overlap_paired(int64_t nrow, int64_t ncol, const int64_t* p, const int32_t* i, const double* x,
                               int sparse, const int32_t* left, const int32_t* right, int64_t npairs,
                               int ngroups, const int64_t* gptr, const int32_t* gcells, double lfc, double* out,
                               double* ties_out, int64_t g0, int64_t g1) {
    ranker r;
    row_source rs;
    if (ranker_init(&r, ngroups, gptr, gcells, ncol)) return 1;
    if (row_source_init(&rs, nrow, ncol, p, i, x, sparse)) return 1;
    double* row = (double*)malloc((size_t)(ncol > 0 ? ncol : 1) * sizeof(double));
    if (g1 <= 0) {
        g0 = 0;
        g1 = nrow;
    }
    for (int64_t g = g0; g < g1; ++g) {
        row_source_get(&rs, g, row);
        ranker_load(&r, row);
        for (int64_t k = 0; k < npairs; ++k) {
            double U, T;
            ranker_contrast(&r, left[k] - 1, right[k] - 1, lfc, &U, &T);
            out[(size_t)g * npairs + k] = U;
            if (ties_out) ties_out[(size_t)g * npairs + k] = T;
        }
    }
    free(row);
    row_source_free(&rs);
    ranker_free(&r);
    return 0;
}

/* overlap_exprs (overlap_exprs.cpp:140-193), flattened: u_out/t_out are
 * [ngroups][ngroups][nrow]; entry (i1, i2), i2 < i1, is 'i1 vs i2 shifted by lfc'; when
 * lfc != 0 entry (i2, i1) is 'i1 vs i2 shifted by -lfc' (:183-187). Pairs involving an empty
 * group stay 0 (:175,177). Entries the reference does not allocate are NaN. */
int oracle_port_overlap(int64_t nrow, int64_t ncol, const int64_t* p, const int32_t* i, const double* x,
                        int sparse, int ngroups, const int64_t* gptr, const int32_t* gcells, double lfc,
                        double* u_out, double* t_out) {
    ranker r;
    row_source rs;
    if (ranker_init(&r, ngroups, gptr, gcells, ncol)) return 1;
    if (row_source_init(&rs, nrow, ncol, p, i, x, sparse)) return 1;
    double* row = (double*)malloc((size_t)(ncol > 0 ? ncol : 1) * sizeof(double));
    const size_t G = (size_t)ngroups;
    for (size_t a = 0; a < G; ++a) {
        for (size_t b = 0; b < G; ++b) {
            const int present = (lfc == 0.0) ? (b < a) : 1;
            for (int64_t g = 0; g < nrow; ++g) {
                u_out[(a * G + b) * nrow + g] = present ? 0.0 : NAN;
                t_out[(a * G + b) * nrow + g] = present ? 0.0 : NAN;
            }
        }
    }
    for (int64_t g = 0; g < nrow; ++g) {
        row_source_get(&rs, g, row);
        ranker_load(&r, row);
        for (size_t a = 0; a < G; ++a) {
            if (gptr[a + 1] == gptr[a]) continue;
            for (size_t b = 0; b < a; ++b) {
                if (gptr[b + 1] == gptr[b]) continue;
                double U, T;
                ranker_contrast(&r, (int)a, (int)b, lfc, &U, &T);
                u_out[(a * G + b) * nrow + g] = U;
                t_out[(a * G + b) * nrow + g] = T;
                if (lfc != 0.0) {
                    ranker_contrast(&r, (int)a, (int)b, -lfc, &U, &T);
                    u_out[(b * G + a) * nrow + g] = U;
                    t_out[(b * G + a) * nrow + g] = T;
                }
            }
        }
    }
    free(row);
    row_source_free(&rs);
    ranker_free(&r);
    return 0;
}

/* ---- combineMarkers' native helpers: src/choose_effect_size.cpp:26-97 ---------------------------
 * define_jump (:26-33): index of the (prop * n)-th smallest of n values, 0-based, 0 when prop == 0. */
static size_t port_define_jump(size_t n, double prop) {
    size_t jump = (size_t)ceil((double)n * prop);
    if (jump) --jump;
    return jump;
}

static int port_cmp_int(const void* a, const void* b) {
    const int x = *(const int*)a, y = *(const int*)b;
    return (x > y) - (x < y);
}

/* compute_Top_statistic_from_ranks (:35-61). ranks: [ncon][ngenes], INT_MIN = NA. */
int oracle_port_top_from_ranks(int ncon, int64_t ngenes, const int* ranks, double prop, int* out) {
    int* col = (int*)malloc((size_t)(ncon > 0 ? ncon : 1) * sizeof(int));
    for (int64_t g = 0; g < ngenes; ++g) {
        size_t nonna = 0;
        for (int c = 0; c < ncon; ++c) {
            const int r = ranks[(size_t)c * ngenes + g];
            if (r != INT_MIN) col[nonna++] = r;
        }
        out[g] = INT_MIN;
        if (nonna) {
            qsort(col, nonna, sizeof(int), port_cmp_int);  /* the element nth_element would put at `jump` */
            out[g] = col[port_define_jump(nonna, prop)];
        }
    }
    free(col);
    return 0;
}

typedef struct {
    double p, e;
} port_pe;
static int port_cmp_pe(const void* a, const void* b) { /* std::pair's operator<: first, then second */
    const port_pe *x = (const port_pe*)a, *y = (const port_pe*)b;
    if (x->p < y->p) return -1;
    if (y->p < x->p) return 1;
    return (x->e > y->e) - (x->e < y->e);
}

/* choose_middle_effect_size (:64-97). pvals / effects: [ncon][ngenes], NaN = NA. */
int oracle_port_choose_middle(int ncon, int64_t ngenes, const double* pvals, const double* effects, double prop, double* out) {
    port_pe* col = (port_pe*)malloc((size_t)(ncon > 0 ? ncon : 1) * sizeof(port_pe));
    for (int64_t g = 0; g < ngenes; ++g) {
        size_t nonna = 0;
        for (int c = 0; c < ncon; ++c) {
            const double p = pvals[(size_t)c * ngenes + g], e = effects[(size_t)c * ngenes + g];
            if (!isnan(p) && !isnan(e)) {
                col[nonna].p = p;
                col[nonna].e = e;
                ++nonna;
            }
        }
        out[g] = NAN;
        if (nonna) {
            qsort(col, nonna, sizeof(port_pe), port_cmp_pe);
            out[g] = col[port_define_jump(nonna, prop)].e;
        }
    }
    free(col);
    return 0;
}
