// TEST SCAFFOLDING: flat-array entry points around glue/scran_b200_glue.cpp, which is compiled here against
// the stand-in Rcpp.h of oracle/stubs (R and Rcpp are not installed). Built by tests/test_glue_cpu.py.
#include <cstring>
#include <string>

#include "Rcpp.h"

Rcpp::List compute_blocked_stats_none(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks);
Rcpp::List compute_blocked_stats_lognorm(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks, Rcpp::NumericVector sf, double pseudo);
Rcpp::NumericMatrix overlap_exprs_paired(Rcpp::RObject exprs, Rcpp::IntegerVector left, Rcpp::IntegerVector right, Rcpp::List groups,
                                         double lfc);
Rcpp::List overlap_exprs(Rcpp::RObject exprs, Rcpp::List groups, double lfc);
Rcpp::List score_markers_b200(Rcpp::RObject x, Rcpp::IntegerVector group_codes, int ngroups, Rcpp::RObject block_codes,
                              Rcpp::IntegerVector des_left, Rcpp::IntegerVector des_right, double lfc, bool full_stats);

namespace {
thread_local std::string g_err;
oracle_stub::MatrixView view(int64_t nrow, int64_t ncol, const int64_t* p, const int* i, const double* x, int sparse) {
    oracle_stub::MatrixView v;
    v.nrow = (size_t)nrow;
    v.ncol = (size_t)ncol;
    v.sparse = sparse != 0;
    v.p = p;
    v.i = i;
    v.x = x;
    return v;
}
Rcpp::List groups_of(int ngroups, const int64_t* gptr, const int* gcells) {
    Rcpp::List groups(ngroups);
    for (int g = 0; g < ngroups; ++g) groups[g].ivec = std::make_shared<std::vector<int>>(gcells + gptr[g], gcells + gptr[g + 1]);
    return groups;
}
}  // namespace

extern "C" {
const char* glue_last_error() { return g_err.c_str(); }

int glue_blocked_stats(int64_t nrow, int64_t ncol, const int64_t* p, const int* i, const double* x, int sparse, const int* block,
                       int nblocks, const double* sf, double pseudo, double* mean, double* var) {
    try {
        oracle_stub::MatrixView v = view(nrow, ncol, p, i, x, sparse);
        Rcpp::IntegerVector b(block, block + ncol);
        Rcpp::List res = sf ? compute_blocked_stats_lognorm(Rcpp::RObject(&v), b, nblocks, Rcpp::NumericVector(sf, sf + ncol), pseudo)
                            : compute_blocked_stats_none(Rcpp::RObject(&v), b, nblocks);
        std::memcpy(mean, res[0].nmat->begin(), (size_t)nrow * nblocks * 8);
        std::memcpy(var, res[1].nmat->begin(), (size_t)nrow * nblocks * 8);
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

int glue_overlap_paired(int64_t nrow, int64_t ncol, const int64_t* p, const int* i, const double* x, int sparse, const int* left,
                        const int* right, int64_t npairs, int ngroups, const int64_t* gptr, const int* gcells, double lfc, double* out) {
    try {
        oracle_stub::MatrixView v = view(nrow, ncol, p, i, x, sparse);
        Rcpp::NumericMatrix res = overlap_exprs_paired(Rcpp::RObject(&v), Rcpp::IntegerVector(left, left + npairs),
                                                       Rcpp::IntegerVector(right, right + npairs), groups_of(ngroups, gptr, gcells), lfc);
        std::memcpy(out, res.begin(), (size_t)npairs * nrow * 8);
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// u / t: [ngroups][ngroups][nrow], NaN where the reference's list of matrices has no column
int glue_overlap(int64_t nrow, int64_t ncol, const int64_t* p, const int* i, const double* x, int sparse, int ngroups,
                 const int64_t* gptr, const int* gcells, double lfc, double* u, double* t) {
    try {
        oracle_stub::MatrixView v = view(nrow, ncol, p, i, x, sparse);
        Rcpp::List res = overlap_exprs(Rcpp::RObject(&v), groups_of(ngroups, gptr, gcells), lfc);
        const size_t total = (size_t)ngroups * ngroups * nrow;
        for (size_t k = 0; k < total; ++k) u[k] = t[k] = R_NaReal;
        for (int which = 0; which < 2; ++which) {
            Rcpp::List& per = *res[which].sub;
            double* dst = which == 0 ? u : t;
            for (int gi = 0; gi < ngroups; ++gi) {
                Rcpp::NumericMatrix& m = *per[gi].nmat;
                for (size_t j = 0; j < m.ncol(); ++j)
                    std::memcpy(dst + ((size_t)gi * ngroups + j) * nrow, m.begin() + j * m.nrow(), (size_t)nrow * 8);
            }
        }
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// groups: 0-based codes per cell; block: codes or null; all ordered pairs. out: [ngroups][19][nrow].
int glue_score_markers(int64_t nrow, int64_t ncol, const int64_t* p, const int* i, const double* x, int sparse, const int* groups,
                       int ngroups, const int* block, double lfc, double* out) {
    try {
        oracle_stub::MatrixView v = view(nrow, ncol, p, i, x, sparse);
        Rcpp::RObject blk;
        if (block) blk.ivec = std::make_shared<std::vector<int>>(block, block + ncol);
        Rcpp::List res = score_markers_b200(Rcpp::RObject(&v), Rcpp::IntegerVector(groups, groups + ncol), ngroups, blk,
                                            Rcpp::IntegerVector(), Rcpp::IntegerVector(), lfc, false);
        std::memcpy(out, res[0].nvec->data(), res[0].nvec->size() * 8);
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}
}
