// TEST INFRASTRUCTURE ONLY. extern "C" driver around the reference's own object code.
//
// Links against /root/reference/src/overlap_exprs.cpp, compute_blocked_stats.cpp and choose_effect_size.cpp,
// compiled unmodified against oracle/stubs (see oracle/Makefile), and exposes flat-array
// entry points so that Python (ctypes) can run the reference arithmetic on CSC / dense
// inputs. Output goes to oracle/_ref/liboracle_ref.so, which is git-ignored.
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load it.
#include "Rcpp.h"
#include "beachmat3/beachmat.h"

#include <cstring>
#include <string>

// Exported by the reference translation units (src/compute_blocked_stats.cpp:150,180,205;
// src/overlap_exprs.cpp:140,196).
Rcpp::List compute_blocked_stats_none(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks);
Rcpp::List compute_blocked_stats_lognorm(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks,
                                         Rcpp::NumericVector sf, double pseudo);
Rcpp::List compute_blocked_stats_norm(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks,
                                      Rcpp::NumericVector sf);
Rcpp::List overlap_exprs(Rcpp::RObject exprs, Rcpp::List groups, double lfc);
Rcpp::NumericMatrix overlap_exprs_paired(Rcpp::RObject exprs, Rcpp::IntegerVector left,
                                         Rcpp::IntegerVector right, Rcpp::List groups, double lfc);

// src/choose_effect_size.cpp:35, 64 (the native helpers of combineMarkers)
Rcpp::IntegerVector compute_Top_statistic_from_ranks(Rcpp::List Ranks, double prop);
Rcpp::NumericVector choose_middle_effect_size(Rcpp::List Pvals, Rcpp::List Effects, double prop);

namespace {
thread_local std::string g_err;

oracle_stub::MatrixView make_view(int64_t nrow, int64_t ncol, const int64_t* p, const int* i,
                                  const double* x, int sparse) {
    oracle_stub::MatrixView v;
    v.nrow = static_cast<size_t>(nrow);
    v.ncol = static_cast<size_t>(ncol);
    v.sparse = sparse != 0;
    v.p = p;
    v.i = i;
    v.x = x;
    return v;
}

Rcpp::List make_groups(int ngroups, const int64_t* gptr, const int* gcells) {
    Rcpp::List groups(ngroups);
    for (int g = 0; g < ngroups; ++g) {
        groups[g].ivec = std::make_shared<std::vector<int>>(gcells + gptr[g], gcells + gptr[g + 1]);
    }
    return groups;
}
}  // namespace

extern "C" {

const char* oracle_ref_last_error() { return g_err.c_str(); }

// transform: 0 none, 1 lognorm (sf, pseudo), 2 norm (sf). mean/var: nrow x nblocks column-major.
int oracle_ref_blocked_stats(int64_t nrow, int64_t ncol, const int64_t* p, const int* i, const double* x,
                             int sparse, const int* block, int nblocks, int transform, const double* sf,
                             double pseudo, double* mean_out, double* var_out) {
    try {
        oracle_stub::MatrixView v = make_view(nrow, ncol, p, i, x, sparse);
        Rcpp::IntegerVector b(block, block + ncol);
        Rcpp::List res;
        if (transform == 0) {
            res = compute_blocked_stats_none(Rcpp::RObject(&v), b, nblocks);
        } else if (transform == 1) {
            res = compute_blocked_stats_lognorm(Rcpp::RObject(&v), b, nblocks, Rcpp::NumericVector(sf, sf + ncol), pseudo);
        } else {
            res = compute_blocked_stats_norm(Rcpp::RObject(&v), b, nblocks, Rcpp::NumericVector(sf, sf + ncol));
        }
        const size_t n = static_cast<size_t>(nrow) * nblocks;
        std::memcpy(mean_out, res[0].nmat->begin(), n * sizeof(double));
        std::memcpy(var_out, res[1].nmat->begin(), n * sizeof(double));
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// left/right are 1-based group indices, as at src/overlap_exprs.cpp:215.
// out: npairs x nrow column-major (pairs contiguous within a gene).
int oracle_ref_overlap_paired(int64_t nrow, int64_t ncol, const int64_t* p, const int* i, const double* x,
                              int sparse, const int* left, const int* right, int64_t npairs, int ngroups,
                              const int64_t* gptr, const int* gcells, double lfc, double* out) {
    try {
        oracle_stub::MatrixView v = make_view(nrow, ncol, p, i, x, sparse);
        Rcpp::NumericMatrix res = overlap_exprs_paired(Rcpp::RObject(&v), Rcpp::IntegerVector(left, left + npairs),
                                                       Rcpp::IntegerVector(right, right + npairs),
                                                       make_groups(ngroups, gptr, gcells), lfc);
        std::memcpy(out, res.begin(), static_cast<size_t>(npairs) * nrow * sizeof(double));
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// Flattened form of the ragged list-of-matrices the reference returns (src/overlap_exprs.cpp:153-166):
// u_out/t_out are [ngroups][ngroups][nrow] (entry (i, j) holds column j of matrix i); entries the
// reference does not have (j >= i when lfc == 0) are set to NaN.
int oracle_ref_overlap(int64_t nrow, int64_t ncol, const int64_t* p, const int* i, const double* x, int sparse,
                       int ngroups, const int64_t* gptr, const int* gcells, double lfc, double* u_out,
                       double* t_out) {
    try {
        oracle_stub::MatrixView v = make_view(nrow, ncol, p, i, x, sparse);
        Rcpp::List res = overlap_exprs(Rcpp::RObject(&v), make_groups(ngroups, gptr, gcells), lfc);
        const size_t total = static_cast<size_t>(ngroups) * ngroups * nrow;
        for (size_t k = 0; k < total; ++k) u_out[k] = t_out[k] = R_NaReal;
        for (int which = 0; which < 2; ++which) {
            Rcpp::List& per_group = *res[which].sub;
            double* dst = which == 0 ? u_out : t_out;
            for (int gi = 0; gi < ngroups; ++gi) {
                Rcpp::NumericMatrix& m = *per_group[gi].nmat;
                for (size_t j = 0; j < m.ncol(); ++j) {
                    std::memcpy(dst + (static_cast<size_t>(gi) * ngroups + j) * nrow, m.begin() + j * m.nrow(),
                                static_cast<size_t>(nrow) * sizeof(double));
                }
            }
        }
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// ranks: [ncon][ngenes] integers (NA_INTEGER = NA), one rank vector per contrast. out: ngenes.
int oracle_ref_top_from_ranks(int ncon, int64_t ngenes, const int* ranks, double prop, int* out) {
    try {
        Rcpp::List lst(ncon);
        for (int c = 0; c < ncon; ++c) lst[c].ivec = std::make_shared<std::vector<int>>(ranks + c * ngenes, ranks + (c + 1) * ngenes);
        Rcpp::IntegerVector res = compute_Top_statistic_from_ranks(lst, prop);
        std::memcpy(out, res.begin(), static_cast<size_t>(ngenes) * sizeof(int));
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// pvals / effects: [ncon][ngenes] doubles (NaN = NA). out: ngenes.
int oracle_ref_choose_middle(int ncon, int64_t ngenes, const double* pvals, const double* effects, double prop, double* out) {
    try {
        Rcpp::List lp(ncon), le(ncon);
        for (int c = 0; c < ncon; ++c) {
            lp[c].nvec = std::make_shared<std::vector<double>>(pvals + c * ngenes, pvals + (c + 1) * ngenes);
            le[c].nvec = std::make_shared<std::vector<double>>(effects + c * ngenes, effects + (c + 1) * ngenes);
        }
        Rcpp::NumericVector res = choose_middle_effect_size(lp, le, prop);
        std::memcpy(out, res.begin(), static_cast<size_t>(ngenes) * sizeof(double));
        return 0;
    } catch (std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

}  // extern "C"
