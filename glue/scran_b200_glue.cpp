// The Rcpp side of the boundary: C++ bodies for the reference's .Call symbols that call the C ABI of
// include/scranb200.h instead of the reference's own native code. Signatures as Rcpp generates them in
// /root/reference/src/RcppExports.cpp:52-89 (compute_blocked_stats_*), :162-186 (overlap_exprs*), so that
// R/RcppExports.R and everything above it stay untouched.
//
// R, Rcpp and beachmat are not installed in this image. The file is compiled in tests/test_glue_cpu.py against
// oracle/stubs/Rcpp.h (the stand-in that also lets the reference's own sources compile), and
// tests/test_gpu_glue.py runs compute_blocked_stats_none, overlap_exprs_paired and overlap_exprs through it on
// the GPU against the reference's object code; with the real Rcpp headers it builds as part of the package
// (INTEGRATION.md section 1).
#include <cstdlib>
#include <stdexcept>
#include <vector>

#include "Rcpp.h"
#include "scranb200.h"

namespace {

// One context per R process, created on first use; the device comes from SCRAN_B200_DEVICE.
scranb200_ctx* sb_ctx() {
    static scranb200_ctx* ctx = nullptr;
    if (!ctx) {
        const char* dev = std::getenv("SCRAN_B200_DEVICE");
        if (scranb200_ctx_create(dev ? std::atoi(dev) : 0, &ctx)) throw std::runtime_error(scranb200_last_error());
    }
    return ctx;
}

void sb_check(int rc) {  // -> an R error through END_RCPP, like the reference's std::runtime_error
    if (rc) throw std::runtime_error(scranb200_last_error());
}

// A device matrix that is freed when the call ends, whichever way it ends.
struct DeviceMat {
    scranb200_mat* m = nullptr;
    int nrow = 0, ncol = 0;
    ~DeviceMat() { scranb200_mat_free(m); }
};

// dgCMatrix or ordinary double matrix -> device matrix: what beachmat::read_lin_block does for the reference
// (src/overlap_exprs.cpp:141, src/compute_blocked_stats.cpp:16).
void sb_upload(Rcpp::RObject x, DeviceMat& out) {
    if (x.isS4() && x.inherits("dgCMatrix")) {
        Rcpp::IntegerVector p = x.slot("p"), i = x.slot("i"), dim = x.slot("Dim");
        Rcpp::NumericVector v = x.slot("x");
        out.nrow = dim[0];
        out.ncol = dim[1];
        sb_check(scranb200_upload_csc32(sb_ctx(), dim[0], dim[1], p.begin(), i.begin(), v.begin(), &out.m));
    } else {
        Rcpp::NumericMatrix d(x);
        out.nrow = (int)d.nrow();
        out.ncol = (int)d.ncol();
        sb_check(scranb200_upload_dense(sb_ctx(), d.nrow(), d.ncol(), d.begin(), &out.m));
    }
}

// The reference's list of 0-based cell-index vectors, flattened.
void flatten_groups(Rcpp::List groups, std::vector<int64_t>& gptr, std::vector<int32_t>& cells) {
    gptr.assign(groups.size() + 1, 0);
    cells.clear();
    for (size_t g = 0; g < groups.size(); ++g) {
        Rcpp::IntegerVector cur(groups[g]);
        cells.insert(cells.end(), cur.begin(), cur.end());
        gptr[g + 1] = (int64_t)cells.size();
    }
    if (cells.empty()) cells.push_back(0);
}

Rcpp::List blocked_stats(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks, int transform, const double* sf, double pseudo) {
    DeviceMat m;
    sb_upload(mat, m);
    if ((int)block.size() != m.ncol) throw std::runtime_error("length of 'block' should be the same as 'ncol(x)'");
    Rcpp::NumericMatrix mean(m.nrow, nblocks), var(m.nrow, nblocks);
    if (transform == 0) {
        sb_check(scranb200_blocked_stats(sb_ctx(), m.m, block.begin(), nblocks, mean.begin(), var.begin(), nullptr));
    } else {
        sb_check(scranb200_blocked_stats_transformed(sb_ctx(), m.m, block.begin(), nblocks, transform, sf, pseudo, mean.begin(),
                                                     var.begin()));
    }
    return Rcpp::List::create(mean, var);  // list(mean, var), like src/compute_blocked_stats.cpp:122
}

}  // namespace

// replaces src/compute_blocked_stats.cpp:205-208
// [[Rcpp::export(rng=false)]]
Rcpp::List compute_blocked_stats_none(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks) {
    return blocked_stats(mat, block, nblocks, 0, nullptr, 1.0);
}

// replaces src/compute_blocked_stats.cpp:150-155
// [[Rcpp::export(rng=false)]]
Rcpp::List compute_blocked_stats_lognorm(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks, Rcpp::NumericVector sf,
                                         double pseudo) {
    return blocked_stats(mat, block, nblocks, 1, sf.begin(), pseudo);
}

// replaces src/compute_blocked_stats.cpp:180-185
// [[Rcpp::export(rng=false)]]
Rcpp::List compute_blocked_stats_norm(Rcpp::RObject mat, Rcpp::IntegerVector block, int nblocks, Rcpp::NumericVector sf) {
    return blocked_stats(mat, block, nblocks, 2, sf.begin(), 1.0);
}

// replaces src/overlap_exprs.cpp:196-220
// [[Rcpp::export(rng=false)]]
Rcpp::NumericMatrix overlap_exprs_paired(Rcpp::RObject exprs, Rcpp::IntegerVector left, Rcpp::IntegerVector right,
                                         Rcpp::List groups, double lfc) {
    std::vector<int64_t> gptr;
    std::vector<int32_t> cells;
    flatten_groups(groups, gptr, cells);
    DeviceMat m;
    sb_upload(exprs, m);
    Rcpp::NumericMatrix out(left.size(), m.nrow);  // npairs x ngenes, like the reference (:206)
    if (left.size() > 0 && m.nrow > 0) {
        sb_check(scranb200_overlap_exprs_paired(sb_ctx(), m.m, left.begin(), right.begin(), (int64_t)left.size(), (int)groups.size(),
                                                gptr.data(), cells.data(), lfc, out.begin(), nullptr));
    }
    return out;
}

// replaces src/overlap_exprs.cpp:140-193: the ragged list of matrices, cut from the flat [ngroups][ngroups][ngenes] arrays
// [[Rcpp::export(rng=false)]]
Rcpp::List overlap_exprs(Rcpp::RObject exprs, Rcpp::List groups, double lfc) {
    std::vector<int64_t> gptr;
    std::vector<int32_t> cells;
    flatten_groups(groups, gptr, cells);
    DeviceMat m;
    sb_upload(exprs, m);
    const size_t G = groups.size(), ngenes = (size_t)m.nrow;
    std::vector<double> u(G * G * ngenes + 1), t(G * G * ngenes + 1);
    sb_check(scranb200_overlap_exprs(sb_ctx(), m.m, (int)G, gptr.data(), cells.data(), lfc, u.data(), t.data()));
    Rcpp::List pout(G), tout(G);
    for (size_t i1 = 0; i1 < G; ++i1) {
        const size_t ncol = lfc == 0 ? i1 : G;  // :153-166
        Rcpp::NumericMatrix pm(ngenes, ncol), tm(ngenes, ncol);
        for (size_t j = 0; j < ncol; ++j) {
            if (lfc != 0 && j == i1) continue;  // the diagonal stays 0
            const size_t at = (i1 * G + j) * ngenes;
            for (size_t g = 0; g < ngenes; ++g) {
                pm.begin()[j * ngenes + g] = u[at + g];
                tm.begin()[j * ngenes + g] = t[at + g];
            }
        }
        pout[i1] = pm;
        tout[i1] = tm;
    }
    return Rcpp::List::create(pout, tout);
}

// The fused scoreMarkers call: replaces the rowBlockApply + computeMinRank section of
// .scoreMarkers (R/scoreMarkers.R:234-256). group_codes = match(groups, universe) - 1; block_codes NULL or
// as.integer(factor(block)) - 1; des_left / des_right = the expanded pairings, 0-based. Returns list(out, full) with
// out = [K][19][ngenes] flat.
// [[Rcpp::export(rng=false)]]
Rcpp::List score_markers_b200(Rcpp::RObject x, Rcpp::IntegerVector group_codes, int ngroups, Rcpp::RObject block_codes,
                              Rcpp::IntegerVector des_left, Rcpp::IntegerVector des_right, double lfc, bool full_stats) {
    scranb200_plan* plan = nullptr;
    Rcpp::IntegerVector blk;
    if (!block_codes.isNULL()) blk = Rcpp::IntegerVector(block_codes);
    sb_check(scranb200_plan_create(sb_ctx(), (int64_t)group_codes.size(), group_codes.begin(), ngroups,
                                   block_codes.isNULL() ? nullptr : blk.begin(), des_left.size() ? des_left.begin() : nullptr,
                                   des_left.size() ? des_right.begin() : nullptr, (int64_t)des_left.size(), lfc, &plan));
    struct PlanGuard {
        scranb200_plan* p;
        ~PlanGuard() { scranb200_plan_destroy(p); }
    } guard{plan};
    int64_t info[5];
    sb_check(scranb200_plan_info(plan, info, 5));
    DeviceMat m;
    sb_upload(x, m);
    Rcpp::NumericVector out((size_t)info[0] * SCRANB200_NCOLS * m.nrow), full(full_stats ? (size_t)3 * info[1] * m.nrow : 0);
    sb_check(scranb200_score_markers(sb_ctx(), m.m, plan, out.begin(), full_stats ? full.begin() : nullptr));
    Rcpp::List res(2);
    res[0] = out;
    res[1] = full;
    return res;
}

// pairwiseWilcox / pairwiseTTests with their p-value layers on the device: everything after
// .setup_groups in R/pairwiseWilcox.R:192-350 / R/pairwiseTTests.R:226-309. One column per ordered pair.
// [[Rcpp::export(rng=false)]]
Rcpp::List pairwise_tests_b200(Rcpp::RObject x, Rcpp::IntegerVector group_codes, int ngroups, Rcpp::RObject block_codes,
                               int nblocks, double lfc, int direction, bool t_test, bool std_lfc) {
    DeviceMat m;
    sb_upload(x, m);
    const size_t npairs = (size_t)ngroups * (ngroups - 1);
    Rcpp::NumericMatrix effect(m.nrow, npairs), logp(m.nrow, npairs), logfdr(m.nrow, npairs);
    Rcpp::IntegerVector blk;
    if (!block_codes.isNULL()) blk = Rcpp::IntegerVector(block_codes);
    const int32_t* bp = block_codes.isNULL() ? nullptr : blk.begin();
    if (t_test) {
        sb_check(scranb200_pairwise_t_tests(sb_ctx(), m.m, group_codes.begin(), ngroups, bp, nblocks, lfc, direction, std_lfc ? 1 : 0,
                                            effect.begin(), logp.begin(), logfdr.begin()));
    } else {
        sb_check(scranb200_pairwise_wilcox(sb_ctx(), m.m, group_codes.begin(), ngroups, bp, nblocks, lfc, direction, effect.begin(),
                                           logp.begin(), logfdr.begin()));
    }
    Rcpp::List res(3);
    res[0] = effect;
    res[1] = logp;
    res[2] = logfdr;
    return res;
}
