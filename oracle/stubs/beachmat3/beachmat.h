// TEST INFRASTRUCTURE ONLY — stand-in for the beachmat3 C++ accessors used by the
// reference's hot path (src/overlap_exprs.cpp:141,171,211; src/compute_blocked_stats.cpp:
// 16,32,54,69). beachmat is a third-party package that is absent from /root/reference and
// from this image; its accessors are value-preserving readers, which is all this provides.
// Row reads go through a row-major copy built once per read_lin_block() call, so the
// timed CPU baseline is not charged beachmat's (slower) per-column binary searches.
#ifndef SCRANB200_ORACLE_STUB_BEACHMAT_H
#define SCRANB200_ORACLE_STUB_BEACHMAT_H

#include "Rcpp.h"

#include <algorithm>
#include <memory>
#include <vector>

namespace beachmat {

struct sparse_index {
    size_t n;
    const double* x;
    const int* i;
};

class lin_matrix {
public:
    explicit lin_matrix(const oracle_stub::MatrixView* m) : view(m) {}
    virtual ~lin_matrix() {}
    size_t get_nrow() const { return view->nrow; }
    size_t get_ncol() const { return view->ncol; }
    bool is_sparse() const { return view->sparse; }

    const double* get_col(size_t c, double* work) {
        if (!view->sparse) return view->x + c * view->nrow;
        std::fill(work, work + view->nrow, 0.0);
        for (int64_t k = view->p[c]; k < view->p[c + 1]; ++k) work[view->i[k]] = view->x[k];
        return work;
    }

    const double* get_row(size_t r, double* work) {
        const size_t nc = view->ncol;
        if (!view->sparse) {
            for (size_t c = 0; c < nc; ++c) work[c] = view->x[c * view->nrow + r];
            return work;
        }
        build_rows();
        std::fill(work, work + nc, 0.0);
        for (int64_t k = rptr[r]; k < rptr[r + 1]; ++k) work[rcol[k]] = rval[k];
        return work;
    }

    const oracle_stub::MatrixView* view;

private:
    void build_rows() {
        if (!rptr.empty()) return;
        const size_t nr = view->nrow, nc = view->ncol;
        rptr.assign(nr + 1, 0);
        const int64_t nnz = view->p[nc];
        for (int64_t k = 0; k < nnz; ++k) ++rptr[view->i[k] + 1];
        for (size_t r = 0; r < nr; ++r) rptr[r + 1] += rptr[r];
        rcol.resize(nnz);
        rval.resize(nnz);
        std::vector<int64_t> fill(rptr.begin(), rptr.end() - 1);
        for (size_t c = 0; c < nc; ++c) {
            for (int64_t k = view->p[c]; k < view->p[c + 1]; ++k) {
                const int64_t at = fill[view->i[k]]++;
                rcol[at] = static_cast<int>(c);
                rval[at] = view->x[k];
            }
        }
    }
    std::vector<int64_t> rptr;
    std::vector<int> rcol;
    std::vector<double> rval;
};

class lin_sparse_matrix : public lin_matrix {
public:
    explicit lin_sparse_matrix(const oracle_stub::MatrixView* m) : lin_matrix(m) {}
    using lin_matrix::get_col;
    sparse_index get_col(size_t c, double* /*xwork*/, int* /*iwork*/) {
        const int64_t lo = view->p[c], hi = view->p[c + 1];
        return sparse_index{static_cast<size_t>(hi - lo), view->x + lo, view->i + lo};
    }
};

inline std::unique_ptr<lin_matrix> read_lin_block(Rcpp::RObject obj) {
    if (obj.mat->sparse) return std::unique_ptr<lin_matrix>(new lin_sparse_matrix(obj.mat));
    return std::unique_ptr<lin_matrix>(new lin_matrix(obj.mat));
}

inline std::unique_ptr<lin_sparse_matrix> promote_to_sparse(std::unique_ptr<lin_matrix>& m) {
    return std::unique_ptr<lin_sparse_matrix>(new lin_sparse_matrix(m->view));
}

}  // namespace beachmat

#endif
