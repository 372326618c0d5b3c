// TEST INFRASTRUCTURE ONLY — not product code, never linked into libscranb200.
//
// A minimal stand-in for the parts of Rcpp that the reference's hot-path translation
// units (src/overlap_exprs.cpp, src/compute_blocked_stats.cpp, src/choose_effect_size.cpp) touch, so that those files
// can be compiled UNMODIFIED, from where they lie under /root/reference, into
// oracle/_ref/liboracle_ref.so (see oracle/Makefile). Rcpp and R are not installed in
// this image. Surface listed in SURVEY.md Appendix B. Everything here is plain
// std::vector-backed storage; the arithmetic stays the reference's own.
#ifndef SCRANB200_ORACLE_STUB_RCPP_H
#define SCRANB200_ORACLE_STUB_RCPP_H

#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#ifndef NA_INTEGER
#define NA_INTEGER INT_MIN
#endif
#define R_NaReal (std::numeric_limits<double>::quiet_NaN())
#define NA_REAL R_NaReal
#define ISNAN(x) (std::isnan(x))
#define ISNA(x) (std::isnan(x))
#ifndef M_LN2
#define M_LN2 0.693147180559945309417232121458
#endif

namespace oracle_stub {
// The matrix an RObject stands for: CSC (dgCMatrix-like) or dense column-major.
struct MatrixView {
    size_t nrow = 0, ncol = 0;
    bool sparse = true;
    const int64_t* p = nullptr;  // ncol + 1
    const int* i = nullptr;      // nnz, 0-based rows, ascending within a column
    const double* x = nullptr;   // nnz (sparse) or nrow*ncol (dense)
};
}  // namespace oracle_stub

namespace Rcpp {

class IntegerVector;
class NumericVector;

// What x.slot("p") yields: convertible to the vector type the caller asks for.
struct SlotProxy {
    const oracle_stub::MatrixView* mat;
    const char* name;
    operator IntegerVector() const;
    operator NumericVector() const;
};

class RObject {
public:
    RObject() {}
    explicit RObject(const oracle_stub::MatrixView* m) : mat(m) {}
    const oracle_stub::MatrixView* mat = nullptr;
    std::shared_ptr<std::vector<int>> ivec;  // an integer vector held as a generic object (glue: block codes)
    // the part of the S4 / SEXP surface that glue/scran_b200_glue.cpp uses: a sparse view is a "dgCMatrix"
    bool isNULL() const { return mat == nullptr && !ivec; }
    bool isS4() const { return mat != nullptr && mat->sparse; }
    bool inherits(const char*) const { return mat != nullptr && mat->sparse; }
    SlotProxy slot(const char* name) const { return SlotProxy{mat, name}; }
};

template <typename T>
class StubVector {
public:
    StubVector() : store(std::make_shared<std::vector<T>>()) {}
    explicit StubVector(size_t n) : store(std::make_shared<std::vector<T>>(n)) {}
    StubVector(size_t n, T fill) : store(std::make_shared<std::vector<T>>(n, fill)) {}
    StubVector(const T* b, const T* e) : store(std::make_shared<std::vector<T>>(b, e)) {}
    size_t size() const { return store->size(); }
    T* begin() { return store->data(); }
    T* end() { return store->data() + store->size(); }
    const T* begin() const { return store->data(); }
    const T* end() const { return store->data() + store->size(); }
    T& operator[](size_t k) { return (*store)[k]; }
    const T& operator[](size_t k) const { return (*store)[k]; }
    std::shared_ptr<std::vector<T>> store;
};

class NumericMatrix {
public:
    typedef double* iterator;
    struct Column {
        double* b;
        double* e;
        double* begin() const { return b; }
        double* end() const { return e; }
    };
    NumericMatrix() : nr(0), nc(0), store(std::make_shared<std::vector<double>>()) {}
    NumericMatrix(size_t r, size_t c) : nr(r), nc(c), store(std::make_shared<std::vector<double>>(r * c, 0.0)) {}
    explicit NumericMatrix(const RObject& x) : nr(0), nc(0), store(std::make_shared<std::vector<double>>()) {
        if (!x.mat || x.mat->sparse) throw std::runtime_error("stub: not an ordinary matrix");
        nr = x.mat->nrow;
        nc = x.mat->ncol;
        store = std::make_shared<std::vector<double>>(x.mat->x, x.mat->x + nr * nc);
    }
    double* begin() { return store->data(); }
    double* end() { return store->data() + store->size(); }
    Column column(size_t j) { return Column{store->data() + j * nr, store->data() + (j + 1) * nr}; }
    size_t nrow() const { return nr; }
    size_t ncol() const { return nc; }
    size_t nr, nc;
    std::shared_ptr<std::vector<double>> store;
};

class List;

// One slot of a List: may hold an integer vector, a numeric matrix or a nested list.
struct ListSlot {
    std::shared_ptr<std::vector<int>> ivec;
    std::shared_ptr<std::vector<double>> nvec;
    std::shared_ptr<NumericMatrix> nmat;
    std::shared_ptr<List> sub;
    ListSlot& operator=(const NumericMatrix& m) {
        nmat = std::make_shared<NumericMatrix>(m);
        return *this;
    }
    ListSlot& operator=(const List& l);
    ListSlot& operator=(const NumericVector& v);
};

class IntegerVector : public StubVector<int> {
public:
    IntegerVector() {}
    explicit IntegerVector(size_t n) : StubVector<int>(n) {}
    IntegerVector(size_t n, int fill) : StubVector<int>(n, fill) {}
    IntegerVector(const int* b, const int* e) : StubVector<int>(b, e) {}
    IntegerVector(const ListSlot& s) {
        if (!s.ivec) throw std::runtime_error("stub: list slot holds no integer vector");
        store = s.ivec;
    }
    explicit IntegerVector(const RObject& x) {
        if (!x.ivec) throw std::runtime_error("stub: object holds no integer vector");
        store = x.ivec;
    }
};

class NumericVector : public StubVector<double> {
public:
    NumericVector() {}
    explicit NumericVector(size_t n) : StubVector<double>(n) {}
    NumericVector(size_t n, double fill) : StubVector<double>(n, fill) {}
    NumericVector(const double* b, const double* e) : StubVector<double>(b, e) {}
    NumericVector(const ListSlot& s) {
        if (!s.nvec) throw std::runtime_error("stub: list slot holds no numeric vector");
        store = s.nvec;
    }
};

class List {
public:
    List() {}
    explicit List(size_t n) : slots(n) {}
    size_t size() const { return slots.size(); }
    ListSlot& operator[](size_t k) { return slots[k]; }
    const ListSlot& operator[](size_t k) const { return slots[k]; }
    static List create(const NumericMatrix& a, const NumericMatrix& b) {
        List out(2);
        out[0] = a;
        out[1] = b;
        return out;
    }
    static List create(const List& a, const List& b) {
        List out(2);
        out[0] = a;
        out[1] = b;
        return out;
    }
    std::vector<ListSlot> slots;
};

inline ListSlot& ListSlot::operator=(const List& l) {
    sub = std::make_shared<List>(l);
    return *this;
}
inline ListSlot& ListSlot::operator=(const NumericVector& v) {
    nvec = v.store;
    return *this;
}

// dgCMatrix slots of a sparse view: @p (converted to R's int), @i, @Dim as integer vectors, @x as numeric.
inline SlotProxy::operator IntegerVector() const {
    if (!mat) throw std::runtime_error("stub: no matrix behind the slot");
    const std::string n(name);
    if (n == "Dim") {
        IntegerVector d(2);
        d[0] = (int)mat->nrow;
        d[1] = (int)mat->ncol;
        return d;
    }
    if (n == "p") {
        IntegerVector p(mat->ncol + 1);
        for (size_t k = 0; k <= mat->ncol; ++k) {
            if (mat->p[k] > INT_MAX) throw std::runtime_error("stub: a dgCMatrix holds at most 2^31 - 1 entries");
            p[k] = (int)mat->p[k];
        }
        return p;
    }
    if (n == "i") return IntegerVector(mat->i, mat->i + mat->p[mat->ncol]);
    throw std::runtime_error("stub: unknown integer slot");
}
inline SlotProxy::operator NumericVector() const {
    if (!mat || std::string(name) != "x") throw std::runtime_error("stub: unknown numeric slot");
    return NumericVector(mat->x, mat->x + mat->p[mat->ncol]);
}

}  // namespace Rcpp

#endif
