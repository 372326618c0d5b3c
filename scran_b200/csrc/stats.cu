// K2: per (gene, combination) mean, unbiased variance and detected count.
//
// Follows src/compute_blocked_stats.cpp:38-120 operation for operation: the running update
// (:57-67) over the stored values of the segment in ascending cell order (the order K1
// produced), then the closed-form zero fill (:96-114) and the division by N - 1 (:116-119).
// The library is compiled with -fmad=false, and fp64 '/' is IEEE-correct on the device, so the
// results are bit-identical to the reference built with R's default flags (no FMA on x86-64).
// One thread owns one segment: 13.4M independent segments at config 3, no atomics, no
// reduction-order question.
#include "kernels.h"

namespace sb {

namespace {

__global__ void __launch_bounds__(256)
k2_stats(const double* __restrict__ vals, const int64_t* __restrict__ row_ptr, const uint32_t* __restrict__ seg_off,
         int ngenes, int C, const int32_t* __restrict__ caller_of_dev, const int64_t* __restrict__ ncells_dev,
         double* __restrict__ mean, double* __restrict__ var, double* __restrict__ ndet) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)ngenes * C) return;
    const int g = (int)(t % ngenes);
    const int dc = (int)(t / ngenes);
    const int64_t base = row_ptr[g];
    const int64_t lo = base + seg_off[(size_t)dc * ngenes + g];
    const int64_t hi = base + seg_off[(size_t)(dc + 1) * ngenes + g];

    double M = 0.0, S = 0.0, nz = 0.0, det = 0.0;
    for (int64_t k = lo; k < hi; ++k) {
        const double v = vals[k];
        nz += 1.0;
        const double delta = v - M;
        M += delta / nz;
        S += delta * (v - M);
        det += (v > 0.0) ? 1.0 : 0.0;
    }

    const int64_t n = ncells_dev[dc];
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    if (n <= 0) M = nan;
    if (n <= 1) {
        S = nan;
    } else {
        const double total = (double)n;
        const double ratio = nz / total;
        S += M * M * ratio * (total - nz);
        M *= ratio;
        S /= (double)(n - 1);
    }
    const size_t at = (size_t)caller_of_dev[dc] * ngenes + g;
    if (mean) mean[at] = M;
    if (var) var[at] = S;
    if (ndet) ndet[at] = det;
}

}  // namespace

void run_stats(scranb200_ctx* ctx, const Transposed& T, const PairLayout& L, double* d_mean, double* d_var,
               double* d_ndet) {
    const int64_t total = (int64_t)T.ngenes * T.C;
    if (total == 0) return;
    SB_LAUNCH(ctx, k2_stats, (unsigned)ceil_div(total, 256), 256, 0, T.vals, T.row_ptr, T.seg_off, T.ngenes, T.C,
              L.d_caller_of_dev, L.d_ncells_dev, d_mean, d_var, d_ndet);
}

}  // namespace sb
