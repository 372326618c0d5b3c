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
         double* __restrict__ mean, double* __restrict__ var, double* __restrict__ ndet, double* __restrict__ vmin,
         double* __restrict__ vmax) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)ngenes * C) return;
    const int g = (int)(t % ngenes);
    const int dc = (int)(t / ngenes);
    const int64_t base = row_ptr[g];
    const int64_t lo = base + seg_off[(size_t)dc * ngenes + g];
    const int64_t hi = base + seg_off[(size_t)(dc + 1) * ngenes + g];

    double M = 0.0, S = 0.0, nz = 0.0, det = 0.0;
    double lo_v = __longlong_as_double(0x7ff0000000000000ll), hi_v = -lo_v;  // range of the non-zero values (for K3)
    auto update = [&](double v) {
        if (v != 0.0) {
            lo_v = fmin(lo_v, v);
            hi_v = fmax(hi_v, v);
        }
        nz += 1.0;
        const double delta = v - M;
        M += delta / nz;
        S += delta * (v - M);
        det += (v > 0.0) ? 1.0 : 0.0;
    };
    int64_t k = lo;
    // The update chain is sequential by definition; the loads are not. Eight values are fetched
    // ahead of the arithmetic so that every thread keeps several memory requests in flight.
    for (; k + 8 <= hi; k += 8) {
        double v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = vals[k + q];
#pragma unroll
        for (int q = 0; q < 8; ++q) update(v[q]);
    }
    for (; k < hi; ++k) update(vals[k]);

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
    vmin[(size_t)dc * ngenes + g] = lo_v;
    vmax[(size_t)dc * ngenes + g] = hi_v;
}

}  // namespace

void run_stats(scranb200_ctx* ctx, Transposed& T, const PairLayout& L, double* d_mean, double* d_var, double* d_ndet) {
    const int64_t total = (int64_t)T.ngenes * T.C;
    if (total == 0) return;
    ctx->vmin.reserve((size_t)total * 8);
    ctx->vmax.reserve((size_t)total * 8);
    T.vmin = ctx->vmin.as<double>();
    T.vmax = ctx->vmax.as<double>();
    SB_LAUNCH(ctx, k2_stats, (unsigned)ceil_div(total, 256), 256, 0, T.vals, T.row_ptr, T.seg_off, T.ngenes, T.C,
              L.d_caller_of_dev, L.d_ncells_dev, d_mean, d_var, d_ndet, ctx->vmin.as<double>(), ctx->vmax.as<double>());
}

}  // namespace sb
