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

// A CTA takes K2_GENES genes x 32 consecutive combinations: warp w owns gene g0 + w and its lanes
// the 32 combinations, whose segments lie next to each other in the gene's row. The 32 streams of
// a warp therefore stay inside one stretch of memory of 32 segments (a gene-major mapping would
// spread them over 32 rows, hundreds of kilobytes apart: a different DRAM page and TLB entry each).
// The results go through shared memory so that each output array is written K2_GENES consecutive
// genes at a time.
constexpr int K2_GENES = 8;

__global__ void __launch_bounds__(K2_GENES * 32)
k2_stats(const double* __restrict__ vals, const int64_t* __restrict__ row_ptr, const uint32_t* __restrict__ seg_off,
         int ngenes, int C, const int32_t* __restrict__ caller_of_dev, const int64_t* __restrict__ ncells_dev,
         double* __restrict__ mean, double* __restrict__ var, double* __restrict__ ndet, double* __restrict__ vmin,
         double* __restrict__ vmax) {
    __shared__ double stage[5][32][K2_GENES + 1];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int g = blockIdx.x * K2_GENES + w;
    const int dc = blockIdx.y * 32 + lane;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    if (g < ngenes && dc < C) {
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
        // The update chain is sequential by definition (a division per value); the loads are not. The
        // next eight values are requested before the current eight are folded in, so that a thread
        // has memory requests in flight all the time and not only between two stretches of arithmetic.
        double v[8];
        bool have = k + 8 <= hi;
        if (have) {
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = vals[k + q];
        }
        while (have) {
            double w[8];
            const bool more = k + 16 <= hi;
            if (more) {
#pragma unroll
                for (int q = 0; q < 8; ++q) w[q] = vals[k + 8 + q];
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) update(v[q]);
            k += 8;
            have = more;
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = w[q];
        }
        for (; k < hi; ++k) update(vals[k]);

        const int64_t n = ncells_dev[dc];
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
        stage[0][lane][w] = M;
        stage[1][lane][w] = S;
        stage[2][lane][w] = det;
        stage[3][lane][w] = lo_v;
        stage[4][lane][w] = hi_v;
    }
    __syncthreads();
    const int g0 = blockIdx.x * K2_GENES;
    for (int idx = threadIdx.x; idx < 5 * 32 * K2_GENES; idx += blockDim.x) {
        const int arr = idx / (32 * K2_GENES), rest = idx % (32 * K2_GENES);
        const int c = rest / K2_GENES, gi = rest % K2_GENES;
        const int odc = blockIdx.y * 32 + c;
        if (odc >= C || g0 + gi >= ngenes) continue;
        const double x = stage[arr][c][gi];
        if (arr < 3) {
            double* out = arr == 0 ? mean : (arr == 1 ? var : ndet);
            if (out) out[(size_t)caller_of_dev[odc] * ngenes + g0 + gi] = x;
        } else {
            (arr == 3 ? vmin : vmax)[(size_t)odc * ngenes + g0 + gi] = x;
        }
    }
}

}  // namespace

void run_stats(scranb200_ctx* ctx, Transposed& T, const PairLayout& L, double* d_mean, double* d_var, double* d_ndet) {
    const int64_t total = (int64_t)T.ngenes * T.C;
    if (total == 0) return;
    ctx->vmin.reserve((size_t)total * 8);
    ctx->vmax.reserve((size_t)total * 8);
    T.vmin = ctx->vmin.as<double>();
    T.vmax = ctx->vmax.as<double>();
    const dim3 grid((unsigned)ceil_div(T.ngenes, K2_GENES), (unsigned)ceil_div(T.C, 32));
    SB_LAUNCH(ctx, k2_stats, grid, K2_GENES * 32, 0, T.vals, T.row_ptr, T.seg_off, T.ngenes, T.C,
              L.d_caller_of_dev, L.d_ncells_dev, d_mean, d_var, d_ndet, ctx->vmin.as<double>(), ctx->vmax.as<double>());
}

}  // namespace sb
