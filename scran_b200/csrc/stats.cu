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

// A warp takes 32 consecutive (gene, combination) segments: they lie next to each other in memory (mostly in
// one gene's row), and consecutive warps take consecutive stretches of memory (a gene-major mapping would spread
// a warp's 32 streams over 32 rows, hundreds of kilobytes apart: a different DRAM page and TLB entry each).
// Segments differ in length by an order of magnitude between genes, so nothing waits for anything: no
// CTA-wide barrier (round 2's staging of the results through shared memory, for coalesced stores, kept seven
// warps of a CTA waiting for the eighth: 67 % of the kernel's stall samples), small CTAs that retire and are
// replaced quickly, results stored straight from the lanes (0.5 GB in 8-byte pieces that L2 merges).
#ifndef SB_K2_WARPS
#define SB_K2_WARPS 2
#endif
#ifndef SB_K2_MINB
#define SB_K2_MINB 1
#endif
constexpr int K2_WARPS = SB_K2_WARPS;

// Four consecutive doubles with one 256-bit load (32-byte aligned address; read-only data).
__device__ __forceinline__ void load4(const double* p, double& a, double& b, double& c, double& d) {
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}

__global__ void __launch_bounds__(K2_WARPS * 32, SB_K2_MINB)
k2_stats(const double* __restrict__ vals, const int64_t* __restrict__ row_ptr, const uint32_t* __restrict__ seg_off,
         int ngenes, int C, const int32_t* __restrict__ caller_of_dev, const int64_t* __restrict__ ncells_dev,
         double* __restrict__ mean, double* __restrict__ var, double* __restrict__ ndet, double* __restrict__ vmin,
         double* __restrict__ vmax) {
    const int lane = threadIdx.x & 31;
    // 32 consecutive segments in memory order (gene-major, then combination): with fewer than 32 combinations a
    // warp spans several genes, so all lanes work (modelGeneVar's blocks: 8 combinations)
    const int64_t sid = ((int64_t)blockIdx.x * K2_WARPS + (threadIdx.x >> 5)) * 32 + lane;
    const int g = (int)(sid / C);
    const int dc = (int)(sid % C);
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
        // The 32 lanes of a warp read 32 different segments, so every load instruction costs the L1 one
        // lookup per lane: the values are therefore fetched two per load (16 bytes, after one value on
        // its own if the segment starts at an odd position), which halves the lookups - the kernel's bound.
        // (Default: four per load, 32 bytes - sm_100's 256-bit global load - after up to three single values.)
#ifdef SB_K2_LD128
        if ((k & 1) && k < hi) {
            update(vals[k]);
            ++k;
        }
        const double2* __restrict__ vals2 = reinterpret_cast<const double2*>(vals);
        double2 v[4];
        bool have = k + 8 <= hi;
        if (have) {
#pragma unroll
            for (int q = 0; q < 4; ++q) v[q] = vals2[(k >> 1) + q];
        }
        while (have) {
            double2 w[4];
            const bool more = k + 16 <= hi;
            if (more) {
#pragma unroll
                for (int q = 0; q < 4; ++q) w[q] = vals2[(k >> 1) + 4 + q];
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                update(v[q].x);
                update(v[q].y);
            }
            k += 8;
            have = more;
#pragma unroll
            for (int q = 0; q < 4; ++q) v[q] = w[q];
        }
#else
        while ((k & 3) && k < hi) {  // up to a 32-byte boundary
            update(vals[k]);
            ++k;
        }
#ifndef SB_K2_BATCH
#define SB_K2_BATCH 8
#endif
        constexpr int NB = SB_K2_BATCH;  // values per batch (a multiple of 4)
        double v[NB];
        bool have = k + NB <= hi;
        if (have) {
#pragma unroll
            for (int q = 0; q < NB; q += 4) load4(vals + k + q, v[q], v[q + 1], v[q + 2], v[q + 3]);
        }
        while (have) {
            double w[NB];
            const bool more = k + 2 * NB <= hi;
            if (more) {
#pragma unroll
                for (int q = 0; q < NB; q += 4) load4(vals + k + NB + q, w[q], w[q + 1], w[q + 2], w[q + 3]);
            }
#pragma unroll
            for (int q = 0; q < NB; ++q) update(v[q]);
            k += NB;
            have = more;
#pragma unroll
            for (int q = 0; q < NB; ++q) v[q] = w[q];
        }
#endif
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
        const size_t o = (size_t)caller_of_dev[dc] * ngenes + g;
        if (mean) mean[o] = M;
        if (var) var[o] = S;
        if (ndet) ndet[o] = det;
        vmin[(size_t)dc * ngenes + g] = lo_v;
        vmax[(size_t)dc * ngenes + g] = hi_v;
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
    const int64_t warps = ceil_div(total, 32);
    SB_LAUNCH(ctx, k2_stats, (unsigned)ceil_div(warps, K2_WARPS), K2_WARPS * 32, 0, T.vals, T.row_ptr, T.seg_off, T.ngenes, T.C,
              L.d_caller_of_dev, L.d_ncells_dev, d_mean, d_var, d_ndet, ctx->vmin.as<double>(), ctx->vmax.as<double>());
}

}  // namespace sb
