// K4: the three effect sizes per block-level pair, folded straight into the block-averaged
// comparison they belong to, and the self./other. columns.
//
//   Cohen's d        R/scoreMarkers.R:417-426   (mL - mR - lfc) / sqrt((vL + vR) / 2), 0 if the
//                                               numerator is exactly 0
//   AUC              R/scoreMarkers.R:442-445   U / (nL * nR)
//   logFC.detected   R/scoreMarkers.R:428-440   log2 of the ratio of shrunken detected fractions
//   block average    R/scoreMarkers.R:479-492   matrixStats::rowWeightedMeans(na.rm=TRUE) with
//                                               w = nL * nR (:369): weights divided by their sum
//                                               first, then sum(w_i * e_i); the reference's row
//                                               sum runs in long double, here in double-double
//   self./other.     R/scoreMarkers.R:395-411   per-group average / detected fraction
//
// The reference materialises genes x block-pairs matrices for all three effects (7.9 GB at
// config 3); here each (gene, comparison) thread consumes its block-level values on the fly.
#include "score.h"

namespace sb {

namespace {

struct DD {
    double hi, lo;
};
__device__ inline void dd_add(DD& s, double x) {  // s += x, error-free (Knuth two-sum)
    const double t = s.hi + x;
    if (isinf(t) || isnan(t)) {  // infinite effects (zero variance) propagate like a plain sum
        s.hi = t;
        s.lo = 0.0;
        return;
    }
    const double bp = t - s.hi;
    const double err = (s.hi - (t - bp)) + (x - bp);
    s.hi = t;
    s.lo += err;
}
__device__ inline double dd_value(const DD& s) { return s.hi + s.lo; }

__device__ inline double effect_cohen(double mL, double mR, double vL, double vR, double lfc) {
    const double delta = mL - mR - lfc;
    if (delta == 0.0) return 0.0;
    return delta / sqrt((vL + vR) / 2.0);
}

__device__ inline double effect_detected(double dL, double dR, double nL, double nR) {
    const double mn = fmin(nL, nR);
    const double pl = 1.0 * nL / mn, pr = 1.0 * nR / mn;
    const double lp = (dL + pl) / (nL + pl * 2.0);
    const double rp = (dR + pr) / (nR + pr * 2.0);
    return log2(lp / rp);
}

struct EffParams {
    int ngenes, Q;
    int64_t ld, row0;
    const double *mean, *var, *ndet, *u;  // [C][ngenes], u: [ngenes][npairs]
    int64_t npairs;
    const int32_t *left, *right, *avg_ptr, *avg_idx;
    const double* ncells;
    double lfc;
    double* eff;  // [3][Q][ld]
};

// (Five CTAs per SM: the kernel waits on loads, so residency beats the few spilled registers:
// 2.0 -> 1.4 ms at the config-3 shard against the compiler's own 114 registers.)
__global__ void __launch_bounds__(256, 5) k4_effects(EffParams P) {
    // 4 consecutive genes per comparison share a 32-byte sector of every [.][ngenes] array.
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t tiles = (P.ngenes + 3) / 4;
    const int64_t tile = t / ((int64_t)4 * P.Q);
    if (tile >= tiles) return;
    const int rem = (int)(t % ((int64_t)4 * P.Q));
    const int q = rem / 4;
    const int g = (int)(tile * 4 + (rem & 3));
    if (g >= P.ngenes) return;

    const int j0 = P.avg_ptr[q], j1 = P.avg_ptr[q + 1];
    double wsum[3] = {0.0, 0.0, 0.0};
    DD acc[3] = {{0.0, 0.0}, {0.0, 0.0}, {0.0, 0.0}};
    auto block_effects = [&](int j, double& w, double& d, double& a, double& f) {
        const int p = P.avg_idx[j];
        const int l = P.left[p], r = P.right[p];
        const double nL = P.ncells[l], nR = P.ncells[r];
        w = nL * nR;
        const size_t al = (size_t)l * P.ngenes + g, ar = (size_t)r * P.ngenes + g;
        d = effect_cohen(P.mean[al], P.mean[ar], P.var[al], P.var[ar], P.lfc);
        a = P.u[(size_t)g * P.npairs + p] / w;
        f = effect_detected(P.ndet[al], P.ndet[ar], nL, nR);
    };
    constexpr int KEEP = 8;  // block levels whose effects stay in registers between the two passes
    if (j1 - j0 <= KEEP) {
        double w[KEEP], d[KEEP], a[KEEP], f[KEEP];
        // Pass 1: the effects, and the sum of weights over the blocks where each effect is defined.
#pragma unroll
        for (int jj = 0; jj < KEEP; ++jj) {
            if (j0 + jj < j1) {
                block_effects(j0 + jj, w[jj], d[jj], a[jj], f[jj]);
                if (!isnan(d[jj])) wsum[0] += w[jj];
                if (!isnan(a[jj])) wsum[1] += w[jj];
                if (!isnan(f[jj])) wsum[2] += w[jj];
            }
        }
        // Pass 2: sum of (w / wsum) * effect.
#pragma unroll
        for (int jj = 0; jj < KEEP; ++jj) {
            if (j0 + jj < j1) {
                if (!isnan(d[jj])) dd_add(acc[0], (w[jj] / wsum[0]) * d[jj]);
                if (!isnan(a[jj])) dd_add(acc[1], (w[jj] / wsum[1]) * a[jj]);
                if (!isnan(f[jj])) dd_add(acc[2], (w[jj] / wsum[2]) * f[jj]);
            }
        }
    } else {  // many block levels: the effects are computed twice
        for (int j = j0; j < j1; ++j) {
            double w, d, a, f;
            block_effects(j, w, d, a, f);
            if (!isnan(d)) wsum[0] += w;
            if (!isnan(a)) wsum[1] += w;
            if (!isnan(f)) wsum[2] += w;
        }
        for (int j = j0; j < j1; ++j) {
            double w, d, a, f;
            block_effects(j, w, d, a, f);
            if (!isnan(d)) dd_add(acc[0], (w / wsum[0]) * d);
            if (!isnan(a)) dd_add(acc[1], (w / wsum[1]) * a);
            if (!isnan(f)) dd_add(acc[2], (w / wsum[2]) * f);
        }
    }
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    for (int e = 0; e < 3; ++e) {
        P.eff[((size_t)e * P.Q + q) * P.ld + P.row0 + g] = wsum[e] > 0.0 ? dd_value(acc[e]) : nan;
    }
}

struct SelfParams {
    int ngenes, K, C, ngroups;
    int64_t ld, row0;
    const double *mean, *ndet;  // [C][ngenes]
    const double* ncells;
    const int32_t *out_group, *proj_ptr, *proj_comb, *other_ptr, *other_group;
    const double* proj_coef;
    int blocked;
    double* logit;    // [C][ngenes]: logit of the shrunken detected fraction (blocked only)
    double* grp_avg;  // [2][ngroups][ngenes]: per-group average and detected fraction
    double* out;      // [K][19][ld]
};

// scuttle::correctGroupSummary(transform="logit"): (p + 0.01) / 1.02, then log(x / (1 - x)).
__global__ void __launch_bounds__(256) k4_logit(SelfParams P) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)P.ngenes * P.C) return;
    const int c = (int)(t / P.ngenes);
    const double prop = P.ndet[t] / P.ncells[c];
    const double x = (prop + 0.01) / (1.0 + 2.0 * 0.01);
    P.logit[t] = log(x / (1.0 - x));
}

// One value per (gene, group): the group's own summary when there is no blocking, else the
// group coefficient of the weighted fit (a fixed projector row applied to the C summaries).
__global__ void __launch_bounds__(256) k4_group_avg(SelfParams P) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)P.ngenes * P.ngroups) return;
    const int g = (int)(t % P.ngenes), grp = (int)(t / P.ngenes);
    const int a = P.proj_ptr[grp], b = P.proj_ptr[grp + 1];
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    double avg = nan, det = nan;
    if (b > a) {
        if (!P.blocked) {
            const int c = P.proj_comb[a];
            avg = P.mean[(size_t)c * P.ngenes + g];
            det = P.ndet[(size_t)c * P.ngenes + g] / P.ncells[c];
        } else {
            DD sa = {0.0, 0.0}, sd = {0.0, 0.0};
            for (int k = a; k < b; ++k) {
                const size_t at = (size_t)P.proj_comb[k] * P.ngenes + g;
                const double w = P.proj_coef[k];
                dd_add(sa, w * P.mean[at]);
                dd_add(sd, w * P.logit[at]);
            }
            avg = dd_value(sa);
            double v = exp(dd_value(sd));
            v = v / (v + 1.0);
            v = v * (1.0 + 2.0 * 0.01) - 0.01;
            det = fmin(fmax(v, 0.0), 1.0);
        }
    }
    P.grp_avg[(size_t)grp * P.ngenes + g] = avg;
    P.grp_avg[((size_t)P.ngroups + grp) * P.ngenes + g] = det;
}

__global__ void __launch_bounds__(256) k4_selfother(SelfParams P) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)P.ngenes * P.K) return;
    const int g = (int)(t % P.ngenes), k = (int)(t / P.ngenes);
    const int grp = P.out_group[k];
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    for (int which = 0; which < 2; ++which) {
        const double* src = P.grp_avg + (size_t)which * P.ngroups * P.ngenes;
        const double self = src[(size_t)grp * P.ngenes + g];
        DD s = {0.0, 0.0};
        const int o0 = P.other_ptr[k], o1 = P.other_ptr[k + 1];
        for (int o = o0; o < o1; ++o) dd_add(s, src[(size_t)P.other_group[o] * P.ngenes + g]);
        const double other = o1 > o0 ? dd_value(s) / (double)(o1 - o0) : nan;
        double* base = P.out + ((size_t)k * SCRANB200_NCOLS + 2 * which) * P.ld + P.row0 + g;
        base[0] = self;
        base[P.ld] = other;
    }
}

}  // namespace

void run_effects(scranb200_ctx* ctx, const scranb200_plan& P, int ngenes, int64_t ld, int64_t row0, const double* d_mean,
                 const double* d_var, const double* d_ndet, const double* d_u, double* d_eff, double* d_out) {
    if (ngenes == 0) return;
    if (P.Q > 0 && d_u != nullptr && d_eff != nullptr) {  // (null: only the self. / other. columns are wanted)
        EffParams ep;
        ep.ngenes = ngenes;
        ep.Q = P.Q;
        ep.ld = ld;
        ep.row0 = row0;
        ep.mean = d_mean;
        ep.var = d_var;
        ep.ndet = d_ndet;
        ep.u = d_u;
        ep.npairs = (int64_t)P.left.size();
        ep.left = P.d_left;
        ep.right = P.d_right;
        ep.avg_ptr = P.d_avg_ptr;
        ep.avg_idx = P.d_avg_idx;
        ep.ncells = P.d_ncells_comb;
        ep.lfc = P.lfc;
        ep.eff = d_eff;
        const int64_t threads = ceil_div(ngenes, 4) * 4 * P.Q;
        SB_LAUNCH(ctx, k4_effects, (unsigned)ceil_div(threads, 256), 256, 0, ep);
    }
    if (P.K > 0) {
        SelfParams sp;
        sp.ngenes = ngenes;
        sp.K = P.K;
        sp.C = P.C;
        sp.ngroups = P.ngroups;
        sp.ld = ld;
        sp.row0 = row0;
        sp.mean = d_mean;
        sp.ndet = d_ndet;
        sp.ncells = P.d_ncells_comb;
        sp.out_group = P.d_out_group;
        sp.proj_ptr = P.d_proj_ptr;
        sp.proj_comb = P.d_proj_comb;
        sp.other_ptr = P.d_other_ptr;
        sp.other_group = P.d_other_group;
        sp.proj_coef = P.d_proj_coef;
        sp.blocked = P.blocked ? 1 : 0;
        ctx->grp_avg.reserve((size_t)2 * P.ngroups * ngenes * 8);
        sp.grp_avg = ctx->grp_avg.as<double>();
        sp.logit = nullptr;
        if (P.blocked) {
            ctx->logit.reserve((size_t)P.C * ngenes * 8);
            sp.logit = ctx->logit.as<double>();
            SB_LAUNCH(ctx, k4_logit, (unsigned)ceil_div((int64_t)ngenes * P.C, 256), 256, 0, sp);
        }
        sp.out = d_out;
        SB_LAUNCH(ctx, k4_group_avg, (unsigned)ceil_div((int64_t)ngenes * P.ngroups, 256), 256, 0, sp);
        SB_LAUNCH(ctx, k4_selfother, (unsigned)ceil_div((int64_t)ngenes * P.K, 256), 256, 0, sp);
    }
}

}  // namespace sb
