// The p-value layer of pairwiseWilcox on the device.
//
//   per block and pair   R/pairwiseWilcox.R:263-345   AUC = U / (nh nt); z = U - nh nt / 2;
//                                                     sigma (.get_sigma, :347-350);
//                                                     pnorm((z +- 0.5) / sigma, log.p=TRUE)
//   across blocks        R/utils_markers.R:52-130     Stouffer's method on the log p-values with
//                                                     weights nh nt (metapod::combineParallelPValues),
//                                                     weighted mean of the effects, choice by direction
//   per pair             R/utils_markers.R:184-190    .logBH over the genes
//
// One thread folds all block levels of one (gene, unordered group pair); the Benjamini-Hochberg step
// sorts every ordered pair's column with the radix sort of sort.cuh.
#include "kernels.h"
#include "sort.cuh"
#include "tdist.cuh"

namespace sb {

namespace {

constexpr double HALF_LOG_2PI = 0.918938533204672741780329736405617639;
constexpr double SQRT_HALF = 0.707106781186547524400844362104849039;

// log of the standard normal distribution function, accurate in both tails (pnorm(x, log.p=TRUE)).
__device__ inline double log_ndtr(double x) {
    if (isnan(x)) return x;
    if (x > 0.0) return log1p(-0.5 * erfc(x * SQRT_HALF));
    if (x > -37.0) return log(0.5 * erfc(-x * SQRT_HALF));
    // erfc underflows: asymptotic series log(phi(x) / -x) + log(1 - 1/x^2 + 3/x^4 - ...)
    const double r = 1.0 / (x * x);
    const double s = 1.0 + r * (-1.0 + r * (3.0 + r * (-15.0 + r * (105.0 + r * (-945.0 + r * (10395.0 - 135135.0 * r))))));
    return -0.5 * x * x - log(-x) - HALF_LOG_2PI + log(s);
}

// Inverse of log_ndtr (qnorm(y, log.p=TRUE)), y <= 0.
__device__ inline double ndtri_exp(double y) {
    if (isnan(y) || y > 0.0) return __longlong_as_double(0x7ff8000000000000ll);
    if (y == 0.0) return __longlong_as_double(0x7ff0000000000000ll);
    if (y > -1.0) return -normcdfinv(-expm1(y));  // 1 - p without cancellation
    if (y > -700.0) return normcdfinv(exp(y));
    if (isinf(y)) return y;
    // exp(y) underflows: Newton steps on log_ndtr(z) = y from the leading asymptotic term
    double z = -sqrt(-2.0 * (y + HALF_LOG_2PI));
    z = -sqrt(-2.0 * (y + HALF_LOG_2PI + log(-z)));
    for (int it = 0; it < 4; ++it) {
        const double ln = log_ndtr(z);
        z -= (ln - y) / exp(-0.5 * z * z - HALF_LOG_2PI - ln);  // d/dz log Phi = phi / Phi
    }
    return z;
}

// Stouffer's method on log p-values (metapod::combineParallelPValues(method = "stouffer", log.p = TRUE),
// called at R/utils_markers.R:96-97): z = sum w_b qnorm(p_b) / sqrt(sum w_b^2). A block with p = 1 (z = +Inf)
// and one with p = 0 (z = -Inf) would add up to NaN; they are kept out of the sum and resolved by weight:
// the heavier side wins, equal weights give p = 1/2. (metapod's source is not in the reference tree: this
// edge is restated from its documented behaviour, parity unpinned.)
struct StoufferAcc {
    double z = 0.0, w_zero = 0.0, w_one = 0.0;
    __device__ void add(double logp, double w) {
        if (logp == 0.0) w_one += w;
        else if (isinf(logp) && logp < 0.0) w_zero += w;
        else z += ndtri_exp(logp) * w;
    }
    __device__ double log_p(double sw2) const {
        if (w_zero > 0.0 || w_one > 0.0) {
            if (isnan(z)) return z;
            if (w_zero > w_one) return -__longlong_as_double(0x7ff0000000000000ll);
            if (w_zero < w_one) return 0.0;
            return -0.693147180559945309417232121458176568;
        }
        return log_ndtr(z / sqrt(sw2));
    }
};

struct WilcoxLayerParams {
    int ngenes, G, nblocks, npairs;  // npairs: block-level pairs (columns of u / t)
    int direction;                   // 0 any, 1 up, 2 down
    int shifted;                     // lfc != 0: u_minus / t_minus hold the run with -lfc
    const double *u, *t;             // [ngenes][npairs], run with +lfc (or lfc == 0)
    const double *u_minus, *t_minus;
    const int32_t* pair_of;          // [nblocks][G][G]: column of (block, i, j), i > j; -1 if a group is empty
    const double* sizes;             // [nblocks][G]
    double *effect, *logp;           // [G * (G - 1)][ngenes], ordered pairs (first, second) in level order
};

__device__ inline double sigma_of(double nh, double nt, double ties) {  // .get_sigma; NaN stays NaN like pmax()
    const double s2 = (nh * nt / 12.0) * ((nh + nt + 1.0) - ties / ((nh + nt) * (nh + nt - 1.0)));
    const double s = sqrt(s2);
    return isnan(s) ? s : fmax(s, 1e-8);
}

__device__ inline double choose_p(double left, double right, int direction) {  // first = the tail to report for "up"
    if (direction == 1) return right;
    if (direction == 2) return left;
    const double m = fmin(left, right);
    if (isnan(left) || isnan(right)) return __longlong_as_double(0x7ff8000000000000ll);
    return fmin(0.0, m + 0.693147180559945309417232121458176568);
}

// position of the ordered pair (a, b), a != b, in the output: levels in order, second index skipping a
__device__ __host__ inline int ordered_index(int a, int b, int G) { return a * (G - 1) + (b < a ? b : b - 1); }

__global__ void __launch_bounds__(256) k6_wilcox_layer(WilcoxLayerParams P) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int upairs = P.G * (P.G - 1) / 2;
    if (t >= (int64_t)P.ngenes * upairs) return;
    const int g = (int)(t % P.ngenes);
    int rem = (int)(t / P.ngenes), host = 1;
    while (rem >= host) {  // unordered pair number -> (host, target), target < host
        rem -= host;
        ++host;
    }
    const int target = rem;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    double sw = 0.0, sw2 = 0.0, fwd = 0.0, rev = 0.0;
    StoufferAcc zl, zr;
    bool any = false;
    for (int b = 0; b < P.nblocks; ++b) {
        const double nh = P.sizes[b * P.G + host], nt = P.sizes[b * P.G + target];
        const double prod = nh * nt;
        if (!(prod > 1.5)) continue;  // not a valid test in this block
        const int col = P.pair_of[((size_t)b * P.G + host) * P.G + target];
        if (col < 0) continue;
        const size_t at = (size_t)g * P.npairs + col;
        double f, r, left, right;
        if (!P.shifted) {
            const double u = P.u[at];
            const double auc = u / prod;
            f = auc;
            r = 1.0 - auc;
            const double z = u - prod / 2.0;
            const double sig = sigma_of(nh, nt, P.t[at]);
            left = log_ndtr((z + 0.5) / sig);
            right = log_ndtr(-(z - 0.5) / sig);
        } else {
            // entry [host][target] was computed with +lfc ("minus": host - lfc > target), entry
            // [target][host] with -lfc ("added"), R/pairwiseWilcox.R:300-345
            const double minus = P.u[at], added = P.u_minus[at];
            if (P.direction == 0) {
                const double auc = (minus / 2.0 + added / 2.0) / prod;
                f = auc;
                r = 1.0 - auc;
            } else if (P.direction == 1) {
                f = minus / prod;
                r = 1.0 - added / prod;
            } else {
                f = added / prod;
                r = 1.0 - minus / prod;
            }
            const double msig = sigma_of(nh, nt, P.t[at]), asig = sigma_of(nh, nt, P.t_minus[at]);
            left = log_ndtr(((added - prod / 2.0) + 0.5) / asig);
            right = log_ndtr(-((minus - prod / 2.0) - 0.5) / msig);
        }
        any = true;
        sw += prod;
        sw2 += prod * prod;
        zl.add(left, prod);
        zr.add(right, prod);
        fwd += f * prod;
        rev += r * prod;
    }
    double fe = nan, re = nan, hp = nan, tp = nan;
    if (any) {
        const double cl = zl.log_p(sw2), cr = zr.log_p(sw2);
        hp = choose_p(cl, cr, P.direction);
        tp = choose_p(cr, cl, P.direction);
        fe = fwd / sw;
        re = rev / sw;
    }
    const size_t o1 = (size_t)ordered_index(host, target, P.G) * P.ngenes + g;
    const size_t o2 = (size_t)ordered_index(target, host, P.G) * P.ngenes + g;
    P.effect[o1] = fe;
    P.logp[o1] = hp;
    P.effect[o2] = re;
    P.logp[o2] = tp;
}

// The same fold for pairwiseTTests (R/pairwiseTTests.R:270-302): per block Welch's t-test from the block's
// means and variances (.get_t_test_stats, :311-332: variances floored at 1e-8, Welch's degrees of freedom),
// log p-values of both tails from the t distribution (.run_t_test, :439-461; with lfc != 0 the two tails are
// tested against -|lfc| and +|lfc|), effect = log-fold change or, with std.lfc, Cohen's d with the
// df-pooled variance (:283-289); blocks weighted by 1 / (1/nh + 1/nt) (:296); a block is valid when both
// groups have two or more cells (:299).
struct TTestLayerParams {
    int ngenes, G, nblocks;
    int direction, std_lfc;
    double thr;                 // |lfc|
    const double *mean, *var;   // [nblocks * G][ngenes]
    const double* sizes;        // [nblocks][G]
    double *effect, *logp;      // [G * (G - 1)][ngenes]
};

__global__ void __launch_bounds__(256) k6_ttest_layer(TTestLayerParams P) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int upairs = P.G * (P.G - 1) / 2;
    if (t >= (int64_t)P.ngenes * upairs) return;
    const int g = (int)(t % P.ngenes);
    int rem = (int)(t / P.ngenes), host = 1;
    while (rem >= host) {
        rem -= host;
        ++host;
    }
    const int target = rem;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    double sw = 0.0, sw2 = 0.0, fwd = 0.0;
    StoufferAcc zl, zr;
    bool any = false;
    for (int b = 0; b < P.nblocks; ++b) {
        const double nh = P.sizes[b * P.G + host], nt = P.sizes[b * P.G + target];
        const double hdf = fmax(0.0, nh - 1.0), tdf = fmax(0.0, nt - 1.0);
        if (!(hdf > 0.0 && tdf > 0.0)) continue;  // df would be NA: no valid test in this block
        const size_t ch = ((size_t)b * P.G + host) * P.ngenes + g, ct = ((size_t)b * P.G + target) * P.ngenes + g;
        const double hv = P.var[ch], tv = P.var[ct];
        const double hs2 = fmax(hv, 1e-8), ts2 = fmax(tv, 1e-8);
        const double herr = hs2 / nh, terr = ts2 / nt;
        const double err = herr + terr;
        const double df = err * err / (herr * herr / hdf + terr * terr / tdf);
        const double cur = P.mean[ch] - P.mean[ct];
        const double se = sqrt(err);
        double left, right;
        if (P.thr == 0.0) {
            const double tt = cur / se;
            left = log_pt(tt, df, true);
            right = log_pt(tt, df, false);
        } else {
            left = log_pt((cur + P.thr) / se, df, true);
            right = log_pt((cur - P.thr) / se, df, false);
        }
        double eff = cur;
        if (P.std_lfc) {
            const double pooled = ((nh - 1.0) * hv + (nt - 1.0) * tv) / (nt + nh - 2.0);
            eff = cur == 0.0 ? 0.0 : cur / sqrt(pooled);
        }
        const double w = 1.0 / (1.0 / nh + 1.0 / nt);
        any = true;
        sw += w;
        sw2 += w * w;
        zl.add(left, w);
        zr.add(right, w);
        fwd += eff * w;
    }
    double fe = nan, hp = nan, tp = nan;
    if (any) {
        const double cl = zl.log_p(sw2), cr = zr.log_p(sw2);
        hp = choose_p(cl, cr, P.direction);
        tp = choose_p(cr, cl, P.direction);
        fe = fwd / sw;
    }
    const size_t o1 = (size_t)ordered_index(host, target, P.G) * P.ngenes + g;
    const size_t o2 = (size_t)ordered_index(target, host, P.G) * P.ngenes + g;
    P.effect[o1] = fe;
    P.logp[o1] = hp;
    P.effect[o2] = -fe;
    P.logp[o2] = tp;
}

// .logBH for every column: sort, add log(n / rank), minimum from the right, back to gene order. A
// column with any NaN is all NaN (R's cummin / numpy's minimum.accumulate propagate it).
struct LogBhParams {
    int n, ncols;
    const double* logp;  // [ncols][n]
    double* logfdr;      // [ncols][n]
    uint64_t *g_ka, *g_kb;  // per-CTA sort buffers, n each
    uint32_t *g_pa, *g_pb;
};

__global__ void __launch_bounds__(SORT_THREADS) k6_log_bh(LogBhParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* whist = reinterpret_cast<uint32_t*>(smem_raw);          // SORT_HIST_WORDS
    double* cmin = reinterpret_cast<double*>(whist + SORT_HIST_WORDS + (SORT_HIST_WORDS & 1));  // [SORT_THREADS]
    uint32_t* misc = reinterpret_cast<uint32_t*>(cmin + SORT_THREADS);
    const int n = P.n, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint64_t* ka = P.g_ka + (size_t)blockIdx.x * n;
    uint64_t* kb = P.g_kb + (size_t)blockIdx.x * n;
    uint32_t* pa = P.g_pa + (size_t)blockIdx.x * n;
    uint32_t* pb = P.g_pb + (size_t)blockIdx.x * n;
    const double inf = __longlong_as_double(0x7ff0000000000000ll), nan = __longlong_as_double(0x7ff8000000000000ll);
    for (int c = blockIdx.x; c < P.ncols; c += gridDim.x) {
        const double* in = P.logp + (size_t)c * n;
        double* out = P.logfdr + (size_t)c * n;
        __syncthreads();
        if (tid == 0) misc[0] = 0;
        __syncthreads();
        uint32_t bad = 0;
        for (int i = tid; i < n; i += blockDim.x) {
            const double v = in[i];
            bad |= isnan(v) ? 1u : 0u;
            ka[i] = key_of(v);
            pa[i] = (uint32_t)i;
        }
        if (bad) atomicOr(&misc[0], 1u);
        __syncthreads();
        if (misc[0]) {
            for (int i = tid; i < n; i += blockDim.x) out[i] = nan;
            continue;
        }
        const int cur = block_sort_keys<uint32_t>(ka, pa, kb, pb, n, whist);
        const uint64_t* k = cur ? kb : ka;
        const uint32_t* p = cur ? pb : pa;
        // minimum from the right of rep[j] = logp_(j) + log(n / (j + 1)): per-thread chunks, then a scan of the chunk minima
        const int per = (n + blockDim.x - 1) / blockDim.x;
        const int lo = min(n, tid * per), hi = min(n, lo + per);
        double m = inf;
        for (int j = lo; j < hi; ++j) m = fmin(m, value_of(k[j]) + log((double)n / (double)(j + 1)));
        cmin[tid] = m;
        __syncthreads();
        if (warp == 0) {  // cmin[t] <- min over chunks t' > t (exclusive suffix minimum): 32 entries per lane
            const int per_lane = SORT_THREADS / 32;
            double local[SORT_THREADS / 32];
            double run = inf;
            for (int q = per_lane - 1; q >= 0; --q) {
                local[q] = run;
                run = fmin(run, cmin[lane * per_lane + q]);
            }
            double excl = inf;  // minimum over the lanes above
            double v = run;
            for (int o = 1; o < 32; o <<= 1) {
                const double y = __shfl_down_sync(0xffffffffu, v, o);
                if (lane + o < 32) v = fmin(v, y);
            }
            excl = __shfl_down_sync(0xffffffffu, v, 1);
            if (lane == 31) excl = inf;
            for (int q = 0; q < per_lane; ++q) cmin[lane * per_lane + q] = fmin(local[q], excl);
        }
        __syncthreads();
        double run = cmin[tid];
        for (int j = hi - 1; j >= lo; --j) {
            run = fmin(run, value_of(k[j]) + log((double)n / (double)(j + 1)));
            out[p[j]] = run;
        }
    }
}

}  // namespace

void run_pairs(scranb200_ctx* ctx, const scranb200_mat* mat, PairLayout& L, DevBuf& lbuf, bool want_stats, double lfc,
               bool want_u, bool want_t);  // api.cu

namespace {

// .logBH over the genes for every ordered pair, then the three result arrays to the host.
void finish_layer(scranb200_ctx* ctx, int64_t ngenes, int nout, double* d_effect, double* d_logp, double* d_logfdr,
                  double* effect, double* logp, double* logfdr) {
    const size_t out_bytes = (size_t)nout * ngenes * 8;
    LogBhParams bp;
    bp.n = (int)ngenes;
    bp.ncols = nout;
    bp.logp = d_logp;
    bp.logfdr = d_logfdr;
    const int grid = std::min(nout, ctx->sm_count);
    ctx->rank_keys.reserve((size_t)grid * ngenes * 16);
    ctx->rank_idx.reserve((size_t)grid * ngenes * 8);
    bp.g_ka = ctx->rank_keys.as<uint64_t>();
    bp.g_kb = bp.g_ka + (size_t)grid * ngenes;
    bp.g_pa = ctx->rank_idx.as<uint32_t>();
    bp.g_pb = bp.g_pa + (size_t)grid * ngenes;
    const size_t smem = (size_t)(SORT_HIST_WORDS + 1) * 4 + (size_t)SORT_THREADS * 8 + 64;
    allow_smem(k6_log_bh, smem);
    SB_LAUNCH(ctx, k6_log_bh, grid, SORT_THREADS, smem, bp);
    SB_CUDA(cudaMemcpyAsync(effect, d_effect, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaMemcpyAsync(logp, d_logp, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaMemcpyAsync(logfdr, d_logfdr, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
}

}  // namespace

// effect / logp / logfdr: host arrays [G (G - 1)][ngenes].
void run_pairwise_wilcox(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* group, int G, const int32_t* block,
                         int nblocks, double lfc, int direction, double* effect, double* logp, double* logfdr) {
    const int64_t ncells = mat->ncol, ngenes = mat->nrow;
    if (G < 2 || ngenes == 0) return;
    std::vector<int32_t> ids(ncells);
    std::vector<double> sizes((size_t)nblocks * G, 0.0);
    for (int64_t c = 0; c < ncells; ++c) {
        const int b = block ? block[c] : 0;
        if (group[c] >= G || (block && b >= nblocks)) throw Error("group or block code out of range");
        ids[c] = group[c] < 0 || b < 0 ? -1 : b * G + group[c];
        if (ids[c] >= 0) sizes[ids[c]] += 1.0;
    }
    std::vector<int32_t> l0, r0, pair_of((size_t)nblocks * G * G, -1);
    for (int b = 0; b < nblocks; ++b) {
        for (int i = 0; i < G; ++i) {
            if (sizes[(size_t)b * G + i] == 0.0) continue;  // src/overlap_exprs.cpp:175,177
            for (int j = 0; j < i; ++j) {
                if (sizes[(size_t)b * G + j] == 0.0) continue;
                pair_of[((size_t)b * G + i) * G + j] = (int32_t)l0.size();
                l0.push_back(b * G + i);
                r0.push_back(b * G + j);
            }
        }
    }
    const int64_t npairs = (int64_t)l0.size();
    const int nout = G * (G - 1);
    const size_t out_bytes = (size_t)nout * ngenes * 8;
    ctx->eff.reserve(out_bytes * 3);
    double* d_effect = ctx->eff.as<double>();
    double* d_logp = d_effect + (size_t)nout * ngenes;
    double* d_logfdr = d_logp + (size_t)nout * ngenes;
    const bool shifted = lfc != 0.0;
    const size_t ubytes = (size_t)std::max<int64_t>(1, npairs) * ngenes * 8;
    if (npairs > 0) {
        PairLayout L;
        build_layout(ids.data(), ncells, nblocks * G, l0.data(), r0.data(), npairs, ctx->sm_count * 4, L);
        upload_layout(ctx, L, ctx->layout_i32);
        if (shifted) {  // the mirrored entries use -lfc (src/overlap_exprs.cpp:179-187): that run first, kept aside
            run_pairs(ctx, mat, L, ctx->layout_i32, false, -lfc, true, true);
            ctx->io_a.reserve(ubytes);
            ctx->io_b.reserve(ubytes);
            SB_CUDA(cudaMemcpyAsync(ctx->io_a.p, ctx->u_pairs.p, ubytes, cudaMemcpyDeviceToDevice, ctx->stream));
            SB_CUDA(cudaMemcpyAsync(ctx->io_b.p, ctx->t_pairs.p, ubytes, cudaMemcpyDeviceToDevice, ctx->stream));
        }
        run_pairs(ctx, mat, L, ctx->layout_i32, false, lfc, true, true);
    }
    // small tables
    ctx->io_c.reserve(pair_of.size() * 4 + sizes.size() * 8 + 64);
    int32_t* d_pair_of = ctx->io_c.as<int32_t>();
    double* d_sizes = reinterpret_cast<double*>(ctx->io_c.as<char>() + ((pair_of.size() * 4 + 7) & ~size_t(7)));
    SB_CUDA(cudaMemcpyAsync(d_pair_of, pair_of.data(), pair_of.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    SB_CUDA(cudaMemcpyAsync(d_sizes, sizes.data(), sizes.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));  // the host vectors go out of use
    WilcoxLayerParams wp;
    wp.ngenes = (int)ngenes;
    wp.G = G;
    wp.nblocks = nblocks;
    wp.npairs = (int)npairs;
    wp.direction = direction;
    wp.shifted = shifted ? 1 : 0;
    wp.u = ctx->u_pairs.as<double>();
    wp.t = ctx->t_pairs.as<double>();
    wp.u_minus = shifted ? ctx->io_a.as<double>() : nullptr;
    wp.t_minus = shifted ? ctx->io_b.as<double>() : nullptr;
    wp.pair_of = d_pair_of;
    wp.sizes = d_sizes;
    wp.effect = d_effect;
    wp.logp = d_logp;
    const int64_t items = ngenes * (int64_t)(G * (G - 1) / 2);
    SB_LAUNCH(ctx, k6_wilcox_layer, (unsigned)ceil_div(items, 256), 256, 0, wp);
    finish_layer(ctx, ngenes, nout, d_effect, d_logp, d_logfdr, effect, logp, logfdr);
}

// pairwiseTTests (block= branch): blocked statistics on the device (K1 + K2 with the reference's own ids,
// group + block * ngroups: R/pairwiseTTests.R:242), then the t-test layer and .logBH.
void run_pairwise_ttests(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* group, int G, const int32_t* block,
                         int nblocks, double lfc, int direction, int std_lfc, double* effect, double* logp, double* logfdr) {
    const int64_t ncells = mat->ncol, ngenes = mat->nrow;
    if (G < 2 || ngenes == 0) return;
    std::vector<int32_t> ids(ncells);
    std::vector<double> sizes((size_t)nblocks * G, 0.0);
    for (int64_t c = 0; c < ncells; ++c) {
        const int b = block ? block[c] : 0;
        if (group[c] >= G || (block && b >= nblocks)) throw Error("group or block code out of range");
        ids[c] = group[c] < 0 || b < 0 ? -1 : b * G + group[c];
        if (ids[c] >= 0) sizes[ids[c]] += 1.0;
    }
    PairLayout L;
    build_layout(ids.data(), ncells, nblocks * G, nullptr, nullptr, 0, ctx->sm_count * 4, L);
    upload_layout(ctx, L, ctx->layout_i32);
    run_pairs(ctx, mat, L, ctx->layout_i32, true, 0.0, false, false);
    const int nout = G * (G - 1);
    const size_t out_bytes = (size_t)nout * ngenes * 8;
    ctx->eff.reserve(out_bytes * 3);
    double* d_effect = ctx->eff.as<double>();
    double* d_logp = d_effect + (size_t)nout * ngenes;
    double* d_logfdr = d_logp + (size_t)nout * ngenes;
    ctx->io_c.reserve(sizes.size() * 8 + 64);
    SB_CUDA(cudaMemcpyAsync(ctx->io_c.p, sizes.data(), sizes.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    TTestLayerParams tp;
    tp.ngenes = (int)ngenes;
    tp.G = G;
    tp.nblocks = nblocks;
    tp.direction = direction;
    tp.std_lfc = std_lfc;
    tp.thr = std::fabs(lfc);
    tp.mean = ctx->mean.as<double>();
    tp.var = ctx->var.as<double>();
    tp.sizes = ctx->io_c.as<double>();
    tp.effect = d_effect;
    tp.logp = d_logp;
    const int64_t items = ngenes * (int64_t)(G * (G - 1) / 2);
    SB_LAUNCH(ctx, k6_ttest_layer, (unsigned)ceil_div(items, 256), 256, 0, tp);
    finish_layer(ctx, ngenes, nout, d_effect, d_logp, d_logfdr, effect, logp, logfdr);
}

}  // namespace sb
