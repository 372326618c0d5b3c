// combineMarkers on the device: what R/combineMarkers.R:232-330 does for the
// contrasts of ONE first group, on log p-values (log.p.in = log.p.out = TRUE), together with the two
// native helpers it calls, src/choose_effect_size.cpp:35-61 (compute_Top_statistic_from_ranks) and
// :64-97 (choose_middle_effect_size).
//
//   per contrast   rank(p, ties.method = "first", na.last = "keep") across the genes (pval.type "any")
//   per gene       the combined p-value over the contrasts (metapod::combineParallelPValues: Simes for "any",
//                  the ceiling(n min.prop)-th Holm-adjusted p-value for "some", the largest for "all"), Top = the
//                  ceiling(n min.prop)-th smallest rank, the summary effect (.choose_effect_size)
//   across genes   .logBH on the combined p-values, and the row order the reference applies last:
//                  order(Top) for "any", order(p) otherwise
//
// One CTA per contrast ranks its column with the stable radix sort of sort.cuh; one thread per gene folds the
// contrasts (its values live in a strided scratch column, sorted by insertion: a gene has at most a few hundred
// contrasts); one CTA does the log BH and one the final ordering.
#include <climits>

#include "kernels.h"
#include "sort.cuh"

namespace sb {

namespace {

constexpr int32_t NA_I32 = INT32_MIN;

struct RankFirstParams {
    int n, ncols;
    const double* x;     // [ncols][n], NaN = NA
    int32_t* rank;       // [ncols][n], NA_I32 for NA
    uint64_t *g_ka, *g_kb;
    uint32_t *g_pa, *g_pb;
};

// rank(x, ties.method = "first", na.last = "keep") per column: position in the stable ascending order.
__global__ void __launch_bounds__(SORT_THREADS) k7_rank_first(RankFirstParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* whist = reinterpret_cast<uint32_t*>(smem_raw);
    const int n = P.n, tid = threadIdx.x;
    uint64_t* ka = P.g_ka + (size_t)blockIdx.x * n;
    uint64_t* kb = P.g_kb + (size_t)blockIdx.x * n;
    uint32_t* pa = P.g_pa + (size_t)blockIdx.x * n;
    uint32_t* pb = P.g_pb + (size_t)blockIdx.x * n;
    for (int c = blockIdx.x; c < P.ncols; c += gridDim.x) {
        const double* in = P.x + (size_t)c * n;
        int32_t* out = P.rank + (size_t)c * n;
        __syncthreads();
        for (int i = tid; i < n; i += blockDim.x) {
            const double v = in[i];
            ka[i] = isnan(v) ? ~0ull : key_of(v);  // NA sorts last and gets no rank
            pa[i] = (uint32_t)i;
        }
        __syncthreads();
        const int cur = block_sort_keys<uint32_t>(ka, pa, kb, pb, n, whist);
        const uint64_t* k = cur ? kb : ka;
        const uint32_t* p = cur ? pb : pa;
        for (int j = tid; j < n; j += blockDim.x) out[p[j]] = k[j] == ~0ull ? NA_I32 : j + 1;
    }
}

struct CombineParams {
    int ngenes, ncon, pval_type;  // 0 any, 1 some, 2 all
    double min_prop;
    const double *logp, *effect;  // [ncon][ngenes]; effect nullable
    const int32_t* rank;          // [ncon][ngenes] (any only)
    double *sp, *se;              // scratch [ncon][ngenes]
    int32_t* sr;                  // scratch [ncon][ngenes]
    int32_t* top;                 // [ngenes] (any only)
    double *p_out, *summary;      // [ngenes]; summary nullable
    double* order_key;            // [ngenes]: what the rows are ordered by (Top or p), NaN = NA
};

__device__ inline int define_jump(int n, double prop) {  // src/choose_effect_size.cpp:26-33
    int jump = (int)ceil((double)n * prop);
    if (jump) --jump;
    return jump;
}

__global__ void __launch_bounds__(256) k7_combine(CombineParams P) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P.ngenes) return;
    const size_t ld = (size_t)P.ngenes;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    // ---- the gene's non-NA p-values in ascending order (insertion into the strided scratch column) ----
    int n = 0;
    for (int c = 0; c < P.ncon; ++c) {
        const double v = P.logp[(size_t)c * ld + g];
        if (isnan(v)) continue;
        int b = n - 1;
        while (b >= 0 && P.sp[(size_t)b * ld + g] > v) {
            P.sp[(size_t)(b + 1) * ld + g] = P.sp[(size_t)b * ld + g];
            --b;
        }
        P.sp[(size_t)(b + 1) * ld + g] = v;
        ++n;
    }
    double pv = nan;
    if (n > 0) {
        if (P.pval_type == 0) {  // Simes: the smallest BH-adjusted p-value
            double best = 0.0;
            for (int k = 0; k < n; ++k) best = fmin(best, P.sp[(size_t)k * ld + g] + log((double)n / (double)(k + 1)));
            pv = best;
        } else if (P.pval_type == 2) {  // Berger: the largest
            pv = P.sp[(size_t)(n - 1) * ld + g];
        } else {  // Holm-min: the (jump + 1)-th smallest Holm-adjusted p-value
            const int jump = define_jump(n, P.min_prop);
            double run = -__longlong_as_double(0x7ff0000000000000ll);
            for (int k = 0; k <= jump; ++k) run = fmax(run, fmin(0.0, P.sp[(size_t)k * ld + g] + log((double)(n - k))));
            pv = run;
        }
    }
    P.p_out[g] = pv;
    double okey = pv;
    // ---- Top: the (jump + 1)-th smallest rank over the contrasts (compute_Top_statistic_from_ranks) ----
    if (P.pval_type == 0) {
        int m = 0;
        for (int c = 0; c < P.ncon; ++c) {
            const int32_t r = P.rank[(size_t)c * ld + g];
            if (r == NA_I32) continue;
            int b = m - 1;
            while (b >= 0 && P.sr[(size_t)b * ld + g] > r) {
                P.sr[(size_t)(b + 1) * ld + g] = P.sr[(size_t)b * ld + g];
                --b;
            }
            P.sr[(size_t)(b + 1) * ld + g] = r;
            ++m;
        }
        const int32_t top = m > 0 ? P.sr[(size_t)define_jump(m, P.min_prop) * ld + g] : NA_I32;
        P.top[g] = top;
        okey = top == NA_I32 ? nan : (double)top;
    }
    P.order_key[g] = okey;
    // ---- summary effect (.choose_effect_size, R/combineMarkers.R:302-330) ----
    if (P.summary) {
        double chosen = nan;
        if (P.pval_type == 2) {  // the effect of the largest p-value; the first contrast wins a tie
            double worst = -__longlong_as_double(0x7ff0000000000000ll);
            for (int c = 0; c < P.ncon; ++c) {
                const double v = P.logp[(size_t)c * ld + g];
                if (!isnan(v) && v > worst) {
                    worst = v;
                    chosen = P.effect[(size_t)c * ld + g];
                }
            }
        } else if (P.pval_type == 0) {  // the effect of the smallest p-value
            double best = __longlong_as_double(0x7ff0000000000000ll);
            for (int c = 0; c < P.ncon; ++c) {
                const double v = P.logp[(size_t)c * ld + g];
                if (!isnan(v) && v < best) {
                    best = v;
                    chosen = P.effect[(size_t)c * ld + g];
                }
            }
        } else {  // choose_middle_effect_size: (p, effect) pairs with neither NA, in std::pair order, the one at `jump`
            int m = 0;
            for (int c = 0; c < P.ncon; ++c) {
                const double v = P.logp[(size_t)c * ld + g], e = P.effect[(size_t)c * ld + g];
                if (isnan(v) || isnan(e)) continue;
                int b = m - 1;
                while (b >= 0) {
                    const double pb = P.sp[(size_t)b * ld + g], eb = P.se[(size_t)b * ld + g];
                    if (!(pb > v || (pb == v && eb > e))) break;
                    P.sp[(size_t)(b + 1) * ld + g] = pb;
                    P.se[(size_t)(b + 1) * ld + g] = eb;
                    --b;
                }
                P.sp[(size_t)(b + 1) * ld + g] = v;
                P.se[(size_t)(b + 1) * ld + g] = e;
                ++m;
            }
            if (m > 0) chosen = P.se[(size_t)define_jump(m, P.min_prop) * ld + g];
        }
        P.summary[g] = chosen;
    }
}

struct OrderParams {
    int n;
    const double* key;  // NaN = NA (last)
    int32_t* order;     // [n] 0-based
    double* logfdr;     // [n]
    const double* logp;
    uint64_t *g_ka, *g_kb;
    uint32_t *g_pa, *g_pb;
};

// One CTA: order(key) (ascending, ties by position, NA last) and .logBH(logp) (R/utils_markers.R:184-190; any NA
// makes every value NA, as cummin does).
__global__ void __launch_bounds__(SORT_THREADS) k7_order_and_bh(OrderParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* whist = reinterpret_cast<uint32_t*>(smem_raw);
    double* cmin = reinterpret_cast<double*>(whist + SORT_HIST_WORDS + (SORT_HIST_WORDS & 1));  // [SORT_THREADS]
    uint32_t* misc = reinterpret_cast<uint32_t*>(cmin + SORT_THREADS);
    const int n = P.n, tid = threadIdx.x;
    const double inf = __longlong_as_double(0x7ff0000000000000ll), nan = __longlong_as_double(0x7ff8000000000000ll);
    // ---- the row order ----
    for (int i = tid; i < n; i += blockDim.x) {
        const double v = P.key[i];
        P.g_ka[i] = isnan(v) ? ~0ull : key_of(v);
        P.g_pa[i] = (uint32_t)i;
    }
    __syncthreads();
    {
        const int cur = block_sort_keys<uint32_t>(P.g_ka, P.g_pa, P.g_kb, P.g_pb, n, whist);
        const uint32_t* p = cur ? P.g_pb : P.g_pa;
        for (int j = tid; j < n; j += blockDim.x) P.order[j] = (int32_t)p[j];
    }
    __syncthreads();
    // ---- log BH ----
    if (tid == 0) misc[0] = 0;
    __syncthreads();
    uint32_t bad = 0;
    for (int i = tid; i < n; i += blockDim.x) {
        const double v = P.logp[i];
        bad |= isnan(v) ? 1u : 0u;
        P.g_ka[i] = key_of(v);
        P.g_pa[i] = (uint32_t)i;
    }
    if (bad) atomicOr(&misc[0], 1u);
    __syncthreads();
    if (misc[0]) {
        for (int i = tid; i < n; i += blockDim.x) P.logfdr[i] = nan;
        return;
    }
    const int cur = block_sort_keys<uint32_t>(P.g_ka, P.g_pa, P.g_kb, P.g_pb, n, whist);
    const uint64_t* k = cur ? P.g_kb : P.g_ka;
    const uint32_t* p = cur ? P.g_pb : P.g_pa;
    // minimum from the right of logp_(j) + log(n / (j + 1)): per-thread chunks, then thread 0 chains the chunk minima
    const int per = (n + blockDim.x - 1) / blockDim.x;
    const int lo = min(n, tid * per), hi = min(n, lo + per);
    double m = inf;
    for (int j = lo; j < hi; ++j) m = fmin(m, value_of(k[j]) + log((double)n / (double)(j + 1)));
    cmin[tid] = m;
    __syncthreads();
    if (tid == 0) {  // cmin[t] <- min over the chunks after t
        double run = inf;
        for (int t = SORT_THREADS - 1; t >= 0; --t) {
            const double own = cmin[t];
            cmin[t] = run;
            run = fmin(run, own);
        }
    }
    __syncthreads();
    double run = cmin[tid];
    for (int j = hi - 1; j >= lo; --j) {
        run = fmin(run, value_of(k[j]) + log((double)n / (double)(j + 1)));
        P.logfdr[p[j]] = run;
    }
}

}  // namespace

void run_combine_markers(scranb200_ctx* ctx, int64_t ngenes, int ncon, const double* logp, const double* effect, int pval_type,
                         double min_prop, int32_t* top, double* logp_out, double* logfdr_out, double* summary_out,
                         int32_t* order_out) {
    if (ngenes == 0) return;
    if (ngenes > 2000000000ll / std::max(1, ncon)) throw Error("too many genes x contrasts");
    const size_t cells = (size_t)std::max(1, ncon) * (size_t)ngenes;
    // device layout: io_a = logp | effect, io_b = scratch p | scratch e, io_c = rank | scratch rank | per-gene outputs
    ctx->io_a.reserve(cells * 16);
    ctx->io_b.reserve(cells * 16);
    ctx->io_c.reserve(cells * 8 + (size_t)ngenes * (4 + 8 * 4 + 4) + 256);
    double* d_logp = ctx->io_a.as<double>();
    double* d_eff = d_logp + cells;
    double* d_sp = ctx->io_b.as<double>();
    double* d_se = d_sp + cells;
    int32_t* d_rank = ctx->io_c.as<int32_t>();
    int32_t* d_sr = d_rank + cells;
    double* d_p = reinterpret_cast<double*>(d_sr + cells);  // (2 * cells int32 = a multiple of 8 bytes)
    double* d_fdr = d_p + ngenes;
    double* d_sum = d_fdr + ngenes;
    double* d_key = d_sum + ngenes;
    int32_t* d_top = reinterpret_cast<int32_t*>(d_key + ngenes);
    int32_t* d_order = d_top + ngenes;
    if (ncon > 0) {
        SB_CUDA(cudaMemcpyAsync(d_logp, logp, cells * 8, cudaMemcpyHostToDevice, ctx->stream));
        if (effect) SB_CUDA(cudaMemcpyAsync(d_eff, effect, cells * 8, cudaMemcpyHostToDevice, ctx->stream));
    }
    const int grid = std::max(1, std::min(ncon, ctx->sm_count));
    ctx->rank_keys.reserve((size_t)grid * ngenes * 16);
    ctx->rank_idx.reserve((size_t)grid * ngenes * 8);
    uint64_t* g_ka = ctx->rank_keys.as<uint64_t>();
    uint64_t* g_kb = g_ka + (size_t)grid * ngenes;
    uint32_t* g_pa = ctx->rank_idx.as<uint32_t>();
    uint32_t* g_pb = g_pa + (size_t)grid * ngenes;
    if (pval_type == 0 && ncon > 0) {
        RankFirstParams rp;
        rp.n = (int)ngenes;
        rp.ncols = ncon;
        rp.x = d_logp;
        rp.rank = d_rank;
        rp.g_ka = g_ka;
        rp.g_kb = g_kb;
        rp.g_pa = g_pa;
        rp.g_pb = g_pb;
        const size_t smem = (size_t)SORT_HIST_WORDS * 4 + 64;
        allow_smem(k7_rank_first, smem);
        SB_LAUNCH(ctx, k7_rank_first, grid, SORT_THREADS, smem, rp);
    }
    CombineParams cp;
    cp.ngenes = (int)ngenes;
    cp.ncon = ncon;
    cp.pval_type = pval_type;
    cp.min_prop = min_prop;
    cp.logp = d_logp;
    cp.effect = effect ? d_eff : nullptr;
    cp.rank = d_rank;
    cp.sp = d_sp;
    cp.se = d_se;
    cp.sr = d_sr;
    cp.top = d_top;
    cp.p_out = d_p;
    cp.summary = effect && summary_out ? d_sum : nullptr;
    cp.order_key = d_key;
    SB_LAUNCH(ctx, k7_combine, (unsigned)ceil_div(ngenes, 256), 256, 0, cp);
    OrderParams op;
    op.n = (int)ngenes;
    op.key = d_key;
    op.order = d_order;
    op.logfdr = d_fdr;
    op.logp = d_p;
    op.g_ka = g_ka;
    op.g_kb = g_kb;
    op.g_pa = g_pa;
    op.g_pb = g_pb;
    const size_t smem = (size_t)(SORT_HIST_WORDS + 1) * 4 + (size_t)SORT_THREADS * 8 + 64;
    allow_smem(k7_order_and_bh, smem);
    SB_LAUNCH(ctx, k7_order_and_bh, 1, SORT_THREADS, smem, op);
    if (top && pval_type == 0) SB_CUDA(cudaMemcpyAsync(top, d_top, (size_t)ngenes * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (logp_out) SB_CUDA(cudaMemcpyAsync(logp_out, d_p, (size_t)ngenes * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (logfdr_out) SB_CUDA(cudaMemcpyAsync(logfdr_out, d_fdr, (size_t)ngenes * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (cp.summary) SB_CUDA(cudaMemcpyAsync(summary_out, d_sum, (size_t)ngenes * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (order_out) SB_CUDA(cudaMemcpyAsync(order_out, d_order, (size_t)ngenes * 4, cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
}

}  // namespace sb

using namespace sb;

extern "C" int scranb200_combine_markers(scranb200_ctx* ctx, int64_t ngenes, int ncon, const double* logp, const double* effect,
                                         int pval_type, double min_prop, int32_t* top_out, double* logp_out,
                                         double* logfdr_out, double* summary_out, int32_t* order_out) {
    try {
        ctx_guard(ctx);
        if (ngenes < 0 || ncon < 0) throw Error("negative number of genes or contrasts");
        if (pval_type < 0 || pval_type > 2) throw Error("pval_type must be 0 (any), 1 (some) or 2 (all)");
        if (!(min_prop >= 0.0 && min_prop <= 1.0)) throw Error("min_prop must lie in [0, 1]");
        run_combine_markers(ctx, ngenes, ncon, logp, effect, pval_type, min_prop, top_out, logp_out, logfdr_out, summary_out,
                            order_out);
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}
