// K5: per-gene summaries of the block-averaged effects.
//
//  k5_summaries  mean / min / median / max over a group's comparisons, NaN-skipping
//                (R/scoreMarkers.R:562-566: rowMeans, rowMins, rowMedians, rowMaxs, na.rm=TRUE).
//                One warp per (gene, output table); the comparison values are staged in shared
//                memory, reductions run on warp shuffles, the median by rank counting.
//  k5_rank       computeMinRank (R/computeMinRank.R:45-50): every comparison column is ranked
//                across ALL genes in decreasing order with ties sharing the smallest rank, and a
//                gene's min-rank is the minimum over its table's columns. One CTA sorts one
//                column (key = effect, payload = gene) and scatters 1 + #(strictly greater) with
//                atomicMin; the same sorted column also serves the mirrored comparison, whose
//                values are REVERSE(x) = -x, or 1 - x for the AUC (R/scoreMarkers.R:373-377).
#include <cfloat>

#include <cstdlib>

#include "score.h"
#include "sort.cuh"

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

struct SumParams {
    int ngenes, K, Q, maxr;
    int64_t ld, row0;
    const double* eff;  // [3][Q][ld]
    const int32_t *right_ptr, *q_direct, *q_flipped;
    double* out;  // [K][19][ld]
};

// One warp per (tile of 32 consecutive genes, output table): lane = gene, so every load of a comparison
// column and every store of a result column is one coalesced 256-byte access. Nothing but a few counters is
// kept per gene: the comparison values are streamed (from L2 after the first pass) once for the mean / min /
// max and then a few more times for the median, which is found by selection through histograms: 32
// equal-width bins over the range that still holds the wanted rank, the bin with that rank either gathered
// (up to SUM_SLOTS values, sorted by insertion) or, if it holds more, narrowed to its exact [min, max] and
// binned again. Equal values end the search at once. A table of hundreds of comparisons costs the same
// 6 KB of shared memory a warp as one of ten, so the SMs stay full.
// (Tables of 65 - 255 comparisons take 128 bins with 8-bit counters - one histogram then usually settles the
// median - at the price of fewer resident warps. Keeping a table's values in shared memory instead of
// streaming them again was measured too: three warps an SM at 199 comparisons, 2.5 times slower.)
constexpr int SUM_WARPS = 8;
constexpr int SUM_SLOTS = 16;

template <int SUM_BINS, typename CountT>
__global__ void __launch_bounds__(SUM_WARPS * 32, 4) k5_summaries(SumParams P) {
    extern __shared__ __align__(16) unsigned char sum_smem[];
    double(*s_slot)[SUM_SLOTS][32] = reinterpret_cast<double(*)[SUM_SLOTS][32]>(sum_smem);
    CountT(*s_hist)[SUM_BINS][32] = reinterpret_cast<CountT(*)[SUM_BINS][32]>(sum_smem + sizeof(double) * SUM_WARPS * SUM_SLOTS * 32);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t ntiles = (P.ngenes + 31) / 32;
    const int64_t item = (int64_t)blockIdx.x * SUM_WARPS + warp;
    if (item >= ntiles * P.K) return;
    const int k = (int)(item / ntiles);
    const int g = (int)(item % ntiles) * 32 + lane;
    const bool live = g < P.ngenes;
    const int j0 = P.right_ptr[k], nr = P.right_ptr[k + 1] - j0;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const double inf = __longlong_as_double(0x7ff0000000000000ll);

    for (int e = 0; e < 3; ++e) {
        const double* col = P.eff + (size_t)e * P.Q * P.ld + P.row0 + (live ? g : 0);
        // f(x) for every comparison value of this lane's gene (NaN = NA), four loads in flight
        auto for_each_value = [&](auto&& f) {
            for (int j = 0; j < nr; j += 4) {
                double x[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    x[u] = nan;
                    if (j + u < nr && live) {
                        const int qd = P.q_direct[j0 + j + u], qf = P.q_flipped[j0 + j + u];
                        if (qd >= 0) {
                            x[u] = col[(size_t)qd * P.ld];
                        } else if (qf >= 0) {
                            const double y = col[(size_t)qf * P.ld];
                            x[u] = e == 1 ? 1.0 - y : -y;
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) f(x[u]);
            }
        };
        DD sum = {0.0, 0.0};
        double mn = inf, mx = -inf, fmn = inf, fmx = -inf;  // over all values / over the finite ones
        int cnt = 0, n_ninf = 0, n_pinf = 0;
        for_each_value([&](double x) {
            if (!isnan(x)) {
                dd_add(sum, x);
                mn = fmin(mn, x);
                mx = fmax(mx, x);
                ++cnt;
                if (isinf(x)) {
                    if (x < 0.0) ++n_ninf; else ++n_pinf;
                } else {
                    fmn = fmin(fmn, x);
                    fmx = fmax(fmx, x);
                }
            }
        });
        // ---- the values at sorted positions r_lo = (cnt - 1) / 2 and r_lo + 1 ----
        double v_lo = nan, v_next = nan;
        const int r_lo = (cnt - 1) / 2;
        const int n_fin = cnt - n_ninf - n_pinf;
        bool searching = false;
        int target = 0;                    // rank among the finite values that are still candidates, counted from `below`
        double lo = fmn, hi = fmx;         // the candidates are the finite values in [lo, hi]
        double above = n_pinf > 0 ? inf : nan;  // the smallest value known to lie above the candidates
        if (cnt > 0) {
            if (r_lo < n_ninf) {
                v_lo = -inf;
                v_next = r_lo + 1 < n_ninf ? -inf : (n_fin > 0 ? fmn : (n_pinf > 0 ? inf : nan));
            } else if (r_lo >= n_ninf + n_fin) {
                v_lo = inf;
                v_next = inf;
            } else {
                searching = true;
                target = r_lo - n_ninf;
            }
        }
        int ncand = n_fin;                 // candidates left
        while (__any_sync(0xffffffffu, searching)) {
            if (searching && lo == hi) {   // all candidates are one value
                v_lo = lo;
                v_next = target + 1 < ncand ? lo : above;
                searching = false;
            }
            // bin = floor((x - lo) / width), on values scaled by 1/32 so that hi - lo cannot overflow
            const double sc = 1.0 / (double)SUM_BINS;
            const double width = (hi * sc - lo * sc) * sc;
            const double inv = searching && width > 0.0 ? 1.0 / width : 0.0;
            const bool by_inv = inv < inf;
            auto bin_of = [&](double x) {
                const double t = x * sc - lo * sc;
                return min(SUM_BINS - 1, (int)(by_inv ? t * inv : t / width));
            };
#pragma unroll
            for (int b = 0; b < SUM_BINS; ++b) s_hist[warp][b][lane] = 0;
            for_each_value([&](double x) {
                if (searching && !isnan(x) && x >= lo && x <= hi) ++s_hist[warp][bin_of(x)][lane];
            });
            int bstar = 0, before = 0, c = 0;
            if (searching) {
                for (bstar = 0; bstar < SUM_BINS; ++bstar) {
                    c = s_hist[warp][bstar][lane];
                    if (target < before + c) break;
                    before += c;
                }
            }
            // second pass: gather the bin (if it is small), or its exact range (if not); the smallest value above it
            const bool gather = searching && c <= SUM_SLOTS;
            int ns = 0;
            double bmin = inf, bmax = -inf, up = above;
            for_each_value([&](double x) {
                if (searching && !isnan(x) && x >= lo && x <= hi) {
                    const int b = bin_of(x);
                    if (b == bstar) {
                        bmin = fmin(bmin, x);
                        bmax = fmax(bmax, x);
                        if (gather) {  // insertion into the lane's sorted slots
                            int i = ns - 1;
                            while (i >= 0 && s_slot[warp][i][lane] > x) {
                                s_slot[warp][i + 1][lane] = s_slot[warp][i][lane];
                                --i;
                            }
                            s_slot[warp][i + 1][lane] = x;
                            ++ns;
                        }
                    } else if (b > bstar) {
                        up = isnan(up) ? x : fmin(up, x);
                    }
                }
            });
            if (searching) {
                if (gather) {
                    const int idx = target - before;
                    v_lo = s_slot[warp][idx][lane];
                    v_next = idx + 1 < ns ? s_slot[warp][idx + 1][lane] : up;
                    searching = false;
                } else {
                    lo = bmin;
                    hi = bmax;
                    target -= before;
                    ncand = c;
                    above = up;
                }
            }
        }
        if (live) {
            double* base = P.out + ((size_t)k * SCRANB200_NCOLS + 4 + 5 * e) * P.ld + P.row0 + g;
            const double total = sum.hi + sum.lo;
            base[0] = cnt > 0 ? total / (double)cnt : nan;
            base[P.ld] = mn;
            base[2 * P.ld] = cnt > 0 ? ((cnt & 1) ? v_lo : (v_lo + v_next) / 2.0) : nan;
            base[3 * P.ld] = mx;
        }
    }
}

// ------------------------------------------------------------------------------------------
// k5_summaries_gene: the same summaries for tables of 65 - 256 comparisons (hundreds of groups), one CTA per
// (gene, effect). The warp-per-(gene tile, table) kernel above streams a table's comparison columns three to five
// times, and with 200 tables of 199 comparisons the resident warps' columns (240 MB) no longer fit L2: every pass
// came from HBM (79 GB of reads for 9.55 GB of effects at config 5). Here the gene's value in EVERY comparison
// column is fetched once into shared memory (Q x 8 bytes: 159 KB at Q = 19,900; the 32-byte sectors are shared
// with the CTAs of the three neighbouring genes through L2), and each warp then serves some of the tables from
// there: a lane holds up to eight of the table's values in registers; sum (double-double), range and counts by
// warp reductions; the median by the same histogram selection as above, with the 32 bins counted by
// shared-memory atomics and the chosen bin (at most 32 values) ranked across the lanes.
// ------------------------------------------------------------------------------------------
#ifndef SB_SG_WARPS
#define SB_SG_WARPS 32
#endif
constexpr int SG_WARPS = SB_SG_WARPS;
constexpr int SG_VPL = 8;  // values per lane: tables of up to 256 comparisons

struct SumGeneParams {
    int ngenes, K, Q;
    int64_t ld, row0;
    const double* eff;  // [3][Q][ld]
    const int32_t *right_ptr, *q_packed;
    double* out;  // [K][19][ld]
};

__device__ __forceinline__ double warp_min_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_max_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(SG_WARPS * 32) k5_summaries_gene(SumGeneParams P) {
    extern __shared__ __align__(16) unsigned char sg_smem[];
    double* vals = reinterpret_cast<double*>(sg_smem);                           // [Q]
    double* slots = vals + ((P.Q + 1) & ~1);                                     // [SG_WARPS][32]
    uint32_t* hist = reinterpret_cast<uint32_t*>(slots + SG_WARPS * 32);         // [SG_WARPS][32]
    uint32_t* nslot = hist + SG_WARPS * 32;                                      // [SG_WARPS]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int e = (int)(blockIdx.x / (unsigned)P.ngenes), g = (int)(blockIdx.x % (unsigned)P.ngenes);
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const double inf = __longlong_as_double(0x7ff0000000000000ll);
    // the comparison indices of a warp's next table are requested before the current table is worked on
    int pk_next[SG_VPL];
    auto fetch_indices = [&](int k) {
#pragma unroll
        for (int u = 0; u < SG_VPL; ++u) pk_next[u] = -1;
        if (k < P.K) {
            const int j0 = P.right_ptr[k], nr = P.right_ptr[k + 1] - j0;
#pragma unroll
            for (int u = 0; u < SG_VPL; ++u)
                if (lane + 32 * u < nr) pk_next[u] = P.q_packed[j0 + lane + 32 * u];
        }
    };
    fetch_indices(warp);
    {
        const double* col = P.eff + (size_t)e * P.Q * P.ld + P.row0 + g;
#pragma unroll 8
        for (int q = threadIdx.x; q < P.Q; q += SG_WARPS * 32) vals[q] = col[(size_t)q * P.ld];
    }
    __syncthreads();
    double* myslots = slots + warp * 32;
    uint32_t* myhist = hist + warp * 32;

    for (int k = warp; k < P.K; k += SG_WARPS) {
        double x[SG_VPL];
        int pks[SG_VPL];
#pragma unroll
        for (int u = 0; u < SG_VPL; ++u) pks[u] = pk_next[u];
        fetch_indices(k + SG_WARPS);
#pragma unroll
        for (int u = 0; u < SG_VPL; ++u) {
            x[u] = nan;
            {
                const int pk = pks[u];  // >= 0: the comparison as it is; <= -2: comparison -2 - pk mirrored; -1: none
                if (pk >= 0) {
                    x[u] = vals[pk];
                } else if (pk <= -2) {
                    const double y = vals[-2 - pk];
                    x[u] = e == 1 ? 1.0 - y : -y;
                }
            }
        }
        // ---- sum, counts, range of the finite values ----
        DD sum = {0.0, 0.0};
        double fmn = inf, fmx = -inf;
        int cnt = 0, n_ninf = 0, n_pinf = 0;
#pragma unroll
        for (int u = 0; u < SG_VPL; ++u) {
            if (!isnan(x[u])) {
                dd_add(sum, x[u]);
                ++cnt;
                if (isinf(x[u])) {
                    if (x[u] < 0.0) ++n_ninf; else ++n_pinf;
                } else {
                    fmn = fmin(fmn, x[u]);
                    fmx = fmax(fmx, x[u]);
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double oh = __shfl_xor_sync(0xffffffffu, sum.hi, o), ol = __shfl_xor_sync(0xffffffffu, sum.lo, o);
            dd_add(sum, oh);
            sum.lo += ol;
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        n_ninf = __reduce_add_sync(0xffffffffu, n_ninf);
        n_pinf = __reduce_add_sync(0xffffffffu, n_pinf);
        fmn = warp_min_f64(fmn);
        fmx = warp_max_f64(fmx);
        const int n_fin = cnt - n_ninf - n_pinf;
        const double mn = n_ninf > 0 ? -inf : (n_fin > 0 ? fmn : inf);
        const double mx = n_pinf > 0 ? inf : (n_fin > 0 ? fmx : -inf);

        // ---- the values at sorted positions r_lo = (cnt - 1) / 2 and r_lo + 1 (everything below is warp-uniform) ----
        double v_lo = nan, v_next = nan;
        const int r_lo = (cnt - 1) / 2;
        bool searching = false;
        int target = 0;
        double lo = fmn, hi = fmx;
        double above = n_pinf > 0 ? inf : nan;
        if (cnt > 0) {
            if (r_lo < n_ninf) {
                v_lo = -inf;
                v_next = r_lo + 1 < n_ninf ? -inf : (n_fin > 0 ? fmn : (n_pinf > 0 ? inf : nan));
            } else if (r_lo >= n_ninf + n_fin) {
                v_lo = inf;
                v_next = inf;
            } else {
                searching = true;
                target = r_lo - n_ninf;
            }
        }
        int ncand = n_fin;
        while (searching) {
            if (lo == hi) {  // all candidates are one value
                v_lo = lo;
                v_next = target + 1 < ncand ? lo : above;
                break;
            }
            const double sc = 1.0 / 32.0;
            const double width = (hi * sc - lo * sc) * sc;
            const double inv = width > 0.0 ? 1.0 / width : 0.0;
            const bool by_inv = inv < inf;
            auto bin_of = [&](double v) {
                const double t = v * sc - lo * sc;
                return min(31, (int)(by_inv ? t * inv : t / width));
            };
            myhist[lane] = 0;
            __syncwarp();
            int bin[SG_VPL];
#pragma unroll
            for (int u = 0; u < SG_VPL; ++u) {
                bin[u] = -1;
                if (!isnan(x[u]) && x[u] >= lo && x[u] <= hi) {
                    bin[u] = bin_of(x[u]);
                    atomicAdd(&myhist[bin[u]], 1u);
                }
            }
            __syncwarp();
            const int c_lane = (int)myhist[lane];
            int incl = c_lane;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            const int bstar = __ffs(__ballot_sync(0xffffffffu, target < incl)) - 1;  // the first bin that reaches the rank
            const int c = __shfl_sync(0xffffffffu, c_lane, bstar);
            const int before = __shfl_sync(0xffffffffu, incl - c_lane, bstar);
            // the smallest candidate above the bin
            double up = inf;
            bool any_up = false;
#pragma unroll
            for (int u = 0; u < SG_VPL; ++u) {
                if (bin[u] > bstar) {
                    up = fmin(up, x[u]);
                    any_up = true;
                }
            }
            up = warp_min_f64(up);
            const double up_total = __any_sync(0xffffffffu, any_up) ? up : above;
            if (c <= 32) {  // rank the bin's values across the lanes
                if (lane == 0) nslot[warp] = 0;
                __syncwarp();
#pragma unroll
                for (int u = 0; u < SG_VPL; ++u)
                    if (bin[u] == bstar) myslots[atomicAdd(&nslot[warp], 1u)] = x[u];
                __syncwarp();
                const double xi = lane < c ? myslots[lane] : inf;
                int rank = 0;
                for (int j = 0; j < c; ++j) {
                    const double xj = __shfl_sync(0xffffffffu, xi, j);
                    rank += (xj < xi || (xj == xi && j < lane)) ? 1 : 0;
                }
                const int idx = target - before;
                const int src_lo = __ffs(__ballot_sync(0xffffffffu, lane < c && rank == idx)) - 1;
                v_lo = __shfl_sync(0xffffffffu, xi, src_lo);
                if (idx + 1 < c) {
                    const int src_next = __ffs(__ballot_sync(0xffffffffu, lane < c && rank == idx + 1)) - 1;
                    v_next = __shfl_sync(0xffffffffu, xi, src_next);
                } else {
                    v_next = up_total;
                }
                __syncwarp();
                break;
            }
            double bmin = inf, bmax = -inf;
#pragma unroll
            for (int u = 0; u < SG_VPL; ++u) {
                if (bin[u] == bstar) {
                    bmin = fmin(bmin, x[u]);
                    bmax = fmax(bmax, x[u]);
                }
            }
            lo = warp_min_f64(bmin);
            hi = warp_max_f64(bmax);
            target -= before;
            ncand = c;
            above = up_total;
        }
        if (lane < 4) {
            const double total = sum.hi + sum.lo;
            const double mean = cnt > 0 ? total / (double)cnt : nan;
            const double med = cnt > 0 ? ((cnt & 1) ? v_lo : (v_lo + v_next) / 2.0) : nan;
            const double w = lane == 0 ? mean : lane == 1 ? mn : lane == 2 ? med : mx;
            P.out[((size_t)k * SCRANB200_NCOLS + 4 + 5 * e + lane) * P.ld + P.row0 + g] = w;
        }
    }
}

// ------------------------------------------------------------------------------------------
struct RankParams {
    const double* x;  // columns of length ngenes, column c at x + col_src[c] * ld
    int64_t ld;
    int ngenes, ncols, cap;
    const int32_t *col_src, *tab_direct, *tab_flipped, *is_auc;
    int32_t* minrank;  // [tables][ngenes]
    uint64_t *g_ka, *g_kb;  // per-CTA global sort buffers (used when ngenes > cap)
    uint32_t *g_pa, *g_pb;
};

__global__ void __launch_bounds__(SORT_THREADS) k5_rank(RankParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* s_ka = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* s_kb = s_ka + P.cap;
    uint32_t* s_pa = reinterpret_cast<uint32_t*>(s_kb + P.cap);
    uint32_t* s_pb = s_pa + P.cap;
    uint32_t* whist = s_pb + P.cap;
    uint32_t* misc = whist + SORT_HIST_WORDS;
    const int n = P.ngenes;
    const bool in_smem = n <= P.cap;
    uint64_t* ka = in_smem ? s_ka : P.g_ka + (size_t)blockIdx.x * n;
    uint64_t* kb = in_smem ? s_kb : P.g_kb + (size_t)blockIdx.x * n;
    uint32_t* pa = in_smem ? s_pa : P.g_pa + (size_t)blockIdx.x * n;
    uint32_t* pb = in_smem ? s_pb : P.g_pb + (size_t)blockIdx.x * n;

    for (int c = blockIdx.x; c < P.ncols; c += gridDim.x) {
        const int td = P.tab_direct[c], tf = P.tab_flipped[c];
        if (td < 0 && tf < 0) continue;
        const double* col = P.x + (size_t)P.col_src[c] * P.ld;
        __syncthreads();
        if (threadIdx.x == 0) misc[0] = 0;
        __syncthreads();
        uint32_t nans = 0;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const double v = col[i];
            const bool bad = isnan(v);
            ka[i] = bad ? ~0ull : key_of(v);
            pa[i] = (uint32_t)i;
            nans += bad;
        }
        if (nans) atomicAdd(&misc[0], nans);
        __syncthreads();
        const int cur = block_sort_keys<uint32_t>(ka, pa, kb, pb, n, whist);
        const uint64_t* k = cur ? kb : ka;
        const uint32_t* p = cur ? pb : pa;
        const int nv = n - (int)misc[0];
        const bool auc = P.is_auc[c] != 0;
        for (int t = threadIdx.x; t < nv; t += blockDim.x) {
            const uint64_t kt = k[t];
            const uint32_t gene = p[t];
            if (td >= 0) {  // 1 + #(x > x_t)
                int lo = t, hi = nv;  // first index with key > kt
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (k[mid] > kt) hi = mid; else lo = mid + 1;
                }
                atomicMin(&P.minrank[(size_t)td * n + gene], 1 + (nv - lo));
            }
            if (tf >= 0) {  // 1 + #(REVERSE(x) > REVERSE(x_t))
                int cntv;
                if (!auc) {  // -x > -x_t  <=>  x < x_t
                    int lo = 0, hi = t;  // first index with key >= kt
                    while (lo < hi) {
                        const int mid = (lo + hi) >> 1;
                        if (k[mid] >= kt) hi = mid; else lo = mid + 1;
                    }
                    cntv = lo;
                } else {  // 1 - x is non-increasing along the sorted column
                    const double rt = 1.0 - value_of(kt);
                    int lo = 0, hi = t;  // first index with 1 - x <= rt
                    while (lo < hi) {
                        const int mid = (lo + hi) >> 1;
                        if (1.0 - value_of(k[mid]) <= rt) hi = mid; else lo = mid + 1;
                    }
                    cntv = lo;
                }
                atomicMin(&P.minrank[(size_t)tf * n + gene], 1 + cntv);
            }
        }
    }
}

__global__ void k5_fill_i32(int32_t* d, int64_t n, int32_t v) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) d[t] = v;
}

// ------------------------------------------------------------------------------------------
// k5_rank_buckets: the same ranks without sorting the column. One CTA per "run" = one comparison
// column seen through one key (x, or 1 - x for the mirrored AUC table, whose rounding may merge
// neighbouring values: R/scoreMarkers.R:373-377) and up to two output tables: 1 + #(key > own key)
// and 1 + #(key < own key).
// The valid (non-NaN) values are spread over as many value buckets as there are values: 1024
// equal-width coarse buckets over [min, max], each divided into as many fine buckets as it has
// values (so the fine buckets follow the density). Count -> scan gives every fine bucket's
// position range [start, end) in the sorted order without sorting; what is left is the order
// INSIDE a gene's own bucket, for which the keys of a window of buckets are scattered to shared
// memory: a bucket of <= 32 is scanned, a larger one (thousands of genes share AUC = 0.5 or d = 0)
// is checked for a single value, else sorted by one warp and searched. 1 + #greater = 1 + (valid -
// end) + greater inside the bucket; 1 + #less = 1 + start + less inside the bucket.
// ------------------------------------------------------------------------------------------
constexpr int RB_COARSE = 1024;
constexpr int RB_SMALL = 32;

struct RankRunParams {
    const double* x;
    int64_t ld;
    int n, nruns, capk;
    const int32_t *src, *flip, *tab_greater, *tab_less;  // per run
    int32_t* minrank;
};

__device__ __forceinline__ uint32_t rb_fine(uint64_t key, const uint32_t* coff, double fmin, double scale) {
    const double x = (value_of(key) - fmin) * scale;  // the conversions saturate; +-inf end up in the end buckets
    const uint32_t d = min(__double2uint_rz(x), (uint32_t)RB_COARSE - 1u);
    const uint32_t lo = coff[d], c = coff[d + 1] - lo;
    return lo + min(__double2uint_rz((x - (double)d) * (double)c), c - 1u);
}

// in-place ascending sort of sz keys in shared memory by one warp (bitonic with flips; positions >= sz
// act as +infinity and are never touched)
__device__ inline void rb_warp_sort(uint64_t* k, int sz, int lane) {
    int n2 = 2;
    while (n2 < sz) n2 <<= 1;
    for (int span = 2; span <= n2; span <<= 1) {
        const int half = span >> 1;
        for (int i = lane; i < (n2 >> 1); i += 32) {
            const int blk = i / half, r = i - blk * half;
            const int a = blk * span + r, b = blk * span + span - 1 - r;
            if (b < sz) {
                const uint64_t ka = k[a], kb = k[b];
                if (kb < ka) {
                    k[a] = kb;
                    k[b] = ka;
                }
            }
        }
        __syncwarp();
        for (int j = span >> 2; j > 0; j >>= 1) {
            for (int i = lane; i < (n2 >> 1); i += 32) {
                const int a = ((i / j) * 2 * j) + (i % j), b = a + j;
                if (b < sz) {
                    const uint64_t ka = k[a], kb = k[b];
                    if (kb < ka) {
                        k[a] = kb;
                        k[b] = ka;
                    }
                }
            }
            __syncwarp();
        }
    }
}

__global__ void __launch_bounds__(1024) k5_rank_buckets(RankRunParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int n = P.n;
    uint64_t* skeys = reinterpret_cast<uint64_t*>(smem_raw);                 // [capk]
    double* red = reinterpret_cast<double*>(skeys + P.capk);                  // [64]
    unsigned long long* kred = reinterpret_cast<unsigned long long*>(red + 64);  // [2]
    uint32_t* coff = reinterpret_cast<uint32_t*>(kred + 2);                   // [RB_COARSE + 1]
    uint32_t* scan = coff + RB_COARSE + 1;                                    // [32]
    uint32_t* misc = scan + 32;                                               // [8]
    uint32_t* biglist = misc + 8;                                             // [n / RB_SMALL + 2]
    uint32_t* bk = biglist + (n / RB_SMALL + 2);                              // [(n + 1) / 2]: 16-bit counters
    uint16_t* bh = reinterpret_cast<uint16_t*>(bk);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const double inf = __longlong_as_double(0x7ff0000000000000ll);

    for (int run = blockIdx.x; run < P.nruns; run += gridDim.x) {
        const double* col = P.x + (size_t)P.src[run] * P.ld;
        const bool flip = P.flip[run] != 0;
        const int tg = P.tab_greater[run], tl = P.tab_less[run];
        auto value = [&](int i) {
            const double v = col[i];
            return flip ? 1.0 - v : v;
        };
        // f(i, value) for every row of the column, four loads in flight per thread (every pass below is a
        // handful of dependent round trips to L2 per thread otherwise: 20 rows a thread at 20,000 genes)
        auto for_each_row = [&](auto&& f) {
            for (int i0 = tid; i0 < n; i0 += 4 * (int)blockDim.x) {
                double v[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int i = i0 + q * (int)blockDim.x;
                    v[q] = i < n ? col[i] : 0.0;
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int i = i0 + q * (int)blockDim.x;
                    if (i < n) f(i, flip ? 1.0 - v[q] : v[q]);
                }
            }
        };
        __syncthreads();
        // ---- valid count and finite range ----
        {
            double lo = inf, hi = -inf;
            uint32_t nans = 0;
            for_each_row([&](int, double v) {
                if (isnan(v)) {
                    ++nans;
                } else if (!isinf(v)) {
                    lo = fmin(lo, v);
                    hi = fmax(hi, v);
                }
            });
            for (int o = 16; o > 0; o >>= 1) {
                lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
                nans += __shfl_xor_sync(0xffffffffu, nans, o);
            }
            if (lane == 0) {
                red[warp] = lo;
                red[32 + warp] = hi;
                scan[warp] = nans;
            }
            for (int i = tid; i <= RB_COARSE; i += blockDim.x) coff[i] = 0;
            for (int i = tid; i < (n + 1) / 2; i += blockDim.x) bk[i] = 0;
            __syncthreads();
            if (warp == 0) {
                double a = lane < nwarps ? red[lane] : inf, b = lane < nwarps ? red[32 + lane] : -inf;
                uint32_t c = lane < nwarps ? scan[lane] : 0u;
                for (int o = 16; o > 0; o >>= 1) {
                    a = fmin(a, __shfl_xor_sync(0xffffffffu, a, o));
                    b = fmax(b, __shfl_xor_sync(0xffffffffu, b, o));
                    c += __shfl_xor_sync(0xffffffffu, c, o);
                }
                if (lane == 0) {
                    red[0] = a;
                    red[1] = b;
                    misc[0] = c;
                    misc[1] = 0;  // number of large buckets
                }
            }
            __syncthreads();
        }
        const double fmin_v = red[0] <= red[1] ? red[0] : 0.0, fmax_v = red[0] <= red[1] ? red[1] : 0.0;
        const double scale = fmax_v > fmin_v ? (double)RB_COARSE / (fmax_v - fmin_v) : 0.0;
        const uint32_t nv = (uint32_t)n - misc[0];
        if (nv == 0) continue;
        // ---- coarse counts -> offsets ----
        for_each_row([&](int, double v) {
            if (!isnan(v)) {  // the same expression as rb_fine
                const double xx = (value_of(key_of(v)) - fmin_v) * scale;
                atomicAdd(&coff[min(__double2uint_rz(xx), (uint32_t)RB_COARSE - 1u)], 1u);
            }
        });
        __syncthreads();
        if (warp == 0) {  // exclusive scan of RB_COARSE counts: 32 per lane
            uint32_t c[RB_COARSE / 32], sum = 0;
#pragma unroll
            for (int q = 0; q < RB_COARSE / 32; ++q) {
                c[q] = coff[lane * (RB_COARSE / 32) + q];
                sum += c[q];
            }
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            uint32_t at = incl - sum;
#pragma unroll
            for (int q = 0; q < RB_COARSE / 32; ++q) {
                coff[lane * (RB_COARSE / 32) + q] = at;
                at += c[q];
            }
            if (lane == 31) coff[RB_COARSE] = at;
        }
        __syncthreads();
        // ---- fine counts -> start offsets (in place), large buckets noted ----
        for_each_row([&](int, double v) {
            if (!isnan(v)) {
                const uint32_t b = rb_fine(key_of(v), coff, fmin_v, scale);
                atomicAdd(&bk[b >> 1], 1u << ((b & 1u) * 16u));
            }
        });
        __syncthreads();
        {
            const uint32_t per = (nv + blockDim.x - 1) / blockDim.x;
            const uint32_t lo = min(nv, (uint32_t)tid * per), hi = min(nv, lo + per);
            uint32_t sum = 0;
            for (uint32_t i = lo; i < hi; ++i) sum += bh[i];
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            if (lane == 31) scan[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                uint32_t v = lane < nwarps ? scan[lane] : 0u;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
                    if (lane >= o) v += y;
                }
                scan[lane] = v;
            }
            __syncthreads();
            uint32_t at = (warp ? scan[warp - 1] : 0u) + incl - sum;
            for (uint32_t i = lo; i < hi; ++i) {
                const uint32_t c = bh[i];
                bh[i] = (uint16_t)at;
                at += c;
                if (c > (uint32_t)RB_SMALL) biglist[atomicAdd(&misc[1], 1u)] = i;
            }
        }
        __syncthreads();
        const uint32_t nbig = misc[1];

        // ---- windows of consecutive fine buckets whose keys fit shared memory ----
        uint32_t b0 = 0;  // first bucket not yet done; bh[b] = start for b >= b0, end for b < b0
        while (b0 < nv) {
            __syncthreads();
            const uint32_t wstart = bh[b0];
            if (tid == 0) {  // last bucket b1 with start[b1] - wstart <= capk (start[nv] = nv)
                uint32_t lo = b0, hi = nv;  // invariant: start[lo] - wstart <= capk
                while (lo < hi) {
                    const uint32_t mid = (lo + hi + 1) >> 1;
                    const uint32_t st = mid == nv ? nv : (uint32_t)bh[mid];
                    if (st - wstart <= (uint32_t)P.capk) lo = mid; else hi = mid - 1;
                }
                misc[2] = lo;
            }
            __syncthreads();
            uint32_t b1 = misc[2];
            const bool giant = b1 == b0;  // bucket b0 alone exceeds the window
            if (giant) b1 = b0 + 1;
            const uint32_t wend = b1 == nv ? nv : (uint32_t)bh[b1];
            __syncthreads();
            if (giant) {
                // One bucket with more keys than shared memory holds: in practice thousands of genes
                // with one and the same value. Its key range tells; if the keys do differ (values
                // clustered beyond the resolution of the two-level buckets) every member counts over
                // the whole column instead.
                if (tid < 2) kred[tid] = tid == 0 ? ~0ull : 0ull;
                __syncthreads();
                unsigned long long klo = ~0ull, khi = 0ull;
                for (int i = tid; i < n; i += blockDim.x) {
                    const double v = value(i);
                    if (isnan(v)) continue;
                    const unsigned long long key = key_of(v);
                    if (rb_fine(key, coff, fmin_v, scale) == b0) {
                        klo = key < klo ? key : klo;
                        khi = key > khi ? key : khi;
                    }
                }
                for (int o = 16; o > 0; o >>= 1) {
                    const unsigned long long a = __shfl_xor_sync(0xffffffffu, klo, o), b = __shfl_xor_sync(0xffffffffu, khi, o);
                    klo = a < klo ? a : klo;
                    khi = b > khi ? b : khi;
                }
                if (lane == 0) {
                    atomicMin(&kred[0], klo);
                    atomicMax(&kred[1], khi);
                }
                __syncthreads();
                const bool uniform = kred[0] == kred[1];
                for (int i = tid; i < n; i += blockDim.x) {
                    const double v = value(i);
                    if (isnan(v)) continue;
                    const uint64_t key = key_of(v);
                    if (rb_fine(key, coff, fmin_v, scale) != b0) continue;
                    uint32_t greater = nv - wend, less = wstart;
                    if (!uniform) {
                        greater = 0;
                        less = 0;
                        for (int j = 0; j < n; ++j) {
                            const double vj = value(j);
                            if (isnan(vj)) continue;
                            const uint64_t kj = key_of(vj);
                            greater += kj > key ? 1u : 0u;
                            less += kj < key ? 1u : 0u;
                        }
                    }
                    if (tg >= 0) atomicMin(&P.minrank[(size_t)tg * n + i], (int)(1u + greater));
                    if (tl >= 0) atomicMin(&P.minrank[(size_t)tl * n + i], (int)(1u + less));
                }
                __syncthreads();
                if (tid == 0) bh[b0] = (uint16_t)wend;
                b0 = b1;
                continue;
            }
            // scatter the window's keys into bucket order: afterwards bh[b] is the END of bucket b
            for_each_row([&](int, double v) {
                if (isnan(v)) return;
                const uint64_t key = key_of(v);
                const uint32_t b = rb_fine(key, coff, fmin_v, scale);
                if (b >= b0 && b < b1) {
                    const uint32_t sh = (b & 1u) * 16u;
                    const uint32_t pos = (atomicAdd(&bk[b >> 1], 1u << sh) >> sh) & 0xffffu;
                    skeys[pos - wstart] = key;
                }
            });
            __syncthreads();
            // large buckets of the window: one warp each; left alone if all keys are equal, else sorted
            for (uint32_t w = warp; w < nbig; w += nwarps) {
                const uint32_t b = biglist[w];
                if (b < b0 || b >= b1) continue;
                const uint32_t start = b ? (uint32_t)bh[b - 1] : 0u, end = bh[b];
                uint64_t* k = skeys + (start - wstart);
                const int sz = (int)(end - start);
                const uint64_t k0 = k[0];
                bool same = true;
                for (int i = lane; i < sz; i += 32) same &= k[i] == k0;
                if (!__all_sync(0xffffffffu, same)) rb_warp_sort(k, sz, lane);
            }
            __syncthreads();
            for_each_row([&](int i, double v) {
                if (isnan(v)) return;
                const uint64_t key = key_of(v);
                const uint32_t b = rb_fine(key, coff, fmin_v, scale);
                if (b < b0 || b >= b1) return;
                const uint32_t start = b ? (uint32_t)bh[b - 1] : 0u, end = bh[b];
                const uint64_t* k = skeys + (start - wstart);
                const uint32_t sz = end - start;
                uint32_t greater = 0, less = 0;
                if (sz <= (uint32_t)RB_SMALL) {
                    for (uint32_t j = 0; j < sz; ++j) {
                        const uint64_t kj = k[j];
                        greater += kj > key ? 1u : 0u;
                        less += kj < key ? 1u : 0u;
                    }
                } else {  // sorted (or all equal): first key >= own, first key > own
                    uint32_t lo = 0, hi = sz;
                    while (lo < hi) {
                        const uint32_t mid = (lo + hi) >> 1;
                        if (k[mid] >= key) hi = mid; else lo = mid + 1;
                    }
                    less = lo;
                    hi = sz;
                    while (lo < hi) {
                        const uint32_t mid = (lo + hi) >> 1;
                        if (k[mid] > key) hi = mid; else lo = mid + 1;
                    }
                    greater = sz - lo;
                }
                if (tg >= 0) atomicMin(&P.minrank[(size_t)tg * n + i], (int)(1u + (nv - end) + greater));
                if (tl >= 0) atomicMin(&P.minrank[(size_t)tl * n + i], (int)(1u + start + less));
            });
            b0 = b1;
        }
    }
}

// minrank: [K][3][ngenes]; rows [row0, row0 + nrows) go to rows [0, nrows) of out [K][19][ld].
__global__ void k5_finish(const int32_t* minrank, int64_t ngenes, int64_t row0, int64_t nrows, int64_t ld, int K, double* out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nrows * K * 3) return;
    const int64_t g = t % nrows;
    const int ke = (int)(t / nrows);
    const int k = ke / 3, e = ke % 3;
    const int32_t r = minrank[(size_t)ke * ngenes + row0 + g];
    const double inf = __longlong_as_double(0x7ff0000000000000ll);
    out[((size_t)k * SCRANB200_NCOLS + 4 + 5 * e + 4) * ld + g] = r == INT32_MAX ? inf : (double)r;
}

void launch_rank(scranb200_ctx* ctx, const double* d_x, int64_t ld, int64_t ngenes, const std::vector<int32_t>& col_src,
                 const std::vector<int32_t>& tab_direct, const std::vector<int32_t>& tab_flipped,
                 const std::vector<int32_t>& is_auc, int32_t* d_minrank) {
    const int ncols = (int)col_src.size();
    if (ncols == 0 || ngenes == 0) return;
    if (ngenes > 2147483647ll / 2) throw Error("too many genes to rank");
    if (ngenes <= 65000) {  // bucket ranking (16-bit positions); longer columns take the sorting kernel below
        std::vector<int32_t> src, flip, tg, tl;
        for (int c = 0; c < ncols; ++c) {
            if (tab_direct[c] < 0 && tab_flipped[c] < 0) continue;
            if (!is_auc[c]) {  // -x > -x_t  <=>  x < x_t: both tables from one run
                src.push_back(col_src[c]); flip.push_back(0); tg.push_back(tab_direct[c]); tl.push_back(tab_flipped[c]);
            } else {
                if (tab_direct[c] >= 0) {
                    src.push_back(col_src[c]); flip.push_back(0); tg.push_back(tab_direct[c]); tl.push_back(-1);
                }
                if (tab_flipped[c] >= 0) {  // ranks of 1 - x, as computed values
                    src.push_back(col_src[c]); flip.push_back(1); tg.push_back(tab_flipped[c]); tl.push_back(-1);
                }
            }
        }
        const int nruns = (int)src.size();
        if (nruns == 0) return;
        std::vector<int32_t> host;
        host.insert(host.end(), src.begin(), src.end());
        host.insert(host.end(), flip.begin(), flip.end());
        host.insert(host.end(), tg.begin(), tg.end());
        host.insert(host.end(), tl.begin(), tl.end());
        ctx->io_c.reserve(host.size() * 4);
        SB_CUDA(cudaMemcpyAsync(ctx->io_c.p, host.data(), host.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
        SB_CUDA(cudaStreamSynchronize(ctx->stream));
        RankRunParams rp;
        rp.x = d_x;
        rp.ld = ld;
        rp.n = (int)ngenes;
        rp.nruns = nruns;
        rp.src = ctx->io_c.as<int32_t>();
        rp.flip = rp.src + nruns;
        rp.tab_greater = rp.flip + nruns;
        rp.tab_less = rp.tab_greater + nruns;
        rp.minrank = d_minrank;
        const size_t fixed = 64 * 8 + 16 + (size_t)(RB_COARSE + 1 + 32 + 8) * 4 + (size_t)(ngenes / RB_SMALL + 2) * 4 +
                             (size_t)((ngenes + 1) / 2) * 4 + 16;
        rp.capk = (int)std::min<size_t>((ctx->smem_optin - fixed) / 8, (size_t)ngenes);
        const size_t smem = (size_t)rp.capk * 8 + fixed;
        allow_smem(k5_rank_buckets, smem);
        SB_LAUNCH(ctx, k5_rank_buckets, std::min(nruns, ctx->sm_count), 1024, smem, rp);
        return;
    }
    std::vector<int32_t> host;
    host.insert(host.end(), col_src.begin(), col_src.end());
    host.insert(host.end(), tab_direct.begin(), tab_direct.end());
    host.insert(host.end(), tab_flipped.begin(), tab_flipped.end());
    host.insert(host.end(), is_auc.begin(), is_auc.end());
    ctx->io_c.reserve(host.size() * 4);
    SB_CUDA(cudaMemcpyAsync(ctx->io_c.p, host.data(), host.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));

    RankParams rp;
    rp.x = d_x;
    rp.ld = ld;
    rp.ngenes = (int)ngenes;
    rp.ncols = ncols;
    const size_t fixed = (size_t)SORT_HIST_WORDS * 4 + 64;
    int cap = (int)((ctx->smem_optin - fixed) / 24);
    cap &= ~31;
    rp.cap = cap;
    rp.col_src = ctx->io_c.as<int32_t>();
    rp.tab_direct = rp.col_src + ncols;
    rp.tab_flipped = rp.tab_direct + ncols;
    rp.is_auc = rp.tab_flipped + ncols;
    rp.minrank = d_minrank;
    const int grid = std::min(ncols, ctx->sm_count);
    rp.g_ka = rp.g_kb = nullptr;
    rp.g_pa = rp.g_pb = nullptr;
    if (ngenes > cap) {
        ctx->rank_keys.reserve((size_t)grid * ngenes * 16);
        ctx->rank_idx.reserve((size_t)grid * ngenes * 8);
        rp.g_ka = ctx->rank_keys.as<uint64_t>();
        rp.g_kb = rp.g_ka + (size_t)grid * ngenes;
        rp.g_pa = ctx->rank_idx.as<uint32_t>();
        rp.g_pb = rp.g_pa + (size_t)grid * ngenes;
    }
    const size_t smem = (size_t)cap * 24 + fixed;
    allow_smem(k5_rank, smem);
    SB_LAUNCH(ctx, k5_rank, grid, SORT_THREADS, smem, rp);
}

}  // namespace

void run_summaries(scranb200_ctx* ctx, const scranb200_plan& P, int ngenes, int64_t ld, int64_t row0, const double* d_eff,
                   double* d_out) {
    if (ngenes == 0 || P.K == 0) return;
    int maxr = 1;
    for (int k = 0; k < P.K; ++k) maxr = std::max(maxr, (int)(P.right_ptr[k + 1] - P.right_ptr[k]));
    SumParams sp;
    sp.ngenes = ngenes;
    sp.K = P.K;
    sp.Q = P.Q;
    sp.maxr = maxr;
    sp.ld = ld;
    sp.row0 = row0;
    sp.eff = d_eff;
    sp.right_ptr = P.d_right_ptr;
    sp.q_direct = P.d_q_direct;
    sp.q_flipped = P.d_q_flipped;
    sp.out = d_out;
    // tables of 65 - 256 comparisons: one CTA per (gene, effect) with the gene's Q values in shared memory
    // (SCRANB200_SUMMARIES=gene / tile forces one or the other kernel: the tests run both on the same inputs)
    {
        const size_t smem = (size_t)((P.Q + 1) & ~1) * 8 + (size_t)SG_WARPS * 32 * 8 + (size_t)SG_WARPS * 32 * 4 + (size_t)SG_WARPS * 4 + 16;
        const char* force = std::getenv("SCRANB200_SUMMARIES");
        const bool can = maxr <= 32 * SG_VPL && smem <= ctx->smem_optin && (int64_t)3 * ngenes < 2147483647ll;
        const bool want = force && force[0] == 'g' ? true : force && force[0] == 't' ? false : maxr > 64;
        if (can && want) {
            SumGeneParams gp;
            gp.ngenes = ngenes;
            gp.K = P.K;
            gp.Q = P.Q;
            gp.ld = ld;
            gp.row0 = row0;
            gp.eff = d_eff;
            gp.right_ptr = P.d_right_ptr;
            gp.q_packed = P.d_q_packed;
            gp.out = d_out;
            allow_smem(k5_summaries_gene, smem);
            SB_LAUNCH(ctx, k5_summaries_gene, (unsigned)(3 * ngenes), SG_WARPS * 32, smem, gp);
            return;
        }
    }
    const int64_t items = ceil_div(ngenes, 32) * P.K;
    const size_t slots = sizeof(double) * SUM_WARPS * SUM_SLOTS * 32;
    if (maxr > 64 && maxr <= 255) {
        const size_t smem = slots + (size_t)SUM_WARPS * 128 * 32;
        allow_smem(k5_summaries<128, uint8_t>, smem);
        SB_LAUNCH(ctx, (k5_summaries<128, uint8_t>), (unsigned)ceil_div(items, SUM_WARPS), SUM_WARPS * 32, smem, sp);
    } else {
        const size_t smem = slots + (size_t)SUM_WARPS * 32 * 32 * 2;
        allow_smem(k5_summaries<32, uint16_t>, smem);
        SB_LAUNCH(ctx, (k5_summaries<32, uint16_t>), (unsigned)ceil_div(items, SUM_WARPS), SUM_WARPS * 32, smem, sp);
    }
}

// d_eff holds the comparison columns [eff_q0, eff_q0 + eff_nq) of every effect: [3][eff_nq][ld].
void run_minrank_cols(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes, int64_t ld, const double* d_eff, int64_t q0,
                      int64_t q1, int64_t eff_q0, int64_t eff_nq, int32_t* d_minrank) {
    if (q0 < eff_q0 || q1 > eff_q0 + eff_nq) throw Error("internal: comparison columns outside the effect buffer");
    std::vector<int32_t> src, td, tf, auc;
    for (int e = 0; e < 3; ++e) {
        for (int64_t q = q0; q < q1; ++q) {
            const int need = P.q_need[q];
            if (!need) continue;
            src.push_back((int32_t)(e * eff_nq + (q - eff_q0)));
            td.push_back((need & 1) ? P.group_table[P.avg_left[q]] * 3 + e : -1);
            tf.push_back((need & 2) ? P.group_table[P.avg_right[q]] * 3 + e : -1);
            auc.push_back(e == 1);
        }
    }
    launch_rank(ctx, d_eff, ld, ngenes, src, td, tf, auc, d_minrank);
}

void run_minrank(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes, int64_t ld, const double* d_eff, int64_t q0,
                 int64_t q1, int32_t* d_minrank) {
    run_minrank_cols(ctx, P, ngenes, ld, d_eff, q0, q1, 0, P.Q, d_minrank);
}

void run_fill_i32(scranb200_ctx* ctx, int32_t* d, int64_t n, int32_t v) {
    if (n > 0) SB_LAUNCH(ctx, k5_fill_i32, (unsigned)ceil_div(n, 256), 256, 0, d, n, v);
}

void run_finish_rows(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes_total, int64_t row0, int64_t nrows,
                     int64_t ld, const int32_t* d_minrank, double* d_out) {
    const int64_t n = nrows * P.K * 3;
    if (n > 0) SB_LAUNCH(ctx, k5_finish, (unsigned)ceil_div(n, 256), 256, 0, d_minrank, ngenes_total, row0, nrows, ld, P.K, d_out);
}

void run_finish(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes, int64_t ld, const int32_t* d_minrank,
                double* d_out) {
    run_finish_rows(ctx, P, ngenes, 0, ngenes, ld, d_minrank, d_out);
}

void run_min_rank_matrix(scranb200_ctx* ctx, const double* d_x, int64_t nrow, int64_t ncol, int32_t* d_out) {
    run_fill_i32(ctx, d_out, nrow, INT32_MAX);
    std::vector<int32_t> src(ncol), td(ncol, 0), tf(ncol, -1), auc(ncol, 0);
    for (int64_t c = 0; c < ncol; ++c) src[c] = (int32_t)c;
    launch_rank(ctx, d_x, nrow, nrow, src, td, tf, auc, d_out);
}

}  // namespace sb

// ------------------------------------------------------------------------------------------
// scranb200_compute_min_rank_static: computeMinRank without any device allocation.
//
// The regular min-rank path takes its scratch (column descriptors and, for columns longer than
// the shared-memory sort, the ping-pong sort buffers) from growable cudaMalloc'ed buffers of the
// context. Two kinds of caller cannot live with that: a host application that has handed the
// whole device to its own memory pool, so that a cudaMalloc issued from inside this library
// fails or fights the pool; and a caller on a latency-sensitive path, where cudaMalloc/cudaFree
// stall the device. For them this variant keeps its scratch in the fixed array below, which is
// part of the library image: every GPU that loads libscranb200.so carries these 16 MiB until the
// process exits, whether or not the entry point is ever called.
// ------------------------------------------------------------------------------------------
namespace sb {
namespace {

constexpr size_t STATIC_SCRATCH_BYTES = (size_t)16 << 20;
constexpr int STATIC_MAX_COLS = 4096;                       // descriptor table: 4 int32 per column
constexpr size_t STATIC_DESC_BYTES = (size_t)STATIC_MAX_COLS * 4 * 4;

static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

// Device address of g_static_scratch per device ordinal. In host code the symbol itself is only a
// handle; the address a kernel can use comes from cudaGetSymbolAddress, asked for once per device
// while that device is current.
constexpr int STATIC_MAX_DEVICES = 64;
unsigned char* g_static_ptr[STATIC_MAX_DEVICES] = {nullptr};

unsigned char* static_scratch_for(scranb200_ctx* ctx) {
    if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES) throw Error("device ordinal too large for the static scratch table");
    if (!g_static_ptr[ctx->device]) {
        void* p = nullptr;
        SB_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));  // ctx_guard has made ctx->device current
        g_static_ptr[ctx->device] = reinterpret_cast<unsigned char*>(p);
    }
    return g_static_ptr[ctx->device];
}

// Every column is ranked as it is and contributes to table 0. Written by the device into the
// static array (device code may name the array directly).
__global__ void static_fill_descriptors(int ncols) {
    int32_t* d = reinterpret_cast<int32_t*>(g_static_scratch);
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < ncols) {
        d[c] = c;                  // col_src
        d[ncols + c] = 0;          // tab_direct
        d[2 * ncols + c] = -1;     // tab_flipped
        d[3 * ncols + c] = 0;      // is_auc
    }
}

}  // namespace
}  // namespace sb

extern "C" int scranb200_compute_min_rank_static(scranb200_ctx* ctx, const double* d_x, int64_t nrow, int64_t ncol,
                                                  int32_t* d_out) {
    using namespace sb;
    try {
        ctx_guard(ctx);
        if (nrow < 0 || ncol < 0) throw Error("negative matrix dimension");
        if (ncol > STATIC_MAX_COLS) throw Error("compute_min_rank_static: more than 4096 columns exceed the static scratch");
        if (nrow > 2147483647ll / 2) throw Error("too many rows to rank");
        if (nrow == 0) return 0;
        unsigned char* base = static_scratch_for(ctx);

        RankParams rp;
        rp.x = d_x;
        rp.ld = nrow;
        rp.ngenes = (int)nrow;
        rp.ncols = (int)ncol;
        const size_t fixed = (size_t)SORT_HIST_WORDS * 4 + 64;
        int cap = (int)((ctx->smem_optin - fixed) / 24);
        cap &= ~31;
        rp.cap = cap;
        rp.col_src = reinterpret_cast<const int32_t*>(base);
        rp.tab_direct = rp.col_src + ncol;
        rp.tab_flipped = rp.tab_direct + ncol;
        rp.is_auc = rp.tab_flipped + ncol;
        rp.minrank = d_out;
        rp.g_ka = rp.g_kb = nullptr;
        rp.g_pa = rp.g_pb = nullptr;
        int grid = (int)std::min<int64_t>(ncol, ctx->sm_count);
        if (nrow > cap) {  // columns are sorted through the static array: 24 bytes per row and CTA
            const size_t per_cta = (size_t)nrow * 24;
            const size_t avail = STATIC_SCRATCH_BYTES - STATIC_DESC_BYTES;
            if (per_cta > avail) {
                throw Error("compute_min_rank_static: matrix too tall for the 16 MiB static scratch (limit about 696,000 rows); "
                            "use scranb200_compute_min_rank");
            }
            grid = (int)std::min<size_t>((size_t)grid, avail / per_cta);
            unsigned char* s = base + STATIC_DESC_BYTES;
            rp.g_ka = reinterpret_cast<uint64_t*>(s);
            rp.g_kb = rp.g_ka + (size_t)grid * nrow;
            rp.g_pa = reinterpret_cast<uint32_t*>(rp.g_kb + (size_t)grid * nrow);
            rp.g_pb = rp.g_pa + (size_t)grid * nrow;
        }
        // The array still holds whatever the previous call left: the descriptors are rewritten
        // here, the sort buffers are fully written before they are read, and the output starts
        // from "no rank yet".
        SB_LAUNCH(ctx, k5_fill_i32, (unsigned)ceil_div(nrow, 256), 256, 0, d_out, nrow, INT32_MAX);
        if (ncol > 0) {
            SB_LAUNCH(ctx, static_fill_descriptors, (unsigned)ceil_div(ncol, 256), 256, 0, (int)ncol);
            const size_t smem = (size_t)cap * 24 + fixed;
            allow_smem(k5_rank, smem);
            SB_LAUNCH(ctx, k5_rank, grid, SORT_THREADS, smem, rp);
        }
        SB_CUDA(cudaStreamSynchronize(ctx->stream));
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}
