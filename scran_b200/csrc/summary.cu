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

constexpr int SUM_WARPS = 8;

struct SumParams {
    int ngenes, K, Q, maxr;
    int64_t ld, row0;
    const double* eff;  // [3][Q][ld]
    const int32_t *right_ptr, *q_direct, *q_flipped;
    double* out;  // [K][19][ld]
};

__global__ void __launch_bounds__(SUM_WARPS * 32) k5_summaries(SumParams P) {
    extern __shared__ double s_vals[];  // [SUM_WARPS][maxr]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t item = (int64_t)blockIdx.x * SUM_WARPS + warp;
    if (item >= (int64_t)P.ngenes * P.K) return;
    const int g = (int)(item % P.ngenes), k = (int)(item / P.ngenes);
    double* v = s_vals + (size_t)warp * P.maxr;
    const int j0 = P.right_ptr[k], nr = P.right_ptr[k + 1] - j0;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const double inf = __longlong_as_double(0x7ff0000000000000ll);

    for (int e = 0; e < 3; ++e) {
        DD sum = {0.0, 0.0};
        double mn = inf, mx = -inf;
        int cnt = 0;
        for (int j = lane; j < nr; j += 32) {
            double x = nan;
            const int qd = P.q_direct[j0 + j], qf = P.q_flipped[j0 + j];
            if (qd >= 0) {
                x = P.eff[((size_t)e * P.Q + qd) * P.ld + P.row0 + g];
            } else if (qf >= 0) {
                const double y = P.eff[((size_t)e * P.Q + qf) * P.ld + P.row0 + g];
                x = e == 1 ? 1.0 - y : -y;
            }
            v[j] = x;
            if (!isnan(x)) {
                dd_add(sum, x);
                mn = fmin(mn, x);
                mx = fmax(mx, x);
                ++cnt;
            }
        }
        __syncwarp();
        // Warp reduction: the partial sums are combined pairwise in double-double.
        for (int o = 16; o > 0; o >>= 1) {
            const double oh = __shfl_down_sync(0xffffffffu, sum.hi, o);
            const double ol = __shfl_down_sync(0xffffffffu, sum.lo, o);
            dd_add(sum, oh);
            if (!isinf(sum.hi) && !isnan(sum.hi)) sum.lo += ol;
            mn = fmin(mn, __shfl_down_sync(0xffffffffu, mn, o));
            mx = fmax(mx, __shfl_down_sync(0xffffffffu, mx, o));
            cnt += __shfl_down_sync(0xffffffffu, cnt, o);
        }
        cnt = __shfl_sync(0xffffffffu, cnt, 0);
        // Median by rank counting: entry i has rank #(x_j < x_i) + #(x_j == x_i, j < i).
        double med_lo = nan, med_hi = nan;
        if (cnt > 0) {
            const int r_lo = (cnt - 1) / 2, r_hi = cnt / 2;
            for (int i = lane; i < nr; i += 32) {
                const double xi = v[i];
                if (isnan(xi)) continue;
                int r = 0;
                for (int j = 0; j < nr; ++j) {
                    const double xj = v[j];
                    r += (xj < xi) || (xj == xi && j < i);
                }
                if (r == r_lo) med_lo = xi;
                if (r == r_hi) med_hi = xi;
            }
            // exactly one lane holds each of them
            for (int o = 16; o > 0; o >>= 1) {
                const double a = __shfl_down_sync(0xffffffffu, med_lo, o);
                const double b = __shfl_down_sync(0xffffffffu, med_hi, o);
                if (isnan(med_lo)) med_lo = a;
                if (isnan(med_hi)) med_hi = b;
            }
        }
        if (lane == 0) {
            double* base = P.out + ((size_t)k * SCRANB200_NCOLS + 4 + 5 * e) * P.ld + P.row0 + g;
            const double total = sum.hi + sum.lo;
            base[0] = cnt > 0 ? total / (double)cnt : nan;
            base[P.ld] = mn;
            base[2 * P.ld] = cnt > 0 ? ((cnt & 1) ? med_lo : (med_lo + med_hi) / 2.0) : nan;
            base[3 * P.ld] = mx;
        }
        __syncwarp();
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

__global__ void k5_finish(const int32_t* minrank, int64_t ngenes, int64_t ld, int K, double* out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ngenes * K * 3) return;
    const int64_t g = t % ngenes;
    const int ke = (int)(t / ngenes);
    const int k = ke / 3, e = ke % 3;
    const int32_t r = minrank[t];
    const double inf = __longlong_as_double(0x7ff0000000000000ll);
    out[((size_t)k * SCRANB200_NCOLS + 4 + 5 * e + 4) * ld + g] = r == INT32_MAX ? inf : (double)r;
}

void launch_rank(scranb200_ctx* ctx, const double* d_x, int64_t ld, int64_t ngenes, const std::vector<int32_t>& col_src,
                 const std::vector<int32_t>& tab_direct, const std::vector<int32_t>& tab_flipped,
                 const std::vector<int32_t>& is_auc, int32_t* d_minrank) {
    const int ncols = (int)col_src.size();
    if (ncols == 0 || ngenes == 0) return;
    if (ngenes > 2147483647ll / 2) throw Error("too many genes to rank");
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
    const size_t smem = (size_t)SUM_WARPS * maxr * 8;
    if (smem > ctx->smem_optin) throw Error("too many comparisons per group for the summary kernel");
    allow_smem(k5_summaries, smem);
    SB_LAUNCH(ctx, k5_summaries, (unsigned)ceil_div((int64_t)ngenes * P.K, SUM_WARPS), SUM_WARPS * 32, smem, sp);
}

void run_minrank(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes, int64_t ld, const double* d_eff, int64_t q0,
                 int64_t q1, int32_t* d_minrank) {
    std::vector<int32_t> src, td, tf, auc;
    for (int e = 0; e < 3; ++e) {
        for (int64_t q = q0; q < q1; ++q) {
            const int need = P.q_need[q];
            if (!need) continue;
            src.push_back((int32_t)(e * P.Q + q));
            td.push_back((need & 1) ? P.group_table[P.avg_left[q]] * 3 + e : -1);
            tf.push_back((need & 2) ? P.group_table[P.avg_right[q]] * 3 + e : -1);
            auc.push_back(e == 1);
        }
    }
    launch_rank(ctx, d_eff, ld, ngenes, src, td, tf, auc, d_minrank);
}

void run_fill_i32(scranb200_ctx* ctx, int32_t* d, int64_t n, int32_t v) {
    if (n > 0) SB_LAUNCH(ctx, k5_fill_i32, (unsigned)ceil_div(n, 256), 256, 0, d, n, v);
}

void run_finish(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes, int64_t ld, const int32_t* d_minrank,
                double* d_out) {
    const int64_t n = ngenes * P.K * 3;
    if (n > 0) SB_LAUNCH(ctx, k5_finish, (unsigned)ceil_div(n, 256), 256, 0, d_minrank, ngenes, ld, P.K, d_out);
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
