// The extern "C" surface declared in include/scranb200.h. Every entry point converts C++
// exceptions into a status code plus a thread-local message, like Rcpp's BEGIN_RCPP/END_RCPP
// does for the reference (src/RcppExports.cpp:16,22).
#include <climits>
#include <cmath>

#include "score.h"

namespace sb {

static thread_local std::string g_last_error;
void set_last_error(const std::string& m) { g_last_error = m; }

static thread_local scranb200_ctx* g_current_ctx = nullptr;

void ctx_guard(scranb200_ctx* ctx) {
    if (!ctx) throw Error("null context");
    if (ctx->owner != getpid()) throw Error("this context was created in another process (fork); create a new one");
    SB_CUDA(cudaSetDevice(ctx->device));
    g_current_ctx = ctx;
}

// Called by DevBuf::reserve when cudaMalloc fails: give back the spare matrix arrays of the context
// this thread is working for; true if there was anything to give back.
bool release_spares_current() {
    scranb200_ctx* ctx = g_current_ctx;
    if (!ctx || ctx->spare.empty()) return false;
    for (auto& sp : ctx->spare) cudaFree(sp.p);
    ctx->spare.clear();
    return true;
}

// A matrix handle is only good with the context that made it, and only while that context lives.
static void mat_guard(scranb200_ctx* ctx, const scranb200_mat* mat) {
    if (!mat) throw Error("null matrix");
    if (mat->ctx != ctx) throw Error("matrix belongs to another context, or its context was destroyed");
}

void phase_collect(scranb200_ctx* ctx) {
    if (ctx->nvtx_open) {
        nvtxRangePop();
        ctx->nvtx_open = false;
    }
    for (int k = 0; k < PH_N; ++k) ctx->timing[k] = 0.0;
    for (int i = 1; i <= ctx->nev; ++i) {
        float ms = 0.f;
        SB_CUDA(cudaEventElapsedTime(&ms, ctx->ev[i - 1], ctx->ev[i]));
        ctx->timing[ctx->ev_phase[i]] += ms;
    }
    if (ctx->nev > 0) {
        float ms = 0.f;
        SB_CUDA(cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[ctx->nev]));
        ctx->timing[PH_TOTAL] = ms;
    }
}

namespace {

// Device array for an uploaded matrix: a spare one of a fitting size if the context has it, else a
// new allocation (after giving the spares back if memory is short).
template <typename T>
T* dev_alloc(scranb200_ctx* ctx, size_t n, size_t& bytes_out) {
    const size_t need = std::max<size_t>(n, 1) * sizeof(T);
    int best = -1;
    for (int k = 0; k < (int)ctx->spare.size(); ++k) {
        const size_t have = ctx->spare[k].bytes;
        if (have >= need && have <= need + need / 4 + (1u << 20) && (best < 0 || have < ctx->spare[best].bytes)) best = k;
    }
    if (best >= 0) {
        void* p = ctx->spare[best].p;
        bytes_out = ctx->spare[best].bytes;
        ctx->spare.erase(ctx->spare.begin() + best);
        return reinterpret_cast<T*>(p);
    }
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, need);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        for (auto& sp : ctx->spare) cudaFree(sp.p);
        ctx->spare.clear();
        e = cudaMalloc(&p, need);
    }
    SB_CUDA(e);
    bytes_out = need;
    return reinterpret_cast<T*>(p);
}

__global__ void widen_i32(const int32_t* in, int64_t n, int64_t* out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) out[t] = in[t];
}

// Dense column-major -> CSC without the exact zeros. One warp per column, rows in order.
__global__ void dense_count(const double* x, int64_t nrow, int64_t ncol, int64_t* counts) {
    const int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (c >= ncol) return;
    int64_t n = 0;
    for (int64_t r = lane; r < nrow; r += 32) n += x[c * nrow + r] != 0.0;
    for (int o = 16; o > 0; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
    if (lane == 0) counts[c] = n;
}
__global__ void dense_fill(const double* x, int64_t nrow, int64_t ncol, const int64_t* p, int32_t* idx, double* val) {
    const int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (c >= ncol) return;
    int64_t at = p[c];
    for (int64_t r0 = 0; r0 < nrow; r0 += 32) {
        const int64_t r = r0 + lane;
        const double v = r < nrow ? x[c * nrow + r] : 0.0;
        const bool keep = r < nrow && v != 0.0;
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int64_t o = at + __popc(m & ((1u << lane) - 1u));
            idx[o] = (int32_t)r;
            val[o] = v;
        }
        at += __popc(m);
    }
}
// counts[0..n) -> exclusive prefix in out[0..n], one CTA.
__global__ void __launch_bounds__(1024) scan_i64(const int64_t* counts, int64_t n, int64_t* out) {
    __shared__ int64_t part[1024];
    const int t = threadIdx.x;
    const int64_t per = (n + 1023) / 1024;
    const int64_t lo = min(n, t * per), hi = min(n, lo + per);
    int64_t s = 0;
    for (int64_t k = lo; k < hi; ++k) s += counts[k];
    part[t] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        const int64_t v = t >= d ? part[t - d] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    int64_t run = part[t] - s;
    for (int64_t k = lo; k < hi; ++k) {
        const int64_t c = counts[k];
        out[k] = run;
        run += c;
    }
    if (t == 1023) out[n] = part[1023];
}

// p_out[k] = base + p_in[k], k < n: a chunk's own column pointers (starting at 0) moved to their place in the
// whole matrix. PT = int64_t or int32_t (R's own @p).
template <typename PT>
__global__ void offset_colptr(const PT* __restrict__ p_in, int64_t n, int64_t base, int64_t* __restrict__ p_out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) p_out[t] = base + (int64_t)p_in[t];
}

scranb200_mat* new_mat(scranb200_ctx* ctx, int64_t nrow, int64_t ncol) {
    if (nrow < 0 || ncol < 0 || nrow > INT32_MAX || ncol > INT32_MAX - 1) throw Error("matrix dimensions out of range");
    scranb200_mat* m = new scranb200_mat();
    m->ctx = ctx;
    m->nrow = nrow;
    m->ncol = ncol;
    m->owned = true;
    ctx->live_mats.push_back(m);
    return m;
}

// Frees a half-built matrix when an upload throws (out of memory on a 25 GB array, a malformed `p`).
struct MatGuard {
    scranb200_mat* m;
    explicit MatGuard(scranb200_mat* mm) : m(mm) {}
    ~MatGuard() {
        if (m) scranb200_mat_free(m);
    }
    scranb200_mat* release() {
        scranb200_mat* r = m;
        m = nullptr;
        return r;
    }
};

// Column pointers of a CSC matrix (dgCMatrix @p): p[0] == 0, non-decreasing, no column longer than
// the matrix has rows, at most 2^32 - 1 stored entries. Everything downstream trusts p[c+1] - p[c]
// as a length (k1_count, the copies), so a malformed @p is refused here with a message.
template <typename PT>
int64_t validate_colptr(const PT* p, int64_t nrow, int64_t ncol) {
    if (p[0] != 0) throw Error("column pointers 'p' must start at 0");
    for (int64_t c = 0; c < ncol; ++c) {
        const int64_t d = (int64_t)p[c + 1] - (int64_t)p[c];
        if (d < 0) throw Error("column pointers 'p' must be non-decreasing");
        if (d > nrow) throw Error("a column has more stored entries than the matrix has rows");
    }
    const int64_t nnz = (int64_t)p[ncol];
    if (nnz > 4294967295ll) throw Error("matrix has more than 2^32-1 stored entries; upload it in row blocks");
    return nnz;
}

__global__ void check_colptr(const int64_t* __restrict__ p, int64_t nrow, int64_t ncol, int64_t nnz, int* bad) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c == 0 && (p[0] != 0 || p[ncol] != nnz)) *bad = 1;
    if (c < ncol) {
        const int64_t d = p[c + 1] - p[c];
        if (d < 0 || d > nrow) *bad = 1;
    }
}

void finish_upload(scranb200_ctx* ctx, scranb200_mat* m, const int32_t* i, const double* x) {
    m->d_i = dev_alloc<int32_t>(ctx, m->nnz, m->bytes_i);
    m->d_x = dev_alloc<double>(ctx, m->nnz, m->bytes_x);
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (m->nnz > 0) {
        h2d_copy(ctx, m->d_i, i, (size_t)m->nnz * 4);
        h2d_copy(ctx, m->d_x, x, (size_t)m->nnz * 8);
    }
}

// The reference's list-of-index-vectors form (src/overlap_exprs.cpp:13-32): its bounds check, message kept.
void check_group_lists(int64_t ncells, int ngroups, const int64_t* gptr, const int32_t* gcells) {
    if (ngroups < 0) throw Error("negative number of groups");
    for (int g = 0; g < ngroups; ++g) {
        if (gptr[g + 1] < gptr[g]) throw Error("group pointers must be non-decreasing");
        for (int64_t k = gptr[g]; k < gptr[g + 1]; ++k) {
            if (gcells[k] < 0 || gcells[k] >= ncells) throw Error("indices in 'groups' out of range");
        }
    }
}

struct PairRun {
    PairLayout L;
    Transposed T;
};

}  // namespace

// transpose (+ optional stats) (+ optional U / ties) for one layout; results stay on the device.
// (Also used by pvalues.cu.)
void run_pairs(scranb200_ctx* ctx, const scranb200_mat* mat, PairLayout& L, DevBuf& lbuf, bool want_stats, double lfc,
               bool want_u, bool want_t) {
    Transposed T;
    run_transpose(ctx, mat, L, T);
    phase_mark(ctx, PH_TRANSPOSE);
    const size_t cells = (size_t)std::max<int64_t>(1, mat->nrow) * std::max(1, L.C);
    if (want_stats || want_u) {  // the sort takes each segment's value range from this pass
        double *pm = nullptr, *pv = nullptr, *pd = nullptr;
        if (want_stats) {
            ctx->mean.reserve(cells * 8);
            ctx->var.reserve(cells * 8);
            ctx->ndet.reserve(cells * 8);
            pm = ctx->mean.as<double>();
            pv = ctx->var.as<double>();
            pd = ctx->ndet.as<double>();
        }
        run_stats(ctx, T, L, pm, pv, pd);
        phase_mark(ctx, PH_STATS);
    }
    if (want_u) {
        const size_t n = (size_t)std::max<int64_t>(1, mat->nrow) * std::max<int64_t>(1, L.npairs);
        ctx->u_pairs.reserve(n * 8);
        if (want_t) ctx->t_pairs.reserve(n * 8);
        run_wilcox(ctx, T, L, lfc, ctx->u_pairs.as<double>(), want_t ? ctx->t_pairs.as<double>() : nullptr);
    }
}

void comm_release(scranb200_ctx* ctx);  // dist.cu

void score_stage1(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan, int64_t ld, int64_t row0,
                  double* d_eff, double* d_out) {
    if (mat->ncol != plan->ncells) throw Error("plan was built for a different number of cells");
    if (plan->ctx != ctx) throw Error("plan was built without or for another context");
    PairLayout& L = const_cast<PairLayout&>(plan->layout);
    run_pairs(ctx, mat, L, const_cast<DevBuf&>(plan->d_layout), true, plan->lfc, true, false);
    run_effects(ctx, *plan, (int)mat->nrow, ld, row0, ctx->mean.as<double>(), ctx->var.as<double>(),
                ctx->ndet.as<double>(), ctx->u_pairs.as<double>(), d_eff, d_out);
    phase_mark(ctx, PH_EFFECTS);
    run_summaries(ctx, *plan, (int)mat->nrow, ld, row0, d_eff, d_out);
    phase_mark(ctx, PH_SUMMARY);
}

namespace {

// The transformers of src/compute_blocked_stats.cpp:129-177 applied to the stored values: one warp per
// column (the size factor belongs to the column). lognorm: log(x / sf + pseudo) / M_LN2 (:136), the
// natural log followed by a division like the reference; norm: x / sf (:168). `shift` is subtracted
// (0 unless the zeros do not stay zero, see scranb200_blocked_stats_transformed).
__global__ void __launch_bounds__(256) transform_values(const int64_t* __restrict__ p, const double* __restrict__ x, int64_t ncol,
                                                        const double* __restrict__ sf, int transform, double pseudo,
                                                        double shift, double* __restrict__ y) {
    const int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= ncol) return;
    const double target = sf[c];
    for (int64_t k = p[c] + (threadIdx.x & 31); k < p[c + 1]; k += 32) {
        const double v = x[k];
        y[k] = transform == 1 ? log(v / target + pseudo) / 0.693147180559945309417232121458176568 - shift : v / target;
    }
}

__global__ void __launch_bounds__(256) shift_means(double* mean, int64_t n, double shift) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) mean[t] += shift;  // NaN (empty block) stays NaN
}


}  // namespace
}  // namespace sb

// A matrix being assembled from column chunks (scranb200_mat_begin / _append_columns / _finish).
struct scranb200_builder {
    scranb200_ctx* ctx = nullptr;
    scranb200_mat* mat = nullptr;  // owns the device arrays while they grow
    int64_t ncol_done = 0, nnz_done = 0;
    size_t cap = 0;                // entries the value / index arrays hold
};

using namespace sb;

namespace {

// Grows the builder's index / value arrays to hold `need` entries, keeping what is there.
void builder_reserve(scranb200_builder* b, size_t need) {
    if (need <= b->cap) return;
    if (need > 4294967295ull) throw Error("matrix has more than 2^32-1 stored entries; upload it in row blocks");
    scranb200_ctx* ctx = b->ctx;
    size_t cap = std::max<size_t>(need, b->cap + b->cap / 2);
    cap = std::min<size_t>(cap, 4294967295ull);
    scranb200_mat* m = b->mat;
    size_t bytes_i = 0, bytes_x = 0;
    int32_t* ni = dev_alloc<int32_t>(ctx, cap, bytes_i);
    double* nx = nullptr;
    try {
        nx = dev_alloc<double>(ctx, cap, bytes_x);
    } catch (...) {
        cudaFree(ni);
        throw;
    }
    if (b->nnz_done > 0) {
        SB_CUDA(cudaMemcpyAsync(ni, m->d_i, (size_t)b->nnz_done * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        SB_CUDA(cudaMemcpyAsync(nx, m->d_x, (size_t)b->nnz_done * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        SB_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    if (m->d_i) cudaFree(m->d_i);
    if (m->d_x) cudaFree(m->d_x);
    m->d_i = ni;
    m->d_x = nx;
    m->bytes_i = bytes_i;
    m->bytes_x = bytes_x;
    b->cap = cap;
}

template <typename PT>
void builder_append(scranb200_builder* b, int64_t ncol_chunk, const PT* p, const int32_t* i, const double* x) {
    if (!b || !b->mat) throw Error("null or finished matrix builder");
    scranb200_ctx* ctx = b->ctx;
    ctx_guard(ctx);
    scranb200_mat* m = b->mat;
    if (ncol_chunk < 0 || b->ncol_done + ncol_chunk > m->ncol) throw Error("more columns appended than the matrix was begun with");
    if (ncol_chunk == 0) return;
    const int64_t nnz_chunk = validate_colptr(p, m->nrow, ncol_chunk);
    builder_reserve(b, (size_t)(b->nnz_done + nnz_chunk));
    // the chunk's column pointers, moved to their place; entry ncol_total is written by _finish
    ctx->io_a.reserve((size_t)(ncol_chunk + 1) * sizeof(PT));
    SB_CUDA(cudaMemcpyAsync(ctx->io_a.p, p, (size_t)ncol_chunk * sizeof(PT), cudaMemcpyHostToDevice, ctx->stream));
    SB_LAUNCH(ctx, offset_colptr<PT>, (unsigned)ceil_div(ncol_chunk, 256), 256, 0, ctx->io_a.as<PT>(), ncol_chunk, b->nnz_done,
              m->d_p + b->ncol_done);
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (nnz_chunk > 0) {  // (synchronous: the caller's buffers are free again when this returns)
        h2d_copy(ctx, m->d_i + b->nnz_done, i, (size_t)nnz_chunk * 4);
        h2d_copy(ctx, m->d_x + b->nnz_done, x, (size_t)nnz_chunk * 8);
    }
    b->ncol_done += ncol_chunk;
    b->nnz_done += nnz_chunk;
}

}  // namespace

#define API_BEGIN try {
#define API_END                           \
    }                                     \
    catch (const std::exception& e) {     \
        sb::set_last_error(e.what());     \
        return 1;                         \
    }                                     \
    catch (...) {                         \
        sb::set_last_error("unknown error"); \
        return 1;                         \
    }                                     \
    return 0;

extern "C" {

const char* scranb200_last_error(void) { return g_last_error.c_str(); }
int scranb200_version(void) { return 100; }

int scranb200_ctx_create(int device, scranb200_ctx** out) {
    API_BEGIN
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        throw Error(std::string("no CUDA device available (there is no CPU fallback): ") + cudaGetErrorString(e));
    }
    if (device < 0 || device >= count) throw Error("device ordinal out of range");
    SB_CUDA(cudaSetDevice(device));
    scranb200_ctx* ctx = new scranb200_ctx();
    ctx->device = device;
    ctx->owner = getpid();
    cudaDeviceProp prop;
    SB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        delete ctx;
        throw Error("this library is built for sm_100a (B200) only");
    }
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    SB_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    for (int k = 0; k < 16; ++k) SB_CUDA(cudaEventCreate(&ctx->ev[k]));
    *out = ctx;
    API_END
}

void scranb200_ctx_destroy(scranb200_ctx* ctx) {
    if (!ctx) return;
    if (ctx->owner == getpid()) {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        sb::DevBuf* bufs[] = {&ctx->table, &ctx->seg_off, &ctx->row_total, &ctx->row_ptr, &ctx->vals, &ctx->stream_lab,
                              &ctx->unit_m, &ctx->sort_keys, &ctx->sort_labs, &ctx->flags, &ctx->mean, &ctx->var,
                              &ctx->ndet, &ctx->u_pairs, &ctx->t_pairs, &ctx->layout_i32, &ctx->layout_f64, &ctx->eff,
                              &ctx->out, &ctx->minrank, &ctx->rank_keys, &ctx->rank_idx, &ctx->plan_tmp, &ctx->io_a,
                              &ctx->io_b, &ctx->io_c, &ctx->grp_avg, &ctx->logit, &ctx->jobs, &ctx->split, &ctx->vmin, &ctx->vmax,
                              &ctx->xchg_part, &ctx->xchg_recv, &ctx->xchg_cols};
        sb::comm_release(ctx);
        sb::stage_release(ctx);
        for (sb::DevBuf* b : bufs) b->release();
        for (auto& sp : ctx->spare) cudaFree(sp.p);
        ctx->spare.clear();
        for (scranb200_mat* m : ctx->live_mats) {  // matrices that outlive their context keep nothing
            if (m->owned) {
                cudaFree(m->d_p);
                cudaFree(m->d_i);
                cudaFree(m->d_x);
            }
            m->d_p = nullptr;
            m->d_i = nullptr;
            m->d_x = nullptr;
        }
        for (scranb200_plan* pl : ctx->live_plans) {
            pl->d_layout.release();
            pl->d_ints.release();
            pl->d_dbls.release();
        }
        for (int k = 0; k < 16; ++k) {
            if (ctx->ev[k]) cudaEventDestroy(ctx->ev[k]);
        }
        cudaStreamDestroy(ctx->stream);
    }
    for (scranb200_mat* m : ctx->live_mats) m->ctx = nullptr;
    for (scranb200_plan* pl : ctx->live_plans) pl->ctx = nullptr;
    delete ctx;
}

int scranb200_ctx_sync(scranb200_ctx* ctx) {
    API_BEGIN
    ctx_guard(ctx);
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    API_END
}

uint64_t scranb200_ctx_stream(scranb200_ctx* ctx) { return ctx ? (uint64_t)(uintptr_t)ctx->stream : 0; }
int64_t scranb200_ctx_launches(scranb200_ctx* ctx) { return ctx ? ctx->launches : 0; }

int scranb200_ctx_last_timing(scranb200_ctx* ctx, double* out_ms, int n) {
    if (!ctx || !out_ms) return 0;
    const int m = n < (int)PH_N ? n : (int)PH_N;
    for (int k = 0; k < m; ++k) out_ms[k] = ctx->timing[k];
    return m;
}

int scranb200_upload_csc(scranb200_ctx* ctx, int64_t nrow, int64_t ncol, const int64_t* p, const int32_t* i,
                         const double* x, scranb200_mat** out) {
    API_BEGIN
    ctx_guard(ctx);
    MatGuard guard(new_mat(ctx, nrow, ncol));
    scranb200_mat* m = guard.m;
    m->nnz = validate_colptr(p, nrow, ncol);
    m->d_p = dev_alloc<int64_t>(ctx, ncol + 1, m->bytes_p);
    SB_CUDA(cudaMemcpyAsync(m->d_p, p, (size_t)(ncol + 1) * 8, cudaMemcpyHostToDevice, ctx->stream));
    finish_upload(ctx, m, i, x);
    *out = guard.release();
    API_END
}

int scranb200_upload_csc32(scranb200_ctx* ctx, int64_t nrow, int64_t ncol, const int32_t* p, const int32_t* i,
                           const double* x, scranb200_mat** out) {
    API_BEGIN
    ctx_guard(ctx);
    MatGuard guard(new_mat(ctx, nrow, ncol));
    scranb200_mat* m = guard.m;
    m->nnz = validate_colptr(p, nrow, ncol);
    m->d_p = dev_alloc<int64_t>(ctx, ncol + 1, m->bytes_p);
    ctx->io_a.reserve((size_t)(ncol + 1) * 4);
    SB_CUDA(cudaMemcpyAsync(ctx->io_a.p, p, (size_t)(ncol + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
    SB_LAUNCH(ctx, widen_i32, (unsigned)ceil_div(ncol + 1, 256), 256, 0, ctx->io_a.as<int32_t>(), ncol + 1, m->d_p);
    finish_upload(ctx, m, i, x);
    *out = guard.release();
    API_END
}

int scranb200_mat_begin(scranb200_ctx* ctx, int64_t nrow, int64_t ncol_total, int64_t nnz_hint, scranb200_builder** out) {
    API_BEGIN
    ctx_guard(ctx);
    if (nnz_hint < 0) nnz_hint = 0;
    MatGuard guard(new_mat(ctx, nrow, ncol_total));
    scranb200_mat* m = guard.m;
    m->d_p = dev_alloc<int64_t>(ctx, ncol_total + 1, m->bytes_p);
    scranb200_builder* b = new scranb200_builder();
    b->ctx = ctx;
    b->mat = m;
    try {
        builder_reserve(b, (size_t)std::max<int64_t>(nnz_hint, 1));
    } catch (...) {
        delete b;
        throw;
    }
    guard.release();
    *out = b;
    API_END
}

int scranb200_mat_append_columns(scranb200_builder* b, int64_t ncol_chunk, const int64_t* p, const int32_t* i, const double* x) {
    API_BEGIN
    builder_append<int64_t>(b, ncol_chunk, p, i, x);
    API_END
}

int scranb200_mat_append_columns32(scranb200_builder* b, int64_t ncol_chunk, const int32_t* p, const int32_t* i, const double* x) {
    API_BEGIN
    builder_append<int32_t>(b, ncol_chunk, p, i, x);
    API_END
}

int scranb200_mat_finish(scranb200_builder* b, scranb200_mat** out) {
    API_BEGIN
    if (!b || !b->mat) throw Error("null or finished matrix builder");
    ctx_guard(b->ctx);
    scranb200_mat* m = b->mat;
    if (b->ncol_done != m->ncol) throw Error("fewer columns appended than the matrix was begun with");
    m->nnz = b->nnz_done;
    SB_CUDA(cudaMemcpyAsync(m->d_p + m->ncol, &m->nnz, 8, cudaMemcpyHostToDevice, b->ctx->stream));
    SB_CUDA(cudaStreamSynchronize(b->ctx->stream));
    b->mat = nullptr;
    delete b;
    *out = m;
    API_END
}

void scranb200_mat_abort(scranb200_builder* b) {
    if (!b) return;
    if (b->mat) scranb200_mat_free(b->mat);
    delete b;
}

int scranb200_upload_dense(scranb200_ctx* ctx, int64_t nrow, int64_t ncol, const double* x, scranb200_mat** out) {
    API_BEGIN
    ctx_guard(ctx);
    MatGuard guard(new_mat(ctx, nrow, ncol));
    scranb200_mat* m = guard.m;
    const size_t cells = (size_t)nrow * ncol;
    ctx->io_a.reserve(std::max<size_t>(cells, 1) * 8);
    ctx->io_b.reserve((size_t)(ncol + 1) * 8);
    if (cells) SB_CUDA(cudaMemcpyAsync(ctx->io_a.p, x, cells * 8, cudaMemcpyHostToDevice, ctx->stream));
    m->d_p = dev_alloc<int64_t>(ctx, ncol + 1, m->bytes_p);
    if (ncol > 0) {
        SB_LAUNCH(ctx, dense_count, (unsigned)ceil_div(ncol * 32, 256), 256, 0, ctx->io_a.as<double>(), nrow, ncol,
                  ctx->io_b.as<int64_t>());
    }
    SB_LAUNCH(ctx, scan_i64, 1, 1024, 0, ctx->io_b.as<int64_t>(), ncol, m->d_p);
    SB_CUDA(cudaMemcpyAsync(&m->nnz, m->d_p + ncol, 8, cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    m->d_i = dev_alloc<int32_t>(ctx, m->nnz, m->bytes_i);
    m->d_x = dev_alloc<double>(ctx, m->nnz, m->bytes_x);
    if (ncol > 0 && m->nnz > 0) {
        SB_LAUNCH(ctx, dense_fill, (unsigned)ceil_div(ncol * 32, 256), 256, 0, ctx->io_a.as<double>(), nrow, ncol, m->d_p,
                  m->d_i, m->d_x);
    }
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = guard.release();
    API_END
}

int scranb200_wrap_device_csc(scranb200_ctx* ctx, int64_t nrow, int64_t ncol, int64_t nnz, const int64_t* d_p,
                              const int32_t* d_i, const double* d_x, scranb200_mat** out) {
    API_BEGIN
    ctx_guard(ctx);
    if (nnz < 0 || nnz > 4294967295ll) throw Error("matrix has more than 2^32-1 stored entries; wrap it in row blocks");
    MatGuard guard(new_mat(ctx, nrow, ncol));
    scranb200_mat* m = guard.m;
    m->owned = false;
    m->nnz = nnz;
    m->d_p = const_cast<int64_t*>(d_p);
    m->d_i = const_cast<int32_t*>(d_i);
    m->d_x = const_cast<double*>(d_x);
    ctx->flags.reserve(64);
    int* bad = ctx->flags.as<int>();
    SB_CUDA(cudaMemsetAsync(bad, 0, 4, ctx->stream));
    SB_LAUNCH(ctx, check_colptr, (unsigned)ceil_div(std::max<int64_t>(1, ncol), 256), 256, 0, d_p, nrow, ncol, nnz, bad);
    int h_bad = 0;
    SB_CUDA(cudaMemcpyAsync(&h_bad, bad, 4, cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (h_bad) throw Error("column pointers 'p' must start at 0, be non-decreasing, keep every column within nrow entries and end at nnz");
    *out = guard.release();
    API_END
}

void scranb200_mat_free(scranb200_mat* mat) {
    if (!mat) return;
    if (mat->ctx) {  // (null: the context went first and has already released what the matrix held)
        auto& live = mat->ctx->live_mats;
        live.erase(std::remove(live.begin(), live.end(), mat), live.end());
    }
    if (mat->owned && mat->ctx && mat->ctx->owner == getpid()) {
        cudaSetDevice(mat->ctx->device);
        // kept for the next upload; at most two matrices' worth
        scranb200_ctx* ctx = mat->ctx;
        void* ptrs[3] = {mat->d_p, mat->d_i, mat->d_x};
        const size_t sizes[3] = {mat->bytes_p, mat->bytes_i, mat->bytes_x};
        for (int k = 0; k < 3; ++k) {
            if (!ptrs[k]) continue;
            if (sizes[k] > 0 && ctx->spare.size() < 6) ctx->spare.push_back({ptrs[k], sizes[k]});
            else cudaFree(ptrs[k]);
        }
    }
    delete mat;
}

int64_t scranb200_mat_nnz(const scranb200_mat* mat) { return mat ? mat->nnz : 0; }

int scranb200_download_csc(scranb200_ctx* ctx, const scranb200_mat* mat, int64_t* p, int32_t* i, double* x) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    SB_CUDA(cudaMemcpyAsync(p, mat->d_p, (size_t)(mat->ncol + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (mat->nnz > 0) {
        SB_CUDA(cudaMemcpyAsync(i, mat->d_i, (size_t)mat->nnz * 4, cudaMemcpyDeviceToHost, ctx->stream));
        SB_CUDA(cudaMemcpyAsync(x, mat->d_x, (size_t)mat->nnz * 8, cudaMemcpyDeviceToHost, ctx->stream));
    }
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    API_END
}

int scranb200_blocked_stats_transformed(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* block, int nblocks,
                                        int transform, const double* sf, double pseudo, double* mean_out,
                                        double* var_out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    if (transform != 1 && transform != 2) throw Error("transform must be 1 (lognorm) or 2 (norm)");
    if (nblocks < 0) throw Error("negative number of blocks");
    for (int64_t c = 0; c < mat->ncol; ++c) {
        if (block[c] >= nblocks) throw Error("block id out of range");
    }
    // lognorm with pseudo != 1 does not keep zeros at zero (the reference then reads dense columns,
    // src/compute_blocked_stats.cpp:69-82,142). Every zero becomes the same constant z0 =
    // log(pseudo) / M_LN2, so the statistics are those of (value - z0), whose zeros are zeros again,
    // with z0 added back to the mean; the variance is unchanged by the shift.
    const double shift = transform == 1 && pseudo != 1.0 ? std::log(pseudo) / 0.693147180559945309417232121458176568 : 0.0;
    ctx->io_a.reserve(std::max<size_t>(1, (size_t)mat->nnz) * 8);
    ctx->io_b.reserve(std::max<size_t>(1, (size_t)mat->ncol) * 8);
    if (mat->ncol) SB_CUDA(cudaMemcpyAsync(ctx->io_b.p, sf, (size_t)mat->ncol * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (mat->ncol) {
        SB_LAUNCH(ctx, transform_values, (unsigned)ceil_div(mat->ncol * 32, 256), 256, 0, mat->d_p, mat->d_x, mat->ncol,
                  ctx->io_b.as<double>(), transform, pseudo, shift, ctx->io_a.as<double>());
    }
    scranb200_mat view = *mat;
    view.d_x = ctx->io_a.as<double>();
    view.owned = false;
    PairLayout L;
    build_layout(block, mat->ncol, nblocks, nullptr, nullptr, 0, ctx->sm_count * 4, L);
    upload_layout(ctx, L, ctx->layout_i32);
    phase_begin(ctx);
    run_pairs(ctx, &view, L, ctx->layout_i32, true, 0.0, false, false);
    const int64_t cells = mat->nrow * (int64_t)nblocks;
    if (cells) {
        if (shift != 0.0) SB_LAUNCH(ctx, shift_means, (unsigned)ceil_div(cells, 256), 256, 0, ctx->mean.as<double>(), cells, shift);
        SB_CUDA(cudaMemcpyAsync(mean_out, ctx->mean.p, (size_t)cells * 8, cudaMemcpyDeviceToHost, ctx->stream));
        SB_CUDA(cudaMemcpyAsync(var_out, ctx->var.p, (size_t)cells * 8, cudaMemcpyDeviceToHost, ctx->stream));
    }
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    phase_collect(ctx);
    API_END
}

int scranb200_blocked_stats(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* block, int nblocks,
                            double* mean_out, double* var_out, double* ndetected_out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    if (nblocks < 0) throw Error("negative number of blocks");
    for (int64_t c = 0; c < mat->ncol; ++c) {
        if (block[c] >= nblocks) throw Error("block id out of range");
    }
    PairLayout L;
    build_layout(block, mat->ncol, nblocks, nullptr, nullptr, 0, ctx->sm_count * 4, L);
    upload_layout(ctx, L, ctx->layout_i32);
    phase_begin(ctx);
    run_pairs(ctx, mat, L, ctx->layout_i32, true, 0.0, false, false);
    const size_t bytes = (size_t)mat->nrow * nblocks * 8;
    if (bytes) {
        SB_CUDA(cudaMemcpyAsync(mean_out, ctx->mean.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        SB_CUDA(cudaMemcpyAsync(var_out, ctx->var.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        if (ndetected_out) SB_CUDA(cudaMemcpyAsync(ndetected_out, ctx->ndet.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    }
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    phase_collect(ctx);
    API_END
}

int scranb200_pairwise_wilcox(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* group, int ngroups,
                              const int32_t* block, int nblocks, double lfc, int direction, double* effect_out,
                              double* logp_out, double* logfdr_out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    if (ngroups < 0 || nblocks < 1) throw Error("bad number of groups or blocks");
    if (direction < 0 || direction > 2) throw Error("direction must be 0 (any), 1 (up) or 2 (down)");
    phase_begin(ctx);
    run_pairwise_wilcox(ctx, mat, group, ngroups, block, nblocks, lfc, direction, effect_out, logp_out, logfdr_out);
    phase_collect(ctx);
    API_END
}

int scranb200_pairwise_t_tests(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* group, int ngroups,
                               const int32_t* block, int nblocks, double lfc, int direction, int std_lfc,
                               double* effect_out, double* logp_out, double* logfdr_out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    if (ngroups < 0 || nblocks < 1) throw Error("bad number of groups or blocks");
    if (direction < 0 || direction > 2) throw Error("direction must be 0 (any), 1 (up) or 2 (down)");
    phase_begin(ctx);
    run_pairwise_ttests(ctx, mat, group, ngroups, block, nblocks, lfc, direction, std_lfc, effect_out, logp_out, logfdr_out);
    phase_collect(ctx);
    API_END
}

int scranb200_overlap_exprs_paired(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* left,
                                   const int32_t* right, int64_t npairs, int ngroups, const int64_t* group_ptr,
                                   const int32_t* group_cells, double lfc, double* u_out, double* ties_out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    check_group_lists(mat->ncol, ngroups, group_ptr, group_cells);
    std::vector<int32_t> l0(npairs), r0(npairs);
    for (int64_t k = 0; k < npairs; ++k) {
        l0[k] = left[k] - 1;  // 1-based, src/overlap_exprs.cpp:215
        r0[k] = right[k] - 1;
        if (l0[k] < 0 || l0[k] >= ngroups || r0[k] < 0 || r0[k] >= ngroups) throw Error("pair index out of range");
    }
    PairLayout L;
    build_layout_members(group_ptr, group_cells, mat->ncol, ngroups, l0.data(), r0.data(), npairs, ctx->sm_count * 4, L);
    upload_layout(ctx, L, ctx->layout_i32);
    phase_begin(ctx);
    run_pairs(ctx, mat, L, ctx->layout_i32, false, lfc, true, ties_out != nullptr);
    const size_t bytes = (size_t)mat->nrow * npairs * 8;
    if (bytes) {
        SB_CUDA(cudaMemcpyAsync(u_out, ctx->u_pairs.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        if (ties_out) SB_CUDA(cudaMemcpyAsync(ties_out, ctx->t_pairs.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    }
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    phase_collect(ctx);
    API_END
}

int scranb200_overlap_exprs(scranb200_ctx* ctx, const scranb200_mat* mat, int ngroups, const int64_t* group_ptr,
                            const int32_t* group_cells, double lfc, double* u_out, double* ties_out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    check_group_lists(mat->ncol, ngroups, group_ptr, group_cells);
    std::vector<int32_t> l0, r0;
    for (int a = 0; a < ngroups; ++a) {
        if (group_ptr[a + 1] == group_ptr[a]) continue;  // src/overlap_exprs.cpp:175
        for (int b = 0; b < a; ++b) {
            if (group_ptr[b + 1] == group_ptr[b]) continue;  // :177
            l0.push_back(a);
            r0.push_back(b);
        }
    }
    const int64_t npairs = (int64_t)l0.size();
    const int64_t ngenes = mat->nrow;
    const size_t G = (size_t)ngroups;
    const double nan = std::nan("");
    for (size_t a = 0; a < G; ++a) {
        for (size_t b = 0; b < G; ++b) {
            const bool present = lfc == 0.0 ? b < a : true;  // :153-166
            for (int64_t g = 0; g < ngenes; ++g) {
                u_out[(a * G + b) * ngenes + g] = present ? 0.0 : nan;
                if (ties_out) ties_out[(a * G + b) * ngenes + g] = present ? 0.0 : nan;
            }
        }
    }
    if (npairs == 0 || ngenes == 0) return 0;
    PairLayout L;
    build_layout_members(group_ptr, group_cells, mat->ncol, ngroups, l0.data(), r0.data(), npairs, ctx->sm_count * 4, L);
    upload_layout(ctx, L, ctx->layout_i32);
    std::vector<double> hu((size_t)ngenes * npairs), ht(ties_out ? (size_t)ngenes * npairs : 0);
    phase_begin(ctx);
    for (int pass = 0; pass < (lfc == 0.0 ? 1 : 2); ++pass) {
        const double shift = pass == 0 ? lfc : -lfc;  // :179-187
        run_pairs(ctx, mat, L, ctx->layout_i32, false, shift, true, ties_out != nullptr);
        SB_CUDA(cudaMemcpyAsync(hu.data(), ctx->u_pairs.p, hu.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
        if (ties_out) SB_CUDA(cudaMemcpyAsync(ht.data(), ctx->t_pairs.p, ht.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
        SB_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int64_t k = 0; k < npairs; ++k) {
            const size_t a = pass == 0 ? l0[k] : r0[k], b = pass == 0 ? r0[k] : l0[k];
            for (int64_t g = 0; g < ngenes; ++g) {
                u_out[(a * G + b) * ngenes + g] = hu[(size_t)g * npairs + k];
                if (ties_out) ties_out[(a * G + b) * ngenes + g] = ht[(size_t)g * npairs + k];
            }
        }
    }
    phase_collect(ctx);
    API_END
}

int scranb200_plan_create(scranb200_ctx* ctx, int64_t ncells, const int32_t* groups, int ngroups,
                          const int32_t* block, const int32_t* des_left, const int32_t* des_right,
                          int64_t ndesired, double lfc, scranb200_plan** out) {
    API_BEGIN
    if (ctx) ctx_guard(ctx);  // a null context gives a host-only plan (bookkeeping tests without a GPU)
    scranb200_plan* P = new scranb200_plan();
    try {
        build_plan(*P, ncells, groups, ngroups, block, des_left, des_right, ndesired, lfc);
        if (ctx) {
            upload_plan(ctx, *P);
            ctx->live_plans.push_back(P);
        }
    } catch (...) {
        delete P;
        throw;
    }
    *out = P;
    API_END
}

void scranb200_plan_destroy(scranb200_plan* plan) {
    if (!plan) return;
    if (plan->ctx) {
        auto& live = plan->ctx->live_plans;
        live.erase(std::remove(live.begin(), live.end(), plan), live.end());
    }
    if (plan->ctx && plan->ctx->owner == getpid()) {
        cudaSetDevice(plan->ctx->device);
        plan->d_layout.release();
        plan->d_ints.release();
        plan->d_dbls.release();
    }
    delete plan;
}

int scranb200_plan_info(const scranb200_plan* plan, int64_t* info, int n) {
    API_BEGIN
    const int64_t v[5] = {plan->K, plan->ndesired, plan->Q, (int64_t)plan->left.size(), plan->C};
    for (int k = 0; k < n && k < 5; ++k) info[k] = v[k];
    API_END
}

int scranb200_plan_tables(const scranb200_plan* plan, int32_t* out_groups, int64_t* right_ptr, int32_t* right_group,
                          double* weight) {
    API_BEGIN
    for (int k = 0; k < plan->K; ++k) {
        if (out_groups) out_groups[k] = plan->out_group[k];
    }
    for (int k = 0; k <= plan->K; ++k) {
        if (right_ptr) right_ptr[k] = plan->right_ptr[k];
    }
    for (int64_t k = 0; k < plan->ndesired; ++k) {
        if (right_group) right_group[k] = plan->right_group[k];
        if (weight) weight[k] = plan->right_weight[k];
    }
    API_END
}

int scranb200_fill_i32(scranb200_ctx* ctx, int32_t* d_ptr, int64_t n, int32_t value) {
    API_BEGIN
    ctx_guard(ctx);
    run_fill_i32(ctx, d_ptr, n, value);
    API_END
}

int scranb200_score_effects_dev(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan,
                                int64_t ld, int64_t row0, double* d_eff, double* d_out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    phase_begin(ctx);
    score_stage1(ctx, mat, plan, ld, row0, d_eff, d_out);
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    phase_collect(ctx);
    API_END
}

int scranb200_score_minrank_dev(scranb200_ctx* ctx, const scranb200_plan* plan, int64_t ngenes, int64_t ld,
                                const double* d_eff, int64_t q0, int64_t q1, int32_t* d_minrank) {
    API_BEGIN
    ctx_guard(ctx);
    if (q0 < 0 || q1 > plan->Q || q0 > q1) throw Error("comparison range out of bounds");
    phase_begin(ctx);
    run_minrank(ctx, *plan, ngenes, ld, d_eff, q0, q1, d_minrank);
    phase_mark(ctx, PH_MINRANK);
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    phase_collect(ctx);
    API_END
}

int scranb200_score_finish_dev(scranb200_ctx* ctx, const scranb200_plan* plan, int64_t ngenes, int64_t ld,
                               const int32_t* d_minrank, double* d_out) {
    API_BEGIN
    ctx_guard(ctx);
    run_finish(ctx, *plan, ngenes, ld, d_minrank, d_out);
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    API_END
}

int scranb200_score_markers(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan, double* out,
                            double* full_out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    const int64_t ngenes = mat->nrow;
    const size_t n_eff = (size_t)3 * std::max(1, plan->Q) * std::max<int64_t>(1, ngenes);
    const size_t n_out = (size_t)std::max(1, plan->K) * SCRANB200_NCOLS * std::max<int64_t>(1, ngenes);
    const size_t n_rank = (size_t)std::max(1, plan->K) * 3 * std::max<int64_t>(1, ngenes);
    ctx->eff.reserve(n_eff * 8);
    ctx->out.reserve(n_out * 8);
    ctx->minrank.reserve(n_rank * 4);
    double* d_eff = ctx->eff.as<double>();
    double* d_out = ctx->out.as<double>();
    int32_t* d_rank = ctx->minrank.as<int32_t>();
    phase_begin(ctx);
    score_stage1(ctx, mat, plan, ngenes, 0, d_eff, d_out);
    run_fill_i32(ctx, d_rank, (int64_t)plan->K * 3 * ngenes, INT32_MAX);
    run_minrank(ctx, *plan, ngenes, ngenes, d_eff, 0, plan->Q, d_rank);
    run_finish(ctx, *plan, ngenes, ngenes, d_rank, d_out);
    phase_mark(ctx, PH_MINRANK);
    if (ngenes > 0 && plan->K > 0) {
        SB_CUDA(cudaMemcpyAsync(out, d_out, (size_t)plan->K * SCRANB200_NCOLS * ngenes * 8, cudaMemcpyDeviceToHost,
                                ctx->stream));
    }
    std::vector<double> heff;
    if (full_out && ngenes > 0 && plan->Q > 0) {
        heff.resize((size_t)3 * plan->Q * ngenes);
        SB_CUDA(cudaMemcpyAsync(heff.data(), d_eff, heff.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
    }
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    phase_collect(ctx);
    if (full_out && ngenes > 0) {  // full.* matrices: R/scoreMarkers.R:536-555, 571-574
        const double nan = std::nan("");
        for (int e = 0; e < 3; ++e) {
            for (int64_t j = 0; j < plan->ndesired; ++j) {
                double* dst = full_out + ((size_t)e * plan->ndesired + j) * ngenes;
                const int qd = plan->q_direct[j], qf = plan->q_flipped[j];
                if (qd >= 0) {
                    memcpy(dst, heff.data() + ((size_t)e * plan->Q + qd) * ngenes, (size_t)ngenes * 8);
                } else if (qf >= 0) {
                    const double* src = heff.data() + ((size_t)e * plan->Q + qf) * ngenes;
                    for (int64_t g = 0; g < ngenes; ++g) dst[g] = e == 1 ? 1.0 - src[g] : -src[g];
                } else {
                    for (int64_t g = 0; g < ngenes; ++g) dst[g] = nan;
                }
            }
        }
    }
    API_END
}

int scranb200_summary_marker_stats(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan, double* out) {
    API_BEGIN
    ctx_guard(ctx);
    mat_guard(ctx, mat);
    if (mat->ncol != plan->ncells) throw Error("plan was built for a different number of cells");
    if (plan->ctx != ctx) throw Error("plan was built without or for another context");
    const int64_t ngenes = mat->nrow;
    if (ngenes == 0 || plan->K == 0) return 0;
    ctx->out.reserve((size_t)plan->K * SCRANB200_NCOLS * ngenes * 8);
    double* d_out = ctx->out.as<double>();
    // only the combinations' statistics are needed: a layout without pairs (no sort, no count)
    PairLayout L;
    build_layout(plan->comb_id.data(), plan->ncells, plan->C, nullptr, nullptr, 0, ctx->sm_count * 4, L);
    upload_layout(ctx, L, ctx->layout_i32);
    phase_begin(ctx);
    run_pairs(ctx, mat, L, ctx->layout_i32, true, 0.0, false, false);
    run_effects(ctx, *plan, (int)ngenes, ngenes, 0, ctx->mean.as<double>(), ctx->var.as<double>(), ctx->ndet.as<double>(), nullptr,
                nullptr, d_out);
    phase_mark(ctx, PH_EFFECTS);
    for (int k = 0; k < plan->K; ++k) {
        SB_CUDA(cudaMemcpyAsync(out + (size_t)k * 4 * ngenes, d_out + (size_t)k * SCRANB200_NCOLS * ngenes, (size_t)4 * ngenes * 8,
                                cudaMemcpyDeviceToHost, ctx->stream));
    }
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    phase_collect(ctx);
    API_END
}

int scranb200_compute_min_rank(scranb200_ctx* ctx, const double* x, int64_t nrow, int64_t ncol, int32_t* out) {
    API_BEGIN
    ctx_guard(ctx);
    if (nrow == 0) return 0;
    ctx->io_a.reserve(std::max<size_t>(1, (size_t)nrow * ncol) * 8);
    ctx->io_b.reserve((size_t)nrow * 4);
    if (ncol > 0) SB_CUDA(cudaMemcpyAsync(ctx->io_a.p, x, (size_t)nrow * ncol * 8, cudaMemcpyHostToDevice, ctx->stream));
    run_min_rank_matrix(ctx, ctx->io_a.as<double>(), nrow, ncol, ctx->io_b.as<int32_t>());
    SB_CUDA(cudaMemcpyAsync(out, ctx->io_b.p, (size_t)nrow * 4, cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    API_END
}

}  // extern "C"

// Host-only view of a plan's block-level pairs and averaging map (bookkeeping tests).
extern "C" int scranb200_plan_pairs(const scranb200_plan* plan, int32_t* comb_group, int32_t* comb_block,
                                    double* ncells_comb, int32_t* left, int32_t* right, int32_t* avg_left,
                                    int32_t* avg_right, int32_t* avg_ptr, int32_t* avg_idx, int32_t* q_direct,
                                    int32_t* q_flipped, int32_t* comb_id) {
    try {
        for (int c = 0; c < plan->C; ++c) {
            if (comb_group) comb_group[c] = plan->comb_group[c];
            if (comb_block) comb_block[c] = plan->comb_block[c];
            if (ncells_comb) ncells_comb[c] = plan->ncells_comb[c];
        }
        for (size_t k = 0; k < plan->left.size(); ++k) {
            if (left) left[k] = plan->left[k];
            if (right) right[k] = plan->right[k];
            if (avg_idx) avg_idx[k] = plan->avg_idx[k];
        }
        for (int q = 0; q < plan->Q; ++q) {
            if (avg_left) avg_left[q] = plan->avg_left[q];
            if (avg_right) avg_right[q] = plan->avg_right[q];
        }
        for (int q = 0; q <= plan->Q; ++q) {
            if (avg_ptr) avg_ptr[q] = plan->avg_ptr[q];
        }
        for (int64_t j = 0; j < plan->ndesired; ++j) {
            if (q_direct) q_direct[j] = plan->q_direct[j];
            if (q_flipped) q_flipped[j] = plan->q_flipped[j];
        }
        if (comb_id) {
            for (int64_t c = 0; c < plan->ncells; ++c) comb_id[c] = plan->comb_id[c];
        }
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}
