// Shared plumbing for libscranb200: error handling, the context and matrix handles, growable
// device buffers, launch accounting. No torch, no third-party headers: CUDA runtime only.
#pragma once

#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>  // header-only; ranges cost nothing unless a profiler is attached
#include <stdint.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/scranb200.h"

namespace sb {

struct Error : public std::runtime_error {
    explicit Error(const std::string& m) : std::runtime_error(m) {}
};

inline void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
    if (e != cudaSuccess) {
        char buf[512];
        snprintf(buf, sizeof(buf), "CUDA error at %s:%d (%s): %s", file, line, what, cudaGetErrorString(e));
        throw Error(buf);
    }
}
#define SB_CUDA(x) ::sb::cuda_check((x), #x, __FILE__, __LINE__)

bool release_spares_current();  // api.cu

// A device allocation that only ever grows; contexts keep a few of these so that repeated
// calls of the same shape do not go back to cudaMalloc.
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void reserve(size_t bytes) {
        if (bytes <= cap) return;
        if (p) SB_CUDA(cudaFree(p));
        p = nullptr;
        cap = 0;
        size_t want = bytes + (bytes >> 4) + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaErrorMemoryAllocation && release_spares_current()) {
            cudaGetLastError();
            e = cudaMalloc(&p, want);
        }
        SB_CUDA(e);
        cap = want;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <typename T>
    T* as() const {
        return reinterpret_cast<T*>(p);
    }
};

enum Phase { PH_TRANSPOSE = 0, PH_STATS, PH_SORT, PH_COUNT, PH_EFFECTS, PH_SUMMARY, PH_MINRANK, PH_TOTAL, PH_N };

}  // namespace sb

struct scranb200_mat;
struct scranb200_plan;

struct scranb200_ctx {
    int device = 0;
    pid_t owner = 0;
    cudaStream_t stream = nullptr;
    int64_t launches = 0;
    int sm_count = 148;
    size_t smem_optin = 0;
    double timing[sb::PH_N] = {0};
    bool nvtx_open = false;
    cudaEvent_t ev[16] = {nullptr};
    int ev_phase[16] = {0};
    int nev = 0;
    // scratch, named by role (sizes at config 3 in DESIGN.md)
    sb::DevBuf table, seg_off, row_total, row_ptr, vals, stream_lab, unit_m, sort_keys, sort_labs, flags;
    sb::DevBuf mean, var, ndet, u_pairs, t_pairs, layout_i32, layout_f64;
    // gene-sharded runs (dist.cu): the NCCL communicator (an ncclComm_t, null for world == 1) and the exchange buffers
    void* comm = nullptr;
    int world = 1, rank = 0;
    sb::DevBuf xchg_part, xchg_recv, xchg_cols;
    sb::DevBuf eff, out, minrank, rank_keys, rank_idx, plan_tmp, io_a, io_b, io_c, grp_avg, logit, jobs, split, vmin, vmax;
    // device arrays of freed matrices, kept for the next upload of a similar size (cudaMalloc / cudaFree of
    // tens of gigabytes costs as much as a tenth of the transfer itself)
    struct Spare {
        void* p;
        size_t bytes;
    };
    std::vector<Spare> spare;
    // staging for host -> device copies from pageable memory (hostcopy.cu): pinned buffers and one stream per worker
    void* stage_pinned = nullptr;
    cudaStream_t stage_streams[8] = {nullptr};
    int stage_workers = 0;
    // live matrices and plans of this context: scranb200_ctx_destroy orphans them (frees their device
    // memory, clears their back-pointer), so that a later mat_free / plan_destroy only deletes the handle
    std::vector<scranb200_mat*> live_mats;
    std::vector<scranb200_plan*> live_plans;
};

struct scranb200_mat {
    scranb200_ctx* ctx = nullptr;
    int64_t nrow = 0, ncol = 0, nnz = 0;
    int64_t* d_p = nullptr;
    int32_t* d_i = nullptr;
    double* d_x = nullptr;
    size_t bytes_p = 0, bytes_i = 0, bytes_x = 0;  // allocation sizes (owned matrices)
    bool owned = false;
};

namespace sb {

void set_last_error(const std::string& m);
void h2d_copy(scranb200_ctx* ctx, void* dst, const void* src, size_t bytes);  // hostcopy.cu: synchronous, pinned or pageable source
void stage_release(scranb200_ctx* ctx);
void ctx_guard(scranb200_ctx* ctx);  // device current + fork check

// Launch with accounting and an immediate launch-error check.
#define SB_LAUNCH(ctx, kern, grid, block, smem, ...)                                  \
    do {                                                                              \
        kern<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                \
        (ctx)->launches++;                                                            \
        SB_CUDA(cudaPeekAtLastError());                                               \
    } while (0)

template <typename F>
inline void allow_smem(F kernel, size_t bytes) {
    SB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
}

// Device-time accounting: an event at the start of a call and one at the end of each phase.
// (Each phase is also an NVTX range on the host timeline - "scranb200:<phase>" from the previous mark to this one -
// so that a timeline profiler groups the launches by phase.)
inline void phase_begin(scranb200_ctx* ctx) {
    ctx->nev = 0;
    SB_CUDA(cudaEventRecord(ctx->ev[0], ctx->stream));
    if (ctx->nvtx_open) nvtxRangePop();
    nvtxRangePushA("scranb200:call");
    ctx->nvtx_open = true;
}
inline void phase_mark(scranb200_ctx* ctx, int k) {
    static const char* const names[PH_N] = {"scranb200:transpose done", "scranb200:stats done", "scranb200:sort done",
                                            "scranb200:count done", "scranb200:effects done", "scranb200:summaries done",
                                            "scranb200:minrank done", "scranb200:total"};
    if (k >= 0 && k < PH_N) nvtxMarkA(names[k]);
    if (ctx->nev >= 15) return;
    ctx->nev++;
    ctx->ev_phase[ctx->nev] = k;
    SB_CUDA(cudaEventRecord(ctx->ev[ctx->nev], ctx->stream));
}
void phase_collect(scranb200_ctx* ctx);  // after a stream sync: fills ctx->timing

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Order-preserving map from double to uint64: a < b  <=>  key(a) < key(b); -0.0 and +0.0 share a
// key so that key equality is exactly the reference's '==' on doubles (NaN is not expected).
__host__ __device__ inline uint64_t key_of(double v) {
#ifdef __CUDA_ARCH__
    uint64_t u = (uint64_t)__double_as_longlong(v);
#else
    uint64_t u;
    memcpy(&u, &v, 8);
#endif
    if ((u << 1) == 0) u = 0;  // -0.0 -> +0.0
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__host__ __device__ inline double value_of(uint64_t k) {
    uint64_t u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double v;
    memcpy(&v, &u, 8);
    return v;
#endif
}

}  // namespace sb
