// Gene-sharded scoreMarkers across the GPUs of one box, inside the library.
//
// The reference's only parallelism is this decomposition: beachmat::rowBlockApply cuts the genes
// into row blocks, one worker per block, the results are rbind-ed (R/scoreMarkers.R:234-248), and
// the min-rank is computed after the join because it ranks across ALL genes (:251-256). Here a
// rank (one context = one GPU) holds the rows [row_starts[r], row_starts[r + 1]) of the matrix for
// all cells. Everything up to the block-averaged effects and the mean / min / median / max
// summaries is local. The one exchange step moves the effect columns from gene shards to
// comparison shards (grouped ncclSend / ncclRecv over NVLink: an all-to-all), every rank ranks its
// own comparison columns over all genes, the per-table minima are combined with
// ncclAllReduce(min), and the finished result rows are gathered on the root (or on every rank).
// All of it is queued on the context's stream; nothing waits on the host between the stages.
//
// NCCL is loaded at run time (dlopen "libnccl.so.2") when a communicator is first asked for, so a
// single-GPU user needs no NCCL at all. Ranks may be processes (one context each; the unique id
// travels by whatever the caller has: MPI, a file, torch.distributed) or the contexts of one
// process (scranb200_ctx_comm_init_all, then one host thread per context).
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>
#include <thread>

#include "score.h"

namespace sb {

namespace {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        api.handle = h;
#define SB_SYM(name) api.name = reinterpret_cast<decltype(api.name)>(dlsym(h, "nccl" #name))
        SB_SYM(GetUniqueId);
        SB_SYM(CommInitRank);
        SB_SYM(CommInitAll);
        SB_SYM(CommDestroy);
        SB_SYM(GroupStart);
        SB_SYM(GroupEnd);
        SB_SYM(Send);
        SB_SYM(Recv);
        SB_SYM(AllReduce);
        SB_SYM(AllGather);
        SB_SYM(GetErrorString);
#undef SB_SYM
    });
    if (!api.handle || !api.GetUniqueId || !api.CommInitRank || !api.CommInitAll || !api.CommDestroy || !api.GroupStart ||
        !api.GroupEnd || !api.Send || !api.Recv || !api.AllReduce || !api.AllGather) {
        throw Error("libnccl.so.2 could not be loaded; multi-GPU scoring needs NCCL");
    }
    return api;
}

void nccl_check(ncclResult_t r, const char* what) {
    if (r != ncclSuccess) {
        const char* msg = nccl().GetErrorString ? nccl().GetErrorString(r) : "?";
        throw Error(std::string("NCCL error in ") + what + ": " + msg);
    }
}
#define SB_NCCL(x) nccl_check((x), #x)

// dst[e][q][g0 + i] = src[e][q][i] for a received block of nq columns x n rows per effect.
__global__ void __launch_bounds__(256) dist_unpack(const double* __restrict__ src, int64_t n, int64_t nq, int64_t g0,
                                                    int64_t ld, double* __restrict__ dst) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 3 * nq * n) return;
    const int64_t i = t % n, col = t / n;  // col = e * nq + q
    dst[col * ld + g0 + i] = src[t];
}

// out[k][c][g0 + i] = part[k][c][i]: rows of one rank into the full table.
__global__ void __launch_bounds__(256) dist_place_rows(const double* __restrict__ part, int64_t n, int64_t ld_part,
                                                        int64_t ncolumns, int64_t g0, int64_t ld, double* __restrict__ out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncolumns * n) return;
    const int64_t i = t % n, col = t / n;
    out[col * ld + g0 + i] = part[col * ld_part + i];
}

}  // namespace

// Who sends what where in the one exchange step: the comparison columns are cut into `world` contiguous,
// equally sized shards (q_starts); rank r sends to rank s, for every effect e, the block
// eff_local[e][q_starts[s] .. q_starts[s+1])[0 .. n_r) - contiguous, send_off / send_len doubles inside effect e's
// plane of [Q][n_r] - and receives from s the block [nq_r][n_s] of every effect into its staging buffer at
// recv_off[s] + e * recv_len[s]. Host arithmetic only: score_sharded uses it, and so does the gloo test on CPU.
void exchange_plan(int64_t Q, int world, int rank, const int64_t* row_starts, int64_t* q_starts, int64_t* send_off,
                   int64_t* send_len, int64_t* recv_off, int64_t* recv_len) {
    const int64_t qper = (Q + world - 1) / world;
    for (int s = 0; s <= world; ++s) q_starts[s] = std::min<int64_t>(Q, s * qper);
    const int64_t n_r = row_starts[rank + 1] - row_starts[rank];
    const int64_t nq_r = q_starts[rank + 1] - q_starts[rank];
    for (int s = 0; s < world; ++s) {
        send_off[s] = q_starts[s] * n_r;
        send_len[s] = (q_starts[s + 1] - q_starts[s]) * n_r;
        recv_off[s] = 3 * nq_r * row_starts[s];
        recv_len[s] = nq_r * (row_starts[s + 1] - row_starts[s]);
    }
}

void comm_release(scranb200_ctx* ctx) {
    if (ctx->comm) {
        try {
            nccl().CommDestroy(reinterpret_cast<ncclComm_t>(ctx->comm));
        } catch (...) {
        }
        ctx->comm = nullptr;
    }
}

// K5b variants used here (summary.cu)
void run_minrank_cols(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes, int64_t ld, const double* d_eff,
                      int64_t q0, int64_t q1, int64_t eff_q0, int64_t eff_nq, int32_t* d_minrank);
void run_finish_rows(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes_total, int64_t row0, int64_t nrows,
                     int64_t ld, const int32_t* d_minrank, double* d_out);

void score_stage1(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan, int64_t ld, int64_t row0,
                  double* d_eff, double* d_out);  // api.cu

// One rank's part of a sharded scoreMarkers pass. Leaves the full [K][19][ngenes] table in ctx->out
// on the root (root < 0: on every rank) and returns its device pointer (null elsewhere).
double* score_sharded(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan, int64_t ngenes,
                      const int64_t* row_starts, int root) {
    const int W = ctx->world, r = ctx->rank;
    if (W < 1) throw Error("the context has no communicator; call scranb200_ctx_comm_init first");
    if (row_starts[0] != 0 || row_starts[W] != ngenes) throw Error("row_starts must run from 0 to the number of genes");
    for (int s = 0; s < W; ++s) {
        if (row_starts[s + 1] < row_starts[s]) throw Error("row_starts must be non-decreasing");
    }
    if (root >= W) throw Error("root rank out of range");
    const int64_t g0 = row_starts[r], n = row_starts[r + 1] - g0;
    if (mat->nrow != n) throw Error("the matrix does not have this rank's number of rows");
    const int K = plan->K, Q = plan->Q;
    const int64_t nmax = [&] {
        int64_t m = 1;
        for (int s = 0; s < W; ++s) m = std::max(m, row_starts[s + 1] - row_starts[s]);
        return m;
    }();
    // comparison shards and block positions
    std::vector<int64_t> q_starts(W + 1), send_off(W), send_len(W), recv_off(W), recv_len(W);
    exchange_plan(Q, W, r, row_starts, q_starts.data(), send_off.data(), send_len.data(), recv_off.data(), recv_len.data());
    const int64_t qa = q_starts[r], qb = q_starts[r + 1], nq = qb - qa;

    const size_t K1 = (size_t)std::max(1, K), Q1 = (size_t)std::max(1, Q);
    ctx->eff.reserve(3 * Q1 * (size_t)std::max<int64_t>(1, n) * 8);                 // local effects [3][Q][n]
    // result rows: one slot of [K][19][nmax] per rank (slot r = this rank's own rows, [K][19][n])
    const size_t part_elems = K1 * SCRANB200_NCOLS * (size_t)nmax;
    ctx->xchg_part.reserve(part_elems * 8 * (size_t)W);
    ctx->minrank.reserve(K1 * 3 * (size_t)std::max<int64_t>(1, ngenes) * 4);
    double* d_eff = ctx->eff.as<double>();
    double* slots = ctx->xchg_part.as<double>();
    double* d_part = slots + (size_t)r * part_elems;
    int32_t* d_rank = ctx->minrank.as<int32_t>();

    if (n > 0) score_stage1(ctx, mat, plan, n, 0, d_eff, d_part);  // d_part: [K][19][n]

    const double* d_cols = d_eff;  // [3][nq][ngenes] after the exchange
    int64_t cols_q0 = 0, cols_nq = Q;
    if (W > 1) {
        NcclApi& N = nccl();
        ncclComm_t comm = reinterpret_cast<ncclComm_t>(ctx->comm);
        ctx->xchg_recv.reserve(3 * (size_t)std::max<int64_t>(1, nq) * (size_t)ngenes * 8);
        ctx->xchg_cols.reserve(3 * (size_t)std::max<int64_t>(1, nq) * (size_t)ngenes * 8);
        double* d_recv = ctx->xchg_recv.as<double>();
        double* d_full = ctx->xchg_cols.as<double>();
        // staging: for source s the block [3][nq][n_s] at 3 * nq * row_starts[s]
        SB_NCCL(N.GroupStart());
        for (int s = 0; s < W; ++s) {
            if (s == r) continue;
            for (int e = 0; e < 3 && send_len[s] > 0; ++e) {
                SB_NCCL(N.Send(d_eff + (size_t)e * Q * n + send_off[s], (size_t)send_len[s], ncclDouble, s, comm, ctx->stream));
            }
            for (int e = 0; e < 3 && recv_len[s] > 0; ++e) {
                SB_NCCL(N.Recv(d_recv + recv_off[s] + (size_t)e * recv_len[s], (size_t)recv_len[s], ncclDouble, s, comm, ctx->stream));
            }
        }
        SB_NCCL(N.GroupEnd());
        for (int e = 0; e < 3 && send_len[r] > 0; ++e) {  // this rank's own rows of its own columns
            SB_CUDA(cudaMemcpyAsync(d_recv + recv_off[r] + (size_t)e * recv_len[r], d_eff + (size_t)e * Q * n + send_off[r],
                                    (size_t)send_len[r] * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        }
        for (int s = 0; s < W && nq > 0; ++s) {
            const int64_t ns = row_starts[s + 1] - row_starts[s];
            if (ns == 0) continue;
            SB_LAUNCH(ctx, dist_unpack, (unsigned)ceil_div(3 * nq * ns, 256), 256, 0, d_recv + recv_off[s], ns, nq, row_starts[s], ngenes,
                      d_full);
        }
        d_cols = d_full;
        cols_q0 = qa;
        cols_nq = nq;
    }
    phase_mark(ctx, PH_SUMMARY);  // (the exchange is booked with the summaries)

    run_fill_i32(ctx, d_rank, (int64_t)K * 3 * ngenes, INT32_MAX);
    if (nq > 0 || W == 1) {
        run_minrank_cols(ctx, *plan, ngenes, W > 1 ? ngenes : n, d_cols, W > 1 ? qa : 0, W > 1 ? qb : Q, cols_q0, cols_nq, d_rank);
    }
    if (W > 1 && (int64_t)K * 3 * ngenes > 0) {
        SB_NCCL(nccl().AllReduce(d_rank, d_rank, (size_t)K * 3 * ngenes, ncclInt32, ncclMin, reinterpret_cast<ncclComm_t>(ctx->comm),
                                 ctx->stream));
    }
    run_finish_rows(ctx, *plan, ngenes, g0, n, n, d_rank, d_part);
    phase_mark(ctx, PH_MINRANK);

    // ---- gather the result rows ----
    if (W == 1) return d_part;
    const bool want = root < 0 || root == r;
    double* d_out = nullptr;
    if (want) {
        ctx->out.reserve(K1 * SCRANB200_NCOLS * (size_t)std::max<int64_t>(1, ngenes) * 8);
        d_out = ctx->out.as<double>();
    }
    if (K > 0) {
        NcclApi& N = nccl();
        ncclComm_t comm = reinterpret_cast<ncclComm_t>(ctx->comm);
        SB_NCCL(N.GroupStart());
        for (int s = 0; s < W; ++s) {
            if (s == r) continue;
            const bool s_wants = root < 0 || root == s;
            if (s_wants && n > 0) SB_NCCL(N.Send(d_part, (size_t)K * SCRANB200_NCOLS * n, ncclDouble, s, comm, ctx->stream));
            const int64_t ns = row_starts[s + 1] - row_starts[s];
            if (want && ns > 0) {
                SB_NCCL(N.Recv(slots + (size_t)s * part_elems, (size_t)K * SCRANB200_NCOLS * ns, ncclDouble, s, comm, ctx->stream));
            }
        }
        SB_NCCL(N.GroupEnd());
        for (int s = 0; s < W && want; ++s) {
            const int64_t ns = row_starts[s + 1] - row_starts[s];
            if (ns == 0) continue;
            SB_LAUNCH(ctx, dist_place_rows, (unsigned)ceil_div((int64_t)K * SCRANB200_NCOLS * ns, 256), 256, 0,
                      slots + (size_t)s * part_elems, ns, ns, (int64_t)K * SCRANB200_NCOLS, row_starts[s], ngenes, d_out);
        }
    }
    return d_out;
}

}  // namespace sb

using namespace sb;

extern "C" {

int scranb200_comm_unique_id(char* id_out) {
    try {
        ncclUniqueId id;
        static_assert(sizeof(ncclUniqueId) == SCRANB200_COMM_ID_BYTES, "ncclUniqueId size");
        SB_NCCL(nccl().GetUniqueId(&id));
        memcpy(id_out, &id, sizeof(id));
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}

int scranb200_ctx_comm_init(scranb200_ctx* ctx, int world, int rank, const char* id_in) {
    try {
        ctx_guard(ctx);
        if (world < 1 || rank < 0 || rank >= world) throw Error("bad world size or rank");
        comm_release(ctx);
        ctx->world = world;
        ctx->rank = rank;
        if (world > 1) {
            ncclUniqueId id;
            memcpy(&id, id_in, sizeof(id));
            ncclComm_t comm;
            SB_NCCL(nccl().CommInitRank(&comm, world, id, rank));
            ctx->comm = comm;
        }
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}

int scranb200_ctx_comm_init_all(scranb200_ctx** ctxs, int n) {
    try {
        if (n < 1) throw Error("need at least one context");
        std::vector<int> devs(n);
        for (int k = 0; k < n; ++k) {
            if (!ctxs[k]) throw Error("null context");
            if (ctxs[k]->owner != getpid()) throw Error("this context was created in another process (fork); create a new one");
            devs[k] = ctxs[k]->device;
            comm_release(ctxs[k]);
            ctxs[k]->world = n;
            ctxs[k]->rank = k;
        }
        if (n > 1) {
            std::vector<ncclComm_t> comms(n);
            SB_NCCL(nccl().CommInitAll(comms.data(), n, devs.data()));
            for (int k = 0; k < n; ++k) ctxs[k]->comm = comms[k];
        }
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}

int scranb200_exchange_plan(int64_t Q, int world, int rank, const int64_t* row_starts, int64_t* q_starts, int64_t* send_off,
                            int64_t* send_len, int64_t* recv_off, int64_t* recv_len) {
    try {
        if (world < 1 || rank < 0 || rank >= world || Q < 0) throw Error("bad world size, rank or number of comparisons");
        exchange_plan(Q, world, rank, row_starts, q_starts, send_off, send_len, recv_off, recv_len);
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}

int scranb200_balance_rows(const double* weight, int64_t nrow, int world, int64_t* row_starts) {
    try {
        if (world < 1 || nrow < 0) throw Error("bad world size or number of rows");
        double total = 0.0;
        for (int64_t g = 0; g < nrow; ++g) total += weight[g] > 0.0 ? weight[g] : 0.0;
        row_starts[0] = 0;
        double run = 0.0;
        int64_t g = 0;
        for (int s = 1; s < world; ++s) {  // cut s where the running weight first reaches s / world of the total
            const double target = total * (double)s / (double)world;
            while (g < nrow && run + 0.5 * (weight[g] > 0.0 ? weight[g] : 0.0) < target) {
                run += weight[g] > 0.0 ? weight[g] : 0.0;
                ++g;
            }
            row_starts[s] = g;
        }
        row_starts[world] = nrow;
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}

int scranb200_score_markers_sharded(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan,
                                    int64_t ngenes_total, const int64_t* row_starts, int root, double* out,
                                    double** d_out) {
    try {
        ctx_guard(ctx);
        if (!mat || mat->ctx != ctx) throw Error("matrix belongs to another context, or its context was destroyed");
        phase_begin(ctx);
        double* d = score_sharded(ctx, mat, plan, ngenes_total, row_starts, root);
        if (out && d && ngenes_total > 0 && plan->K > 0) {
            SB_CUDA(cudaMemcpyAsync(out, d, (size_t)plan->K * SCRANB200_NCOLS * ngenes_total * 8, cudaMemcpyDeviceToHost, ctx->stream));
        }
        SB_CUDA(cudaStreamSynchronize(ctx->stream));
        phase_collect(ctx);
        if (d_out) *d_out = d;
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}

}  // extern "C"
