// K1: cell-major CSC -> gene-major values bucketed by (gene, device combination), stable in cell
// order. Replaces the reference's per-gene dense row extraction and per-group gather
// (src/overlap_exprs.cpp:171,211 and :37-47) and the column walk of
// src/compute_blocked_stats.cpp:53-56 with one O(nnz) counting sort.
//
// Why it is stable without global atomics: the cells are visited in (combination, cell) order,
// cut into chunks of columns. The counting pass (k1_count) gives, per chunk, every gene's number
// of stored entries; scanned in (gene, chunk) order that is where each chunk's part of every
// (gene, combination) segment starts. The scatter (k1_scatter_warp) gives every warp one chunk
// and one sub-range of 256 genes: it walks the chunk's columns in order, and inside one column
// every gene occurs at most once (dgCMatrix invariant, validated in the counting pass), so the
// per-gene write cursors in shared memory are bumped with plain loads and stores; a __syncwarp
// between two columns orders the updates of consecutive cells. Every segment therefore holds its
// gene's stored values of that combination in ascending cell order: exactly the sequence the
// reference's running mean / variance update consumes.
#include "kernels.h"

namespace sb {

namespace {

constexpr int K1_THREADS = 1024;
constexpr int K1_SUB_SHIFT = 8;  // gene sub-ranges of 256: one scatter warp each
constexpr int K1_SUB = 1 << K1_SUB_SHIFT;
constexpr int K1_MAXCOLS = MAX_CHUNK_COLS;
#ifndef SB_K1C_LOADS
#define SB_K1C_LOADS 16
#endif
constexpr int K1C_LOADS = SB_K1C_LOADS;  // columns per chunk (their bounds are staged in shared memory)

// Stages the bounds of the chunk's columns in shared memory. Returns the number of columns.
__device__ inline int stage_bounds(const int64_t* __restrict__ p, const int32_t* __restrict__ perm, int c_lo, int c_hi,
                                   int64_t* s_lo, int* s_n) {
    const int ncols = c_hi - c_lo;
    for (int j = threadIdx.x; j < ncols; j += blockDim.x) {
        const int c = perm[c_lo + j];
        const int64_t a = p[c], b = p[c + 1];
        s_lo[j] = a;
        s_n[j] = (int)(b - a);
    }
    return ncols;
}

// Pass A: table[chunk][gene] = stored entries of `gene` in the chunk's columns. Counting does not
// care about order, so the warps of a CTA take columns independently (32 columns in flight per
// CTA) and bump the shared-memory counters with atomics; no barrier between columns.
__global__ void __launch_bounds__(K1_THREADS, 1)
k1_count(const int64_t* __restrict__ p, const int32_t* __restrict__ idx, const int32_t* __restrict__ perm,
         const int32_t* __restrict__ chunk_start, int ngenes, int g0, int glen, uint32_t* __restrict__ table,
         int* __restrict__ bad, uint32_t* __restrict__ sub_split, int nsub) {
    extern __shared__ __align__(16) unsigned char k1_smem[];
    int64_t* s_lo = reinterpret_cast<int64_t*>(k1_smem);
    int* s_n = reinterpret_cast<int*>(s_lo + K1_MAXCOLS);
    uint32_t* cnt = reinterpret_cast<uint32_t*>(s_n + K1_MAXCOLS);
    for (int g = threadIdx.x; g < glen; g += blockDim.x) cnt[g] = 0;
    const int chunk = blockIdx.x;
    const int c_lo = chunk_start[chunk], c_hi = chunk_start[chunk + 1];
    const int ncols = stage_bounds(p, perm, c_lo, c_hi, s_lo, s_n);
    __syncthreads();
    int flag = 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (int j = warp; j < ncols; j += nwarps) {
        const int64_t lo = s_lo[j];
        const int n = s_n[j];
        // sub_split (optional): for the column at permuted position c_lo + j, where the entries with
        // gene >= r * K1_SUB start, r = 0 .. nsub: the rows are sorted, so the boundaries are where the
        // sub-range of an entry differs from that of its left neighbour
        uint32_t* srow = sub_split ? sub_split + (size_t)(c_lo + j) * (nsub + 1) : nullptr;
        if (srow && n == 0) {
            for (int r = lane; r <= nsub; r += 32) srow[r] = 0u;
        }
        // K1C_LOADS loads in flight per lane: with 32 warps per SM the bytes in flight decide the bandwidth
        for (int k0 = 0; k0 < n; k0 += 32 * K1C_LOADS) {
            int g[K1C_LOADS];
#pragma unroll
            for (int q = 0; q < K1C_LOADS; ++q) {
                const int k = k0 + 32 * q + lane;
                g[q] = k < n ? idx[lo + k] : -2;
            }
#pragma unroll
            for (int q = 0; q < K1C_LOADS; ++q) {
                const int k = k0 + 32 * q + lane;
                // strictly increasing rows inside a column: compare with the left neighbour
                int prev = __shfl_up_sync(0xffffffffu, g[q], 1);
                const int before = q > 0 ? __shfl_sync(0xffffffffu, g[q - 1], 31) : -1;  // all lanes take part
                if (lane == 0) prev = q > 0 ? before : ((k0 > 0 && k < n) ? idx[lo + k - 1] : -1);
                if (k < n) {
                    if (g[q] < 0 || g[q] >= ngenes || prev >= g[q]) {
                        flag = 1;
                    } else {
                        const unsigned gl = (unsigned)(g[q] - g0);
                        if (gl < (unsigned)glen) atomicAdd(&cnt[gl], 1u);
                        const int sub = g[q] >> K1_SUB_SHIFT, sub_prev = prev >> K1_SUB_SHIFT;  // (-1 >> s = -1)
                        if (srow && sub != sub_prev) {
                            srow[sub] = (uint32_t)k;  // usually the only sub-range that starts here
#pragma unroll 1
                            for (int r = sub_prev + 1; r < sub; ++r) srow[r] = (uint32_t)k;  // (sub-ranges without entries)
                        }
                    }
                }
            }
        }
        if (srow && n > 0) {  // the sub-ranges behind the column's last entry: all lanes
            const int last = idx[lo + n - 1];
            if (last >= 0 && last < ngenes) {
                for (int r = (last >> K1_SUB_SHIFT) + 1 + lane; r <= nsub; r += 32) srow[r] = (uint32_t)n;
            }
        }
    }
    __syncthreads();
    if (flag) *bad = 1;
    for (int g = threadIdx.x; g < glen; g += blockDim.x) table[(size_t)chunk * ngenes + g0 + g] = cnt[g];
}

// Exclusive scan of the table along chunks for every gene; also records where every
// combination's segment starts inside the gene's row (seg_off: [C + 1][ngenes]).
__global__ void k1_scan_table(uint32_t* __restrict__ table, int nchunks, int ngenes, int C,
                              const int32_t* __restrict__ first_chunk, uint32_t* __restrict__ seg_off,
                              uint32_t* __restrict__ row_total) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ngenes) return;
    uint32_t run = 0;
    int dc = 0;
    int next_first = first_chunk[0];
    for (int k0 = 0; k0 < nchunks; k0 += 8) {  // eight loads in flight: one thread per gene is all the parallelism there is
        uint32_t v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = k0 + q < nchunks ? table[(size_t)(k0 + q) * ngenes + g] : 0u;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int k = k0 + q;
            if (k < nchunks) {
                while (dc <= C && next_first == k) {
                    seg_off[(size_t)dc * ngenes + g] = run;
                    ++dc;
                    next_first = dc <= C ? first_chunk[dc] : -1;
                }
                table[(size_t)k * ngenes + g] = run;
                run += v[q];
            }
        }
    }
    while (dc <= C) {
        seg_off[(size_t)dc * ngenes + g] = run;
        ++dc;
    }
    row_total[g] = run;
}

// row_ptr[g] = sum of row_total[0..g); one CTA. Also reports nnz of the used cells.
__global__ void __launch_bounds__(1024) k1_scan_rows(const uint32_t* __restrict__ row_total, int ngenes,
                                                      int64_t* __restrict__ row_ptr) {
    __shared__ int64_t part[1024];
    const int t = threadIdx.x;
    const int per = (ngenes + 1023) / 1024;
    const int lo = min(ngenes, t * per), hi = min(ngenes, lo + per);
    int64_t s = 0;
    for (int g = lo; g < hi; ++g) s += row_total[g];
    part[t] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        int64_t v = t >= d ? part[t - d] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    int64_t run = part[t] - s;
    for (int g = lo; g < hi; ++g) {
        row_ptr[g] = run;
        run += row_total[g];
    }
    if (t == 1023) row_ptr[ngenes] = part[1023];
}

// Pass B, one warp per (chunk, sub-range of K1_SUB genes). A gene's entries must be appended in
// column order, but different gene ranges are independent of each other: with the positions where
// every sub-range starts inside every column (sub_split, written by the counting pass) a warp can
// walk the chunk's columns on its own, one short slice (its genes' entries of that column) a step,
// with no barrier shared with other warps. The slices of the next columns are fetched while the
// current one is applied (K1W_DEPTH in flight). Per gene the write cursor and the last up-to-4 values
// are staged in shared memory, so that HBM sees whole 32-byte sectors.
constexpr int K1W_WARPS = 4;
constexpr int K1W_DEPTH = 8;  // (6, 12, 16 measured: 35.2, 39.9, 42.2 ms against 33.0 at config 3)

__global__ void __launch_bounds__(K1W_WARPS * 32)
k1_scatter_warp(const int64_t* __restrict__ p, const int32_t* __restrict__ idx, const double* __restrict__ x,
                const int32_t* __restrict__ perm, const int32_t* __restrict__ chunk_start, int ngenes, int nsub,
                const uint32_t* __restrict__ sub_split, const uint32_t* __restrict__ table,
                const int64_t* __restrict__ row_ptr, double* __restrict__ vals) {
    __shared__ double s_stage[K1W_WARPS][K1_SUB * 3];  // the fourth value of a sector never waits: it completes the sector
    __shared__ uint32_t s_cursor[K1W_WARPS][K1_SUB], s_first[K1W_WARPS][K1_SUB];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int blocks_per_chunk = (nsub + K1W_WARPS - 1) / K1W_WARPS;
    const int chunk = blockIdx.x / blocks_per_chunk, sr = (blockIdx.x % blocks_per_chunk) * K1W_WARPS + warp;
    if (sr >= nsub) return;  // (no block-wide barrier anywhere in this kernel)
    double* stage = s_stage[warp];
    uint32_t* cursor = s_cursor[warp];
    uint32_t* first = s_first[warp];
    const int g0 = sr * K1_SUB, glen = min(K1_SUB, ngenes - g0);
    for (int g = lane; g < glen; g += 32) {
        const uint32_t at = (uint32_t)(row_ptr[g0 + g] + table[(size_t)chunk * ngenes + g0 + g]);
        cursor[g] = at;
        first[g] = at;
    }
    __syncwarp();
    const int c_lo = chunk_start[chunk], c_hi = chunk_start[chunk + 1];
    const int ncols = c_hi - c_lo;

    auto put = [&](int g, double v) {
        const unsigned gl = (unsigned)(g - g0);
        if (gl >= (unsigned)glen) return;  // only reachable with an invalid (unsorted) matrix, reported by pass A
        const uint32_t at = cursor[gl];
        cursor[gl] = at + 1;
        double* sg = stage + (size_t)gl * 3;
        if ((at & 3u) != 3u) {
            sg[at & 3u] = v;
        } else {
            const uint32_t base = at - 3u, f = first[gl];
            if (base >= f) {  // a whole 32-byte sector
                double2* dst = reinterpret_cast<double2*>(vals + base);
                dst[0] = make_double2(sg[0], sg[1]);
                dst[1] = make_double2(sg[2], v);
            } else {          // the sector started in an earlier chunk: write only our part
                for (uint32_t s = f; s < at; ++s) vals[s] = sg[s & 3u];
                vals[at] = v;
            }
        }
    };
    struct Slice {
        int32_t g;
        double v;
        int64_t lo;
        int n;
    };
    for (int j0 = 0; j0 < ncols; j0 += 32) {
        // bounds of my sub-range inside the next 32 columns: one column per lane
        int64_t my_lo = 0;
        int my_n = 0;
        if (j0 + lane < ncols) {
            const int pos = c_lo + j0 + lane;
            const uint32_t* srow = sub_split + (size_t)pos * (nsub + 1);
            const uint32_t a = srow[sr], b = srow[sr + 1];
            my_lo = p[perm[pos]] + a;
            my_n = (int)(b - a);
        }
        const int nb = min(32, ncols - j0);
        Slice ring[K1W_DEPTH];
        auto load = [&](Slice& sl, int q) {
            sl.lo = __shfl_sync(0xffffffffu, my_lo, q & 31);
            sl.n = q < nb ? __shfl_sync(0xffffffffu, my_n, q & 31) : 0;
            sl.g = -1;
            if (lane < sl.n) {
                sl.g = idx[sl.lo + lane];
                sl.v = x[sl.lo + lane];
            }
        };
#pragma unroll
        for (int d = 0; d < K1W_DEPTH; ++d) load(ring[d], d);
        for (int q0 = 0; q0 < nb; q0 += K1W_DEPTH) {
#pragma unroll
            for (int d = 0; d < K1W_DEPTH; ++d) {
                Slice& sl = ring[d];
                if (sl.g >= 0) put(sl.g, sl.v);
                for (int k = 32 + lane; k < sl.n; k += 32) put(idx[sl.lo + k], x[sl.lo + k]);
                __syncwarp();  // the next column may touch the same genes from other lanes
                load(sl, q0 + d + K1W_DEPTH);
            }
        }
    }
    __syncwarp();
    // flush the incomplete last sector of every gene
    for (int g = lane; g < glen; g += 32) {
        const uint32_t c = cursor[g], f = first[g];
        if ((c & 3u) != 0u && c > f) {
            const uint32_t from = max(f, c & ~3u);
            for (uint32_t s = from; s < c; ++s) vals[s] = stage[(size_t)g * 3 + (s & 3u)];
        }
    }
}

}  // namespace

void run_transpose(scranb200_ctx* ctx, const scranb200_mat* mat, const PairLayout& L, Transposed& T) {
    const int ngenes = (int)mat->nrow;
    const int C = L.C;
    if (mat->nnz > 4294967295ll) throw Error("matrix has more than 2^32-1 stored entries; split it into row blocks");
    T.ngenes = ngenes;
    T.C = C;
    ctx->table.reserve((size_t)std::max(1, L.nchunks) * ngenes * 4);
    ctx->seg_off.reserve((size_t)(C + 1) * ngenes * 4);
    ctx->row_total.reserve((size_t)ngenes * 4 + 4);
    ctx->row_ptr.reserve((size_t)(ngenes + 1) * 8);
    ctx->vals.reserve((size_t)std::max<int64_t>(1, mat->nnz) * 8);
    ctx->flags.reserve(64);
    T.seg_off = ctx->seg_off.as<uint32_t>();
    T.row_ptr = ctx->row_ptr.as<int64_t>();
    T.vals = ctx->vals.as<double>();
    uint32_t* table = ctx->table.as<uint32_t>();
    int* bad = ctx->flags.as<int>();
    SB_CUDA(cudaMemsetAsync(bad, 0, 64, ctx->stream));

    // Counting pass: genes in ranges whose counters fit in shared memory (one range for <= ~50k genes).
    const size_t bounds_bytes = (size_t)K1_MAXCOLS * 12;
    const int max_glen = (int)((ctx->smem_optin - bounds_bytes - 1024) / 4);
    const int nranges = std::max(1, (int)ceil_div(ngenes, max_glen));
    const int glen_all = (int)ceil_div(ngenes, nranges);
    allow_smem(k1_count, (size_t)glen_all * 4 + bounds_bytes);

    // the scatter needs, per column, where every sub-range of K1_SUB genes starts
    const int nsub = (int)ceil_div(ngenes, K1_SUB);
    uint32_t* sub_split = nullptr;
    if (L.nchunks > 0 && ngenes > 0) {
        ctx->split.reserve((size_t)(nsub + 1) * L.perm.size() * 4 + 64);
        sub_split = ctx->split.as<uint32_t>();
        for (int r = 0; r < nranges; ++r) {
            const int g0 = r * glen_all, glen = std::min(glen_all, ngenes - g0);
            SB_LAUNCH(ctx, k1_count, L.nchunks, K1_THREADS, (size_t)glen * 4 + bounds_bytes, mat->d_p, mat->d_i, L.d_perm,
                      L.d_chunk_start, ngenes, g0, glen, table, bad, r == 0 ? sub_split : nullptr, nsub);
        }
    }
    if (ngenes > 0) {
        SB_LAUNCH(ctx, k1_scan_table, (int)ceil_div(ngenes, 128), 128, 0, table, L.nchunks, ngenes, C, L.d_first_chunk,
                  ctx->seg_off.as<uint32_t>(), ctx->row_total.as<uint32_t>());
    }
    SB_LAUNCH(ctx, k1_scan_rows, 1, 1024, 0, ctx->row_total.as<uint32_t>(), ngenes, ctx->row_ptr.as<int64_t>());
    if (L.nchunks > 0 && ngenes > 0) {
        const int64_t grid = (int64_t)L.nchunks * ceil_div(nsub, K1W_WARPS);
        if (grid > 2147483647ll) throw Error("too many transpose tiles");
        SB_LAUNCH(ctx, k1_scatter_warp, (unsigned)grid, K1W_WARPS * 32, 0, mat->d_p, mat->d_i, mat->d_x, L.d_perm, L.d_chunk_start,
                  ngenes, nsub, sub_split, table, T.row_ptr, T.vals);
    }
    int h_bad = 0;
    SB_CUDA(cudaMemcpyAsync(&h_bad, bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (h_bad) throw Error("row indices of the sparse matrix must be in range and strictly increasing within a column");
}

}  // namespace sb
