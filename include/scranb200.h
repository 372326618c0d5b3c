/* scranb200.h — C ABI of the B200-native marker-scoring path.
 *
 * This is the boundary a maintainer of the reference binds instead of the reference's own
 * native routines (INTEGRATION.md shows the Rcpp-side stubs). Plain C: pointers, sizes, int
 * status codes. Unless a parameter is marked DEVICE, every pointer is HOST memory owned by the
 * caller; outputs are written into caller-allocated buffers and nothing allocated here is
 * ever handed to the caller except the two opaque handles below.
 *
 * Every function returns 0 on success. On failure it returns non-zero and
 * scranb200_last_error() (thread-local) describes the problem; messages that the reference's
 * tests match are kept verbatim ("indices in 'groups' out of range").
 *
 * There is no CPU fallback: without a CUDA device every compute call fails.
 *
 * Citations are to /root/reference (scran 1.41.1).
 */
#ifndef SCRANB200_H
#define SCRANB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct scranb200_ctx scranb200_ctx; /* one device, its stream and its scratch memory */
typedef struct scranb200_mat scranb200_mat; /* a sparse genes x cells matrix resident in HBM  */

#define SCRANB200_NA_INT INT32_MIN /* R's NA_integer_ (src/utils.h:8-10) */

const char* scranb200_last_error(void);
int scranb200_version(void);

/* Context: binds CUDA device `device` (cudaSetDevice ordinal). Not shareable across fork(). */
int scranb200_ctx_create(int device, scranb200_ctx** out);
void scranb200_ctx_destroy(scranb200_ctx* ctx);
/* Blocks until all work queued by this context has finished. */
int scranb200_ctx_sync(scranb200_ctx* ctx);
/* The cudaStream_t this context launches on, as an integer (for callers that own events). */
uint64_t scranb200_ctx_stream(scranb200_ctx* ctx);
/* Number of kernels launched by this context since creation (bench.py's gpu_launches). */
int64_t scranb200_ctx_launches(scranb200_ctx* ctx);
/* Device-time breakdown of the last score_markers/overlap/blocked_stats call in milliseconds,
 * measured with CUDA events on the context's stream: out[0..7] = transpose, stats, sort, count,
 * effects, summaries, minrank, total. Returns how many entries were written. */
int scranb200_ctx_last_timing(scranb200_ctx* ctx, double* out_ms, int n);

/* ---- matrix upload: replaces beachmat::read_lin_block on the reference side ----------------
 * (src/overlap_exprs.cpp:141, src/compute_blocked_stats.cpp:16). A dgCMatrix maps directly:
 * p = @p (ncol + 1), i = @i (0-based rows, strictly increasing within a column), x = @x.
 * 64-bit column pointers because 33,538 x 1.3M at 7 % has more than 2^31 non-zeros; the _csc32
 * form takes R's own int @p. A device matrix is limited to 2^32 - 1 stored entries. */
int scranb200_upload_csc(scranb200_ctx* ctx, int64_t nrow, int64_t ncol, const int64_t* p, const int32_t* i,
                         const double* x, scranb200_mat** out);
int scranb200_upload_csc32(scranb200_ctx* ctx, int64_t nrow, int64_t ncol, const int32_t* p, const int32_t* i,
                           const double* x, scranb200_mat** out);
/* The same matrix in column chunks: for matrices that do not exist as one object on
 * the caller's side. An R dgCMatrix holds at most 2^31 - 1 entries (its @p is int), so the 1.3M-cell matrix
 * of config 3 (3.1e9 entries) can only arrive as several; HDF5-backed matrices (TENxMatrix) are read block by
 * block anyway. scranb200_mat_begin fixes the shape (nnz_hint: expected stored entries, 0 = unknown; the
 * device arrays grow when it is exceeded); every scranb200_mat_append_columns adds the next ncol_chunk
 * columns as a CSC piece of its own (p: ncol_chunk + 1 entries starting at 0; _columns32 takes R's int @p);
 * the host buffers may be pageable or pinned and are free again when the call returns.
 * scranb200_mat_finish (all ncol_total columns appended) yields the matrix and ends the builder;
 * scranb200_mat_abort drops a builder that will not be finished. */
typedef struct scranb200_builder scranb200_builder;
int scranb200_mat_begin(scranb200_ctx* ctx, int64_t nrow, int64_t ncol_total, int64_t nnz_hint, scranb200_builder** out);
int scranb200_mat_append_columns(scranb200_builder* b, int64_t ncol_chunk, const int64_t* p, const int32_t* i,
                                 const double* x);
int scranb200_mat_append_columns32(scranb200_builder* b, int64_t ncol_chunk, const int32_t* p, const int32_t* i,
                                   const double* x);
int scranb200_mat_finish(scranb200_builder* b, scranb200_mat** out);
void scranb200_mat_abort(scranb200_builder* b);
/* Ordinary column-major double matrix; exact zeros are not stored (the reference's dense branch
 * skips them too: src/compute_blocked_stats.cpp:74, src/overlap_exprs.cpp:43-47). */
int scranb200_upload_dense(scranb200_ctx* ctx, int64_t nrow, int64_t ncol, const double* x, scranb200_mat** out);
/* Borrow CSC arrays that already live in this device's memory (DEVICE pointers; not copied, not
 * freed). Used by the multi-GPU harness and by bench.py's device-resident timing. */
int scranb200_wrap_device_csc(scranb200_ctx* ctx, int64_t nrow, int64_t ncol, int64_t nnz, const int64_t* d_p,
                              const int32_t* d_i, const double* d_x, scranb200_mat** out);
/* Frees the handle. The device arrays of an uploaded matrix are kept by its context (at most two
 * matrices' worth) and reused by the next upload of a similar size, because allocating and freeing
 * tens of gigabytes costs a tenth of the transfer; they are released by scranb200_ctx_destroy, or
 * earlier when a device allocation of the library would otherwise fail. Free matrices before their
 * context. */
void scranb200_mat_free(scranb200_mat* mat);
int64_t scranb200_mat_nnz(const scranb200_mat* mat);

/* ---- compute_blocked_stats_none: src/compute_blocked_stats.cpp:205-208 (template :15-123) ---
 * Per gene and block: mean and unbiased variance over ALL cells of the block, zeros included,
 * with the reference's operation order (running update over stored values in cell order, then
 * the closed-form zero fill), so results are bit-identical to the reference's sparse branch.
 * block: one id per cell, 0-based, SCRANB200_NA_INT (or any negative) = skip the cell (:39-42).
 * mean_out, var_out: nrow x nblocks column-major, like the reference's NumericMatrix pair.
 * ndetected_out (nullable, same shape): number of values > 0, i.e. what
 * scuttle::summarizeAssayByGroup(statistics="num.detected") gives at R/scoreMarkers.R:361-363.
 * Empty block: mean and variance NaN (:86-88); one cell: variance NaN (:91-94). */
int scranb200_blocked_stats(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* block, int nblocks,
                            double* mean_out, double* var_out, double* ndetected_out);

/* ---- compute_blocked_stats_lognorm / _norm: src/compute_blocked_stats.cpp:129-185 ------------
 * (.Call symbols _scran_compute_blocked_stats_lognorm, src/RcppExports.cpp:52-64, and
 * _scran_compute_blocked_stats_norm, :66-77.) The same statistics after a per-cell transformation
 * of the values: transform = 1: log(x / sf[cell] + pseudo) / M_LN2 (:136); transform = 2:
 * x / sf[cell] (:168; pseudo ignored). sf: one size factor per cell. norm is bit-identical to the
 * reference (IEEE division); lognorm agrees to the accuracy of the device's log (<= 1 ulp per
 * value; tests: 1e-12 relative). With pseudo != 1 the reference leaves its sparse branch (:142);
 * here the zeros' common value log(pseudo) / M_LN2 is subtracted before and added to the mean
 * after, which is the same mean and variance up to rounding (tests: 1e-12 relative, and 1e-12
 * absolute on means that cancel to ~0). */
int scranb200_blocked_stats_transformed(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* block, int nblocks,
                                        int transform, const double* sf, double pseudo, double* mean_out,
                                        double* var_out);

/* ---- pairwiseWilcox, statistics AND p-value layer: R/pairwiseWilcox.R:192-350 ----------------
 *  The R function calls overlap_exprs once per block level
 * (R/pairwiseWilcox.R:221-228) and then, in R closures per (block, host, target): AUC, the normal
 * approximation with continuity and tie correction (:263-350), Stouffer's combination over the
 * block levels with weights n_host n_target and the weighted mean of the AUCs
 * (R/utils_markers.R:52-130), and .logBH over the genes (:184-190). This entry point does all of
 * it on the device, log.p = TRUE.
 * group: one code per cell, 0 .. ngroups-1 (negative = NA: the cell is ignored); block: one code
 * per cell, 0 .. nblocks-1, or NULL (one block); direction: 0 "any", 1 "up", 2 "down".
 * Outputs: host arrays [ngroups (ngroups - 1)][nrow]: ordered pair (first, second), first in level
 * order, second in level order skipping first, the order of the reference's `pairs` DataFrame;
 * effect = AUC (overlap proportion), logp, logfdr natural logs. Pairs without a valid block: NaN. */
int scranb200_pairwise_wilcox(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* group, int ngroups,
                              const int32_t* block, int nblocks, double lfc, int direction, double* effect_out,
                              double* logp_out, double* logfdr_out);

/* ---- pairwiseTTests (block= branch), statistics AND p-value layer: R/pairwiseTTests.R:226-332, 439-461
 *  The R function takes the per-(block, group) means and variances
 * from compute_blocked_stats_none (ids group + block * ngroups, :242) and then, in R closures per
 * (block, host, target): Welch's t-test (.get_t_test_stats, :311-332), stats::pt(log.p = TRUE) for both
 * tails (.run_t_test, :439-461; against -|lfc| / +|lfc| when lfc != 0), the log-fold change or with
 * std_lfc Cohen's d (:283-289), Stouffer over the block levels with weights 1 / (1/nh + 1/nt)
 * (R/utils_markers.R:52-130), and .logBH over the genes (:184-190). All of it runs on the device here.
 * Arguments and outputs as for scranb200_pairwise_wilcox; effect = logFC (host - target). */
int scranb200_pairwise_t_tests(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* group, int ngroups,
                               const int32_t* block, int nblocks, double lfc, int direction, int std_lfc,
                               double* effect_out, double* logp_out, double* logfdr_out);

/* ---- combineMarkers: R/combineMarkers.R:182-330 with src/choose_effect_size.cpp:35-97 ----------
 *  What .combine_markers_internal does for the contrasts of ONE first group,
 * on log p-values (log.p.in = log.p.out = TRUE), including the two native helpers it calls
 * (.Call symbols _scran_compute_Top_statistic_from_ranks, src/RcppExports.cpp, and
 * _scran_choose_middle_effect_size). logp, effect: host [ncon][ngenes], NaN = NA; effect nullable.
 * pval_type: 0 "any" (Simes; Top = the ceiling(n min_prop)-th smallest rank over the contrasts, ranks by
 * rank(ties.method = "first") across the genes), 1 "some" (the ceiling(n min_prop)-th Holm-adjusted
 * p-value), 2 "all" (the largest). The reference's defaults: min_prop = 0 for "any", 0.5 otherwise.
 * Outputs, host arrays of ngenes in the ORIGINAL gene order, any of them nullable: top_out (int32,
 * SCRANB200_NA_INT = NA; written for "any" only), logp_out, logfdr_out (.logBH), summary_out
 * (.choose_effect_size), and order_out: the 0-based row order the reference applies last
 * (order(Top) for "any", order(p) otherwise; ties by position, NA last). */
int scranb200_combine_markers(scranb200_ctx* ctx, int64_t ngenes, int ncon, const double* logp, const double* effect,
                              int pval_type, double min_prop, int32_t* top_out, double* logp_out, double* logfdr_out,
                              double* summary_out, int32_t* order_out);

/* ---- overlap_exprs_paired: src/overlap_exprs.cpp:196-220 ------------------------------------
 * groups: the reference's list of 0-based cell-index vectors, flattened: group g owns
 * group_cells[group_ptr[g] .. group_ptr[g+1]). As in the reference, a cell may be listed by several
 * groups, or several times by one (it then counts that many times).
 * left/right: 1-based group indices, exactly as the reference takes them (:215).
 * u_out: npairs x nrow column-major (pairs contiguous within a gene), U = #(a - lfc > b) +
 * 0.5 #(a - lfc == b) over a in left group, b in right group; exact half-integers.
 * ties_out (nullable, same shape): the tie term sum t(t^2 - 1) the reference computes alongside
 * (:121-124) but only returns from overlap_exprs. */
int scranb200_overlap_exprs_paired(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* left,
                                   const int32_t* right, int64_t npairs, int ngroups, const int64_t* group_ptr,
                                   const int32_t* group_cells, double lfc, double* u_out, double* ties_out);

/* ---- overlap_exprs: src/overlap_exprs.cpp:140-193 -------------------------------------------
 * All pairs i2 < i1 of the given groups; the reference's ragged list-of-matrices is returned
 * flattened as [ngroups][ngroups][nrow]: entry (i1, i2) holds column i2 of matrix i1. With
 * lfc != 0 entry (i2, i1) holds the "-lfc" contrast (:183-187). Entries the reference does not
 * allocate (i2 >= i1 when lfc == 0) are NaN; pairs with an empty group stay 0 (:175,177). */
int scranb200_overlap_exprs(scranb200_ctx* ctx, const scranb200_mat* mat, int ngroups, const int64_t* group_ptr,
                            const int32_t* group_cells, double lfc, double* u_out, double* ties_out);

/* ---- fused scoreMarkers: replaces .compute_all_effect_sizes + computeMinRank ----------------
 * (R/scoreMarkers.R:351-414 and :251-256, R/computeMinRank.R:45-50), so that R never
 * materialises the genes x block-pairs matrices.
 *
 * groups: per cell, the index of the cell's group in sort(unique(groups)) (R/scoreMarkers.R:192);
 *         negative = NA (cell ignored). block: nullable; per cell block code, any non-negative ints.
 * des_left/des_right: the desired comparisons as group indices, in the order
 *         .expand_pairings gives them (R/expandPairings.R:1-58); nullable = all ordered pairs
 *         of the ngroups groups in expand.grid order (first index fastest).
 * A "score plan" holds the bookkeeping of R/scoreMarkers.R:196-217, 293-345, 463-521.
 * ctx may be null: the plan is then host-only (plan_info / plan_tables work, compute calls refuse it). */
typedef struct scranb200_plan scranb200_plan;

int scranb200_plan_create(scranb200_ctx* ctx, int64_t ncells, const int32_t* groups, int ngroups,
                          const int32_t* block, const int32_t* des_left, const int32_t* des_right,
                          int64_t ndesired, double lfc, scranb200_plan** out);
void scranb200_plan_destroy(scranb200_plan* plan);
/* info[0] = number of output groups K (distinct left groups, ascending), [1] = ndesired,
 * [2] = number of block-averaged comparisons Q, [3] = block-level pairs, [4] = combinations. */
int scranb200_plan_info(const scranb200_plan* plan, int64_t* info, int n);
/* out_groups[K]: group index of each output table; right_ptr[K+1], right_group[ndesired]: the
 * compared-against groups of each table in column order; weight[ndesired]: combined weights
 * (mcols(full)$weight, R/scoreMarkers.R:487, 571-574). Any pointer may be null. */
int scranb200_plan_tables(const scranb200_plan* plan, int32_t* out_groups, int64_t* right_ptr,
                          int32_t* right_group, double* weight);

/* The plan's internals, for checking the bookkeeping against R/scoreMarkers.R:293-345, 463-521:
 * combinations (group, block index, cells), block-level pairs (0-based combination ids), the
 * averaged comparisons with their fold lists, per desired comparison the direct / flipped
 * averaged comparison (-1 = none), and each cell's combination (-1 = dropped). Sizes from
 * scranb200_plan_info; any pointer may be null. */
int scranb200_plan_pairs(const scranb200_plan* plan, int32_t* comb_group, int32_t* comb_block, double* ncells_comb,
                         int32_t* left, int32_t* right, int32_t* avg_left, int32_t* avg_right, int32_t* avg_ptr,
                         int32_t* avg_idx, int32_t* q_direct, int32_t* q_flipped, int32_t* comb_id);

#define SCRANB200_NCOLS 19 /* self.average, other.average, self.detected, other.detected, then
                              {mean,min,median,max,rank} x {logFC.cohen, AUC, logFC.detected}:
                              the reference's column order (R/scoreMarkers.R:370,392,401-411,562-576) */

/* out: [K][19][nrow] doubles (rank columns hold integers). full_out (nullable):
 * [3][ndesired][nrow], effect-major in the order cohen, AUC, detected, then comparisons grouped
 * by output table in column order: the full.* matrices of full.stats=TRUE. */
int scranb200_score_markers(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan,
                            double* out, double* full_out);

/* ---- summaryMarkerStats: R/summaryMarkerStats.R:62-95 ---------------------
 * Per group the mean expression and the proportion of cells with a value > 0 (self.average, self.detected)
 * and their mean over the other groups (other.*): the first four columns of the scoreMarkers tables, from
 * the blocked statistics alone (no sort, no pair counting). plan: scranb200_plan_create with the groups
 * (block = NULL for the reference's function; with a block the columns are scoreMarkers' block-corrected
 * ones). out: host [K][4][nrow] in the order self.average, other.average, self.detected, other.detected. */
int scranb200_summary_marker_stats(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan, double* out);

/* ---- staged form for gene-sharded multi-GPU runs (all buffers are DEVICE pointers) ----------
 * stage 1 on each gene shard: effects d_eff [3][Q][ld] (gene-contiguous columns, this shard's
 * rows at [row0, row0 + nrow_shard)), and the non-rank columns of d_out [K][19][ld].
 * stage 2 after the gather, on columns [q0, q1): min-ranks over all ngenes rows, combined into
 * d_minrank [K][3][ngenes] (int32, pre-filled with INT32_MAX by scranb200_fill_i32) with atomicMin.
 * stage 3: write d_minrank into the rank columns of d_out. */
int scranb200_score_effects_dev(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan,
                                int64_t ld, int64_t row0, double* d_eff, double* d_out);
int scranb200_score_minrank_dev(scranb200_ctx* ctx, const scranb200_plan* plan, int64_t ngenes, int64_t ld,
                                const double* d_eff, int64_t q0, int64_t q1, int32_t* d_minrank);
int scranb200_score_finish_dev(scranb200_ctx* ctx, const scranb200_plan* plan, int64_t ngenes, int64_t ld,
                               const int32_t* d_minrank, double* d_out);
int scranb200_fill_i32(scranb200_ctx* ctx, int32_t* d_ptr, int64_t n, int32_t value);

/* ---- gene-sharded scoreMarkers across the GPUs of one box -------------------------------------
 * Replaces what beachmat::rowBlockApply + BiocParallel do for the reference (R/scoreMarkers.R:
 * 234-248: one worker per row block, results rbind-ed) together with the min-rank over ALL genes
 * after the join (:251-256). A rank is one context (one GPU) holding the rows
 * [row_starts[r], row_starts[r+1]) of the matrix for all cells; ranks are either processes (one
 * context each) or the contexts of one process (one host thread per context while scoring).
 * The exchange (effect columns from gene shards to comparison shards, MIN-combination of the
 * per-table ranks, gather of the result rows) runs inside the library over NCCL, which is loaded
 * at run time (libnccl.so.2) on the first of these calls.
 *
 * scranb200_comm_unique_id: rank 0 makes the 128-byte id and hands it to the others by any means.
 * scranb200_ctx_comm_init:  collective over the world ranks. world == 1 needs no NCCL.
 * scranb200_ctx_comm_init_all: the same for n contexts of THIS process (ncclCommInitAll).
 * scranb200_balance_rows: contiguous row ranges of equal total weight (weight = stored entries per
 *   gene; housekeeping genes are ten times heavier than the average): row_starts[world + 1].
 * scranb200_score_markers_sharded: collective. mat = this rank's row block (row_starts[rank+1] -
 *   row_starts[rank] rows), plan = the same plan on every rank. root >= 0: only that rank
 *   receives the result; root < 0: every rank. out (nullable, host): [K][19][ngenes_total] as
 *   scranb200_score_markers writes it; d_out (nullable): receives the DEVICE address of the same
 *   table inside the context's scratch (valid until the context's next call; null on ranks that
 *   do not receive the result).
 * scranb200_exchange_plan: host arithmetic only, no GPU needed - the block positions of the one exchange step
 *   as scranb200_score_markers_sharded uses them (comparison shards q_starts[world + 1]; for `rank`, per peer s:
 *   offset / length in doubles of the block sent to s inside one effect's [Q][nrows_rank] plane, offset /
 *   length of the block received from s in the staging buffer, effect e at offset + e * length). Exposed so
 *   that the plan can be exercised by a multi-process test on CPUs (tests/test_distributed_cpu.py, gloo). */
#define SCRANB200_COMM_ID_BYTES 128
int scranb200_comm_unique_id(char* id_out);
int scranb200_ctx_comm_init(scranb200_ctx* ctx, int world, int rank, const char* id_in);
int scranb200_ctx_comm_init_all(scranb200_ctx** ctxs, int n);
int scranb200_balance_rows(const double* weight, int64_t nrow, int world, int64_t* row_starts);
int scranb200_exchange_plan(int64_t Q, int world, int rank, const int64_t* row_starts, int64_t* q_starts, int64_t* send_off,
                            int64_t* send_len, int64_t* recv_off, int64_t* recv_len);
int scranb200_score_markers_sharded(scranb200_ctx* ctx, const scranb200_mat* mat, const scranb200_plan* plan,
                                    int64_t ngenes_total, const int64_t* row_starts, int root, double* out,
                                    double** d_out);

/* ---- computeMinRank: R/computeMinRank.R:45-50 (ties.method="min", decreasing=TRUE) ----------
 * x: nrow x ncol column-major; out[nrow]: min over columns of (1 + number of non-NaN entries of
 * the column strictly greater), NaN entries ignored; INT32_MAX if a row has none. */
int scranb200_compute_min_rank(scranb200_ctx* ctx, const double* x, int64_t nrow, int64_t ncol, int32_t* out);

/* ---- computeMinRank without device allocation ------------------------------------------------
 * No counterpart in the reference: this exists for callers of this library that must not, or
 * cannot, allocate device memory while it runs. Typical cases: the host application has given
 * the whole GPU to its own memory pool, so a cudaMalloc from inside a library fails or competes
 * with the pool; or the call sits on a latency-sensitive path where the device-wide stall of
 * cudaMalloc / cudaFree is not acceptable. Same result as scranb200_compute_min_rank, but:
 *   - d_x (nrow x ncol column-major, NaN = NA) and d_out (int32[nrow]) are DEVICE pointers that
 *     the caller owns; nothing is copied to or from the host.
 *   - all scratch lives in a fixed 16 MiB array inside the library image. Every GPU that loads
 *     libscranb200.so holds those 16 MiB until the process exits. Limits: ncol <= 4096, and
 *     nrow * 24 bytes must fit the array (about 696,000 rows) when a column is too long for the
 *     shared-memory sort; beyond that the call fails with a message and computes nothing.
 *   - the call issues no cudaMalloc / cudaFree and uses no allocating library routine. The first
 *     call on a device looks up the array's device address (cudaGetSymbolAddress); make that
 *     first call outside any stream capture.
 *   - all calls on a device share the one array and it keeps the previous call's contents; the
 *     call re-initialises what it reads. Two calls running on the same GPU at the same time
 *     (from two contexts or threads) race on it: serialise them.
 * The call returns after its work on the context's stream has finished; inputs written on
 * another stream must be complete before it is made. */
int scranb200_compute_min_rank_static(scranb200_ctx* ctx, const double* d_x, int64_t nrow, int64_t ncol,
                                      int32_t* d_out);

/* ---- synthetic input (bench.py), generated on the device -------------
 * Writes a CSC matrix of genes [g0, g1) x ncells log-counts into device memory owned by the
 * returned matrix. density < 0: natural Poisson density (config 1). */
int scranb200_synth_csc(scranb200_ctx* ctx, int64_t ngenes_total, int64_t g0, int64_t g1, int64_t ncells,
                        const int32_t* groups, int ngroups, double density, uint64_t seed, scranb200_mat** out);
/* Expected stored entries per gene of that matrix (weights for scranb200_balance_rows). */
int scranb200_synth_row_weights(int64_t ngenes_total, int64_t ncells, double density, uint64_t seed, double* weight);
/* Copy a device matrix's CSC arrays to host buffers (sizes from nrow/ncol/scranb200_mat_nnz). */
int scranb200_download_csc(scranb200_ctx* ctx, const scranb200_mat* mat, int64_t* p, int32_t* i, double* x);

#ifdef __cplusplus
}
#endif
#endif
