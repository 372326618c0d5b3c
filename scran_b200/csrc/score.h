// The scoreMarkers plan: everything .scoreMarkers works out before it touches the matrix
// (R/scoreMarkers.R:191-217 with helpers :272-345, :463-476, :495-521), restated in C++ on the
// host, plus the device copies the effect / summary kernels read.
#pragma once

#include "kernels.h"

struct scranb200_plan {
    scranb200_ctx* ctx = nullptr;
    int64_t ncells = 0;
    int ngroups = 0;
    double lfc = 0.0;
    bool collapse = true;  // lfc == 0: one orientation per pair (R/scoreMarkers.R:210-211)
    bool blocked = false;

    // Combinations, sorted by (group, block) like .uniquify_DataFrame (:272-290).
    int C = 0;
    std::vector<int32_t> comb_group, comb_block;  // block = index among sorted distinct block codes
    std::vector<int32_t> comb_id;                 // per cell, -1 = dropped (:196-202)
    std::vector<double> ncells_comb;

    // Block-level pairs (:304-339), 0-based combination ids.
    std::vector<int32_t> left, right;

    // Block-averaged comparisons (:463-476): Q sorted (left group, right group) keys and the
    // block-level pairs folded into each; combined weights (:487).
    int Q = 0;
    std::vector<int32_t> avg_left, avg_right, avg_ptr, avg_idx;
    std::vector<double> avg_weight;

    // Output tables (:495-521): K left groups ascending; for each its desired right groups in
    // order with the direct / flipped averaged comparison (-1 = none).
    int K = 0;
    int64_t ndesired = 0;
    std::vector<int32_t> out_group, right_ptr, right_group, q_direct, q_flipped;
    std::vector<int32_t> q_packed;  // q_direct, or -2 - q_flipped, or -1: one load per entry
    std::vector<double> right_weight;

    // self./other. averages (:395-411): avg[group] = sum_c proj[group][c] * y[c]; sparse rows.
    std::vector<int32_t> proj_ptr, proj_comb;
    std::vector<double> proj_coef;
    std::vector<int32_t> other_ptr, other_group;  // per output table: distinct right groups that have cells

    sb::PairLayout layout;
    sb::DevBuf d_layout, d_ints, d_dbls;
    // device views
    int32_t *d_left = nullptr, *d_right = nullptr, *d_avg_ptr = nullptr, *d_avg_idx = nullptr, *d_out_group = nullptr,
            *d_right_ptr = nullptr, *d_q_direct = nullptr, *d_q_flipped = nullptr, *d_proj_ptr = nullptr,
            *d_proj_comb = nullptr, *d_other_ptr = nullptr, *d_other_group = nullptr, *d_need = nullptr,
            *d_q_lgroup = nullptr, *d_q_rgroup = nullptr, *d_group_table = nullptr, *d_q_packed = nullptr;
    double *d_ncells_comb = nullptr, *d_proj_coef = nullptr;
    std::vector<int32_t> q_need;       // bit 0: ranked as is (direct), bit 1: ranked reversed (flipped)
    std::vector<int32_t> group_table;  // group -> output table index or -1
};

namespace sb {

void build_plan(scranb200_plan& P, int64_t ncells, const int32_t* groups, int ngroups, const int32_t* block,
                const int32_t* des_left, const int32_t* des_right, int64_t ndesired, double lfc);
void upload_plan(scranb200_ctx* ctx, scranb200_plan& P);

// K4: block-level effects -> block-averaged effects d_eff [3][Q][ld] (rows [row0, row0+ngenes)),
// and the self/other columns of d_out [K][19][ld].
void run_effects(scranb200_ctx* ctx, const scranb200_plan& P, int ngenes, int64_t ld, int64_t row0, const double* d_mean,
                 const double* d_var, const double* d_ndet, const double* d_u, double* d_eff, double* d_out);
// K5a: mean/min/median/max columns of d_out from d_eff, rows [row0, row0+ngenes).
void run_summaries(scranb200_ctx* ctx, const scranb200_plan& P, int ngenes, int64_t ld, int64_t row0, const double* d_eff,
                   double* d_out);
// K5b: min-rank contributions of comparisons [q0, q1) over all ngenes rows into d_minrank [K][3][ngenes].
void run_minrank(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes, int64_t ld, const double* d_eff, int64_t q0,
                 int64_t q1, int32_t* d_minrank);
void run_finish(scranb200_ctx* ctx, const scranb200_plan& P, int64_t ngenes, int64_t ld, const int32_t* d_minrank,
                double* d_out);
void run_fill_i32(scranb200_ctx* ctx, int32_t* d, int64_t n, int32_t v);
// Generic computeMinRank over a column-major device matrix.
void run_min_rank_matrix(scranb200_ctx* ctx, const double* d_x, int64_t nrow, int64_t ncol, int32_t* d_out);

}  // namespace sb
