// Internal interfaces between the translation units of libscranb200.
#pragma once

#include <algorithm>
#include <cstring>

#include "common.cuh"
#include "layout.h"

namespace sb {

// Output of K1 (device pointers into the context's scratch).
struct Transposed {
    int ngenes = 0, C = 0;
    const uint32_t* seg_off = nullptr;  // [C + 1][ngenes]: start of (gene, device combination) inside the gene's row
    const int64_t* row_ptr = nullptr;   // [ngenes + 1]
    const double* vmin = nullptr;       // [C][ngenes] (device combination order): smallest / largest stored non-zero
    const double* vmax = nullptr;       // value of the segment (+inf / -inf when it has none); filled by run_stats
    double* vals = nullptr;             // [nnz of the used cells], ordered (gene, device combination, cell)
};

void run_transpose(scranb200_ctx* ctx, const scranb200_mat* mat, const PairLayout& L, Transposed& T);

// K2: per (gene, combination) mean / variance / number of values > 0, written in the CALLER's
// combination numbering as [C][ngenes] (= ngenes x C column-major). Any output may be null.
void run_stats(scranb200_ctx* ctx, Transposed& T, const PairLayout& L, double* d_mean, double* d_var, double* d_ndet);

// K3: U statistic (and optionally the tie term) of every pair of the layout for every gene.
// d_u / d_t: [ngenes][npairs] (pairs contiguous within a gene, caller's pair order).
void run_wilcox(scranb200_ctx* ctx, const Transposed& T, const PairLayout& L, double lfc, double* d_u, double* d_t);

// pvalues.cu: pairwiseWilcox's p-value layer on the device; outputs are host arrays [G (G - 1)][ngenes].
void run_pairwise_wilcox(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* group, int G, const int32_t* block,
                         int nblocks, double lfc, int direction, double* effect, double* logp, double* logfdr);

// pvalues.cu: pairwiseTTests (block= branch) with its p-value layer on the device.
void run_pairwise_ttests(scranb200_ctx* ctx, const scranb200_mat* mat, const int32_t* group, int G, const int32_t* block,
                         int nblocks, double lfc, int direction, int std_lfc, double* effect, double* logp, double* logfdr);

}  // namespace sb
