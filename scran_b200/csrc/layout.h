// Host-side bookkeeping for one pairwise call: how cells map to combinations, how combinations
// are ordered on the device, which of them are compared with each other ("units"), and how the
// cell axis is cut into chunks for the transpose. All of it is O(ncells + npairs) integer work.
#pragma once

#include <cstdint>
#include <vector>

#include "common.cuh"

namespace sb {

constexpr int MAX_CHUNK_COLS = 2048;  // the transpose stages a chunk's column bounds in shared memory

struct PairLayout {
    int64_t ncells = 0;
    int C = 0;  // combinations as the caller numbers them

    // Device order of combinations: the members of unit 0 (ascending caller id), of unit 1, ...,
    // then the combinations that are in no pair. A unit is a connected component of the
    // "is compared with" graph; for scoreMarkers it is one block level.
    std::vector<int32_t> dev_of_caller, caller_of_dev;
    int U = 0;
    std::vector<int32_t> unit_first, unit_ncomb;  // in device order
    int maxG = 0;

    // Cells that belong to a combination, stably sorted by device combination, cut into chunks
    // of at most chunk_cols columns that never straddle two combinations.
    std::vector<int32_t> perm;
    std::vector<int32_t> chunk_start;  // nchunks + 1 offsets into perm
    std::vector<int32_t> first_chunk;  // C + 1: first chunk of each device combination
    int nchunks = 0;

    std::vector<int64_t> ncells_dev;  // cells per device combination

    // Pairs grouped by unit; l/r are positions inside the unit, out = index in the caller's list.
    std::vector<int32_t> pair_ptr, pair_l, pair_r, pair_out;
    int64_t npairs = 0;

    // Device mirror (one allocation): offsets in int32 elements.
    int32_t* d_base = nullptr;
    int32_t *d_perm = nullptr, *d_chunk_start = nullptr, *d_first_chunk = nullptr, *d_caller_of_dev = nullptr,
            *d_unit_first = nullptr, *d_unit_ncomb = nullptr, *d_pair_ptr = nullptr, *d_pair_l = nullptr,
            *d_pair_r = nullptr, *d_pair_out = nullptr;
    int64_t* d_ncells_dev = nullptr;
};

// comb_id: per cell, caller combination id or negative for "in no combination".
// left/right: 0-based caller combination ids.
void build_layout(const int32_t* comb_id, int64_t ncells, int C, const int32_t* left, const int32_t* right,
                  int64_t npairs, int target_chunks, PairLayout& L);

// The same from the reference's list form (src/overlap_exprs.cpp:13-32): combination c holds the cells
// gcells[gptr[c] .. gptr[c + 1]). A cell may be listed by several combinations, or more than once.
void build_layout_members(const int64_t* gptr, const int32_t* gcells, int64_t ncells, int C, const int32_t* left,
                          const int32_t* right, int64_t npairs, int target_chunks, PairLayout& L);

// Copies the arrays above into buf (grown as needed) on the context's stream.
void upload_layout(scranb200_ctx* ctx, PairLayout& L, DevBuf& buf);

}  // namespace sb
