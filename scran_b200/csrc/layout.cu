#include "layout.h"

#include <algorithm>
#include <cstring>
#include <numeric>

namespace sb {

namespace {
int uf_find(std::vector<int32_t>& parent, int a) {
    while (parent[a] != a) {
        parent[a] = parent[parent[a]];
        a = parent[a];
    }
    return a;
}
}  // namespace

void build_layout(const int32_t* comb_id, int64_t ncells, int C, const int32_t* left, const int32_t* right,
                  int64_t npairs, int target_chunks, PairLayout& L) {
    // cell -> combination becomes the list form (cells of every combination in ascending order)
    std::vector<int64_t> gptr(C + 1, 0);
    for (int64_t k = 0; k < ncells; ++k) {
        const int c = comb_id[k];
        if (c < 0) continue;
        if (c >= C) throw Error("combination id out of range");
        gptr[c + 1]++;
    }
    for (int c = 0; c < C; ++c) gptr[c + 1] += gptr[c];
    std::vector<int32_t> gcells(std::max<int64_t>(1, gptr[C]));
    {
        std::vector<int64_t> fill(gptr.begin(), gptr.end() - 1);
        for (int64_t k = 0; k < ncells; ++k) {
            const int c = comb_id[k];
            if (c >= 0) gcells[fill[c]++] = (int32_t)k;
        }
    }
    build_layout_members(gptr.data(), gcells.data(), ncells, C, left, right, npairs, target_chunks, L);
}

void build_layout_members(const int64_t* gptr, const int32_t* gcells, int64_t ncells, int C, const int32_t* left,
                          const int32_t* right, int64_t npairs, int target_chunks, PairLayout& L) {
    L = PairLayout();
    L.ncells = ncells;
    L.C = C;
    L.npairs = npairs;

    // Units = connected components over the pairs.
    std::vector<int32_t> parent(C);
    std::iota(parent.begin(), parent.end(), 0);
    std::vector<char> paired(C, 0);
    for (int64_t k = 0; k < npairs; ++k) {
        const int a = left[k], b = right[k];
        if (a < 0 || a >= C || b < 0 || b >= C) throw Error("pair refers to a combination that does not exist");
        paired[a] = paired[b] = 1;
        const int ra = uf_find(parent, a), rb = uf_find(parent, b);
        if (ra != rb) parent[std::max(ra, rb)] = std::min(ra, rb);
    }
    // Number components by their smallest member so that the order is deterministic.
    std::vector<int32_t> unit_of(C, -1);
    L.U = 0;
    for (int c = 0; c < C; ++c) {
        if (!paired[c]) continue;
        const int r = uf_find(parent, c);
        if (unit_of[r] < 0) unit_of[r] = L.U++;
        unit_of[c] = unit_of[r];
    }
    L.unit_ncomb.assign(L.U, 0);
    for (int c = 0; c < C; ++c) {
        if (unit_of[c] >= 0) L.unit_ncomb[unit_of[c]]++;
    }
    L.unit_first.assign(L.U, 0);
    int acc = 0;
    for (int u = 0; u < L.U; ++u) {
        L.unit_first[u] = acc;
        acc += L.unit_ncomb[u];
        L.maxG = std::max(L.maxG, (int)L.unit_ncomb[u]);
    }
    L.dev_of_caller.assign(C, -1);
    L.caller_of_dev.assign(C, -1);
    {
        std::vector<int32_t> fill(L.unit_first);
        int tail = acc;
        for (int c = 0; c < C; ++c) {
            const int d = unit_of[c] >= 0 ? fill[unit_of[c]]++ : tail++;
            L.dev_of_caller[c] = d;
            L.caller_of_dev[d] = c;
        }
    }

    // Cells per device combination and the column list: the members of device combination 0 in the order
    // given, then those of 1, ... A cell that belongs to several combinations is simply listed several times.
    L.ncells_dev.assign(C, 0);
    for (int c = 0; c < C; ++c) L.ncells_dev[L.dev_of_caller[c]] = gptr[c + 1] - gptr[c];
    std::vector<int64_t> start(C + 1, 0);
    for (int d = 0; d < C; ++d) start[d + 1] = start[d] + L.ncells_dev[d];
    if (start[C] > 2147483647ll) throw Error("too many group memberships");
    L.perm.resize(start[C]);
    for (int d = 0; d < C; ++d) {
        const int c = L.caller_of_dev[d];
        for (int64_t k = gptr[c]; k < gptr[c + 1]; ++k) {
            if (gcells[k] < 0 || gcells[k] >= ncells) throw Error("indices in 'groups' out of range");
            L.perm[start[d] + (k - gptr[c])] = gcells[k];
        }
    }

    // Chunks: about target_chunks of them, none straddling a combination.
    const int64_t used = start[C];
    int64_t chunk_cols = std::max<int64_t>(1, ceil_div(used, std::max(1, target_chunks)));
    chunk_cols = std::min<int64_t>(chunk_cols, MAX_CHUNK_COLS);
    L.first_chunk.assign(C + 1, 0);
    L.chunk_start.clear();
    for (int d = 0; d < C; ++d) {
        L.first_chunk[d] = (int32_t)L.chunk_start.size();
        for (int64_t s = start[d]; s < start[d + 1]; s += chunk_cols) L.chunk_start.push_back((int32_t)s);
    }
    L.first_chunk[C] = (int32_t)L.chunk_start.size();
    L.nchunks = (int)L.chunk_start.size();
    L.chunk_start.push_back((int32_t)used);

    // Pairs by unit.
    L.pair_ptr.assign(L.U + 1, 0);
    for (int64_t k = 0; k < npairs; ++k) L.pair_ptr[unit_of[left[k]] + 1]++;
    for (int u = 0; u < L.U; ++u) L.pair_ptr[u + 1] += L.pair_ptr[u];
    L.pair_l.resize(npairs);
    L.pair_r.resize(npairs);
    L.pair_out.resize(npairs);
    {
        std::vector<int32_t> fill(L.pair_ptr.begin(), L.pair_ptr.end() - 1);
        for (int64_t k = 0; k < npairs; ++k) {
            const int u = unit_of[left[k]];
            const int at = fill[u]++;
            L.pair_l[at] = L.dev_of_caller[left[k]] - L.unit_first[u];
            L.pair_r[at] = L.dev_of_caller[right[k]] - L.unit_first[u];
            L.pair_out[at] = (int32_t)k;
        }
    }
}

void upload_layout(scranb200_ctx* ctx, PairLayout& L, DevBuf& buf) {
    auto pad2 = [](size_t n) { return (n + 1) & ~size_t(1); };  // keep 8-byte alignment for the int64 tail
    const size_t n_perm = pad2(L.perm.size()), n_cs = pad2(L.chunk_start.size()), n_fc = pad2(L.first_chunk.size()),
                 n_cod = pad2(L.caller_of_dev.size()), n_uf = pad2(L.unit_first.size()),
                 n_un = pad2(L.unit_ncomb.size()), n_pp = pad2(L.pair_ptr.size()), n_pl = pad2(L.pair_l.size());
    const size_t n_i32 = n_perm + n_cs + n_fc + n_cod + n_uf + n_un + n_pp + 3 * n_pl;
    const size_t bytes = n_i32 * 4 + L.ncells_dev.size() * 8 + 64;
    buf.reserve(bytes);
    std::vector<int32_t> host(n_i32 + 2 * L.ncells_dev.size() + 16, 0);
    size_t at = 0;
    auto put = [&](const std::vector<int32_t>& v, size_t padded, int32_t*& dptr) {
        if (!v.empty()) memcpy(host.data() + at, v.data(), v.size() * 4);
        dptr = buf.as<int32_t>() + at;
        at += padded;
    };
    put(L.perm, n_perm, L.d_perm);
    put(L.chunk_start, n_cs, L.d_chunk_start);
    put(L.first_chunk, n_fc, L.d_first_chunk);
    put(L.caller_of_dev, n_cod, L.d_caller_of_dev);
    put(L.unit_first, n_uf, L.d_unit_first);
    put(L.unit_ncomb, n_un, L.d_unit_ncomb);
    put(L.pair_ptr, n_pp, L.d_pair_ptr);
    put(L.pair_l, n_pl, L.d_pair_l);
    put(L.pair_r, n_pl, L.d_pair_r);
    put(L.pair_out, n_pl, L.d_pair_out);
    if (!L.ncells_dev.empty()) memcpy(host.data() + at, L.ncells_dev.data(), L.ncells_dev.size() * 8);
    L.d_ncells_dev = reinterpret_cast<int64_t*>(buf.as<int32_t>() + at);
    at += 2 * L.ncells_dev.size();
    L.d_base = buf.as<int32_t>();
    SB_CUDA(cudaMemcpyAsync(buf.p, host.data(), at * 4, cudaMemcpyHostToDevice, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));  // 'host' goes out of scope
}

}  // namespace sb
