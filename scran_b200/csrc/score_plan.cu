// Host side of scoreMarkers: restates the bookkeeping of R/scoreMarkers.R:191-217 (and helpers
// :272-345, :463-476, :495-521) in C++. Integer work, O(ncells + comparisons).
#include <cmath>
#include <map>
#include <set>

#include "score.h"

namespace sb {

namespace {

// Weighted least-squares projector for scuttle::correctGroupSummary's blocked case as called at
// R/scoreMarkers.R:395-399: model ~0 + group + block on the per-combination summaries with
// weights = cells per combination; returns, for each group that has cells, the row that maps
// the C summaries to the group coefficient. Aliased columns are dropped in column order (as
// R's lm.wfit pivots them out). PARITY UNPINNED: scuttle's source is not available here.
void wls_projector(const scranb200_plan& P, const std::vector<int32_t>& groups_present, int nblocks,
                   std::vector<std::vector<double>>& rows) {
    const int C = P.C;
    const int ng = (int)groups_present.size();
    const int p = ng + nblocks - 1;
    std::map<int, int> gpos;
    for (int k = 0; k < ng; ++k) gpos[groups_present[k]] = k;
    typedef long double LD;
    std::vector<std::vector<LD>> A(p, std::vector<LD>(C, 0.0L));  // columns of sqrt(w) * design
    std::vector<LD> sw(C);
    for (int c = 0; c < C; ++c) {
        sw[c] = std::sqrt((LD)P.ncells_comb[c]);
        A[gpos[P.comb_group[c]]][c] = sw[c];
        if (P.comb_block[c] > 0) A[ng + P.comb_block[c] - 1][c] = sw[c];
    }
    // Modified Gram-Schmidt with dropping of dependent columns.
    std::vector<int> kept;
    std::vector<std::vector<LD>> Qm;            // orthonormal columns
    std::vector<std::vector<LD>> Rm;            // Rm[j] = coefficients of kept column j on Q columns 0..j
    for (int j = 0; j < p; ++j) {
        std::vector<LD> v = A[j];
        LD norm0 = 0;
        for (int c = 0; c < C; ++c) norm0 += v[c] * v[c];
        std::vector<LD> r;
        for (size_t q = 0; q < Qm.size(); ++q) {
            LD dot = 0;
            for (int c = 0; c < C; ++c) dot += Qm[q][c] * v[c];
            for (int c = 0; c < C; ++c) v[c] -= dot * Qm[q][c];
            r.push_back(dot);
        }
        LD norm1 = 0;
        for (int c = 0; c < C; ++c) norm1 += v[c] * v[c];
        if (norm0 <= 0 || std::sqrt(norm1) < 1e-7L * std::sqrt(norm0)) continue;  // aliased
        const LD nr = std::sqrt(norm1);
        for (int c = 0; c < C; ++c) v[c] /= nr;
        r.push_back(nr);
        Qm.push_back(v);
        Rm.push_back(r);
        kept.push_back(j);
    }
    // beta = R^{-1} Q' (sw * y)  =>  row for kept column j: e_j' R^{-1} Q' diag(sw).
    const int pk = (int)kept.size();
    rows.assign(ng, std::vector<double>());
    for (int jj = 0; jj < pk; ++jj) {
        if (kept[jj] >= ng) continue;
        // Solve x' R = e_jj'  (R upper triangular, column k of R = Rm[k]).
        std::vector<LD> xr(pk, 0.0L);
        for (int k = jj; k < pk; ++k) {
            LD s = (k == jj) ? 1.0L : 0.0L;
            for (int i = jj; i < k; ++i) s -= xr[i] * Rm[k][i];
            xr[k] = s / Rm[k][k];
        }
        // xr holds row jj of R^{-1}: beta_jj = sum_k xr[k] * (Q_k . sw*y).
        std::vector<double> row(C);
        for (int c = 0; c < C; ++c) {
            LD s = 0;
            for (int k = jj; k < pk; ++k) s += xr[k] * Qm[k][c];
            row[c] = (double)(s * sw[c]);
        }
        rows[kept[jj]] = row;
    }
}

}  // namespace

void build_plan(scranb200_plan& P, int64_t ncells, const int32_t* groups, int ngroups, const int32_t* block,
                const int32_t* des_left, const int32_t* des_right, int64_t ndesired, double lfc) {
    P.ncells = ncells;
    P.ngroups = ngroups;
    P.lfc = lfc;
    P.collapse = (lfc == 0.0);
    if (ngroups < 1) throw Error("need at least one group");

    // Desired comparisons; default = every ordered pair, first index fastest (expand.grid).
    std::vector<int32_t> dl, dr;
    if (des_left && des_right) {
        for (int64_t k = 0; k < ndesired; ++k) {
            if (des_left[k] < 0 || des_left[k] >= ngroups || des_right[k] < 0 || des_right[k] >= ngroups) {
                throw Error("desired comparison refers to a group that does not exist");
            }
            dl.push_back(des_left[k]);
            dr.push_back(des_right[k]);
        }
    } else {
        for (int b = 0; b < ngroups; ++b) {
            for (int a = 0; a < ngroups; ++a) {
                if (a != b) {
                    dl.push_back(a);
                    dr.push_back(b);
                }
            }
        }
    }
    P.ndesired = (int64_t)dl.size();
    std::vector<char> used(ngroups, 0);
    for (size_t k = 0; k < dl.size(); ++k) used[dl[k]] = used[dr[k]] = 1;

    // Block codes -> dense indices in sorted order.
    std::map<int32_t, int32_t> bindex;
    if (block) {
        std::set<int32_t> seen;
        for (int64_t c = 0; c < ncells; ++c) {
            const int g = groups[c];
            if (g >= 0 && g < ngroups && used[g] && block[c] >= 0) seen.insert(block[c]);
        }
        int32_t k = 0;
        for (int32_t b : seen) bindex[b] = k++;
    }
    const int nblocks = block ? (int)bindex.size() : 1;
    P.blocked = block != nullptr && nblocks > 1;

    // Combinations sorted by (group, block).
    std::vector<int64_t> count((size_t)ngroups * std::max(1, nblocks), 0);
    for (int64_t c = 0; c < ncells; ++c) {
        const int g = groups[c];
        if (g >= ngroups) throw Error("group index out of range");
        if (g < 0 || !used[g]) continue;
        int b = 0;
        if (block) {
            if (block[c] < 0) continue;
            b = bindex[block[c]];
        }
        count[(size_t)g * nblocks + b]++;
    }
    std::vector<int32_t> comb_of((size_t)ngroups * std::max(1, nblocks), -1);
    P.comb_group.clear();
    P.comb_block.clear();
    P.ncells_comb.clear();
    for (int g = 0; g < ngroups; ++g) {
        for (int b = 0; b < nblocks; ++b) {
            if (count[(size_t)g * nblocks + b] > 0) {
                comb_of[(size_t)g * nblocks + b] = (int32_t)P.comb_group.size();
                P.comb_group.push_back(g);
                P.comb_block.push_back(b);
                P.ncells_comb.push_back((double)count[(size_t)g * nblocks + b]);
            }
        }
    }
    P.C = (int)P.comb_group.size();
    P.comb_id.assign(ncells, -1);
    for (int64_t c = 0; c < ncells; ++c) {
        const int g = groups[c];
        if (g < 0 || !used[g]) continue;
        int b = 0;
        if (block) {
            if (block[c] < 0) continue;
            b = bindex[block[c]];
        }
        P.comb_id[c] = comb_of[(size_t)g * nblocks + b];
    }

    // Block-level pairs: per block, the desired comparisons whose two groups are present,
    // de-duplicated in first-seen order; one orientation only when lfc == 0 (:304-339).
    P.left.clear();
    P.right.clear();
    for (int b = 0; b < nblocks; ++b) {
        std::set<std::pair<int32_t, int32_t>> seen;
        for (size_t k = 0; k < dl.size(); ++k) {
            const int32_t a = comb_of[(size_t)dl[k] * nblocks + b], c = comb_of[(size_t)dr[k] * nblocks + b];
            if (a < 0 || c < 0) continue;
            std::pair<int32_t, int32_t> key = P.collapse ? std::make_pair(std::max(a, c), std::min(a, c))
                                                         : std::make_pair(a, c);
            if (seen.insert(key).second) {
                P.left.push_back(key.first);
                P.right.push_back(key.second);
            }
        }
    }

    // Averaged comparisons (:463-476).
    std::map<std::pair<int32_t, int32_t>, std::vector<int32_t>> fold;
    for (size_t k = 0; k < P.left.size(); ++k) {
        fold[std::make_pair(P.comb_group[P.left[k]], P.comb_group[P.right[k]])].push_back((int32_t)k);
    }
    P.Q = (int)fold.size();
    P.avg_left.clear();
    P.avg_right.clear();
    P.avg_ptr.assign(1, 0);
    P.avg_idx.clear();
    P.avg_weight.clear();
    std::map<std::pair<int32_t, int32_t>, int32_t> qpos;
    for (auto& kv : fold) {
        qpos[kv.first] = (int32_t)P.avg_left.size();
        P.avg_left.push_back(kv.first.first);
        P.avg_right.push_back(kv.first.second);
        double w = 0;
        for (int32_t k : kv.second) {
            P.avg_idx.push_back(k);
            w += P.ncells_comb[P.left[k]] * P.ncells_comb[P.right[k]];
        }
        P.avg_weight.push_back(w);
        P.avg_ptr.push_back((int32_t)P.avg_idx.size());
    }

    // Output tables (:495-521).
    std::vector<char> is_left(ngroups, 0);
    for (int32_t a : dl) is_left[a] = 1;
    P.out_group.clear();
    P.group_table.assign(ngroups, -1);
    for (int g = 0; g < ngroups; ++g) {
        if (is_left[g]) {
            P.group_table[g] = (int32_t)P.out_group.size();
            P.out_group.push_back(g);
        }
    }
    P.K = (int)P.out_group.size();
    P.right_ptr.assign(P.K + 1, 0);
    for (int32_t a : dl) P.right_ptr[P.group_table[a] + 1]++;
    for (int k = 0; k < P.K; ++k) P.right_ptr[k + 1] += P.right_ptr[k];
    P.right_group.assign(dl.size(), 0);
    P.q_direct.assign(dl.size(), -1);
    P.q_flipped.assign(dl.size(), -1);
    P.right_weight.assign(dl.size(), 0.0);
    P.q_need.assign(std::max(1, P.Q), 0);
    {
        std::vector<int32_t> fill(P.right_ptr.begin(), P.right_ptr.end() - 1);
        for (size_t k = 0; k < dl.size(); ++k) {
            const int at = fill[P.group_table[dl[k]]]++;
            P.right_group[at] = dr[k];
            auto d = qpos.find(std::make_pair(dl[k], dr[k]));
            if (d != qpos.end()) {
                P.q_direct[at] = d->second;
                P.right_weight[at] = P.avg_weight[d->second];
                P.q_need[d->second] |= 1;
            } else if (P.collapse) {
                auto f = qpos.find(std::make_pair(dr[k], dl[k]));
                if (f != qpos.end()) {
                    P.q_flipped[at] = f->second;
                    P.right_weight[at] = P.avg_weight[f->second];
                    P.q_need[f->second] |= 2;
                }
            }
        }
    }

    // self./other. projector rows and "other" group lists (:395-411).
    std::vector<int32_t> groups_present;
    {
        std::vector<char> has(ngroups, 0);
        for (int c = 0; c < P.C; ++c) has[P.comb_group[c]] = 1;
        for (int g = 0; g < ngroups; ++g) {
            if (has[g]) groups_present.push_back(g);
        }
    }
    P.proj_ptr.assign(ngroups + 1, 0);
    P.proj_comb.clear();
    P.proj_coef.clear();
    std::vector<std::vector<double>> rows;
    if (P.blocked) wls_projector(P, groups_present, nblocks, rows);
    {
        size_t gp = 0;
        for (int g = 0; g < ngroups; ++g) {
            if (gp < groups_present.size() && groups_present[gp] == g) {
                if (P.blocked) {
                    const std::vector<double>& row = rows[gp];
                    if (row.empty()) {
                        P.proj_comb.push_back(0);
                        P.proj_coef.push_back(NAN);
                    } else {
                        for (int c = 0; c < P.C; ++c) {
                            if (row[c] != 0.0) {
                                P.proj_comb.push_back(c);
                                P.proj_coef.push_back(row[c]);
                            }
                        }
                    }
                } else {
                    for (int c = 0; c < P.C; ++c) {
                        if (P.comb_group[c] == g) {
                            P.proj_comb.push_back(c);
                            P.proj_coef.push_back(1.0);
                        }
                    }
                }
                ++gp;
            }
            P.proj_ptr[g + 1] = (int32_t)P.proj_comb.size();
        }
    }
    P.other_ptr.assign(P.K + 1, 0);
    P.other_group.clear();
    for (int k = 0; k < P.K; ++k) {
        std::set<int32_t> others;
        for (int j = P.right_ptr[k]; j < P.right_ptr[k + 1]; ++j) {
            const int32_t r = P.right_group[j];
            if (P.proj_ptr[r + 1] > P.proj_ptr[r]) others.insert(r);
        }
        for (int32_t r : others) P.other_group.push_back(r);
        P.other_ptr[k + 1] = (int32_t)P.other_group.size();
    }

    build_layout(P.comb_id.data(), ncells, P.C, P.left.data(), P.right.data(), (int64_t)P.left.size(), 148 * 4, P.layout);
}

void upload_plan(scranb200_ctx* ctx, scranb200_plan& P) {
    P.ctx = ctx;
    upload_layout(ctx, P.layout, P.d_layout);
    std::vector<int32_t> ints;
    auto put = [&](const std::vector<int32_t>& v) {
        const size_t at = ints.size();
        ints.insert(ints.end(), v.begin(), v.end());
        if (ints.size() & 1) ints.push_back(0);
        return at;
    };
    P.q_packed.assign(P.q_direct.size(), -1);
    for (size_t j = 0; j < P.q_direct.size(); ++j)
        P.q_packed[j] = P.q_direct[j] >= 0 ? P.q_direct[j] : (P.q_flipped[j] >= 0 ? -2 - P.q_flipped[j] : -1);
    const size_t o_packed = put(P.q_packed);
    const size_t o_left = put(P.left), o_right = put(P.right), o_avg_ptr = put(P.avg_ptr), o_avg_idx = put(P.avg_idx),
                 o_out_group = put(P.out_group), o_right_ptr = put(P.right_ptr), o_qd = put(P.q_direct),
                 o_qf = put(P.q_flipped), o_pp = put(P.proj_ptr), o_pc = put(P.proj_comb), o_op = put(P.other_ptr),
                 o_og = put(P.other_group), o_need = put(P.q_need), o_ql = put(P.avg_left), o_qr = put(P.avg_right),
                 o_gt = put(P.group_table);
    ints.push_back(0);
    ints.push_back(0);
    P.d_ints.reserve(ints.size() * 4);
    SB_CUDA(cudaMemcpyAsync(P.d_ints.p, ints.data(), ints.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    int32_t* b = P.d_ints.as<int32_t>();
    P.d_left = b + o_left;
    P.d_right = b + o_right;
    P.d_avg_ptr = b + o_avg_ptr;
    P.d_avg_idx = b + o_avg_idx;
    P.d_out_group = b + o_out_group;
    P.d_right_ptr = b + o_right_ptr;
    P.d_q_direct = b + o_qd;
    P.d_q_flipped = b + o_qf;
    P.d_q_packed = b + o_packed;
    P.d_proj_ptr = b + o_pp;
    P.d_proj_comb = b + o_pc;
    P.d_other_ptr = b + o_op;
    P.d_other_group = b + o_og;
    P.d_need = b + o_need;
    P.d_q_lgroup = b + o_ql;
    P.d_q_rgroup = b + o_qr;
    P.d_group_table = b + o_gt;
    std::vector<double> dbls;
    dbls.insert(dbls.end(), P.ncells_comb.begin(), P.ncells_comb.end());
    const size_t o_coef = dbls.size();
    dbls.insert(dbls.end(), P.proj_coef.begin(), P.proj_coef.end());
    dbls.push_back(0.0);
    P.d_dbls.reserve(dbls.size() * 8);
    SB_CUDA(cudaMemcpyAsync(P.d_dbls.p, dbls.data(), dbls.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    P.d_ncells_comb = P.d_dbls.as<double>();
    P.d_proj_coef = P.d_dbls.as<double>() + o_coef;
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
}

}  // namespace sb
