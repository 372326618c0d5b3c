// K3: exact pairwise Wilcoxon U statistics for every gene and every requested pair.
//
// Reference semantics: src/overlap_exprs.cpp:34-128 (class wilcoxer). There, every gene gathers
// the cells of every group from a dense row, sorts each group and runs one two-pointer sweep per
// pair: O(ncells) + O(nnz_gene * (G-1)) scalar steps per gene. Here the work is O(nnz) and the
// zeros are handled in closed form:
//
//  k3_sort  one CTA per (gene, unit). The unit's stored non-zero values (all groups of one
//           block) plus one "probe" element at 0 are sorted by value with their group label as
//           payload (shared-memory radix sort, or the same sort streamed through L2 for units
//           beyond the shared-memory capacity). Output: a 2-byte-per-entry label stream in value
//           order with a "ties with previous" bit. With lfc != 0 the unit is sorted twice over:
//           every value once as b ("R role") and once as a - lfc ("L role"), R before L in a tie.
//  k3_walk  one CTA per (gene, unit): lane R keeps the running count c_R of label-R entries seen
//           so far (cs_R = the count at the start of the current tie run) and adds it into
//           row label(i) of a (G+1) x (G+1) table in shared memory while stepping through the
//           stream. That yields B[L][R] = sum over a in L of #(b in R : b < a) for all ordered
//           pairs at once; the probe row/column gives #(a > 0) and #(b < 0), from which the
//           implicit zeros enter as  B[L][R] + B[L][probe] z_R + z_L B[probe][R]  (z = zero
//           count). Then 2U = nL nR + Bfull[L][R] - Bfull[R][L] in int64, exact.
//           (lfc != 0: the table accumulates cs_R + c_R at L-role entries and
//           2U = B[L][R] + z_R B[L][p] + z_L B[p][R] + z_L z_R B[p][p].)
//  k3_ties  optional: sum over tie runs of t(t^2-1) per pair, accumulated in double in ascending
//           value order like src/overlap_exprs.cpp:121-124.
#include "kernels.h"
#include "sort.cuh"

namespace sb {

namespace {

constexpr uint16_t LAB_MASK = 0x3fff;
constexpr uint16_t ROLE_L = 0x4000;
constexpr uint16_t TIE_BIT = 0x8000;
constexpr int MAX_LABELS = 16383;

struct UnitTables {
    const int64_t* row_ptr;
    const uint32_t* seg_off;
    const int32_t* unit_first;
    const int32_t* unit_ncomb;
    const int64_t* ncells_dev;
    int ngenes, U, mult;  // mult = 1 (lfc == 0) or 2
};

// Offset of (gene, unit) in the label stream / sort scratch, and its value range.
__device__ inline void unit_range(const UnitTables& ut, int g, int u, int64_t& vbase, uint32_t& n, int64_t& soff) {
    const int first = ut.unit_first[u], G = ut.unit_ncomb[u];
    const uint32_t a = ut.seg_off[(size_t)first * ut.ngenes + g];
    const uint32_t b = ut.seg_off[(size_t)(first + G) * ut.ngenes + g];
    vbase = ut.row_ptr[g] + a;
    n = b - a;
    soff = (int64_t)ut.mult * (vbase + (int64_t)g * ut.U + u);
}

// ------------------------------------------------------------------------------------------
// k3_sort
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SORT_THREADS)
k3_sort(UnitTables ut, const double* __restrict__ vals, double shift, int cap, int maxG, uint16_t* stream,
        uint32_t* unit_m, uint64_t* g_keys_a, uint64_t* g_keys_b, uint16_t* g_labs_a, uint16_t* g_labs_b,
        unsigned long long* task_counter) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* s_ka = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* s_kb = s_ka + cap;
    uint16_t* s_la = reinterpret_cast<uint16_t*>(s_kb + cap);
    uint16_t* s_lb = s_la + cap;
    uint32_t* whist = reinterpret_cast<uint32_t*>(s_lb + cap);
    uint32_t* segb = whist + SORT_HIST_WORDS;
    uint32_t* misc = segb + (maxG + 2);  // [0] task, [1] dropped

    const int64_t ntasks = (int64_t)ut.ngenes * ut.U;
    while (true) {
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned long long t = atomicAdd(task_counter, 1ull);
            misc[0] = (uint32_t)(t & 0xffffffffu);
            misc[2] = (uint32_t)(t >> 32);
            misc[1] = 0;
        }
        __syncthreads();
        const int64_t task = (int64_t)(((uint64_t)misc[2] << 32) | misc[0]);
        if (task >= ntasks) break;
        const int g = (int)(task / ut.U), u = (int)(task % ut.U);
        const int first = ut.unit_first[u], G = ut.unit_ncomb[u];
        int64_t vbase, soff;
        uint32_t n;
        unit_range(ut, g, u, vbase, n, soff);
        const uint32_t seg0 = ut.seg_off[(size_t)first * ut.ngenes + g];
        for (int k = threadIdx.x; k <= G; k += blockDim.x) {
            segb[k] = ut.seg_off[(size_t)(first + k) * ut.ngenes + g] - seg0;
        }
        __syncthreads();

        const int64_t per_role = (int64_t)n + 1;
        const int64_t total = per_role * ut.mult;
        const bool in_smem = total <= cap;
        uint64_t* ka = in_smem ? s_ka : g_keys_a + soff;
        uint64_t* kb = in_smem ? s_kb : g_keys_b + soff;
        uint16_t* la = in_smem ? s_la : g_labs_a + soff;
        uint16_t* lb = in_smem ? s_lb : g_labs_b + soff;

        // Build the element list: R role first, then (lfc != 0) the shifted L role.
        uint32_t dropped = 0;
        for (int64_t e = threadIdx.x; e < total; e += blockDim.x) {
            const int role = e >= per_role;
            const uint32_t k = (uint32_t)(e - (role ? per_role : 0));
            double v = 0.0;
            uint16_t lab = (uint16_t)G;
            bool keep = true;
            if (k < n) {
                v = vals[vbase + k];
                keep = v != 0.0;  // stored zeros count as zeros (src/overlap_exprs.cpp:43-47)
                int lo = 0, hi = G;  // label = last boundary <= k
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (segb[mid] <= k) lo = mid; else hi = mid;
                }
                lab = (uint16_t)lo;
            }
            if (role) {
                v = v - shift;
                lab |= ROLE_L;
            }
            ka[e] = keep ? key_of(v) : ~0ull;
            la[e] = lab;
            dropped += keep ? 0u : 1u;
        }
        if (dropped) atomicAdd(&misc[1], dropped);
        __syncthreads();

        const int cur = block_sort_keys<uint16_t>(ka, la, kb, lb, (int)total, whist);
        const uint64_t* rk = cur ? kb : ka;
        const uint16_t* rl = cur ? lb : la;
        const uint32_t m = (uint32_t)total - misc[1];
        for (uint32_t j = threadIdx.x; j < m; j += blockDim.x) {
            uint16_t e = rl[j];
            if (j > 0 && rk[j] == rk[j - 1]) e |= TIE_BIT;
            stream[soff + j] = e;
        }
        if (threadIdx.x == 0) unit_m[task] = m;
    }
}

// ------------------------------------------------------------------------------------------
// k3_walk
// ------------------------------------------------------------------------------------------
constexpr int WALK_THREADS = 512;
constexpr int WALK_WARPS = WALK_THREADS / 32;
constexpr int WALK_MAX_GROUPS = 16;

struct WalkParams {
    UnitTables ut;
    const uint16_t* stream;
    const uint32_t* unit_m;
    const int32_t *pair_ptr, *pair_l, *pair_r, *pair_out;
    int64_t npairs;
    double* u_out;
    unsigned long long* bfull;  // per-CTA scratch: (maxG + 1)^2 entries each
    unsigned long long* task_counter;
    int maxG;
    int tiles_per_batch;  // R tiles whose accumulators fit in shared memory at once
    int groups;           // position chunks walked concurrently (private accumulator copies)
    uint32_t m_lo, m_hi;  // this launch handles units with m_lo < m <= m_hi
};

template <typename AccT, bool SHIFT>
__global__ void __launch_bounds__(WALK_THREADS) k3_walk(WalkParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Gmax1 = P.maxG + 1;
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem_raw);        // [groups + 1][Gmax1]
    uint32_t* nnzv = hist + (WALK_MAX_GROUPS + 1) * Gmax1;          // [Gmax1]
    int* bounds = reinterpret_cast<int*>(nnzv + Gmax1);             // [WALK_MAX_GROUPS + 1]
    uint32_t* misc = reinterpret_cast<uint32_t*>(bounds + WALK_MAX_GROUPS + 1);  // [4]
    size_t off = (size_t)((unsigned char*)(misc + 4) - smem_raw);
    off = (off + 15) & ~size_t(15);
    AccT* acc = reinterpret_cast<AccT*>(smem_raw + off);            // [groups][Gt][tiles_per_batch * 32]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long* bfull = P.bfull + (size_t)blockIdx.x * Gmax1 * Gmax1;
    const int64_t ntasks = (int64_t)P.ut.ngenes * P.ut.U;

    while (true) {
        __syncthreads();
        if (tid == 0) {
            const unsigned long long t = atomicAdd(P.task_counter, 1ull);
            misc[0] = (uint32_t)(t & 0xffffffffu);
            misc[1] = (uint32_t)(t >> 32);
        }
        __syncthreads();
        const int64_t task = (int64_t)(((uint64_t)misc[1] << 32) | misc[0]);
        if (task >= ntasks) break;
        const uint32_t m = P.unit_m[task];
        if (!(m > P.m_lo && m <= P.m_hi)) continue;
        const int g = (int)(task / P.ut.U), u = (int)(task % P.ut.U);
        const int first = P.ut.unit_first[u], G = P.ut.unit_ncomb[u];
        const int Gt = G + 1;
        int64_t vbase, soff;
        uint32_t n;
        unit_range(P.ut, g, u, vbase, n, soff);
        const uint16_t* st = P.stream + soff;

        // Position chunks aligned to tie-run starts.
        const int ntiles = (Gt + 31) / 32;
        const int tpb = min(P.tiles_per_batch, ntiles);
        int groups = max(1, min(P.groups, WALK_WARPS / tpb));
        if (m < 2048u) groups = 1;
        if (tid <= groups) {
            int64_t b = tid == groups ? (int64_t)m : ((int64_t)m * tid) / groups;
            while (b < (int64_t)m && b > 0 && (st[b] & TIE_BIT)) ++b;
            bounds[tid] = (int)b;
        }
        for (int k = tid; k < (groups + 1) * Gmax1; k += blockDim.x) hist[k] = 0;
        __syncthreads();
        for (uint32_t i = tid; i < m; i += blockDim.x) {
            const uint16_t e = st[i];
            if (SHIFT && (e & ROLE_L)) continue;
            int cg = 0;
            while (cg + 1 < groups && (int)i >= bounds[cg + 1]) ++cg;
            atomicAdd(&hist[(cg + 1) * Gmax1 + (e & LAB_MASK)], 1u);
        }
        __syncthreads();
        // hist[cg][R] becomes the count of label R before chunk cg; nnzv the total.
        for (int R = tid; R < Gt; R += blockDim.x) {
            uint32_t run = 0;
            for (int cg = 0; cg <= groups; ++cg) {
                const uint32_t v = hist[cg * Gmax1 + R];
                run += v;
                hist[cg * Gmax1 + R] = run;
            }
            nnzv[R] = run;
        }
        __syncthreads();

        for (int tb = 0; tb < ntiles; tb += tpb) {
            const int bt = min(tpb, ntiles - tb);
            const int rpad = bt * 32;
            for (int k = tid; k < groups * Gt * rpad; k += blockDim.x) acc[k] = 0;
            __syncthreads();
            if (warp < groups * bt) {
                const int cg = warp / bt, tile = tb + warp % bt;
                const int R = tile * 32 + lane;
                const bool live = R < Gt;
                AccT* row0 = acc + (size_t)cg * Gt * rpad + (R - tb * 32);
                uint32_t c = live ? hist[cg * Gmax1 + R] : 0u;
                uint32_t cs = c;
                const int i_hi = bounds[cg + 1];
                for (int i = bounds[cg]; i < i_hi; ++i) {
                    const uint16_t e = st[i];
                    const int lab = e & LAB_MASK;
                    if (!(e & TIE_BIT)) cs = c;
                    if (SHIFT) {
                        if (e & ROLE_L) {
                            if (live) row0[(size_t)lab * rpad] += (AccT)cs + (AccT)c;
                        } else {
                            c += (lab == R) ? 1u : 0u;
                        }
                    } else {
                        if (live) row0[(size_t)lab * rpad] += (AccT)cs;
                        c += (lab == R) ? 1u : 0u;
                    }
                }
            }
            __syncthreads();
            for (int k = tid; k < Gt * rpad; k += blockDim.x) {
                const int L = k / rpad, R = tb * 32 + k % rpad;
                if (R < Gt) {
                    unsigned long long s = 0;
                    for (int cg = 0; cg < groups; ++cg) s += (unsigned long long)acc[(size_t)cg * Gt * rpad + k];
                    bfull[(size_t)L * Gt + R] = s;
                }
            }
            __syncthreads();
        }

        // Pairs of this unit.
        const int pg = G;  // probe label
        for (int pp = P.pair_ptr[u] + tid; pp < P.pair_ptr[u + 1]; pp += blockDim.x) {
            const int l = P.pair_l[pp], r = P.pair_r[pp];
            const long long nL = P.ut.ncells_dev[first + l], nR = P.ut.ncells_dev[first + r];
            const long long zL = nL - (long long)nnzv[l], zR = nR - (long long)nnzv[r];
            long long two;
            if (SHIFT) {
                two = (long long)bfull[(size_t)l * Gt + r] + zR * (long long)bfull[(size_t)l * Gt + pg] +
                      zL * (long long)bfull[(size_t)pg * Gt + r] + zL * zR * (long long)bfull[(size_t)pg * Gt + pg];
            } else {
                const long long a = (long long)bfull[(size_t)l * Gt + r] + (long long)bfull[(size_t)l * Gt + pg] * zR +
                                    zL * (long long)bfull[(size_t)pg * Gt + r];
                const long long b = (long long)bfull[(size_t)r * Gt + l] + (long long)bfull[(size_t)r * Gt + pg] * zL +
                                    zR * (long long)bfull[(size_t)pg * Gt + l];
                two = nL * nR + a - b;
            }
            if (nL == 0 || nR == 0) two = 0;
            P.u_out[(size_t)g * P.npairs + P.pair_out[pp]] = 0.5 * (double)two;
        }
    }
}

// ------------------------------------------------------------------------------------------
// k3_ties: tie term per pair. One CTA per (gene, unit); one thread per pair walks the tie runs.
// A run contributes t(t^2-1) with t = (entries of the left label in L role) + (entries of the
// right label in R role); the zeros of a group join the run at value 0 (R role) / -lfc (L role),
// which is where the probe entries sit. Runs where t <= 1 add exactly 0 and are skipped, so the
// double accumulation visits the same non-zero terms in the same order as the reference.
// ------------------------------------------------------------------------------------------
struct TieParams {
    UnitTables ut;
    const uint16_t* stream;
    const uint32_t* unit_m;
    const int32_t *pair_ptr, *pair_l, *pair_r, *pair_out;
    int64_t npairs;
    double* t_out;
    uint32_t* run_scratch;  // per CTA: compacted tie-run table
    uint32_t run_cap;       // entries per CTA
    int maxG;
};

// Compacted representation per CTA: a list of "events" (run id, label|role, count) for every
// (run, label, role) with a non-zero count, restricted to runs that can contribute (size >= 2 or
// containing a probe). Built by one pass over the stream; then each pair thread scans the events.
template <bool SHIFT>
__global__ void __launch_bounds__(256) k3_ties(TieParams P) {
    const int64_t task = blockIdx.x;
    const int g = (int)(task / P.ut.U), u = (int)(task % P.ut.U);
    const int first = P.ut.unit_first[u], G = P.ut.unit_ncomb[u];
    int64_t vbase, soff;
    uint32_t n;
    unit_range(P.ut, g, u, vbase, n, soff);
    const uint16_t* st = P.stream + soff;
    const uint32_t m = P.unit_m[task];
    __shared__ uint32_t s_nnz[1];
    // zeros per label need the stored non-zero count: count R-role entries per label on the fly
    // (per pair thread, below) - cheap relative to the run scan because it is fused into it.
    (void)s_nnz;
    for (int pp = P.pair_ptr[u] + threadIdx.x; pp < P.pair_ptr[u + 1]; pp += blockDim.x) {
        const int l = P.pair_l[pp], r = P.pair_r[pp];
        const long long nL = P.ut.ncells_dev[first + l], nR = P.ut.ncells_dev[first + r];
        // First pass: stored non-zero counts of the two labels.
        long long cl = 0, cr = 0;
        for (uint32_t i = 0; i < m; ++i) {
            const uint16_t e = st[i];
            if (SHIFT && (e & ROLE_L)) continue;
            const int lab = e & LAB_MASK;
            cl += lab == l;
            cr += lab == r;
        }
        const long long zL = nL - cl, zR = nR - cr;
        double tie = 0.0;
        long long t = 0;
        bool open = false;
        for (uint32_t i = 0; i <= m; ++i) {
            const bool boundary = i == m || !(st[i] & TIE_BIT);
            if (boundary && open) {
                if (t > 0) {
                    const double nt = (double)t;
                    tie += nt * (nt * nt - 1.0);
                }
                t = 0;
            }
            if (i == m) break;
            open = true;
            const uint16_t e = st[i];
            const int lab = e & LAB_MASK;
            const bool roleL = SHIFT ? (e & ROLE_L) != 0 : false;
            if (lab == G) {  // probe: the zeros of the pair's groups sit here
                if (SHIFT) {
                    t += roleL ? zL : zR;
                } else {
                    t += zL + zR;
                }
            } else if (SHIFT) {
                if (roleL && lab == l) t += 1;
                if (!roleL && lab == r) t += 1;
            } else {
                if (lab == l || lab == r) t += 1;
            }
        }
        if (nL == 0 || nR == 0) tie = 0.0;
        P.t_out[(size_t)g * P.npairs + P.pair_out[pp]] = tie;
    }
}

}  // namespace

void run_wilcox(scranb200_ctx* ctx, const Transposed& T, const PairLayout& L, double lfc, double* d_u, double* d_t) {
    if (L.U == 0 || T.ngenes == 0 || L.npairs == 0) return;
    if (L.maxG > MAX_LABELS) throw Error("more than 16383 groups compared inside one block are not supported");
    const bool shift = lfc != 0.0;
    const int mult = shift ? 2 : 1;
    const int64_t ntasks = (int64_t)T.ngenes * L.U;

    int64_t nnz_used = 0;
    SB_CUDA(cudaMemcpyAsync(&nnz_used, T.row_ptr + T.ngenes, 8, cudaMemcpyDeviceToHost, ctx->stream));
    SB_CUDA(cudaStreamSynchronize(ctx->stream));
    const int64_t stream_len = mult * (nnz_used + ntasks);

    UnitTables ut;
    ut.row_ptr = T.row_ptr;
    ut.seg_off = T.seg_off;
    ut.unit_first = L.d_unit_first;
    ut.unit_ncomb = L.d_unit_ncomb;
    ut.ncells_dev = L.d_ncells_dev;
    ut.ngenes = T.ngenes;
    ut.U = L.U;
    ut.mult = mult;

    ctx->stream_lab.reserve((size_t)stream_len * 2 + 64);
    ctx->unit_m.reserve((size_t)ntasks * 4 + 64);
    ctx->flags.reserve(64);
    unsigned long long* counter = reinterpret_cast<unsigned long long*>(ctx->flags.as<char>() + 16);

    // ---- sort ----
    const size_t fixed = (size_t)SORT_HIST_WORDS * 4 + (size_t)(L.maxG + 2) * 4 + 64;
    int cap = (int)((ctx->smem_optin - fixed) / 20);
    cap &= ~31;
    const size_t sort_smem = (size_t)cap * 20 + fixed;
    // Units beyond the shared-memory capacity are sorted through global scratch.
    int64_t max_cells = 0;
    for (int u = 0; u < L.U; ++u) {
        int64_t s = 0;
        for (int k = 0; k < L.unit_ncomb[u]; ++k) s += L.ncells_dev[L.unit_first[u] + k];
        max_cells = std::max(max_cells, s);
    }
    const bool may_spill = mult * (std::min<int64_t>(max_cells, nnz_used) + 1) > cap;
    uint64_t *gka = nullptr, *gkb = nullptr;
    uint16_t *gla = nullptr, *glb = nullptr;
    if (may_spill) {
        ctx->sort_keys.reserve((size_t)stream_len * 16 + 64);
        ctx->sort_labs.reserve((size_t)stream_len * 4 + 64);
        gka = ctx->sort_keys.as<uint64_t>();
        gkb = gka + stream_len;
        gla = ctx->sort_labs.as<uint16_t>();
        glb = gla + stream_len;
    }
    allow_smem(k3_sort, sort_smem);
    SB_CUDA(cudaMemsetAsync(counter, 0, 8, ctx->stream));
    const int sort_grid = (int)std::min<int64_t>(ntasks, (int64_t)ctx->sm_count);
    SB_LAUNCH(ctx, k3_sort, sort_grid, SORT_THREADS, sort_smem, ut, T.vals, lfc, cap, L.maxG, ctx->stream_lab.as<uint16_t>(),
              ctx->unit_m.as<uint32_t>(), gka, gkb, gla, glb, counter);
    phase_mark(ctx, PH_SORT);

    // ---- walk ----
    WalkParams wp;
    wp.ut = ut;
    wp.stream = ctx->stream_lab.as<uint16_t>();
    wp.unit_m = ctx->unit_m.as<uint32_t>();
    wp.pair_ptr = L.d_pair_ptr;
    wp.pair_l = L.d_pair_l;
    wp.pair_r = L.d_pair_r;
    wp.pair_out = L.d_pair_out;
    wp.npairs = L.npairs;
    wp.u_out = d_u;
    wp.task_counter = counter;
    wp.maxG = L.maxG;
    const int Gt = L.maxG + 1;
    const int ntiles = (Gt + 31) / 32;
    const size_t walk_fixed = ((size_t)(WALK_MAX_GROUPS + 1) * Gt + Gt) * 4 + (WALK_MAX_GROUPS + 1) * 4 + 16 + 32;
    const int walk_grid_max = ctx->sm_count * 2;
    ctx->plan_tmp.reserve((size_t)walk_grid_max * Gt * Gt * 8 + 64);
    wp.bfull = ctx->plan_tmp.as<unsigned long long>();

    const int64_t max_m = mult * (std::min<int64_t>(max_cells, nnz_used) + 1);
    for (int pass = 0; pass < 2; ++pass) {
        const size_t acc_size = pass == 0 ? 4 : 8;
        if (pass == 1 && max_m <= 65536) break;
        wp.m_lo = pass == 0 ? 0u : 65536u;
        wp.m_hi = pass == 0 ? 65536u : 0xffffffffu;
        // Aim for two CTAs per SM; use what is left for private accumulator copies.
        const size_t budget = std::min<size_t>(ctx->smem_optin, (size_t)110 * 1024) - walk_fixed;
        const size_t big_budget = ctx->smem_optin - walk_fixed;
        const size_t per_tile = (size_t)Gt * 32 * acc_size;
        int tpb, groups;
        size_t smem;
        if (per_tile * ntiles <= budget) {
            tpb = ntiles;
            groups = (int)std::min<size_t>(budget / (per_tile * ntiles), (size_t)WALK_MAX_GROUPS);
            groups = std::max(1, std::min(groups, WALK_WARPS / std::max(1, tpb)));
            smem = walk_fixed + per_tile * ntiles * groups;
        } else {
            tpb = (int)std::max<size_t>(1, std::min<size_t>(big_budget / per_tile, (size_t)ntiles));
            if (per_tile > big_budget) throw Error("too many groups inside one block for the shared-memory count table");
            tpb = std::min(tpb, WALK_WARPS);
            groups = 1;
            smem = walk_fixed + per_tile * tpb;
        }
        wp.tiles_per_batch = tpb;
        wp.groups = groups;
        SB_CUDA(cudaMemsetAsync(counter, 0, 8, ctx->stream));
        const int grid = (int)std::min<int64_t>(ntasks, walk_grid_max);
        if (pass == 0) {
            if (shift) {
                allow_smem(k3_walk<uint32_t, true>, smem);
                SB_LAUNCH(ctx, (k3_walk<uint32_t, true>), grid, WALK_THREADS, smem, wp);
            } else {
                allow_smem(k3_walk<uint32_t, false>, smem);
                SB_LAUNCH(ctx, (k3_walk<uint32_t, false>), grid, WALK_THREADS, smem, wp);
            }
        } else {
            if (shift) {
                allow_smem(k3_walk<unsigned long long, true>, smem);
                SB_LAUNCH(ctx, (k3_walk<unsigned long long, true>), grid, WALK_THREADS, smem, wp);
            } else {
                allow_smem(k3_walk<unsigned long long, false>, smem);
                SB_LAUNCH(ctx, (k3_walk<unsigned long long, false>), grid, WALK_THREADS, smem, wp);
            }
        }
    }

    if (d_t) {
        TieParams tp;
        tp.ut = ut;
        tp.stream = wp.stream;
        tp.unit_m = wp.unit_m;
        tp.pair_ptr = L.d_pair_ptr;
        tp.pair_l = L.d_pair_l;
        tp.pair_r = L.d_pair_r;
        tp.pair_out = L.d_pair_out;
        tp.npairs = L.npairs;
        tp.t_out = d_t;
        tp.run_scratch = nullptr;
        tp.run_cap = 0;
        tp.maxG = L.maxG;
        if (ntasks > 2147483647ll) throw Error("too many (gene, block) units for the tie-term kernel");
        if (shift) {
            SB_LAUNCH(ctx, k3_ties<true>, (unsigned)ntasks, 256, 0, tp);
        } else {
            SB_LAUNCH(ctx, k3_ties<false>, (unsigned)ntasks, 256, 0, tp);
        }
    }
    phase_mark(ctx, PH_COUNT);
}

}  // namespace sb
