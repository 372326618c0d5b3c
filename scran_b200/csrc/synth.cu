// Synthetic sparse log-count matrices for bench.py, generated on the device (BASELINE.md
// section 5 / SURVEY.md section 8d). Counter-based: every (gene, cell) entry is a pure function of
// (seed, gene, cell), so any row block [g0, g1) can be produced independently on any GPU and the
// pieces of a gene-sharded run together equal the single-GPU matrix.
//
//   cell size factor  s_c = exp(0.25 * N(0,1))
//   gene rate         lambda_g = alpha * 2^U(-6,3), alpha chosen so that the expected non-zero
//                     fraction is `density`
//   group effect      2^(0.5 * N(0,1)) per (gene, group) on 10 % of the genes
//   count             k ~ Poisson(lambda_g * s_c * effect);  value = log2(k / s_c + 1)
//   degenerate genes  1 % all zero; genes 7, 1007, 2007 expressed in every cell; every 97th gene
//                     integer-valued (heavy ties); gene 13 alternates sign by cell
#include <cmath>

#include "kernels.h"

namespace sb {

namespace {

__host__ __device__ inline uint64_t mix64(uint64_t z) {
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
__host__ __device__ inline uint64_t hash3(uint64_t seed, uint64_t a, uint64_t b) {
    return mix64(mix64(seed ^ (a * 0xd1342543de82ef95ull)) ^ (b * 0xaf251af3b0f025b5ull));
}
__host__ __device__ inline double unit(uint64_t h) { return ((h >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
__host__ __device__ inline double normal(uint64_t h) {
    const double u1 = unit(h), u2 = unit(mix64(h));
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

enum GeneKind { G_NORMAL = 0, G_ZERO = 1, G_ALWAYS = 2, G_INTEGER = 3, G_NEGATIVE = 4 };

inline int gene_kind(uint64_t seed, int64_t g) {
    if (g == 13) return G_NEGATIVE;
    if (g == 7 || g == 1007 || g == 2007) return G_ALWAYS;
    if (g % 97 == 5) return G_INTEGER;
    if (hash3(seed, 0x5a5a, (uint64_t)g) % 100 == 0) return G_ZERO;
    return G_NORMAL;
}

struct SynthParams {
    int64_t g0, ngenes, ncells;
    const double* rate;       // [ngenes] already scaled
    const uint8_t* kind;      // [ngenes]
    const uint8_t* effect;    // [ngenes] 1 = has group effects
    const int32_t* groups;    // [ncells]
    uint64_t seed;
};

__device__ inline double entry_value(const SynthParams& P, int64_t gl, int64_t c, double s, int grp) {
    const int kind = P.kind[gl];
    if (kind == G_ZERO) return 0.0;
    const int64_t g = P.g0 + gl;
    double lam = P.rate[gl] * s;
    if (P.effect[gl] && grp >= 0) lam *= exp2(0.5 * normal(hash3(P.seed, 0xeffec7ull + (uint64_t)grp, (uint64_t)g)));
    if (kind == G_ALWAYS) lam = 12.0 * s;
    const double u = unit(hash3(P.seed, (uint64_t)g, (uint64_t)c));
    double p = exp(-lam), cdf = p;
    int k = 0;
    while (u > cdf && k < 4096) {
        ++k;
        p *= lam / k;
        cdf += p;
    }
    if (kind == G_ALWAYS && k == 0) k = 1;
    if (k == 0) return 0.0;
    if (kind == G_INTEGER) return (double)k;
    double v = log2((double)k / s + 1.0);
    if (kind == G_NEGATIVE && (c & 1)) v = -v;
    return v;
}

// One warp per cell; lanes stride the genes, so stored rows come out ascending.
template <bool FILL>
__global__ void __launch_bounds__(256) synth_kernel(SynthParams P, int64_t* counts, const int64_t* colptr, int32_t* idx,
                                                     double* val) {
    const int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (c >= P.ncells) return;
    const double s = exp(0.25 * normal(hash3(P.seed, 0x5f5f, (uint64_t)c)));
    const int grp = P.groups ? P.groups[c] : -1;
    int64_t at = FILL ? colptr[c] : 0;
    for (int64_t g0 = 0; g0 < P.ngenes; g0 += 32) {
        const int64_t gl = g0 + lane;
        const double v = gl < P.ngenes ? entry_value(P, gl, c, s, grp) : 0.0;
        const unsigned m = __ballot_sync(0xffffffffu, v != 0.0);
        if (FILL && v != 0.0) {
            const int64_t o = at + __popc(m & ((1u << lane) - 1u));
            idx[o] = (int32_t)gl;
            val[o] = v;
        }
        at += __popc(m);
    }
    if (!FILL && lane == 0) counts[c] = at;
}

__global__ void __launch_bounds__(1024) synth_scan(const int64_t* counts, int64_t n, int64_t* out) {
    __shared__ int64_t part[1024];
    const int t = threadIdx.x;
    const int64_t per = (n + 1023) / 1024;
    const int64_t lo = min(n, t * per), hi = min(n, lo + per);
    int64_t s = 0;
    for (int64_t k = lo; k < hi; ++k) s += counts[k];
    part[t] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        const int64_t v = t >= d ? part[t - d] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    int64_t run = part[t] - s;
    for (int64_t k = lo; k < hi; ++k) {
        const int64_t cnt = counts[k];
        out[k] = run;
        run += cnt;
    }
    if (t == 1023) out[n] = part[1023];
}

}  // namespace
}  // namespace sb

using namespace sb;

// Expected number of stored entries per gene of the synthetic matrix (the same rates as
// scranb200_synth_csc uses), for cutting gene shards of equal weight before anything is generated.
extern "C" int scranb200_synth_row_weights(int64_t ngenes_total, int64_t ncells, double density, uint64_t seed, double* weight) {
    try {
        if (density <= 0.0 || density >= 1.0) density = 0.63;
        std::vector<double> base(ngenes_total);
        std::vector<uint8_t> kind(ngenes_total);
        for (int64_t g = 0; g < ngenes_total; ++g) {
            base[g] = std::exp2(-6.0 + 9.0 * unit(hash3(seed, 0x7a7e, (uint64_t)g)));
            kind[g] = (uint8_t)gene_kind(seed, g);
        }
        auto frac = [&](double alpha) {
            double s = 0;
            for (int64_t g = 0; g < ngenes_total; ++g) {
                if (kind[g] == G_ZERO) continue;
                s += kind[g] == G_ALWAYS ? 1.0 : 1.0 - std::exp(-alpha * base[g]);
            }
            return s / (double)std::max<int64_t>(1, ngenes_total);
        };
        double lo = 1e-6, hi = 1e6;
        for (int it = 0; it < 200; ++it) {
            const double mid = std::sqrt(lo * hi);
            if (frac(mid) < density) lo = mid; else hi = mid;
        }
        const double alpha = std::sqrt(lo * hi);
        for (int64_t g = 0; g < ngenes_total; ++g) {
            const double d = kind[g] == G_ZERO ? 0.0 : kind[g] == G_ALWAYS ? 1.0 : 1.0 - std::exp(-alpha * base[g]);
            weight[g] = d * (double)ncells;
        }
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}

extern "C" int scranb200_synth_csc(scranb200_ctx* ctx, int64_t ngenes_total, int64_t g0, int64_t g1, int64_t ncells,
                                    const int32_t* groups, int ngroups, double density, uint64_t seed,
                                    scranb200_mat** out) {
    try {
        ctx_guard(ctx);
        if (g0 < 0 || g1 > ngenes_total || g0 > g1) throw Error("gene range out of bounds");
        if (density <= 0.0 || density >= 1.0) density = 0.63;
        (void)ngroups;
        // Base rates for ALL genes (so that the scale is the same whichever shard asks).
        std::vector<double> base(ngenes_total);
        std::vector<uint8_t> kind(ngenes_total);
        for (int64_t g = 0; g < ngenes_total; ++g) {
            base[g] = std::exp2(-6.0 + 9.0 * unit(hash3(seed, 0x7a7e, (uint64_t)g)));
            kind[g] = (uint8_t)gene_kind(seed, g);
        }
        auto frac = [&](double alpha) {
            double s = 0;
            for (int64_t g = 0; g < ngenes_total; ++g) {
                if (kind[g] == G_ZERO) continue;
                s += kind[g] == G_ALWAYS ? 1.0 : 1.0 - std::exp(-alpha * base[g]);
            }
            return s / (double)std::max<int64_t>(1, ngenes_total);
        };
        double lo = 1e-6, hi = 1e6;
        for (int it = 0; it < 200; ++it) {
            const double mid = std::sqrt(lo * hi);
            if (frac(mid) < density) lo = mid; else hi = mid;
        }
        const double alpha = std::sqrt(lo * hi);
        const int64_t n = g1 - g0;
        std::vector<double> rate(std::max<int64_t>(1, n));
        std::vector<uint8_t> kk(std::max<int64_t>(1, n)), eff(std::max<int64_t>(1, n));
        for (int64_t g = 0; g < n; ++g) {
            rate[g] = alpha * base[g0 + g];
            kk[g] = kind[g0 + g];
            eff[g] = hash3(seed, 0xe0e0, (uint64_t)(g0 + g)) % 10 == 0;
        }
        ctx->io_a.reserve(rate.size() * 8 + 64);
        ctx->io_b.reserve(kk.size() * 2 + 64);
        ctx->io_c.reserve((size_t)std::max<int64_t>(1, ncells) * 4 + 64);
        SB_CUDA(cudaMemcpyAsync(ctx->io_a.p, rate.data(), rate.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
        SB_CUDA(cudaMemcpyAsync(ctx->io_b.p, kk.data(), kk.size(), cudaMemcpyHostToDevice, ctx->stream));
        SB_CUDA(cudaMemcpyAsync(ctx->io_b.as<uint8_t>() + kk.size(), eff.data(), eff.size(), cudaMemcpyHostToDevice,
                                ctx->stream));
        if (groups && ncells > 0) {
            SB_CUDA(cudaMemcpyAsync(ctx->io_c.p, groups, (size_t)ncells * 4, cudaMemcpyHostToDevice, ctx->stream));
        }
        SynthParams sp;
        sp.g0 = g0;
        sp.ngenes = n;
        sp.ncells = ncells;
        sp.rate = ctx->io_a.as<double>();
        sp.kind = ctx->io_b.as<uint8_t>();
        sp.effect = ctx->io_b.as<uint8_t>() + kk.size();
        sp.groups = groups ? ctx->io_c.as<int32_t>() : nullptr;
        sp.seed = seed;

        scranb200_mat* m = new scranb200_mat();
        m->ctx = ctx;
        m->nrow = n;
        m->ncol = ncells;
        m->owned = true;
        ctx->live_mats.push_back(m);
        void* pp = nullptr;
        SB_CUDA(cudaMalloc(&pp, (size_t)(ncells + 1) * 8));
        m->d_p = (int64_t*)pp;
        DevBuf counts;
        counts.reserve((size_t)std::max<int64_t>(1, ncells) * 8);
        const unsigned grid = (unsigned)ceil_div(std::max<int64_t>(1, ncells) * 32, 256);
        SB_LAUNCH(ctx, synth_kernel<false>, grid, 256, 0, sp, counts.as<int64_t>(), nullptr, nullptr, nullptr);
        SB_LAUNCH(ctx, synth_scan, 1, 1024, 0, counts.as<int64_t>(), ncells, m->d_p);
        SB_CUDA(cudaMemcpyAsync(&m->nnz, m->d_p + ncells, 8, cudaMemcpyDeviceToHost, ctx->stream));
        SB_CUDA(cudaStreamSynchronize(ctx->stream));
        counts.release();
        SB_CUDA(cudaMalloc(&pp, std::max<size_t>(1, (size_t)m->nnz) * 4));
        m->d_i = (int32_t*)pp;
        SB_CUDA(cudaMalloc(&pp, std::max<size_t>(1, (size_t)m->nnz) * 8));
        m->d_x = (double*)pp;
        SB_LAUNCH(ctx, synth_kernel<true>, grid, 256, 0, sp, nullptr, m->d_p, m->d_i, m->d_x);
        SB_CUDA(cudaStreamSynchronize(ctx->stream));
        *out = m;
    } catch (const std::exception& e) {
        sb::set_last_error(e.what());
        return 1;
    }
    return 0;
}
