// A CTA-wide stable LSD radix sort of (uint64 key, payload) pairs. The ping-pong buffers may
// live in shared memory (small inputs) or in global memory (inputs beyond the shared-memory
// capacity; one CTA streams them through L2). Used for the value order of a gene's non-zeros
// inside a unit (K3) and for ranking effect columns across genes (K5).
//
// Stability per pass: every warp owns a contiguous slice of the input, walks it 32 elements at a
// time and ranks equal digits inside the 32 by lane order (match.any + popc); the per-warp digit
// histograms are scanned in (digit, warp) order. Passes whose digit is constant over the whole
// input are detected from the histogram and skipped.
#pragma once

#include <stdint.h>

namespace sb {

constexpr int SORT_THREADS = 1024;
constexpr int SORT_WARPS = SORT_THREADS / 32;
#ifndef SORT_UNROLL
#define SORT_UNROLL 1
#endif
constexpr int SORT_HIST_WORDS = SORT_WARPS * 256 + 64 + 1024;  // + aux words + scan exchange area

// whist: SORT_HIST_WORDS uint32 of shared memory. Returns 0 if the result is in (ka, pa), 1 if
// in (kb, pb). Must be called by all SORT_THREADS threads with identical arguments.
// Turns the per-warp digit histograms whist[warp][digit] into the position where each warp's
// first entry of each digit goes: exclusive scan in (digit, warp) order, by all 1024 threads
// (eight entries each). aux: >= 40 words after the histograms; aux[32] is set when a single
// digit holds all n entries. dbase (nullable, 257 words): where each digit starts (+ total).
// Ends with a barrier. The kernels that use it run 1024 threads = 32 warps.
__device__ __forceinline__ void scan_digit_histograms(uint32_t* whist, uint32_t* aux, int n, uint32_t* dbase,
                                                      uint32_t offset) {
    static_assert(SORT_THREADS == 1024, "the scan maps 8 (digit, warp) cells to each of 1024 threads");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // thread = (digit d, group of eight warps wg): consecutive lanes read consecutive digits (no bank conflicts)
    const int d = tid & 255, wg = tid >> 8;
    uint32_t* xch = aux + 64;  // [4][256] sums of the warp groups, then the digit bases in xch[0][*]
    uint32_t v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = whist[(wg * 8 + q) * 256 + d];
    uint32_t run = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const uint32_t t = v[q];
        v[q] = run;
        run += t;
    }
    xch[wg * 256 + d] = run;
    __syncthreads();
    uint32_t tot = 0, inc = 0;
    if (tid < 256) {  // one thread per digit: total, then an exclusive scan over the 256 digits
        tot = xch[d] + xch[256 + d] + xch[512 + d] + xch[768 + d];
        inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += y;
        }
        if (lane == 31) aux[warp] = inc;
        if (tot == (uint32_t)n) aux[32] = 1;
    }
    __syncthreads();
    if (tid < 256) {
        uint32_t wbase = 0;
        for (int w = 0; w < warp; ++w) wbase += aux[w];
        const uint32_t base = offset + wbase + inc - tot;
        if (dbase) {
            dbase[d] = base;
            if (d == 255) dbase[256] = base + tot;
        }
        // bases of the four warp groups of this digit
        const uint32_t s0 = xch[d], s1 = xch[256 + d], s2 = xch[512 + d];
        xch[d] = base;
        xch[256 + d] = base + s0;
        xch[512 + d] = base + s0 + s1;
        xch[768 + d] = base + s0 + s1 + s2;
    }
    __syncthreads();
    const uint32_t base = xch[wg * 256 + d];
#pragma unroll
    for (int q = 0; q < 8; ++q) whist[(wg * 8 + q) * 256 + d] = base + v[q];
    __syncthreads();
}

// Lanes with equal 8-bit digits, from eight ballots. `valid` lanes only; others get themselves.
__device__ __forceinline__ unsigned match_digit(unsigned d, bool valid, int lane) {
#ifdef SORT_USE_MATCH
    return __match_any_sync(0xffffffffu, valid ? d : 256u + (unsigned)lane);
#else
    unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return valid ? peers : (1u << lane);
#endif
}

template <typename P, bool FROM_PAYLOAD = false>
__device__ int block_radix_sort(uint64_t* ka, P* pa, uint64_t* kb, P* pb, int n, int bit_lo, int bit_hi,
                                uint32_t* whist, uint64_t sub = 0, uint32_t* dbase = nullptr) {
    constexpr int U = SORT_UNROLL;  // 32-element groups in flight per warp: their loads and matches overlap
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t* aux = whist + SORT_WARPS * 256;  // [0..7] warp sums, [32] skip flag
    const int per = (((n + SORT_WARPS - 1) / SORT_WARPS) + 31) & ~31;
    const int s_lo = min(n, warp * per), s_hi = min(n, s_lo + per);
    uint32_t* mine = whist + warp * 256;
    const unsigned lt = (1u << lane) - 1u;
    int cur = 0;
    for (int bit = bit_lo; bit < bit_hi; bit += 8) {
        uint64_t* sk = cur ? kb : ka;
        P* sp = cur ? pb : pa;
        uint64_t* dk = cur ? ka : kb;
        P* dp = cur ? pa : pb;
        for (int i = lane; i < 256; i += 32) mine[i] = 0;
        if (tid == 0) aux[32] = 0;
        __syncwarp();
        // ---- per-warp digit histogram of the warp's slice -----------------------------------
        for (int t = s_lo; t < s_hi; t += 32 * U) {
            unsigned d[U], peers[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = t + 32 * u + lane;
                // lanes past the end get a digit nobody else has, so that they match only themselves
                d[u] = 256u + (unsigned)lane;
                if (i < s_hi) d[u] = FROM_PAYLOAD ? ((unsigned)(sp[i] >> bit) & 255u) : ((unsigned)((sk[i] - sub) >> bit) & 255u);
            }
#ifdef SORT_HIST_BALLOT
#pragma unroll
            for (int u = 0; u < U; ++u) peers[u] = match_digit(d[u], d[u] < 256u, lane);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (d[u] < 256u && (peers[u] & lt) == 0u) mine[d[u]] += __popc(peers[u]);
                __syncwarp();
            }
#else
            // counting needs no order: shared-memory atomics on the warp's own counters
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (d[u] < 256u) atomicAdd(&mine[d[u]], 1u);
            }
            (void)peers;
#endif
        }
        __syncwarp();
        __syncthreads();
        scan_digit_histograms(whist, aux, dbase ? -1 : n, dbase, 0u);  // with dbase the pass is never skipped
        const bool skip = aux[32] != 0;
        if (!skip) {
            // ---- stable reorder: rank inside the 32 by lane order, base from the warp's counters ---
            for (int t = s_lo; t < s_hi; t += 32 * U) {
                uint64_t k[U];
                P pl[U];
                unsigned d[U], peers[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int i = t + 32 * u + lane;
                    d[u] = 256u + (unsigned)lane;
                    if (i < s_hi) {
                        k[u] = sk[i];
                        pl[u] = sp[i];
                        d[u] = FROM_PAYLOAD ? ((unsigned)(pl[u] >> bit) & 255u) : ((unsigned)((k[u] - sub) >> bit) & 255u);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) peers[u] = match_digit(d[u], d[u] < 256u, lane);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const bool live = d[u] < 256u;
                    const bool leader = (peers[u] & lt) == 0u;
                    uint32_t base = 0;
                    if (live && leader) {  // the first lane of every digit group advances the counter
                        base = mine[d[u]];
                        mine[d[u]] = base + __popc(peers[u]);
                    }
                    base = __shfl_sync(0xffffffffu, base, __ffs(peers[u]) - 1);
                    if (live) {
                        const uint32_t at = base + __popc(peers[u] & lt);
                        dk[at] = k[u];
                        dp[at] = pl[u];
                    }
                    __syncwarp();
                }
            }
            cur ^= 1;
        }
        __syncthreads();
    }
    return cur;
}

// After a sort on the high 32 key bits: order every run of equal high words by the full key,
// stably. Runs of up to 32 entries are insertion-sorted by the thread at their start. A longer
// run is fine as long as all its keys are equal (heavy ties); if one holds different keys,
// *overflow is raised and the caller falls back to sorting all eight digits.
template <typename P>
__device__ void fixup_low_bits(uint64_t* k, P* p, int n, uint32_t* overflow, uint64_t sub = 0, int rs = 32) {
    // a "run" = entries with equal (key - sub) >> rs, i.e. equal in the digits already sorted
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
        const uint64_t kj = k[j];
        const uint64_t h = (kj - sub) >> rs;
        if (j > 0 && ((k[j - 1] - sub) >> rs) == h) {
            // inside a run: entries 32 or more past the run start report unsorted neighbours
            if (j >= 32 && ((k[j - 32] - sub) >> rs) == h && k[j - 1] != kj) *overflow = 1;
            continue;
        }
        int e = j + 1;
        bool differ = false;
        while (e < n && e - j <= 32 && ((k[e] - sub) >> rs) == h) {
            differ |= k[e] != kj;
            ++e;
        }
        const int len = e - j;
        if (len > 32) {
            if (differ) *overflow = 1;
        } else if (len >= 2 && differ) {
            for (int a = j + 1; a < e; ++a) {
                const uint64_t ka = k[a];
                const P pa = p[a];
                int b = a - 1;
                while (b >= j && k[b] > ka) {
                    k[b + 1] = k[b];
                    p[b + 1] = p[b];
                    --b;
                }
                k[b + 1] = ka;
                p[b + 1] = pa;
            }
        }
    }
}

// Full 64-bit stable sort: high word first with the low-bit fix-up, falling back to all digits
// when the keys share long runs of equal high words. `top_bit`: the caller's bound on the highest
// key bit that differs between any two keys (64 = unknown); digits above it are not visited.
// Returns the buffer index (0 / 1).
template <typename P>
__device__ int block_sort_keys(uint64_t* ka, P* pa, uint64_t* kb, P* pb, int n, uint32_t* whist, int top_bit = 64,
                               uint64_t sub = 0) {
    // Orders by (key - sub); `top_bit` bounds the highest bit of (key - sub) that still has to be
    // sorted (bits above it are equal, or already in order and left alone by the stable passes).
    uint32_t* overflow = whist + SORT_WARPS * 256 + 33;
    const int bit_hi = min(64, (top_bit + 7) & ~7);
    if (bit_hi <= 32) {  // nothing to do above the low word: plain passes over the low digits
        const int c = block_radix_sort<P>(ka, pa, kb, pb, n, 0, bit_hi, whist, sub);
        __syncthreads();
        return c;
    }
    if (threadIdx.x == 0) *overflow = 0;
    int cur = block_radix_sort<P>(ka, pa, kb, pb, n, 32, bit_hi, whist, sub);
    uint64_t* k = cur ? kb : ka;
    P* p = cur ? pb : pa;
    fixup_low_bits<P>(k, p, n, overflow, sub, 32);
    __syncthreads();
    if (*overflow) {
        // (k, p) is the current input; the other pair is scratch.
        uint64_t* ok = cur ? ka : kb;
        P* op = cur ? pa : pb;
        const int c2 = block_radix_sort<P>(k, p, ok, op, n, 0, bit_hi, whist, sub);
        cur = c2 ? (cur ^ 1) : cur;
    }
    __syncthreads();
    return cur;
}

}  // namespace sb
