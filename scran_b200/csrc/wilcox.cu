// K3: exact pairwise Wilcoxon U statistics for every gene and every requested pair.
//
// Reference semantics: src/overlap_exprs.cpp:34-128 (class wilcoxer). There, every gene gathers
// the cells of every group from a dense row, sorts each group and runs one two-pointer sweep per
// pair: O(ncells) + O(nnz_gene * (G-1)) scalar steps per gene. Here the work is O(nnz) and the
// zeros are handled in closed form.
//
// A "unit" is one gene x one set of mutually compared combinations (for scoreMarkers: one gene
// in one block level). Its stored non-zero values, plus one "probe" entry at 0, are put in value
// order; what the count needs from that order is only the sequence of group labels with a
// "ties with the previous entry" bit: the 2-byte label stream.
//
//  k3_plan  one CTA: cuts the (gene, unit) list into jobs: runs of consecutive small units that
//           together fit the shared-memory sort, or one larger unit on its own.
//  k3_sort  one CTA per job, no radix passes on the main path: the values go to as many equal-width
//           VALUE buckets as there are elements (count, scan, scatter in shared memory), after
//           which an element ranks itself inside its small bucket by counting (which also yields
//           the tie bit) and writes its label to its final stream position; large buckets (many
//           equal values) are finished by one warp each. A unit beyond the shared-memory
//           capacity is first split by value into 256 coarse buckets in global scratch; its fine
//           buckets then follow the coarse counts. With lfc != 0 every value enters twice: as b
//           ("R role") and as a - lfc ("L role"), R before L inside a tie.
//  k3_walk  lane R keeps the running count c_R of label-R entries seen so far (cs_R = the count
//           at the start of the current tie run) and adds it into row label(i) of a
//           (G+1) x (G+1) table in shared memory while stepping through the stream: that yields
//           B[L][R] = sum over a in L of #(b in R : b < a) for all ordered pairs at once. The
//           probe row/column gives #(a > 0) and #(b < 0), so the implicit zeros enter as
//           Bfull[L][R] = B[L][R] + B[L][probe] z_R + z_L B[probe][R]  (z = zero count), and
//           2U = nL nR + Bfull[L][R] - Bfull[R][L] in int64, exact. (lfc != 0: the table
//           accumulates cs_R + c_R at L-role entries and
//           2U = B[L][R] + z_R B[L][p] + z_L B[p][R] + z_L z_R B[p][p].)
//           Small units are walked several at a time (one warp group each, k3_walk_small);
//           large ones by all warps of a CTA over position chunks into one shared table with
//           atomic adds (k3_walk), the chunks being combined through their label counts.
//  k3_ties  optional: sum over tie runs of t(t^2-1) per pair, accumulated in double in ascending
//           value order like src/overlap_exprs.cpp:121-124.
#include "kernels.h"
#include "sort.cuh"

namespace sb {


namespace {

constexpr uint32_t LAB_MASK = 0x3fff;
constexpr uint32_t ROLE_L = 0x4000;
constexpr uint32_t TIE_BIT = 0x8000;
constexpr int MAX_LABELS = 16383;
constexpr int MAXB = 256;      // units per job at most (also: the 256 value buckets of a large unit)
constexpr int MAXSEG = 4096;   // segment boundaries staged per job at most

// L2 eviction hints (createpolicy + .L2::cache_hint). The scratch of a unit beyond the shared-memory capacity is
// written once and read back twice within a few hundred microseconds: it should stay in L2 (evict_last) while the
// values that stream through the same L2 on their last read should go first (evict_first), or the scratch is
// pushed out to HBM and fetched back (round 1 / early round 2: 40 GB of the sort's DRAM traffic).
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void st_hint(uint64_t* a, uint64_t v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.b64 [%0], %1, %2;" ::"l"(a), "l"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_hint(uint32_t* a, uint32_t v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.b32 [%0], %1, %2;" ::"l"(a), "r"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ uint64_t ld_hint(const uint64_t* a, uint64_t pol) {
    uint64_t v;
    asm volatile("ld.global.L2::cache_hint.b64 %0, [%1], %2;" : "=l"(v) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ uint32_t ld_hint(const uint32_t* a, uint64_t pol) {
    uint32_t v;
    asm volatile("ld.global.L2::cache_hint.b32 %0, [%1], %2;" : "=r"(v) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ double ld_hint(const double* a, uint64_t pol) {
    double v;
    asm volatile("ld.global.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(a), "l"(pol));
    return v;
}

struct UnitTables {
    const int64_t* row_ptr;
    const uint32_t* seg_off;
    const int32_t* unit_first;
    const int32_t* unit_ncomb;
    const int64_t* ncells_dev;
    int ngenes, U, mult;  // mult = 1 (lfc == 0) or 2
};

struct UnitInfo {
    int64_t vbase, soff;
    uint32_t n, total;
    int first, G, g, u;
};

__device__ inline UnitInfo unit_info(const UnitTables& ut, int64_t task) {
    UnitInfo r;
    r.g = (int)(task / ut.U);
    r.u = (int)(task % ut.U);
    r.first = ut.unit_first[r.u];
    r.G = ut.unit_ncomb[r.u];
    const uint32_t a = ut.seg_off[(size_t)r.first * ut.ngenes + r.g];
    const uint32_t b = ut.seg_off[(size_t)(r.first + r.G) * ut.ngenes + r.g];
    r.vbase = ut.row_ptr[r.g] + a;
    r.n = b - a;
    r.soff = (int64_t)ut.mult * (r.vbase + task);
    r.total = (uint32_t)ut.mult * (r.n + 1);
    return r;
}

// ------------------------------------------------------------------------------------------
// k3_plan
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k3_plan(UnitTables ut, uint32_t small_max, uint32_t window, int maxb,
                                                 int32_t* jobs, int32_t* njobs) {
    __shared__ int part[1024];
    const int64_t ntasks = (int64_t)ut.ngenes * ut.U;
    const int t = threadIdx.x;
    const int64_t per = (ntasks + 1023) / 1024;
    const int64_t lo = min(ntasks, t * per), hi = min(ntasks, lo + per);
    auto is_start = [&](int64_t task, const UnitInfo& cur, const UnitInfo& prev) {
        if (task == 0) return true;
        if (cur.total > small_max || prev.total > small_max) return true;
        if (cur.soff / window != prev.soff / window) return true;
        return task / maxb != (task - 1) / maxb;
    };
    for (int phase = 0; phase < 2; ++phase) {
        int cnt = 0;
        UnitInfo prev = lo > 0 && lo < ntasks ? unit_info(ut, lo - 1) : UnitInfo();
        int at = 0;
        if (phase == 1) at = t == 0 ? 0 : part[t - 1];
        for (int64_t task = lo; task < hi; ++task) {
            const UnitInfo cur = unit_info(ut, task);
            if (is_start(task, cur, prev)) {
                if (phase == 1) jobs[at] = (int32_t)task;
                ++at;
                ++cnt;
            }
            prev = cur;
        }
        if (phase == 0) {
            part[t] = cnt;
            __syncthreads();
            for (int d = 1; d < 1024; d <<= 1) {
                const int v = t >= d ? part[t - d] : 0;
                __syncthreads();
                part[t] += v;
                __syncthreads();
            }
        }
    }
    if (t == 1023) {
        *njobs = part[1023];
        jobs[part[1023]] = (int32_t)ntasks;
    }
}

// Job order for the persistent kernels: the large single-unit jobs first and among them the larger
// size classes first (powers of two above `big_total`), so that the long jobs do not end up in the
// tail of the launch (longest-processing-time-first over ORDER_CLASSES + 1 classes).
constexpr int ORDER_CLASSES = 6;
__global__ void __launch_bounds__(1024) k3_order(UnitTables ut, const int32_t* jobs, const int32_t* njobs_ptr,
                                                  uint32_t big_total, int32_t* order, int2* order_range,
                                                  int32_t* nbig_out) {
    __shared__ int part[ORDER_CLASSES + 1][1024];
    __shared__ int class_base[ORDER_CLASSES + 2];
    const int njobs = *njobs_ptr;
    const int t = threadIdx.x;
    const int per = (njobs + 1023) / 1024;
    const int lo = min(njobs, t * per), hi = min(njobs, lo + per);
    auto class_of = [&](int j) {  // 0 = largest ... ORDER_CLASSES - 1 = just above big_total; ORDER_CLASSES = the rest
        if (jobs[j + 1] - jobs[j] != 1) return ORDER_CLASSES;
        const uint32_t total = unit_info(ut, jobs[j]).total;
        if (total <= big_total) return ORDER_CLASSES;
        const int above = 31 - __clz(total / big_total);  // floor(log2(total / big_total)) >= 0
        return ORDER_CLASSES - 1 - min(above, ORDER_CLASSES - 1);
    };
    int cnt[ORDER_CLASSES + 1];
#pragma unroll
    for (int c = 0; c <= ORDER_CLASSES; ++c) cnt[c] = 0;
    for (int j = lo; j < hi; ++j) {
        const int cl = class_of(j);
#pragma unroll
        for (int c = 0; c <= ORDER_CLASSES; ++c) cnt[c] += cl == c ? 1 : 0;
    }
#pragma unroll
    for (int c = 0; c <= ORDER_CLASSES; ++c) part[c][t] = cnt[c];
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        int v[ORDER_CLASSES + 1];
#pragma unroll
        for (int c = 0; c <= ORDER_CLASSES; ++c) v[c] = t >= d ? part[c][t - d] : 0;
        __syncthreads();
#pragma unroll
        for (int c = 0; c <= ORDER_CLASSES; ++c) part[c][t] += v[c];
        __syncthreads();
    }
    if (t == 0) {
        int run = 0;
        for (int c = 0; c <= ORDER_CLASSES; ++c) {
            class_base[c] = run;
            run += part[c][1023];
        }
        class_base[ORDER_CLASSES + 1] = run;
        *nbig_out = class_base[ORDER_CLASSES];
    }
    __syncthreads();
    int at[ORDER_CLASSES + 1];
#pragma unroll
    for (int c = 0; c <= ORDER_CLASSES; ++c) at[c] = class_base[c] + part[c][t] - cnt[c];
    for (int j = lo; j < hi; ++j) {
        const int cl = class_of(j);
#pragma unroll
        for (int c = 0; c <= ORDER_CLASSES; ++c)
            if (cl == c) {
                order_range[at[c]] = make_int2(jobs[j], jobs[j + 1]);  // what k3_sort needs of the job, in one load
                order[at[c]++] = j;
            }
    }
}

// ------------------------------------------------------------------------------------------
// k3_sort
// ------------------------------------------------------------------------------------------
struct SortParams {
    UnitTables ut;
    const double* vals;
    const double *vmin, *vmax;  // per (device combination, gene) value range from K2
    double shift;
    int cap, segcap, maxG;
    uint16_t* stream;
    uint32_t* unit_m;
    uint64_t *g_ka, *g_kb;   // global scratch of the units beyond cap: one region of `scratch_len` entries per CTA,
    uint32_t *g_pa, *g_pb;   // reused unit after unit (it stays in L2 instead of streaming through HBM)
    int64_t scratch_len;
    const int32_t* jobs;
    const int32_t* njobs;
    const int32_t* order;
    const int2* order_range;  // (first task, end task) of the jobs in `order`
    unsigned long long* job_counter;
};

// Shared memory of k3_sort. The big "region" of 12 bytes per element holds key[cap] (8 B) |
// lab[cap] (2 B) | bk[cap] (2 B): one 16-bit counter per bucket. The radix fallback for a value
// bucket beyond the capacity works in global memory and borrows the region for its histograms.
struct SortSmem {
    uint64_t* key;
    uint16_t* lab;
    uint32_t* bk;        // cap / 2 words: two 16-bit bucket counters a word
    uint32_t* whist;     // SORT_HIST_WORDS at the start of the region: radix passes of the fallback sort
    uint32_t* seg;       // segcap + MAXB + 8 staged segment boundaries
    uint32_t* segptr;    // MAXB + 1
    uint32_t* off;       // MAXB + 1: compact element offset of each segment of the job
    uint32_t* drop;      // MAXB: stored zeros of the segment (they are not part of the stream)
    uint32_t* n;         // MAXB
    int64_t* vbase;      // MAXB
    int64_t* soff;       // MAXB: stream position of the segment's first entry
    double* fmin;        // MAXB: value range of the segment (ValueMap)
    double* fmax;        // MAXB
    double* scale;       // MAXB
    uint32_t* boff;      // 258: bucket offsets of the value partition
    unsigned long long* bmin;  // 2: smallest / largest key of a bucket that takes the radix fallback
    uint32_t* gcur;      // 256: write cursor of every bucket in global scratch
    uint32_t* biglist;   // cap / BUCKET_RANK_MAX + 2: buckets too large for all-pairs ranking
    uint32_t* scan;      // 40: block scan exchange
    uint32_t* misc;      // 16
};

constexpr int BUCKET_RANK_MAX = 32;  // buckets up to this size are ranked by all-pairs counting

constexpr size_t sort_fixed_bytes(int cap, int segcap) {
    return (size_t)(segcap + MAXB + 8) * 4 + (size_t)(MAXB + 1) * 4 * 2 + (size_t)MAXB * 4 * 2 + (size_t)MAXB * 8 * 5 + 2 * 8 +
           258 * 4 + 256 * 4 + (size_t)(cap / BUCKET_RANK_MAX + 2) * 4 + 40 * 4 + 16 * 4 + 128;
}

__device__ inline SortSmem carve_sort_smem(unsigned char* raw, int cap, int segcap) {
    SortSmem s;
    s.key = reinterpret_cast<uint64_t*>(raw);
    s.lab = reinterpret_cast<uint16_t*>(s.key + cap);
    s.bk = reinterpret_cast<uint32_t*>(s.lab + cap);
    s.whist = reinterpret_cast<uint32_t*>(raw);
    unsigned char* p = raw + (size_t)cap * 12;
    s.vbase = reinterpret_cast<int64_t*>(p);
    s.soff = s.vbase + MAXB;
    s.fmin = reinterpret_cast<double*>(s.soff + MAXB);
    s.fmax = s.fmin + MAXB;
    s.scale = s.fmax + MAXB;
    s.bmin = reinterpret_cast<unsigned long long*>(s.scale + MAXB);
    s.seg = reinterpret_cast<uint32_t*>(s.bmin + 2);
    s.segptr = s.seg + (segcap + MAXB + 8);
    s.off = s.segptr + (MAXB + 1);
    s.drop = s.off + (MAXB + 1);
    s.n = s.drop + MAXB;
    s.boff = s.n + MAXB;
    s.gcur = s.boff + 258;
    s.biglist = s.gcur + 256;
    s.scan = s.biglist + (cap / BUCKET_RANK_MAX + 2);
    s.misc = s.scan + 40;
    return s;
}

// Key of element e of a unit's list, without its label (enough for the bucket counts). Dropped
// (stored zero) entries get key ~0.
__device__ inline void unit_key(const double* __restrict__ vals, int64_t vbase, uint32_t n, double shift, uint32_t e,
                                uint64_t& key, bool& keep) {
    const uint32_t per_role = n + 1;
    const bool role = e >= per_role;
    const uint32_t k = role ? e - per_role : e;
    double v = 0.0;
    keep = true;
    if (k < n) {
        v = vals[vbase + k];
        keep = v != 0.0;
    }
    if (role) v = v - shift;
    key = keep ? key_of(v) : ~0ull;
}

// Value bucket of a key: `nbuckets` buckets of equal width in VALUE space over [fmin, fmax] (equal
// width in key space would follow the exponent and put a whole binade range into one bucket).
// Monotone in the value, so the buckets are ordered and equal values share a bucket; values outside
// the range (the probes) clamp to the ends; dropped entries go to the last bucket, where their
// all-ones key sorts them last.
struct ValueMap {
    double fmin, fmax;  // range of the stored non-zero values (both roles), probes excluded
    double scale;       // (nbuckets - 2) / (fmax - fmin); 0 if there is one value only
};
// The first and the last bucket are kept for what lies outside the range: the probes (0, -lfc) and the
// dropped entries. Otherwise a probe would share the bucket of the smallest values, and with
// integer-valued data that bucket (thousands of copies of one value) would no longer be uniform.
__device__ __forceinline__ unsigned value_bucket(uint64_t key, bool keep, const ValueMap& vm, unsigned nbuckets) {
    if (!keep) return nbuckets - 1u;
    if (nbuckets < 3u) return 0u;
    const double v = value_of(key);
    if (v < vm.fmin) return 0u;
    if (v > vm.fmax) return nbuckets - 1u;
    return 1u + min(__double2uint_rz((v - vm.fmin) * vm.scale), nbuckets - 3u);  // the conversion saturates
}

// Segment of compact element / bucket index p: last s with off[s] <= p.
__device__ __forceinline__ int find_segment(const uint32_t* off, int nb, uint32_t p) {
    if (nb == 1) return 0;
    int lo = 0, hi = nb;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (off[mid] <= p) lo = mid; else hi = mid;
    }
    return lo;
}

// (key, role) order: inside a tie the R role (bit clear) comes before the L role.
__device__ __forceinline__ bool elem_before(uint64_t ka, uint32_t la, uint64_t kb, uint32_t lb) {
    return ka < kb || (ka == kb && (la & ROLE_L) < (lb & ROLE_L));
}

// In-place sort of sz (key, label) pairs in shared memory by one warp: a bitonic network whose
// merges start with a "flip" so that every compare-exchange is ascending; positions >= sz act as
// +infinity and are simply never touched.
__device__ inline void warp_bitonic(uint64_t* k, uint16_t* l, int sz, int lane) {
    int n2 = 2;
    while (n2 < sz) n2 <<= 1;
    auto cmpx = [&](int a, int b) {
        const uint64_t ka = k[a], kb = k[b];
        const uint16_t la = l[a], lb = l[b];
        if (elem_before(kb, lb, ka, la)) {
            k[a] = kb;
            k[b] = ka;
            l[a] = lb;
            l[b] = la;
        }
    };
    for (int span = 2; span <= n2; span <<= 1) {
        const int half = span >> 1;
        for (int i = lane; i < (n2 >> 1); i += 32) {
            const int blk = i / half, r = i - blk * half;
            const int a = blk * span + r, b = blk * span + span - 1 - r;
            if (b < sz) cmpx(a, b);
        }
        __syncwarp();
        for (int j = span >> 2; j > 0; j >>= 1) {
            for (int i = lane; i < (n2 >> 1); i += 32) {
                const int a = ((i / j) * 2 * j) + (i % j), b = a + j;
                if (b < sz) cmpx(a, b);
            }
            __syncwarp();
        }
    }
}

// Sorts the `nb` segments of a job (together `total` <= cap elements) and writes their label
// streams. No radix passes: an element goes straight to one of as many equal-width value buckets
// as its segment has elements (one counting pass, one scan, one scatter, all in shared memory),
// after which almost every bucket holds a handful of elements or copies of one value:
//  * buckets of up to BUCKET_RANK_MAX elements: every element counts the elements of its bucket
//    that precede it (all pairs) and so knows its final position and whether it ties with its
//    predecessor;
//  * larger buckets (many equal values, or a very uneven value distribution) go to a list and
//    are finished by one warp each: nothing to do if all keys are equal, else a bitonic sort.
// The bucket index space is the element index space: bucket off[s] + b of segment s.
// MODE 0: segments are whole units, elements are read from K1's values;
// MODE 1: one run of consecutive coarse buckets of ONE large unit, which the 256-way split left in
//         global memory (gk, gp: the run's first element). There are no segments: the fine
//         bucket of a value follows from the unit-wide coarse offsets (fine_bucket());
// MODE 2: one value range of ONE large unit, which the sample split left in global memory: as many
//         equal-width buckets over the range (run.vm) as it has entries; no dropped entries.
struct ScratchRun {
    uint32_t start;  // position of the run inside the unit (= its first fine bucket)
    uint32_t lim;    // entries of the run that belong to the stream (the stored zeros at the very end do not)
    int64_t out;     // stream position of the run's first entry
    ValueMap vm;     // the unit's coarse bucket map (256 buckets)
};

// Fine bucket of a key inside a large unit: coarse bucket d (256 of equal width over the unit's value
// range, the same expression as the split used) is divided into as many fine buckets as it has
// elements, so the fine buckets follow the density of the values and number exactly the unit's
// elements: fine = boff[d] + floor(frac * count_d). Monotone in the value.
__device__ __forceinline__ uint32_t fine_bucket(uint64_t key, const uint32_t* boff, const ValueMap& vm) {
    if (key == ~0ull) return boff[256] - 1u;  // stored zeros: the very last bucket
    const double v = value_of(key);
    uint32_t d = 0;
    double frac = 0.0;
    if (v > vm.fmax) {
        d = 255u;
    } else if (!(v < vm.fmin)) {  // the same expression as value_bucket(..., 256)
        const double x = (v - vm.fmin) * vm.scale;
        const uint32_t di = min(__double2uint_rz(x), 253u);
        d = 1u + di;
        frac = x - (double)di;
    }
    const uint32_t lo = boff[d], c = boff[d + 1] - lo;
    return lo + min(__double2uint_rz(frac * (double)c), c - 1u);
}

template <int MODE>
__device__ void bucket_sort_job(const SortParams& P, const SortSmem& S, int nb, uint32_t total, const uint64_t* gk,
                                const uint32_t* gp, const ScratchRun run) {
    constexpr bool SCRATCH = MODE != 0;
    auto scratch_bucket = [&](uint64_t key) -> uint32_t {
        if constexpr (MODE == 1) {
            return fine_bucket(key, S.boff, run.vm) - run.start;
        } else {
            return value_bucket(key, true, run.vm, total);
        }
    };
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool two_roles = P.ut.mult == 2;
    uint16_t* bh = reinterpret_cast<uint16_t*>(S.bk);
    // Every warp owns a contiguous chunk of the job's elements and goes through it unit by unit, 32
    // elements a step: what belongs to the unit (offsets, value range, group boundaries) is loaded
    // once and is the same for all lanes, and the group label of a lane's elements only ever moves
    // forward, so a cursor replaces the binary search.
    constexpr uint32_t nwarps = SORT_THREADS / 32;  // (the kernel is launched with SORT_THREADS threads)
    const uint32_t chunk = ((total + nwarps * 128u - 1u) / (nwarps * 128u)) * 128u;
    const uint32_t c_lo = min(total, (uint32_t)warp * chunk), c_hi = min(total, c_lo + chunk);

    // body(key, label | role, keep, bucket, segment) for every element of the warp's chunk
    auto for_each_element = [&](bool want_label, auto&& body) {
        if constexpr (SCRATCH) {
            const uint64_t pol_keep = l2_policy_evict_last();
            for (uint32_t t = c_lo + lane; t < c_hi; t += 128) {  // four loads in flight per thread
                uint64_t key[4];
                uint32_t lab[4];
                bool valid[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const uint32_t e = t + q * 32;
                    valid[q] = e < c_hi;
                    key[q] = valid[q] ? ld_hint(gk + e, pol_keep) : 0ull;
                    lab[q] = valid[q] && want_label ? (ld_hint(gp + e, pol_keep) & 0xffffu) : 0u;
                }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (valid[q]) body(key[q], lab[q], key[q] != ~0ull, scratch_bucket(key[q]), 0);
            }
        } else {
            if (c_lo >= c_hi) return;
            int s = find_segment(S.off, nb, c_lo);
            uint32_t pos = c_lo;
            while (pos < c_hi) {
                while (pos >= S.off[s + 1]) ++s;
                const uint32_t off_lo = S.off[s], off_hi = S.off[s + 1], cnt = off_hi - off_lo;
                const uint32_t n = S.n[s];
                const double* vals = P.vals + S.vbase[s];
                const uint32_t* segb = S.seg + S.segptr[s];
                const uint32_t G = S.segptr[s + 1] - S.segptr[s] - 1u;
                const ValueMap vm = {S.fmin[s], S.fmax[s], S.scale[s]};
                const uint32_t el_lo = pos - off_lo, el_hi = min(c_hi, off_hi) - off_lo;
                for (int role = 0; role < (two_roles ? 2 : 1); ++role) {
                    // this role's part of the unit's list: n values, then the probe
                    const uint32_t base = role ? n + 1u : 0u;
                    const uint32_t part_lo = max(el_lo, base), part_hi = min(el_hi, base + n + 1u);
                    if (part_lo >= part_hi) continue;
                    const uint32_t k_lo = part_lo - base, k_hi = min(part_hi - base, n);
                    const uint32_t rolebit = role ? ROLE_L : 0u;
                    uint32_t lab = 0, next_b = want_label && G > 1u ? segb[1] : 0xffffffffu;  // first position of the next group
                    for (uint32_t k = k_lo + lane; k < k_hi; k += 128) {  // four loads in flight per thread
                        double v[4];
                        bool ok[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            ok[q] = k + 32 * q < k_hi;
                            v[q] = ok[q] ? vals[k + 32 * q] : 1.0;
                        }
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            if (ok[q]) {
                                const uint32_t kk = k + 32 * q;
                                const bool keep = v[q] != 0.0;  // stored zeros count as zeros (src/overlap_exprs.cpp:43-47)
                                while (kk >= next_b) {  // label = last boundary <= kk
                                    ++lab;
                                    next_b = lab + 1u < G ? segb[lab + 1u] : 0xffffffffu;
                                }
                                const double vv = role ? v[q] - P.shift : v[q];
                                const uint64_t key = keep ? key_of(vv) : ~0ull;
                                body(key, lab | rolebit, keep, off_lo + value_bucket(key, keep, vm, cnt), s);
                            }
                        }
                    }
                    if (part_hi - base == n + 1u && lane == 0) {  // the probe: 0, or 0 - lfc in the L role
                        const double vv = role ? 0.0 - P.shift : 0.0;
                        const uint64_t key = key_of(vv);
                        body(key, G | rolebit, true, off_lo + value_bucket(key, true, vm, cnt), s);
                    }
                }
                pos = off_lo + el_hi;
            }
        }
    };

    // ---- counts ----
    for (uint32_t i = tid; i < ((total + 1) >> 1); i += blockDim.x) S.bk[i] = 0;
    if (tid == 0) S.misc[5] = 0;
    __syncthreads();
    for_each_element(false, [&](uint64_t, uint32_t, bool keep, uint32_t b, int s) {
        atomicAdd(&S.bk[b >> 1], 1u << ((b & 1u) * 16u));
        if (!SCRATCH && !keep) atomicAdd(&S.drop[s], 1u);
    });
    __syncthreads();

    // ---- exclusive scan of the counts (in place), large buckets noted ----
    {
        // a thread owns a multiple of 8 counters = whole 16-byte words (counters past `total` are zero
        // up to the end of the last 32-bit word; the words beyond that are never read)
        const uint32_t per = ((total + SORT_THREADS * 8u - 1u) / (SORT_THREADS * 8u)) * 8u;
        const uint32_t lo = min(total, (uint32_t)tid * per), hi = min(total, lo + per);
        uint32_t sum = 0;
        for (uint32_t i = lo; i < hi; i += 8) {
            if (i + 8 <= hi) {
                const uint4 w = *reinterpret_cast<const uint4*>(bh + i);
                sum += (w.x & 0xffffu) + (w.x >> 16) + (w.y & 0xffffu) + (w.y >> 16) + (w.z & 0xffffu) + (w.z >> 16) +
                       (w.w & 0xffffu) + (w.w >> 16);
            } else {
                for (uint32_t k = i; k < hi; ++k) sum += bh[k];
            }
        }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) S.scan[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t v = lane < (int)nwarps ? S.scan[lane] : 0u;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += y;
            }
            S.scan[lane] = v;
        }
        __syncthreads();
        uint32_t at = (warp ? S.scan[warp - 1] : 0u) + incl - sum;
        for (uint32_t i = lo; i < hi; i += 8) {
            if (i + 8 <= hi) {  // eight counters = one 16-byte word in, one out
                const uint4 w = *reinterpret_cast<const uint4*>(bh + i);
                const uint32_t in[4] = {w.x, w.y, w.z, w.w};
                uint32_t outw[4];
                bool big = false;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const uint32_t c0 = in[q] & 0xffffu, c1 = in[q] >> 16;
                    outw[q] = (at & 0xffffu) | ((at + c0) << 16);
                    at += c0 + c1;
                    big |= c0 > (uint32_t)BUCKET_RANK_MAX || c1 > (uint32_t)BUCKET_RANK_MAX;
                }
                *reinterpret_cast<uint4*>(bh + i) = make_uint4(outw[0], outw[1], outw[2], outw[3]);
                if (big) {
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const uint32_t c = (in[q >> 1] >> ((q & 1) * 16)) & 0xffffu;
                        if (c > (uint32_t)BUCKET_RANK_MAX) S.biglist[atomicAdd(&S.misc[5], 1u)] = i + q;
                    }
                }
            } else {
                for (uint32_t k = i; k < hi; ++k) {
                    const uint32_t c = bh[k];
                    bh[k] = (uint16_t)at;
                    at += c;
                    if (c > (uint32_t)BUCKET_RANK_MAX) S.biglist[atomicAdd(&S.misc[5], 1u)] = k;
                }
            }
        }
    }
    __syncthreads();

    // ---- scatter into bucket order: afterwards bh[b] is the END of bucket b ----
    for_each_element(true, [&](uint64_t key, uint32_t lab, bool, uint32_t b, int) {
        const uint32_t sh = (b & 1u) * 16u;
        const uint32_t pos = (atomicAdd(&S.bk[b >> 1], 1u << sh) >> sh) & 0xffffu;
        S.key[pos] = key;
        S.lab[pos] = (uint16_t)lab;
    });
    __syncthreads();

    // ---- small buckets: rank by counting, one thread per element ----
    auto rank_element = [&](uint32_t p, uint64_t key, uint32_t b, uint32_t seg_lo, uint32_t lim, int64_t out) {
        const uint32_t start = b ? (uint32_t)bh[b - 1] : 0u, end = bh[b];
        if (end - start > (uint32_t)BUCKET_RANK_MAX) return;
        const uint32_t lp = S.lab[p];
        uint32_t rank = 0;
        bool tie = false;
        if (end - start > 1) {
            if (two_roles) {
#pragma unroll 1
                for (uint32_t j = start; j < end; ++j) {
                    const uint64_t kj = S.key[j];
                    const uint32_t lj = S.lab[j];
                    const bool eq = kj == key;
                    const bool before = kj < key || (eq && ((lj & ROLE_L) < (lp & ROLE_L) || ((lj & ROLE_L) == (lp & ROLE_L) && j < p)));
                    rank += before ? 1u : 0u;
                    tie |= eq && before;
                }
            } else {
#pragma unroll 1
                for (uint32_t j = start; j < end; ++j) {
                    const uint64_t kj = S.key[j];
                    const bool eq = kj == key;
                    const bool before = kj < key || (eq && j < p);
                    rank += before ? 1u : 0u;
                    tie |= eq && before;
                }
            }
        }
        const uint32_t q = start + rank - seg_lo;
        if (q < lim) P.stream[out + q] = (uint16_t)((lp & 0x7fffu) | (tie ? TIE_BIT : 0u));
    };
    if constexpr (SCRATCH) {
        for (uint32_t p = c_lo + lane; p < c_hi; p += 32) {
            const uint64_t key = S.key[p];
            rank_element(p, key, scratch_bucket(key), 0u, run.lim, run.out);
        }
    } else if (c_lo < c_hi) {
        int s = find_segment(S.off, nb, c_lo);
        uint32_t pos = c_lo;
        while (pos < c_hi) {
            while (pos >= S.off[s + 1]) ++s;
            const uint32_t off_lo = S.off[s], off_hi = S.off[s + 1], cnt = off_hi - off_lo;
            const ValueMap vm = {S.fmin[s], S.fmax[s], S.scale[s]};
            const uint32_t lim = cnt - S.drop[s];
            const int64_t out = S.soff[s];
            const uint32_t hi = min(c_hi, off_hi);
            for (uint32_t p = pos + lane; p < hi; p += 32) {
                const uint64_t key = S.key[p];
                rank_element(p, key, off_lo + value_bucket(key, key != ~0ull, vm, cnt), off_lo, lim, out);
            }
            pos = hi;
        }
    }

    // ---- large buckets: one warp each ----
    const uint32_t nbig = S.misc[5];
    for (uint32_t w = warp; w < nbig; w += blockDim.x / 32) {
        const uint32_t b = S.biglist[w];
        const uint32_t start = b ? (uint32_t)bh[b - 1] : 0u, end = bh[b];
        const int sz = (int)(end - start);
        const int s = SCRATCH ? 0 : find_segment(S.off, nb, b);
        const uint32_t lim = SCRATCH ? run.lim : (S.off[s + 1] - S.off[s]) - S.drop[s];
        const int64_t out = SCRATCH ? run.out : S.soff[s];
        uint64_t* k = S.key + start;
        uint16_t* l = S.lab + start;
        const uint64_t k0 = k[0];
        bool same = true;
        for (int i = lane; i < sz; i += 32) same &= k[i] == k0;
        same = __all_sync(0xffffffffu, same);
        const uint32_t base = start - (SCRATCH ? 0u : S.off[s]);
        if (same && two_roles) {
            // one value in both roles: R-role entries first, each role in the order found
            uint32_t n_r = 0;
            for (int i = lane; i < sz; i += 32) n_r += (l[i] & ROLE_L) ? 0u : 1u;
            for (int o = 16; o > 0; o >>= 1) n_r += __shfl_xor_sync(0xffffffffu, n_r, o);
            uint32_t seen_r = 0, seen_l = 0;
            for (int i0 = 0; i0 < sz; i0 += 32) {
                const int i = i0 + lane;
                const bool valid = i < sz;
                const uint32_t lab = valid ? (uint32_t)l[i] : 0u;
                const bool is_l = valid && (lab & ROLE_L);
                const unsigned ml = __ballot_sync(0xffffffffu, is_l), mr = __ballot_sync(0xffffffffu, valid && !is_l);
                const unsigned below = (1u << lane) - 1u;
                if (valid) {
                    const uint32_t q = is_l ? n_r + seen_l + __popc(ml & below) : seen_r + __popc(mr & below);
                    if (base + q < lim) P.stream[out + base + q] = (uint16_t)((lab & 0x7fffu) | (q > 0 ? TIE_BIT : 0u));
                }
                seen_l += __popc(ml);
                seen_r += __popc(mr);
            }
            continue;
        }
        if (!same) warp_bitonic(k, l, sz, lane);
        for (int i = lane; i < sz; i += 32) {
            if (base + i < lim) {
                const bool tie = i > 0 && k[i] == k[i - 1];
                P.stream[out + base + i] = (uint16_t)(((uint32_t)l[i] & 0x7fffu) | (tie ? TIE_BIT : 0u));
            }
        }
    }
}

// A unit beyond the shared-memory capacity, first choice: split it by value into R ranges of about
// three quarters of the capacity each, with ONE pass over its values. The range boundaries are the
// quantiles of a sample (SPLIT_SAMPLES per range, taken at equal strides over the unit, so every group
// is represented in proportion), read off a histogram of the sample over SPLIT_BINS equal-width value
// bins: a boundary is the lower edge of a bin; an entry finds its range by binary search over the R - 1 splitter
// keys (integers: no floating point), a warp appends what it has for a range with one shared-memory
// atomic, and every range has room for `cap` entries in the CTA's scratch. Each range is then sorted in
// shared memory like a small unit (as many equal-width value buckets as it has entries). Equal keys
// share a range, so ties never straddle one. Returns false - and has written nothing that matters - if
// some range came out larger than the capacity (a heavily tied value, or an unlucky sample): the caller
// then takes the exact 256-way split below.
constexpr uint32_t SPLIT_SAMPLES = 512;    // per range; the counts of the ranges then vary by about 1/22
constexpr uint32_t SPLIT_BINS = 8192;      // value bins of the sample histogram
constexpr uint32_t SPLIT_MAX_RANGES = 64;

__device__ bool sample_split_unit(const SortParams& P, const SortSmem& S, uint32_t total, int64_t vbase, uint32_t n,
                                  int G, int64_t soff, uint64_t* ga, uint32_t* gp, double fmin, double fmax, uint32_t R,
                                  uint32_t* m_out) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t cap = (uint32_t)P.cap;
    uint32_t Rp = 2;
    while (Rp < R) Rp <<= 1;
    const uint32_t ns = Rp * SPLIT_SAMPLES;
    uint64_t* sp = reinterpret_cast<uint64_t*>(S.scale);     // splitters: sp[j] = smallest key of range j (j >= 1)

    // ---- histogram of the sample over SPLIT_BINS equal-width value bins, prefix sums, quantiles ----
    uint16_t* bh = reinterpret_cast<uint16_t*>(S.bk);
    const ValueMap vs = {fmin, fmax, (double)(SPLIT_BINS - 2u) / (fmax - fmin)};
    for (uint32_t i = tid; i < SPLIT_BINS / 2u; i += blockDim.x) S.bk[i] = 0;
    __syncthreads();
    for (uint32_t i = tid; i < ns; i += blockDim.x) {
        uint64_t key;
        bool keep;
        unit_key(P.vals, vbase, n, P.shift, (uint32_t)(((uint64_t)i * total) / ns), key, keep);
        if (keep) {  // stored zeros are not part of the stream
            const uint32_t b = value_bucket(key, true, vs, SPLIT_BINS);
            atomicAdd(&S.bk[b >> 1], 1u << ((b & 1u) * 16u));
        }
    }
    __syncthreads();
    {  // exclusive prefix sums in place: eight bins per thread
        static_assert(SPLIT_BINS == 8u * SORT_THREADS, "one 16-byte word of bins per thread");
        const uint4 w = *reinterpret_cast<const uint4*>(bh + tid * 8);
        const uint32_t in[4] = {w.x, w.y, w.z, w.w};
        uint32_t sum = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) sum += (in[q] & 0xffffu) + (in[q] >> 16);
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) S.scan[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t v = S.scan[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += y;
            }
            S.scan[lane] = v;
        }
        __syncthreads();
        uint32_t at = (warp ? S.scan[warp - 1] : 0u) + incl - sum;
        uint32_t outw[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t c0 = in[q] & 0xffffu, c1 = in[q] >> 16;
            outw[q] = (at & 0xffffu) | ((at + c0) << 16);
            at += c0 + c1;
        }
        *reinterpret_cast<uint4*>(bh + tid * 8) = make_uint4(outw[0], outw[1], outw[2], outw[3]);
    }
    const uint32_t nv = S.scan[31];  // sampled entries that are part of the stream
    __syncthreads();
    if (nv < 8u * R) return false;
    if (tid < 256) S.gcur[tid] = 0;
    if ((uint32_t)tid < Rp) {
        // splitter j = lower edge of the bin that holds the sample of rank j nv / R: the last bin whose prefix is <= that rank
        uint64_t key = tid == 0 ? 0ull : ~0ull;
        if (tid > 0 && (uint32_t)tid < R) {
            const uint32_t want = (uint32_t)(((uint64_t)tid * nv) / R);
            uint32_t lo = 0, hi = SPLIT_BINS;
            while (hi - lo > 1u) {
                const uint32_t mid = (lo + hi) >> 1;
                if ((uint32_t)bh[mid] <= want) lo = mid; else hi = mid;
            }
            key = lo == 0 ? 0ull : key_of(fmin + (double)(lo - 1u) / vs.scale);
        }
        sp[tid] = key;
    }
    __syncthreads();

    // ---- the one pass: every entry to the next free place of its range ----
    {
#ifndef SB_SPLIT_LOADS
#define SB_SPLIT_LOADS 4
#endif
        constexpr int U = SB_SPLIT_LOADS;  // loads in flight per thread
        const uint64_t pol_keep = l2_policy_evict_last();
#ifdef SB_SPLIT_EVICT_FIRST
        const uint64_t pol_stream = l2_policy_evict_first();
#endif
        constexpr uint32_t nwarps = SORT_THREADS / 32;
        const uint32_t chunk = ((total + nwarps * 32u * U - 1u) / (nwarps * 32u * U)) * 32u * U;
        const uint32_t c_lo = min(total, (uint32_t)warp * chunk), c_hi = min(total, c_lo + chunk);
        const unsigned below = (1u << lane) - 1u;
        const bool two_roles = P.ut.mult == 2;
        // group label of the lane's entries: moves forward only; next_b = first position of the next group
        uint32_t cur_lab = 0, next_b = G > 1 ? S.seg[1] : 0xffffffffu;
        bool cur_role = false;
        for (uint32_t t = c_lo; t < c_hi; t += 32u * U) {  // the whole warp takes every step: match_any needs all lanes
            double v[U];
#pragma unroll
            for (int q = 0; q < U; ++q) {
                const uint32_t e = t + lane + q * 32;
                const uint32_t k = e > n ? e - (n + 1u) : e;
#ifdef SB_SPLIT_EVICT_FIRST
                v[q] = e < c_hi && k < n ? ld_hint(P.vals + vbase + k, pol_stream) : 1.0;
#else
                v[q] = e < c_hi && k < n ? P.vals[vbase + k] : 1.0;  // (the probes are not stored zeros)
#endif
            }
#pragma unroll
            for (int q = 0; q < U; ++q) {
                const uint32_t e = t + lane + q * 32;
                const bool valid = e < c_hi;
                const bool role = e > n;
                const uint32_t k = role ? e - (n + 1u) : e;
                uint32_t lab = (uint32_t)G;
                double x = 0.0;
                if (two_roles && valid && role != cur_role) {
                    cur_role = role;
                    cur_lab = 0;
                    next_b = G > 1 ? S.seg[1] : 0xffffffffu;
                }
                if (valid && k < n) {
                    while (k >= next_b) {
                        ++cur_lab;
                        next_b = cur_lab + 1u < (uint32_t)G ? S.seg[cur_lab + 1u] : 0xffffffffu;
                    }
                    lab = cur_lab;
                    x = v[q];
                }
                const bool keep = valid && v[q] != 0.0;  // stored zeros are not part of the stream
                if (role) {
                    x = x - P.shift;
                    lab |= ROLE_L;
                }
                const uint64_t key = key_of(x);
                uint32_t r = 0xffffffffu;
                if (keep) {
                    r = 0;
                    for (uint32_t step = Rp >> 1; step > 0; step >>= 1)
                        if (sp[r + step] <= key) r += step;
                    r = min(r, R - 1u);
                }
                const unsigned peers = __match_any_sync(0xffffffffu, r);
                const int leader = __ffs(peers) - 1;
                uint32_t base = 0;
                if (lane == leader && keep) base = atomicAdd(&S.gcur[r], (uint32_t)__popc(peers));
                base = __shfl_sync(0xffffffffu, base, leader);
                const uint32_t pos = base + (uint32_t)__popc(peers & below);
                if (keep && pos < cap) {
                    st_hint(ga + (size_t)r * cap + pos, key, pol_keep);
                    st_hint(gp + (size_t)r * cap + pos, lab, pol_keep);
                }
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t run = 0, ok = 1;
        for (uint32_t r = 0; r < R; ++r) {
            const uint32_t c = S.gcur[r];
            S.boff[r] = run;
            run += c;
            ok &= c <= cap ? 1u : 0u;
        }
        S.boff[R] = run;
        S.misc[7] = ok;
    }
    __syncthreads();
    if (!S.misc[7]) return false;
    *m_out = S.boff[R];

    // ---- every range through shared memory ----
    for (uint32_t r = 0; r < R; ++r) {
        const uint32_t cnt = S.gcur[r];
        if (cnt == 0) continue;
        const double lo = r == 0 ? fmin : value_of(sp[r]), hi = r + 1u == R ? fmax : value_of(sp[r + 1u]);
        ScratchRun run;
        run.start = 0;
        run.lim = cnt;
        run.out = soff + S.boff[r];
        run.vm = {lo, hi, hi > lo && cnt >= 3u ? (double)(cnt - 2u) / (hi - lo) : 0.0};
        __syncthreads();
        bucket_sort_job<2>(P, S, 1, cnt, ga + (size_t)r * cap, gp + (size_t)r * cap, run);
    }
    return true;
}

#ifndef SB_SORT_JOB_BATCH
#define SB_SORT_JOB_BATCH 4
#endif
constexpr int SORT_JOB_BATCH = SB_SORT_JOB_BATCH;

__global__ void __launch_bounds__(SORT_THREADS) k3_sort(SortParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const SortSmem S = carve_sort_smem(smem_raw, P.cap, P.segcap);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int njobs = *P.njobs;
    const UnitTables& ut = P.ut;
    const double inf = __longlong_as_double(0x7ff0000000000000ll);

    // Jobs are taken SORT_JOB_BATCH at a time (one round trip to the global counter per batch), and a job's
    // task range comes with its place in the order (one load instead of order -> jobs).
    if (tid == 0) S.misc[8] = S.misc[9] = 0;  // next job of the batch, end of the batch
    while (true) {
        __syncthreads();
        if (tid == 0) {
            uint32_t b = S.misc[8], e = S.misc[9];
            if (b == e) {
                b = (uint32_t)min(atomicAdd(P.job_counter, (unsigned long long)SORT_JOB_BATCH), (unsigned long long)njobs);
                e = min(b + (uint32_t)SORT_JOB_BATCH, (uint32_t)njobs);
                S.misc[9] = e;
            }
            S.misc[0] = b;
            S.misc[8] = b < e ? b + 1u : b;
        }
        __syncthreads();
        if ((int)S.misc[0] >= njobs) break;
        const int2 jr = P.order_range[S.misc[0]];
        const int64_t t0 = jr.x, t1 = jr.y;
        const int nb = (int)(t1 - t0);

        // ---- per-unit tables -------------------------------------------------------------
        if (tid < nb) {
            const UnitInfo ui = unit_info(ut, t0 + tid);
            S.vbase[tid] = ui.vbase;
            S.soff[tid] = ui.soff;
            S.n[tid] = ui.n;
            S.drop[tid] = 0;
            S.off[tid + 1] = ui.total;
            S.segptr[tid + 1] = (uint32_t)ui.G + 1;
        }
        if (tid == 0) {
            S.off[0] = 0;
            S.segptr[0] = 0;
        }
        __syncthreads();
        if (warp == 0) {  // inclusive scans of off[1..nb] and segptr[1..nb] by one warp
            uint32_t run_a = 0, run_b = 0;
            for (int base = 0; base < nb; base += 32) {
                const int i = base + lane;
                uint32_t a = i < nb ? S.off[i + 1] : 0u, b = i < nb ? S.segptr[i + 1] : 0u;
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t ya = __shfl_up_sync(0xffffffffu, a, o), yb = __shfl_up_sync(0xffffffffu, b, o);
                    if (lane >= o) {
                        a += ya;
                        b += yb;
                    }
                }
                if (i < nb) {
                    S.off[i + 1] = run_a + a;
                    S.segptr[i + 1] = run_b + b;
                }
                run_a += __shfl_sync(0xffffffffu, a, 31);
                run_b += __shfl_sync(0xffffffffu, b, 31);
            }
        }
        __syncthreads();
        const uint32_t total = S.off[nb];
        const uint32_t nseg = S.segptr[nb];
        for (uint32_t e = tid; e < nseg; e += blockDim.x) {  // stage the segment boundaries
            int lo = 0, hi = nb;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (S.segptr[mid] <= e) lo = mid; else hi = mid;
            }
            const int64_t task = t0 + lo;
            const int g = (int)(task / ut.U), u = (int)(task % ut.U);
            const int first = ut.unit_first[u];
            const uint32_t seg0 = ut.seg_off[(size_t)first * ut.ngenes + g];
            S.seg[e] = ut.seg_off[(size_t)(first + (e - S.segptr[lo])) * ut.ngenes + g] - seg0;
        }
        // value range of every unit (both roles, probes excluded), from the per-segment ranges that
        // K2 left behind: no pass over the values
        for (int s = warp; s < nb; s += blockDim.x / 32) {
            const int64_t task = t0 + s;
            const int gene = (int)(task / ut.U), first = ut.unit_first[(int)(task % ut.U)];
            const int G = (int)(S.segptr[s + 1] - S.segptr[s]) - 1;
            double fmin = inf, fmax = -inf;
            for (int k = lane; k < G; k += 32) {
                const double lo_v = P.vmin[(size_t)(first + k) * ut.ngenes + gene];
                const double hi_v = P.vmax[(size_t)(first + k) * ut.ngenes + gene];
                if (lo_v <= hi_v) {  // the segment has non-zero values
                    fmin = fmin < lo_v ? fmin : lo_v;
                    fmax = fmax > hi_v ? fmax : hi_v;
                    if (ut.mult == 2) {
                        fmin = fmin < lo_v - P.shift ? fmin : lo_v - P.shift;
                        fmax = fmax > hi_v - P.shift ? fmax : hi_v - P.shift;
                    }
                }
            }
            for (int o = 16; o > 0; o >>= 1) {
                const double a = __shfl_xor_sync(0xffffffffu, fmin, o), b = __shfl_xor_sync(0xffffffffu, fmax, o);
                fmin = fmin < a ? fmin : a;
                fmax = fmax > b ? fmax : b;
            }
            if (lane == 0) {
                const uint32_t cnt = S.off[s + 1] - S.off[s];
                if (!(fmin <= fmax)) fmin = fmax = 0.0;  // no stored non-zero value at all
                S.fmin[s] = fmin;
                S.fmax[s] = fmax;
                S.scale[s] = fmax > fmin && cnt >= 3u ? (double)(cnt - 2u) / (fmax - fmin) : 0.0;
                if (nb == 1) {  // the plain range, for the 256-way split of a unit beyond the capacity
                    double* red = reinterpret_cast<double*>(S.scan);
                    red[0] = fmin;
                    red[1] = fmax;
                }
            }
        }
        __syncthreads();

        if (total <= (uint32_t)P.cap) {
            // ---- shared-memory path: all units of the job at once -----------------------------
            bucket_sort_job<0>(P, S, nb, total, nullptr, nullptr, ScratchRun());
            if (tid < nb) P.unit_m[t0 + tid] = (S.off[tid + 1] - S.off[tid]) - S.drop[tid];
            continue;
        }

        // ---- one unit beyond the shared-memory capacity ------------------------------------
        // (nb == 1 by construction.) Split by value into 256 ordered buckets with one stable pass
        // into global scratch; consecutive buckets that fit shared memory together are then sorted
        // there, every bucket as a segment with its own value range.
        const int64_t vbase = S.vbase[0], soff = S.soff[0];
        const uint32_t n = S.n[0];
        const int G = (int)nseg - 1;
        uint64_t* ga = P.g_ka + (size_t)blockIdx.x * P.scratch_len;
        uint32_t* gp = P.g_pa + (size_t)blockIdx.x * P.scratch_len;
        uint64_t* gb = P.g_kb + (size_t)blockIdx.x * P.scratch_len;
        uint32_t* gq = P.g_pb + (size_t)blockIdx.x * P.scratch_len;
        const double fmin = reinterpret_cast<double*>(S.scan)[0], fmax = reinterpret_cast<double*>(S.scan)[1];
        {
#ifndef SB_SPLIT_SLACK_DIV
#define SB_SPLIT_SLACK_DIV 4
#endif
            const uint32_t per = (uint32_t)P.cap - (uint32_t)P.cap / (unsigned)SB_SPLIT_SLACK_DIV;
            const uint32_t R = (total + per - 1u) / per;
            uint32_t m_split = 0;
            if (R <= SPLIT_MAX_RANGES && (int64_t)R * P.cap <= P.scratch_len && fmax > fmin && (uint32_t)P.cap >= SPLIT_BINS &&
                sample_split_unit(P, S, total, vbase, n, G, soff, ga, gp, fmin, fmax, R, &m_split)) {
                if (tid == 0) P.unit_m[t0] = m_split;
                continue;
            }
            __syncthreads();
        }
        const ValueMap vm256 = {fmin, fmax, fmax > fmin ? 254.0 / (fmax - fmin) : 0.0};
        if (tid == 0) S.misc[3] = 0;  // dropped (stored zero) entries, counted in pass 1
        __syncthreads();

        // pass 1: the unit-wide bucket counts, then every element straight to the next free place of
        // its bucket's region (the order inside a bucket is free; each region is filled front to back
        // within microseconds, so L2 collects the 12-byte writes into whole lines)
        {
            for (int i = tid; i < 256; i += blockDim.x) S.gcur[i] = 0;
            __syncthreads();
            uint32_t ndrop = 0;
            for (uint32_t t = tid; t < total; t += 4 * blockDim.x) {  // four loads in flight per thread
                uint64_t key[4];
                bool keep[4], valid[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const uint32_t e = t + q * blockDim.x;
                    valid[q] = e < total;
                    key[q] = 0;
                    keep[q] = false;
                    if (valid[q]) unit_key(P.vals, vbase, n, P.shift, e, key[q], keep[q]);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (valid[q]) {
                        const unsigned d = value_bucket(key[q], keep[q], vm256, 256u);
                        atomicAdd(&S.gcur[d], 1u);
                        ndrop += keep[q] ? 0u : 1u;
                    }
                }
            }
            for (int o = 16; o > 0; o >>= 1) ndrop += __shfl_xor_sync(0xffffffffu, ndrop, o);
            if (lane == 0 && ndrop) atomicAdd(&S.misc[3], ndrop);
            __syncthreads();
            if (warp == 0) {  // exclusive scan of the 256 bucket counts: eight per lane
                uint32_t c[8], run = 0;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    c[q] = S.gcur[lane * 8 + q];
                    run += c[q];
                }
                uint32_t inc = run;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += y;
                }
                uint32_t at = inc - run;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    S.boff[lane * 8 + q] = at;
                    S.gcur[lane * 8 + q] = at;
                    at += c[q];
                }
                if (lane == 31) S.boff[256] = at;
            }
            __syncthreads();
            {  // warp chunks, so that the group label of a lane's elements only moves forward
                const uint64_t pol_keep = l2_policy_evict_last(), pol_stream = l2_policy_evict_first();
                const uint32_t nwarps = blockDim.x / 32;
                const uint32_t chunk = ((total + nwarps * 128u - 1u) / (nwarps * 128u)) * 128u;
                const uint32_t c_lo = min(total, (uint32_t)warp * chunk), c_hi = min(total, c_lo + chunk);
                uint32_t cur_lab = 0;
                bool cur_role = false;
                for (uint32_t t = c_lo + lane; t < c_hi; t += 128) {
                    uint64_t key[4];
                    uint32_t lab[4];
                    bool keep[4], valid[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t e = t + q * 32;
                        valid[q] = e < c_hi;
                        key[q] = 0;
                        lab[q] = 0;
                        keep[q] = false;
                        if (valid[q]) {
                            const bool role = e > n;
                            const uint32_t k = role ? e - (n + 1u) : e;
                            if (role != cur_role) {
                                cur_role = role;
                                cur_lab = 0;
                            }
                            double v = 0.0;
                            keep[q] = true;
                            lab[q] = (uint32_t)G;
                            if (k < n) {
                                v = ld_hint(P.vals + vbase + k, pol_stream);  // the last read of these values
                                keep[q] = v != 0.0;
                                while (cur_lab + 1u < (uint32_t)G && S.seg[cur_lab + 1u] <= k) ++cur_lab;
                                lab[q] = cur_lab;
                            }
                            if (role) {
                                v = v - P.shift;
                                lab[q] |= ROLE_L;
                            }
                            key[q] = keep[q] ? key_of(v) : ~0ull;
                        }
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (valid[q]) {
                            const unsigned d = value_bucket(key[q], keep[q], vm256, 256u);
                            const uint32_t dst = atomicAdd(&S.gcur[d], 1u);
                            st_hint(ga + dst, key[q], pol_keep);
                            st_hint(gp + dst, lab[q], pol_keep);
                        }
                    }
                }
            }
            __syncthreads();
        }
        const uint32_t ndropped = S.misc[3];
        const uint32_t m = total - ndropped;

        // pass 2: runs of consecutive buckets that fit shared memory
        int d0 = 0;
        while (d0 < 256) {
            __syncthreads();
            if (tid == 0) {  // extend the run of buckets while it fits
                const uint32_t st0 = S.boff[d0];
                int d = d0 + 1;
                while (d < 256 && S.boff[d + 1] - st0 <= (uint32_t)P.cap) ++d;
                S.misc[1] = (uint32_t)d;
            }
            __syncthreads();
            const uint32_t start = S.boff[d0];
            const int d1 = (int)S.misc[1];
            const uint32_t cnt = S.boff[d1] - start;
            d0 = d1;
            if (cnt == 0) continue;
            if (cnt <= (uint32_t)P.cap) {
                ScratchRun run;
                run.start = start;
                run.lim = m > start ? min(cnt, m - start) : 0u;
                run.out = soff + start;
                run.vm = vm256;
                bucket_sort_job<1>(P, S, 1, cnt, ga + start, gp + start, run);
                continue;
            }
            // a single coarse bucket larger than shared memory (e.g. one heavily tied value): radix sort
            // in global memory, skipping the digits above the highest bit in which the smallest and the
            // largest key differ (dropped entries, whose key is not a value, switch that off)
            if (tid < 2) S.bmin[tid] = tid == 0 ? ~0ull : 0ull;
            __syncthreads();
            {
                unsigned long long lo_k = ~0ull, hi_k = 0ull;
                for (uint32_t j = tid; j < cnt; j += blockDim.x) {
                    const unsigned long long k = ga[start + j];
                    lo_k = k < lo_k ? k : lo_k;
                    hi_k = k > hi_k ? k : hi_k;
                }
                for (int o = 16; o > 0; o >>= 1) {
                    const unsigned long long a = __shfl_xor_sync(0xffffffffu, lo_k, o), b = __shfl_xor_sync(0xffffffffu, hi_k, o);
                    lo_k = a < lo_k ? a : lo_k;
                    hi_k = b > hi_k ? b : hi_k;
                }
                if (lane == 0) {
                    atomicMin(&S.bmin[0], lo_k);
                    atomicMax(&S.bmin[1], hi_k);
                }
            }
            __syncthreads();
            int top_bit = 64;
            {
                const unsigned long long lo_k = S.bmin[0], hi_k = S.bmin[1];
                if (hi_k != ~0ull) top_bit = lo_k == hi_k ? 0 : 64 - __clzll((long long)(lo_k ^ hi_k));
            }
            __syncthreads();
            if (S.bmin[0] == S.bmin[1]) {
                // one value only (integer-valued data): any order inside a role will do, R role first
                if (tid < 2) S.misc[6 + tid] = 0;
                __syncthreads();
                uint32_t n_r = 0;
                for (uint32_t j = tid; j < cnt; j += blockDim.x) n_r += (gp[start + j] & ROLE_L) ? 0u : 1u;
                for (int o = 16; o > 0; o >>= 1) n_r += __shfl_xor_sync(0xffffffffu, n_r, o);
                if (lane == 0 && n_r) atomicAdd(&S.misc[6], n_r);
                __syncthreads();
                n_r = S.misc[6];
                __syncthreads();
                if (tid == 0) S.misc[6] = 0;
                __syncthreads();
                for (uint32_t j0 = 0; j0 < cnt; j0 += blockDim.x) {  // positions by warp-aggregated counters
                    const uint32_t j = j0 + tid;
                    const bool valid = j < cnt;
                    const uint32_t lab = valid ? gp[start + j] & 0xffffu : 0u;
                    const bool is_l = valid && (lab & ROLE_L);
                    const unsigned ml = __ballot_sync(0xffffffffu, is_l), mr = __ballot_sync(0xffffffffu, valid && !is_l);
                    uint32_t base_l = 0, base_r = 0;
                    if (lane == 0) {
                        base_r = atomicAdd(&S.misc[6], (uint32_t)__popc(mr));
                        base_l = atomicAdd(&S.misc[7], (uint32_t)__popc(ml));
                    }
                    base_r = __shfl_sync(0xffffffffu, base_r, 0);
                    base_l = __shfl_sync(0xffffffffu, base_l, 0);
                    if (valid) {
                        const unsigned below = (1u << lane) - 1u;
                        const uint32_t q = is_l ? n_r + base_l + __popc(ml & below) : base_r + __popc(mr & below);
                        if (start + q < m) P.stream[soff + start + q] = (uint16_t)((lab & 0x7fffu) | (q > 0 ? TIE_BIT : 0u));
                    }
                }
                continue;
            }
            if (blockDim.x != SORT_THREADS) __trap();  // the radix passes below are written for SORT_THREADS threads
            // the split left the bucket in no particular order; with two roles a stable pass on the
            // payload byte that holds the role bit comes first, so that R entries precede L entries
            // among equal keys after the (stable) key passes
            int cur = 0;
            if (ut.mult == 2) cur = block_radix_sort<uint32_t, true>(ga + start, gp + start, gb + start, gq + start, (int)cnt, 8, 16, S.whist);
            {
                uint64_t* k0 = cur ? gb + start : ga + start;
                uint32_t* p0 = cur ? gq + start : gp + start;
                uint64_t* k1 = cur ? ga + start : gb + start;
                uint32_t* p1 = cur ? gp + start : gq + start;
                const int c2 = block_sort_keys<uint32_t>(k0, p0, k1, p1, (int)cnt, S.whist, top_bit, 0ull);
                cur = c2 ? (cur ^ 1) : cur;
            }
            const uint64_t* rk = cur ? gb + start : ga + start;
            const uint32_t* rp = cur ? gq + start : gp + start;
            for (uint32_t j = tid; j < cnt; j += blockDim.x) {
                const uint32_t pos = start + j;
                if (pos < m) {
                    uint32_t e = rp[j] & 0x7fffu;
                    if (j > 0 && rk[j] == rk[j - 1]) e |= TIE_BIT;
                    P.stream[soff + pos] = (uint16_t)e;
                }
            }
        }
        if (tid == 0) P.unit_m[t0] = m;
    }
}

// ------------------------------------------------------------------------------------------
// k3_walk
// ------------------------------------------------------------------------------------------
constexpr int WALK_THREADS = 1024;
constexpr int WALK_WARPS = WALK_THREADS / 32;
constexpr int WALK_MAX_GROUPS = 32;
constexpr int WALK_MAX_CHUNKS = 256;  // position chunks of one unit over all rounds of the 32-bit pass

struct WalkParams {
    UnitTables ut;
    const uint16_t* stream;
    uint32_t* unit_m;   // entries of every unit's stream; bit 31: the unit has been walked by the 32-bit pass of k3_walk
    const int32_t *pair_ptr, *pair_l, *pair_r, *pair_out;
    int64_t npairs;
    double* u_out;
    unsigned long long* bfull;  // per-CTA scratch: (maxG + 1)^2 entries each
    unsigned long long* job_counter;
    const int32_t* jobs;
    const int32_t* njobs;
    const int32_t* order;
    int maxG;
    int tiles_per_batch;  // solo mode: R tiles whose accumulators fit in shared memory at once
    int groups;           // solo mode: position chunks walked concurrently (private tables)
    int nslots;           // batch mode: warp groups (0 = batch mode unavailable: tables too large)
    int slot_tiles;       // batch mode: warps per slot = ceil((maxG + 1) / 32)
    int slot_chunks;      // small-unit kernel: position chunks per unit (2 when one warp covers all labels)
    int pass;             // 0: everything with m <= 65536 (uint32 tables); 1: only m > 65536 (uint64 tables)
    const int32_t* nbig;  // number of large single-unit jobs at the head of `order`
    uint32_t big_total;   // a unit with more entries than this is "large": walked by a whole CTA
    int hist_groups;      // large-unit kernel: rows of per-chunk counts kept in shared memory
    int groups0;          // large-unit kernel: position chunks of the 32-bit pass (the 64-bit pass needs them to tell its units)
    int shared_table;     // large-unit kernel: the position chunks add into ONE table (32-bit cells, atomic adds)
    int by_task;          // large-unit kernel: walk EVERY unit (used when the per-warp mode has no room for its tables)
    int shared0;          // large-unit kernel: the 32-bit pass runs with one shared table (and therefore in rounds)
};

// 2U from the count table (AccT entries, row stride `rpad`) for one pair.
template <typename AccT, bool SHIFT>
__device__ inline double pair_u(const AccT* B, size_t rpad, int l, int r, int pg, long long nL, long long nR,
                                long long zL, long long zR) {
    long long two;
    if (SHIFT) {
        two = (long long)B[(size_t)l * rpad + r] + zR * (long long)B[(size_t)l * rpad + pg] +
              zL * (long long)B[(size_t)pg * rpad + r] + zL * zR * (long long)B[(size_t)pg * rpad + pg];
    } else {
        const long long a = (long long)B[(size_t)l * rpad + r] + (long long)B[(size_t)l * rpad + pg] * zR +
                            zL * (long long)B[(size_t)pg * rpad + r];
        const long long b = (long long)B[(size_t)r * rpad + l] + (long long)B[(size_t)r * rpad + pg] * zL +
                            zR * (long long)B[(size_t)pg * rpad + l];
        two = nL * nR + a - b;
    }
    if (nL == 0 || nR == 0) two = 0;
    return 0.5 * (double)two;
}

__device__ inline void slot_barrier(int slot, int slot_tiles) {
    if (slot_tiles == 1) {
        __syncwarp();
    } else {
        asm volatile("bar.sync %0, %1;" ::"r"(slot + 1), "r"(slot_tiles * 32) : "memory");
    }
}

// The inner loop shared by both kernels: positions [i_lo, i_hi) of one unit's stream. One warp covers
// a "wide tile" of W = 32 * J labels: lane l owns labels R0 + 32 j (j < J), i.e. J running counts
// in registers and J read-modify-writes per entry that share the entry's decode. The table has
// whole wide tiles of columns, so lanes past the last label update padding that nobody reads.
// The running counts start at zero: what the entries BEFORE i_lo contribute is linear in the
// counts this function returns (cnt_r: entries per label in the counted role; cnt_l: entries per
// label in the adding role, which is the same thing unless SHIFT) and is added by the caller.
// One cell of a count table += x. 32-bit cells: a shared-memory atomic add is a single instruction
// that nothing waits for (no load / add / store chain through shared memory); 64-bit cells have no
// native shared-memory add (it compiles to a compare-and-swap loop), so they keep the plain form.
template <typename AccT>
__device__ __forceinline__ void add_cell(AccT* p, AccT x) {
    if constexpr (sizeof(AccT) == 4) atomicAdd(p, x); else *p += x;
}

template <typename AccT, bool SHIFT, int J>
__device__ __forceinline__ void walk_range(const uint16_t* st, int i_lo, int i_hi, int R0, AccT* col, size_t rpad,
                                           uint32_t* cnt_r, uint32_t* cnt_l, int Gt, int lane) {
    uint32_t c[J], cs[J], cl[J];
#pragma unroll
    for (int j = 0; j < J; ++j) c[j] = cs[j] = cl[j] = 0u;
    // byte addressing of the table: one multiply-add per entry
    unsigned char* colb = reinterpret_cast<unsigned char*>(col);
    const uint32_t rowbytes = (uint32_t)(rpad * sizeof(AccT));
    int Rj[J];
#pragma unroll
    for (int j = 0; j < J; ++j) Rj[j] = R0 + 32 * j;
    auto bump = [&](uint32_t& cnt, int lab, int R) {  // cnt += (lab == R): compare + predicated add
        asm("{\n\t.reg .pred p;\n\tsetp.eq.s32 p, %1, %2;\n\t@p add.u32 %0, %0, 1;\n\t}" : "+r"(cnt) : "r"(lab), "r"(R));
    };
    auto step = [&](uint32_t e) {
        const int lab = (int)(e & LAB_MASK);
        AccT* row = reinterpret_cast<AccT*>(colb + (uint32_t)lab * rowbytes);
        if (!(e & TIE_BIT)) {
#pragma unroll
            for (int j = 0; j < J; ++j) cs[j] = c[j];
        }
        if (SHIFT) {
            if (e & ROLE_L) {
#pragma unroll
                for (int j = 0; j < J; ++j) {
                    add_cell(&row[32 * j], (AccT)cs[j] + (AccT)c[j]);
                    bump(cl[j], lab, Rj[j]);
                }
            } else {
#pragma unroll
                for (int j = 0; j < J; ++j) bump(c[j], lab, Rj[j]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < J; ++j) {
                add_cell(&row[32 * j], (AccT)cs[j]);
                bump(c[j], lab, Rj[j]);
            }
        }
    };
    // No entry of the 32 ties with its predecessor (the usual case for continuous data): every entry
    // starts its own run, so the count at the run start is the running count itself, and without a
    // role bit the entry IS its label.
    auto step_untied = [&](uint32_t lab) {
        AccT* row = reinterpret_cast<AccT*>(colb + lab * rowbytes);
#pragma unroll
        for (int j = 0; j < J; ++j) {
            add_cell(&row[32 * j], (AccT)c[j]);
            bump(c[j], (int)lab, Rj[j]);
        }
    };
    // Four untied entries at once. A table update is load / add / store through shared memory and
    // the next one may hit the same row, so taken one by one they form a chain of shared-memory
    // round trips. The labels are warp-uniform: when the four differ (checked without divergence)
    // their rows are disjoint, all loads go out first and the adds and stores follow.
    // (QN entries at once: four while the loaded table words fit in 16 registers, else two, else one.)
    constexpr int QN = sizeof(AccT) == 4 ? 1 : (J * sizeof(AccT) <= 32 ? 2 : 1);  // (32-bit: atomics, nothing to overlap)
    auto multi_untied = [&](const uint32_t (&l)[QN]) {
        bool dup = false;
#pragma unroll
        for (int a = 0; a < QN; ++a)
#pragma unroll
            for (int b = a + 1; b < QN; ++b) dup |= l[a] == l[b];
        if (QN == 1 || dup) {
#pragma unroll
            for (int a = 0; a < QN; ++a) step_untied(l[a]);
            return;
        }
        if constexpr (QN > 1) {
        AccT* r[QN];
        AccT v[QN][J];
#pragma unroll
        for (int a = 0; a < QN; ++a) r[a] = reinterpret_cast<AccT*>(colb + l[a] * rowbytes);
#pragma unroll
        for (int a = 0; a < QN; ++a)
#pragma unroll
            for (int j = 0; j < J; ++j) v[a][j] = r[a][32 * j];
#pragma unroll
        for (int a = 0; a < QN; ++a)
#pragma unroll
            for (int j = 0; j < J; ++j) {
                v[a][j] += (AccT)c[j];
                bump(c[j], (int)l[a], Rj[j]);
            }
#pragma unroll
        for (int a = 0; a < QN; ++a)
#pragma unroll
            for (int j = 0; j < J; ++j) r[a][32 * j] = v[a][j];
        }
    };
    // Long runs of equal values (integer-valued data): inside a run every entry adds the SAME vector
    // (the counts at the run start) to its label's row, and the running counts only matter again when
    // the run ends. Blocks of 32 entries that lie wholly inside a run are therefore only counted
    // per label (cd, in registers: lane R holds the count of its own labels), and the rows are
    // brought up to date once, when the run ends: row[L] += cd[L] * cs. Needs the warp to own every
    // label of the unit (one wide tile), since the count of label L lives in L's lane.
    const bool can_defer = !SHIFT && R0 == lane && Gt <= 32 * J;
    uint32_t cd[J];
#pragma unroll
    for (int j = 0; j < J; ++j) cd[j] = 0u;
    bool deferred = false;
    auto flush = [&]() {
#pragma unroll
        for (int j = 0; j < J; ++j) {
            for (int src = 0; src < 32; ++src) {
                const int L = 32 * j + src;
                if (L >= Gt) break;
                const uint32_t h = __shfl_sync(0xffffffffu, cd[j], src);
                if (h != 0u) {
                    AccT* row = reinterpret_cast<AccT*>(colb + (uint32_t)L * rowbytes);
#pragma unroll
                    for (int q = 0; q < J; ++q) add_cell(&row[32 * q], (AccT)h * (AccT)cs[q]);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < J; ++j) {
            c[j] += cd[j];
            cd[j] = 0u;
        }
        deferred = false;
    };
    // Up to 32 entries the slow way: one entry per lane, handed round with shuffles. Used for the few
    // entries before the first 16-byte boundary of the stream and for the tail of the range.
    auto shuffled_block = [&](int i0, int cnt) {
        const uint32_t ent = lane < cnt ? (uint32_t)st[i0 + lane] : 0u;
        if (deferred) flush();
        for (int k = 0; k < cnt; ++k) step(__shfl_sync(0xffffffffu, ent, k));
    };
    int i = i_lo;
    {
        const int to_boundary = (int)(((16u - (unsigned)(reinterpret_cast<uintptr_t>(st + i) & 15u)) & 15u) >> 1);
        const int cnt = min(to_boundary, i_hi - i);
        if (cnt > 0) {
            shuffled_block(i, cnt);
            i += cnt;
        }
    }
    // Whole blocks of 32 entries: EVERY lane loads all 64 bytes (four 16-byte loads from one address:
    // a broadcast) and picks the entries out of its registers, instead of 32 shuffles a block: the
    // walk is bound by the load/store pipe (two atomic adds per entry), which the shuffles share.
    for (; i + 32 <= i_hi; i += 32) {
        uint32_t u[16];
        {
            const uint4* src = reinterpret_cast<const uint4*>(st + i);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint4 w = src[q];
                u[4 * q] = w.x;
                u[4 * q + 1] = w.y;
                u[4 * q + 2] = w.z;
                u[4 * q + 3] = w.w;
            }
        }
        uint32_t any = 0u, all = 0xffffffffu;
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            any |= u[q];
            all &= u[q];
        }
        const uint32_t tie2 = TIE_BIT | (TIE_BIT << 16);
        auto entry = [&](int k) { return (k & 1) ? u[k >> 1] >> 16 : u[k >> 1] & 0xffffu; };
        if (can_defer && (all & tie2) == tie2) {  // wholly inside a run: count only
            deferred = true;
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                const int lab = (int)(entry(k) & LAB_MASK);
#pragma unroll
                for (int j = 0; j < J; ++j) bump(cd[j], lab, Rj[j]);
            }
            continue;
        }
        if (deferred) flush();
        if (!SHIFT && (any & tie2) == 0u) {  // no tie, no role bit: the entry is its label
#pragma unroll
            for (int k = 0; k < 28; k += QN) {
                uint32_t l[QN];
#pragma unroll
                for (int a2 = 0; a2 < QN; ++a2) l[a2] = entry(k + a2);
                multi_untied(l);
            }
            step_untied(entry(28));
            step_untied(entry(29));
            step_untied(entry(30));
#pragma unroll
            for (int j = 0; j < J; ++j) cs[j] = c[j];  // the last entry's run starts here; the next block may tie with it
            step_untied(entry(31));
        } else {
#pragma unroll
            for (int k = 0; k < 32; ++k) step(entry(k));
        }
    }
    if (i < i_hi) shuffled_block(i, i_hi - i);
    if (deferred) flush();
#pragma unroll
    for (int j = 0; j < J; ++j) {
        if (Rj[j] < Gt) {
            cnt_r[Rj[j]] = c[j];
            if (SHIFT && cnt_l != nullptr) cnt_l[Rj[j]] = cl[j];
        }
    }
}

// Units of up to `big_total` entries: every warp group ("slot": the warps that cover one unit's
// labels, one warp for up to 256 labels) takes units on its own from a global counter, with its own
// count table. There is no CTA-wide barrier: the slots of a CTA never wait for each other.
template <bool SHIFT, int J>
__global__ void __launch_bounds__(WALK_THREADS) k3_walk_small(WalkParams P) {
    constexpr int W = 32 * J;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Gmax1 = P.maxG + 1;
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem_raw);                  // [2][WALK_WARPS][Gmax1]: counted / adding role, per (slot, chunk)
    unsigned long long* slot_task = reinterpret_cast<unsigned long long*>(hist + (size_t)2 * WALK_WARPS * Gmax1);
    int* slot_mid = reinterpret_cast<int*>(slot_task + WALK_WARPS);         // [WALK_WARPS]
    uint32_t* acc = reinterpret_cast<uint32_t*>(slot_mid + WALK_WARPS);     // [nslots][Gmax1][stl * W]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const UnitTables& ut = P.ut;
    // A slot = the warps of one unit: `stl` tiles of labels x `nch` position chunks (two when one warp
    // covers all labels: the halves of the unit are walked at the same time into the slot's one
    // table, with atomic adds, and combined through their label counts as in k3_walk).
    const int stl = P.slot_tiles, nch = P.slot_chunks, swarps = stl * nch;
    const int slot = warp / swarps, ws = warp % swarps, wtile = ws % stl, chunk = ws / stl;
    if (slot >= P.nslots) return;
    const size_t rpad = (size_t)stl * W;
    uint32_t* sacc = acc + (size_t)slot * Gmax1 * rpad;
    uint32_t* hist_r = hist + (size_t)(slot * nch) * Gmax1;                   // rows chunk 0 .. nch-1
    uint32_t* hist_l = hist + (size_t)(WALK_WARPS + slot * nch) * Gmax1;
    const int sthreads = swarps * 32, stid = ws * 32 + lane;
    const int64_t ntasks = (int64_t)ut.ngenes * ut.U;
    while (true) {
        unsigned long long t = 0;
        if (swarps == 1) {
            if (lane == 0) t = atomicAdd(P.job_counter, 1ull);
            t = __shfl_sync(0xffffffffu, t, 0);
        } else {
            if (stid == 0) slot_task[slot] = atomicAdd(P.job_counter, 1ull);
            slot_barrier(slot, swarps);
            t = slot_task[slot];
            slot_barrier(slot, swarps);
        }
        if ((int64_t)t >= ntasks) break;
        const int64_t task = (int64_t)t;
        const UnitInfo ui = unit_info(ut, task);
        if (ui.total > P.big_total) continue;  // walked by k3_walk
        const int Gt = ui.G + 1;
        const uint32_t m = P.unit_m[task] & 0x7fffffffu;
        const uint16_t* st = P.stream + ui.soff;
        for (int k = stid; k < Gt * (int)rpad; k += sthreads) sacc[k] = 0;
        if (nch == 2 && ws == 0) {  // second chunk: from the first run start at or after the middle
            int b = m >= 512u ? (int)(m / 2) : (int)m;
            while (b < (int)m) {
                const bool stop = b + lane >= (int)m || !(st[b + lane] & TIE_BIT);
                const unsigned found = __ballot_sync(0xffffffffu, stop);
                if (found) {
                    b += __ffs(found) - 1;
                    break;
                }
                b += 32;
            }
            if (lane == 0) slot_mid[slot] = min(b, (int)m);
        }
        slot_barrier(slot, swarps);
        const int mid = nch == 2 ? slot_mid[slot] : (int)m;
        const int i_lo = chunk == 0 ? 0 : mid, i_hi = chunk == nch - 1 ? (int)m : mid;
        const int R0 = wtile * W + lane;
        // the label counts of the chunk (counted role / adding role) are left behind by the walk
        walk_range<uint32_t, SHIFT, J>(st, i_lo, i_hi, R0, sacc + R0, rpad, hist_r + chunk * Gmax1,
                                       SHIFT ? hist_l + chunk * Gmax1 : nullptr, Gt, lane);
        slot_barrier(slot, swarps);
        if (nch == 2) {  // what the first chunk's entries add to the second chunk's rows
            const uint32_t* add_l = SHIFT ? hist_l + Gmax1 : hist_r + Gmax1;
            for (int k = stid; k < Gt * Gt; k += sthreads) {
                const int L = k / Gt, R = k % Gt;
                const uint32_t extra = add_l[L] * hist_r[R];
                if (extra) sacc[(size_t)L * rpad + R] += SHIFT ? 2u * extra : extra;
            }
            slot_barrier(slot, swarps);
        }
        for (int pp = P.pair_ptr[ui.u] + stid; pp < P.pair_ptr[ui.u + 1]; pp += sthreads) {
            const int l = P.pair_l[pp], r = P.pair_r[pp];
            const long long nL = ut.ncells_dev[ui.first + l], nR = ut.ncells_dev[ui.first + r];
            const long long sl = nch == 2 ? (long long)hist_r[l] + hist_r[Gmax1 + l] : (long long)hist_r[l];
            const long long sr = nch == 2 ? (long long)hist_r[r] + hist_r[Gmax1 + r] : (long long)hist_r[r];
            P.u_out[(size_t)ui.g * P.npairs + P.pair_out[pp]] = pair_u<uint32_t, SHIFT>(sacc, rpad, l, r, ui.G, nL, nR, nL - sl, nR - sr);
        }
        slot_barrier(slot, swarps);
    }
}

template <typename AccT, bool SHIFT, int J>
__global__ void __launch_bounds__(WALK_THREADS) k3_walk(WalkParams P) {
    constexpr int W = 32 * J;  // labels per warp
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Gmax1 = P.maxG + 1;
    const int HG = P.hist_groups;  // chunk rows kept: enough for the groups of either pass
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem_raw);                 // [2 * HG][Gmax1]: per-chunk counts, counted / adding role
    uint32_t* nnzv = hist + (size_t)(2 * HG) * Gmax1;                         // [Gmax1]
    int* bounds = reinterpret_cast<int*>(nnzv + Gmax1);                      // [WALK_MAX_CHUNKS + 1]
    int* cand = bounds + WALK_MAX_CHUNKS + 1;                                // [WALK_MAX_CHUNKS]: run start / end at every cut target
    uint32_t* misc = reinterpret_cast<uint32_t*>(cand + WALK_MAX_CHUNKS);    // [4]
    size_t off = (size_t)((unsigned char*)(misc + 4) - smem_raw);
    off = (off + 15) & ~size_t(15);
    AccT* acc = reinterpret_cast<AccT*>(smem_raw + off);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long* bfull = P.bfull + (size_t)blockIdx.x * Gmax1 * Gmax1;
    const int nbig = *P.nbig;
    const UnitTables& ut = P.ut;

    while (true) {
        __syncthreads();
        if (tid == 0) misc[0] = (uint32_t)atomicAdd(P.job_counter, 1ull);
        __syncthreads();
        int64_t t0;
        if (P.by_task) {
            if ((int64_t)misc[0] >= (int64_t)ut.ngenes * ut.U) break;
            t0 = (int64_t)misc[0];
        } else {
            if ((int)misc[0] >= nbig) break;
            t0 = P.jobs[P.order[misc[0]]];
        }
        // ---- one large unit, all warps: position chunks with private tables -----------------
        const uint32_t m_raw = P.unit_m[t0];
        const uint32_t m = m_raw & 0x7fffffffu;
        if (P.pass == 1 && ((m_raw >> 31) || m <= 65536u)) continue;  // walked by the 32-bit pass / always fits 32-bit cells
        const UnitInfo ui = unit_info(ut, t0);
        const int G = ui.G, Gt = G + 1;
        const uint16_t* st = P.stream + ui.soff;
        const int ntiles = (Gt + W - 1) / W;  // wide tiles
        const int tpb = min(P.tiles_per_batch, ntiles);
        int groups = max(1, min(P.groups, WALK_WARPS / tpb));
        // chunk starts: the first run start at or after the even split (a warp scans 32 entries a step)
        auto chunk_bounds = [&](int ng) {
            for (int cgb = warp; cgb <= ng; cgb += WALK_WARPS) {
                int64_t b = cgb == ng ? (int64_t)m : ((int64_t)m * cgb) / ng;
                while (b < (int64_t)m && b > 0) {
                    const bool stop = b + lane >= (int64_t)m || !(st[b + lane] & TIE_BIT);
                    const unsigned found = __ballot_sync(0xffffffffu, stop);
                    if (found) {
                        b += __ffs(found) - 1;
                        break;
                    }
                    b += 32;
                }
                if (lane == 0) bounds[cgb] = (int)b;
            }
        };
        // Which pass owns the unit. Every chunk is walked from zero counts, so a cell receives at most
        // the sum over the chunks of (entries / 2)^2 (twice that with a shift); the 32-bit pass takes
        // the unit if that stays below 2^32 for ITS chunks (hundreds of thousands of entries do, unless
        // one giant run of equal values keeps most of them in one chunk), the 64-bit pass what is left.
        auto fits32 = [&](int ng) {
            chunk_bounds(ng);
            __syncthreads();
            if (tid == 0) {
                unsigned long long sum = 0;
                for (int c = 0; c < ng; ++c) {
                    const unsigned long long len = (unsigned long long)(bounds[c + 1] - bounds[c]);
                    sum += len * len;
                }
                misc[1] = (sum / 4) * (SHIFT ? 2ull : 1ull) < 4000000000ull ? 1u : 0u;
            }
            __syncthreads();
            return misc[1] != 0u;
        };
        // The 32-bit pass with ONE shared table walks a long unit in rounds: `g0` position chunks at a time, the
        // table being added into the 64-bit totals (bfull) after every round, so that a unit of any length keeps
        // all warps busy on 32-bit cells as long as no single round can overflow one (only a giant run of equal
        // values, which no chunk boundary may cut, can). nrounds_for: the rounds it takes with g0 concurrent
        // chunks, 0 if the unit has to take the 64-bit pass; leaves the chunk bounds behind.
        // Start of the run (of equal values) that holds position b, and the first run start at or after b: a warp
        // looks at 128 entries a step, four loads in flight per lane (a run of 150,000 equal values - an
        // integer-valued gene - is crossed in a millisecond instead of five).
        auto run_start_at_or_before = [&](int64_t b) {
            while (b > 0) {
                unsigned best = 0u;  // 1 + offset from (b - 127) of the last entry without the tie bit
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int64_t i = b - 127 + (lane * 4 + q);
                    if (i >= 0 && i <= b && !(st[i] & TIE_BIT)) best = (unsigned)(lane * 4 + q) + 1u;
                }
                best = __reduce_max_sync(0xffffffffu, best);
                if (best) return b - 127 + (int64_t)(best - 1u);
                b -= 128;
            }
            return (int64_t)0;
        };
        auto run_start_at_or_after = [&](int64_t b) {
            while (b < (int64_t)m) {
                unsigned best = 0xffffffffu;
#pragma unroll
                for (int q = 3; q >= 0; --q) {
                    const int64_t i = b + (lane * 4 + q);
                    if (i < (int64_t)m && !(st[i] & TIE_BIT)) best = (unsigned)(lane * 4 + q);
                }
                best = __reduce_min_sync(0xffffffffu, best);
                if (best != 0xffffffffu) return b + (int64_t)best;
                b += 128;
            }
            return (int64_t)m;
        };
        // The 32-bit pass with ONE shared table walks a long unit in rounds: `g0` position chunks at a time, the
        // table being added into the 64-bit totals (bfull) after every round, so that a unit of any length keeps
        // all warps busy on 32-bit cells as long as no single round can overflow one. Chunks start where a run of
        // equal values starts. A LONG run gets a chunk of its own (boundaries at its start and at its end):
        // walked from zero counts it adds nothing to the table (every entry adds the counts at the run start, zero),
        // however long it is; what it owes comes from the counts before it, in 64 bits. (With a shift a run is
        // not free of additions, so there it counts in full.) nrounds_for: the rounds it takes with g0
        // concurrent chunks, 0 if the unit has to take the 64-bit pass; leaves the chunk bounds behind.
        auto nrounds_for = [&](int g0) {
            constexpr int LONG_RUN = 2048;
            // entries a chunk may hold: g0 chunks of this length, each stretched by up to LONG_RUN entries when its
            // end moves to the next run start, stay below 2^32 in any cell (sum over the chunks of (entries / 2)^2)
            const double per_chunk = 2.0 * sqrt(3.0e9 / ((double)g0 * (SHIFT ? 2.0 : 1.0)));
            const long long per_round = (long long)(per_chunk * g0);
            const long long rounds = ((long long)m + per_round - 1) / per_round;
            const long long ntargets = rounds * g0 - 1;  // interior cuts of an even split into rounds * g0 chunks
            if (2 * (ntargets + 1) + g0 > WALK_MAX_CHUNKS) return 0;
            if (tid == 0) misc[3] = 0u;
            __syncthreads();
            for (int k = warp; k < (int)ntargets; k += WALK_WARPS) {
                const int64_t b = ((int64_t)m * (k + 1)) / (ntargets + 1);
                const int64_t s0 = run_start_at_or_before(b);
                const int64_t e0 = s0 == b ? b : run_start_at_or_after(b);
                if (lane == 0) {
                    cand[2 * k] = (int)s0;
                    cand[2 * k + 1] = (int)e0;
                    if (e0 - s0 > LONG_RUN) misc[3] = 1u;
                }
            }
            __syncthreads();
            if (misc[3] == 0u) {
                // no long run at any cut (the usual case): chunk k ends at the first run start at or after its even
                // cut; by the choice of per_chunk no round can overflow
                for (int k = tid; k <= (int)ntargets + 1; k += blockDim.x) {
                    bounds[k] = k == 0 ? 0 : (k == (int)ntargets + 1 ? (int)m : cand[2 * (k - 1) + 1]);
                }
                __syncthreads();
                return (int)rounds;
            }
            if (tid == 0) {
                int n = 0;  // chunks so far; bounds[n] = start of the next one
                bounds[0] = 0;
                unsigned long long sum = 0;
                uint32_t ok = 1u;
                auto close_chunk = [&](int end, bool isrun) {
                    if (end <= bounds[n]) return;
                    const unsigned long long len = (unsigned long long)(end - bounds[n]);
                    if (!(isrun && !SHIFT)) sum += len * len;
                    bounds[++n] = end;
                    if (n % g0 == 0) {  // a round is complete
                        if ((sum / 4) * (SHIFT ? 2ull : 1ull) >= 4000000000ull) ok = 0u;
                        sum = 0;
                    }
                };
                for (int k = 0; k < (int)ntargets; ++k) {
                    const int s0 = cand[2 * k], e0 = cand[2 * k + 1];
                    if (e0 - s0 > LONG_RUN) {
                        close_chunk(s0, false);
                        close_chunk(e0, true);
                    } else {
                        close_chunk(e0, false);
                    }
                }
                close_chunk((int)m, false);
                if ((sum / 4) * (SHIFT ? 2ull : 1ull) >= 4000000000ull) ok = 0u;
                const int R = (n + g0 - 1) / g0;
                for (int c = n + 1; c <= R * g0; ++c) bounds[c] = (int)m;  // the last round's spare chunks are empty
                misc[1] = ok;
                misc[2] = (uint32_t)R;
            }
            __syncthreads();
            return misc[1] != 0u ? (int)misc[2] : 0;
        };
        if (P.pass == 0 && P.shared_table) {
            const int R = nrounds_for(groups);
            if (R == 0) continue;
            const size_t rpad = (size_t)ntiles * W;  // (shared_table: all tiles of the table are resident)
            uint32_t* hist_l = hist + (size_t)HG * Gmax1;
            for (int k = tid; k < Gt * Gt; k += blockDim.x) bfull[k] = 0ull;
            for (int k = tid; k < Gt; k += blockDim.x) nnzv[k] = 0u;  // label counts of the rounds before
            for (int r = 0; r < R; ++r) {
                for (int k = tid; k < Gt * (int)rpad; k += blockDim.x) acc[k] = 0;
                __syncthreads();
                if (warp < groups * ntiles) {
                    const int cg = warp / ntiles, tile = warp % ntiles;
                    const int R0 = tile * W + lane;
                    walk_range<AccT, SHIFT, J>(st, bounds[r * groups + cg], bounds[r * groups + cg + 1], R0, acc + R0, rpad,
                                               hist + cg * Gmax1, hist_l + cg * Gmax1, Gt, lane);
                }
                __syncthreads();
                // per-chunk counts -> counts before each chunk (earlier rounds included); the round's totals move on
                for (int Rr = tid; Rr < Gt; Rr += blockDim.x) {
                    uint32_t run = nnzv[Rr];
                    for (int cg = 0; cg < groups; ++cg) {
                        const uint32_t here = hist[cg * Gmax1 + Rr];
                        hist[cg * Gmax1 + Rr] = run;
                        if (!SHIFT) hist_l[cg * Gmax1 + Rr] = here;
                        run += here;
                    }
                    nnzv[Rr] = run;
                }
                __syncthreads();
                for (int k = tid; k < Gt * Gt; k += blockDim.x) {
                    const int L = k / Gt, Rr = k % Gt;
                    unsigned long long s2 = 0;  // what the entries before each chunk add to its rows
                    for (int cg = 0; cg < groups; ++cg)
                        s2 += (unsigned long long)hist_l[cg * Gmax1 + L] * (unsigned long long)hist[cg * Gmax1 + Rr];
                    bfull[k] += (unsigned long long)acc[(size_t)L * rpad + Rr] + (SHIFT ? 2ull * s2 : s2);
                }
                __syncthreads();
            }
            for (int pp = P.pair_ptr[ui.u] + tid; pp < P.pair_ptr[ui.u + 1]; pp += blockDim.x) {
                const int l = P.pair_l[pp], r = P.pair_r[pp];
                const long long nL = ut.ncells_dev[ui.first + l], nR = ut.ncells_dev[ui.first + r];
                const long long zL = nL - (long long)nnzv[l], zR = nR - (long long)nnzv[r];
                P.u_out[(size_t)ui.g * P.npairs + P.pair_out[pp]] =
                    pair_u<unsigned long long, SHIFT>(bfull, (size_t)Gt, l, r, G, nL, nR, zL, zR);
            }
            if (tid == 0) P.unit_m[t0] = m | 0x80000000u;  // the 64-bit pass skips it
            continue;
        }
        if (P.pass == 0) {
            if (!fits32(groups)) continue;
            if (tid == 0) P.unit_m[t0] = m | 0x80000000u;
        } else {
            chunk_bounds(groups);  // (whatever the 32-bit pass took carries the flag and never gets here)
        }
        for (int tb = 0; tb < ntiles; tb += tpb) {
            const int bt = min(tpb, ntiles - tb);
            const size_t rpad = (size_t)bt * W;
            const int copies = P.shared_table ? 1 : groups;
            for (int k = tid; k < copies * Gt * (int)rpad; k += blockDim.x) acc[k] = 0;
            __syncthreads();
            if (warp < groups * bt) {
                const int cg = warp / bt, tile = tb + warp % bt;
                const int R0 = tile * W + lane;
                walk_range<AccT, SHIFT, J>(st, bounds[cg], bounds[cg + 1], R0,
                                           acc + (size_t)(P.shared_table ? 0 : cg) * Gt * rpad + (R0 - tb * W), rpad, hist + cg * Gmax1,
                                           hist + (HG + cg) * Gmax1, Gt, lane);
            }
            __syncthreads();
            for (int k = tid; k < Gt * (int)rpad; k += blockDim.x) {
                const int L = k / (int)rpad, R = tb * W + k % (int)rpad;
                if (R < Gt) {
                    unsigned long long s = 0;
                    for (int cg = 0; cg < copies; ++cg) s += (unsigned long long)acc[(size_t)cg * Gt * rpad + k];
                    bfull[(size_t)L * Gt + R] = s;
                }
            }
            __syncthreads();
        }
        // Every chunk was walked with its running counts starting at zero. What the earlier chunks
        // add to cell (L, R) is (entries of L in the chunk, adding role) x (entries of R before the
        // chunk, counted role) -- twice that with a shift, where an adding entry takes the count
        // before its run plus the count after it. The walkers left both per-chunk counts in `hist`.
        uint32_t* hist_l = hist + (size_t)HG * Gmax1;
        for (int R = tid; R < Gt; R += blockDim.x) {
            uint32_t run = 0;
            for (int cg = 0; cg < groups; ++cg) {
                const uint32_t here = hist[cg * Gmax1 + R];
                hist[cg * Gmax1 + R] = run;
                if (!SHIFT) hist_l[cg * Gmax1 + R] = here;
                run += here;
            }
            nnzv[R] = run;
        }
        __syncthreads();
        for (int k = tid; k < Gt * Gt; k += blockDim.x) {
            const int L = k / Gt, R = k % Gt;
            unsigned long long s = 0;
            for (int cg = 1; cg < groups; ++cg)
                s += (unsigned long long)hist_l[cg * Gmax1 + L] * (unsigned long long)hist[cg * Gmax1 + R];
            bfull[k] += SHIFT ? 2ull * s : s;
        }
        __syncthreads();
        for (int pp = P.pair_ptr[ui.u] + tid; pp < P.pair_ptr[ui.u + 1]; pp += blockDim.x) {
            const int l = P.pair_l[pp], r = P.pair_r[pp];
            const long long nL = ut.ncells_dev[ui.first + l], nR = ut.ncells_dev[ui.first + r];
            const long long zL = nL - (long long)nnzv[l], zR = nR - (long long)nnzv[r];
            P.u_out[(size_t)ui.g * P.npairs + P.pair_out[pp]] =
                pair_u<unsigned long long, SHIFT>(bfull, (size_t)Gt, l, r, G, nL, nR, zL, zR);
        }
    }
}

// ------------------------------------------------------------------------------------------
// k3_ties: tie term per pair. One CTA per (gene, unit); one thread per pair walks the stream.
// A run contributes t(t^2-1) with t = (entries of the left label in L role) + (entries of the
// right label in R role); the zeros of a group join the run at value 0 (R role) / -lfc (L role),
// which is where the probe entries sit. Runs where t <= 1 add exactly 0, so the double
// accumulation visits the same non-zero terms in the same order as the reference.
// ------------------------------------------------------------------------------------------
struct TieParams {
    UnitTables ut;
    const uint16_t* stream;
    const uint32_t* unit_m;
    const int32_t *pair_ptr, *pair_l, *pair_r, *pair_out;
    int64_t npairs;
    double* t_out;
    int maxG;
};

constexpr int TIE_EVENTS = 8192;  // compacted stream entries kept in shared memory per unit

// Adds one stream entry to the running tie count of pair (l, r) of a unit with probe label G.
template <bool SHIFT>
__device__ __forceinline__ void tie_entry(uint32_t e, int l, int r, int G, long long zL, long long zR, long long& t) {
    const int lab = (int)(e & LAB_MASK);
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

template <bool SHIFT>
__global__ void __launch_bounds__(256) k3_ties(TieParams P) {
    extern __shared__ __align__(16) unsigned char tie_smem[];
    uint32_t* hist = reinterpret_cast<uint32_t*>(tie_smem);                 // [maxG + 2] stored non-zero entries per label
    uint32_t* wsum = hist + (P.maxG + 2);                                   // [8] + [1] count + [1] overflow
    uint16_t* events = reinterpret_cast<uint16_t*>(wsum + 12);              // [TIE_EVENTS]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t task = blockIdx.x;
    const UnitInfo ui = unit_info(P.ut, task);
    const int G = ui.G;
    const uint16_t* st = P.stream + ui.soff;
    const uint32_t m = P.unit_m[task] & 0x7fffffffu;
    for (int k = tid; k <= G; k += blockDim.x) hist[k] = 0;
    if (tid == 0) {
        wsum[8] = 0;
        wsum[9] = 0;
    }
    __syncthreads();
    // Only entries of runs with two or more entries, and the probes (where the zeros sit), can add
    // to any pair's tie term. They are compacted, in order, into shared memory; bit 15 of a
    // compacted entry marks the first entry of its run.
    for (uint32_t base = 0; base < m; base += blockDim.x) {
        const uint32_t i = base + tid;
        uint32_t e = 0;
        bool rel = false;
        if (i < m) {
            e = st[i];
            const bool tied_next = i + 1 < m && (st[i + 1] & TIE_BIT);
            rel = (e & TIE_BIT) || tied_next || (int)(e & LAB_MASK) == G;
            if (!(SHIFT && (e & ROLE_L))) atomicAdd(&hist[e & LAB_MASK], 1u);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, rel);
        if (lane == 0) wsum[warp] = __popc(bal);
        __syncthreads();
        uint32_t off = wsum[8];
        for (int w = 0; w < warp; ++w) off += wsum[w];
        if (rel) {
            const uint32_t at = off + __popc(bal & ((1u << lane) - 1u));
            // a compacted entry starts a run unless it ties with its predecessor, which is then relevant too
            if (at < TIE_EVENTS) events[at] = (uint16_t)((e & 0x7fffu) | ((e & TIE_BIT) ? 0u : 0x8000u));
        }
        __syncthreads();
        if (tid == 0) {
            uint32_t tot = wsum[8];
            for (int w = 0; w < 8; ++w) tot += wsum[w];
            wsum[8] = tot;
            if (tot > TIE_EVENTS) wsum[9] = 1;
        }
        __syncthreads();
    }
    const uint32_t nev = wsum[8];
    const bool overflow = wsum[9] != 0;
    for (int pp = P.pair_ptr[ui.u] + tid; pp < P.pair_ptr[ui.u + 1]; pp += blockDim.x) {
        const int l = P.pair_l[pp], r = P.pair_r[pp];
        const long long nL = P.ut.ncells_dev[ui.first + l], nR = P.ut.ncells_dev[ui.first + r];
        const long long zL = nL - (long long)hist[l], zR = nR - (long long)hist[r];
        double tie = 0.0;
        long long t = 0;
        auto close_run = [&]() {
            if (t > 0) {
                const double nt = (double)t;
                tie += nt * (nt * nt - 1.0);
            }
            t = 0;
        };
        if (!overflow) {
            for (uint32_t k = 0; k < nev; ++k) {
                const uint32_t e = events[k];
                if (e & 0x8000u) close_run();
                tie_entry<SHIFT>(e, l, r, G, zL, zR, t);
            }
        } else {  // too many tied entries for shared memory: walk the stream itself
            for (uint32_t i = 0; i < m; ++i) {
                const uint32_t e = st[i];
                if (!(e & TIE_BIT)) close_run();
                tie_entry<SHIFT>(e, l, r, G, zL, zR, t);
            }
        }
        close_run();
        if (nL == 0 || nR == 0) tie = 0.0;
        P.t_out[(size_t)ui.g * P.npairs + P.pair_out[pp]] = tie;
    }
}

}  // namespace


void run_wilcox(scranb200_ctx* ctx, const Transposed& T, const PairLayout& L, double lfc, double* d_u, double* d_t) {
    if (L.U == 0 || T.ngenes == 0 || L.npairs == 0) return;
    if (L.maxG > MAX_LABELS) throw Error("more than 16383 groups compared inside one block are not supported");
    if (L.maxG + 1 > MAXSEG) throw Error("more than 4095 groups compared inside one block are not supported");
    const bool shift = lfc != 0.0;
    const int mult = shift ? 2 : 1;
    const int64_t ntasks = (int64_t)T.ngenes * L.U;
    if (ntasks > 2000000000ll) throw Error("too many (gene, block) units");

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
    ctx->jobs.reserve((size_t)(ntasks + 2) * 16 + 64);  // jobs | order | (first, end) task of the jobs in order
    ctx->flags.reserve(64);
    unsigned long long* counter = reinterpret_cast<unsigned long long*>(ctx->flags.as<char>() + 16);
    int32_t* njobs = reinterpret_cast<int32_t*>(ctx->flags.as<char>() + 32);
    int32_t* nbig = reinterpret_cast<int32_t*>(ctx->flags.as<char>() + 40);
    int32_t* jobs = ctx->jobs.as<int32_t>();
    int32_t* order = jobs + (ntasks + 2);

    // ---- plan ----
    // capacity of the shared-memory sort: 12 bytes per element plus the fixed tables (which hold a
    // list with one entry per BUCKET_RANK_MAX elements)
    const int segcap = std::max(2048, L.maxG + 1);  // staged segment boundaries of a job
    auto cap_for = [&](size_t budget) {
        int c = (int)((budget - sort_fixed_bytes(0, segcap)) * 8 / 97);
        c = std::min(c, 65280) & ~63;
        while (c > 64 && (size_t)c * 12 + sort_fixed_bytes(c, segcap) > budget) c -= 64;
        return c;
    };
    // (Two CTAs of 512 threads per SM with half the capacity each were tried for the small jobs, so
    // that one computes while the other waits at a barrier: slower, 10.2 vs 9.7 ms at config 2.)
    const int cap_large = cap_for(ctx->smem_optin);
    if ((size_t)cap_large * 12 < (size_t)SORT_HIST_WORDS * 4) throw Error("internal: shared memory too small for the sort");
    const uint32_t small_max = (uint32_t)cap_large / 2, window = (uint32_t)cap_large - small_max;
    const int maxb = std::max(1, std::min(MAXB, segcap / (L.maxG + 1)));
    SB_LAUNCH(ctx, k3_plan, 1, 1024, 0, ut, small_max, window, maxb, jobs, njobs);
    int2* order_range = reinterpret_cast<int2*>(order + (ntasks + 2));
    SB_LAUNCH(ctx, k3_order, 1, 1024, 0, ut, jobs, njobs, 16384u, order, order_range, nbig);

    // ---- sort ----
    int64_t max_cells = 0;
    for (int u = 0; u < L.U; ++u) {
        int64_t s = 0;
        for (int k = 0; k < L.unit_ncomb[u]; ++k) s += L.ncells_dev[L.unit_first[u] + k];
        max_cells = std::max(max_cells, s);
    }
    const int64_t max_m = mult * (std::min<int64_t>(max_cells, nnz_used) + 1);
    if (max_m > 2000000000ll) throw Error("a (gene, block) unit with more than 2e9 entries is not supported");
    SortParams sp;
    sp.ut = ut;
    sp.vals = T.vals;
    sp.vmin = T.vmin;
    sp.vmax = T.vmax;
    if (!T.vmin || !T.vmax) throw Error("internal: the value ranges of the segments are missing");
    sp.shift = lfc;
    sp.segcap = segcap;
    sp.maxG = L.maxG;
    sp.stream = ctx->stream_lab.as<uint16_t>();
    sp.unit_m = ctx->unit_m.as<uint32_t>();
    sp.g_ka = sp.g_kb = nullptr;
    sp.g_pa = sp.g_pb = nullptr;
    const int sort_grid = (int)std::min<int64_t>(ntasks, (int64_t)ctx->sm_count);
    // room for the sample split: up to SPLIT_MAX_RANGES ranges of `cap` entries, three quarters full on average
    sp.scratch_len = (std::max(max_m, std::min(max_m + max_m / 3 + 2 * (int64_t)cap_large, (int64_t)(SPLIT_MAX_RANGES + 1) * cap_large)) + 63) & ~int64_t(63);
    if (max_m > cap_large) {  // some unit may not fit shared memory: a scratch region per CTA, sized for the largest unit
        const size_t all = (size_t)sort_grid * (size_t)sp.scratch_len;
        ctx->sort_keys.reserve(all * 16 + 64);
        ctx->sort_labs.reserve(all * 8 + 64);
        sp.g_ka = ctx->sort_keys.as<uint64_t>();
        sp.g_kb = sp.g_ka + all;
        sp.g_pa = ctx->sort_labs.as<uint32_t>();
        sp.g_pb = sp.g_pa + all;
    }
    sp.jobs = jobs;
    sp.njobs = njobs;
    sp.order = order;
    sp.order_range = order_range;
    sp.job_counter = counter;
    allow_smem(k3_sort, (size_t)cap_large * 12 + sort_fixed_bytes(cap_large, segcap));
    sp.cap = cap_large;
    SB_CUDA(cudaMemsetAsync(counter, 0, 8, ctx->stream));
    SB_LAUNCH(ctx, k3_sort, sort_grid, SORT_THREADS, (size_t)cap_large * 12 + sort_fixed_bytes(cap_large, segcap), sp);
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
    wp.job_counter = counter;
    wp.jobs = jobs;
    wp.njobs = njobs;
    wp.order = order;
    wp.maxG = L.maxG;
    const int Gt = L.maxG + 1;
    // Labels per warp: 32 * J with J in {1, 2, 4, 8}; beyond 256 labels several warps share a unit.
    const int narrow_tiles = (Gt + 31) / 32;
    const int J = narrow_tiles <= 1 ? 1 : narrow_tiles <= 2 ? 2 : narrow_tiles <= 4 ? 4 : 8;
    const int W = 32 * J;
    const int ntiles = (Gt + W - 1) / W;  // wide tiles = warps per unit
    // fixed part of the large-unit kernel's shared memory, with rows for hg chunk groups
    auto walk_fixed_for = [&](int hg) { return ((size_t)(2 * hg) * Gt + Gt) * 4 + (2 * WALK_MAX_CHUNKS + 1) * 4 + 16 + 32; };
    const int walk_grid_max = ctx->sm_count * 2;
    ctx->plan_tmp.reserve((size_t)walk_grid_max * Gt * Gt * 8 + 64);
    wp.bfull = ctx->plan_tmp.as<unsigned long long>();

    const uint32_t big_total = 16384u;  // same threshold as k3_order's "large job"
    wp.big_total = big_total;
    wp.nbig = nbig;
    wp.by_task = 0;
    wp.shared0 = 0;

    // ---- units of up to big_total entries: one warp group per unit, no CTA-wide barriers ----
    {
        const size_t table = (size_t)Gt * W * 4 * ntiles;  // one unit's uint32 count table
        const size_t fixed_small = ((size_t)2 * WALK_WARPS * Gt) * 4 + WALK_WARPS * 8 + WALK_WARPS * 4 + 32;
        // halves of a unit walked at the same time, if shared memory holds fewer tables than the CTA has warps
        const int nch = ntiles == 1 && fixed_small + table * WALK_WARPS > ctx->smem_optin ? 2 : 1;
        int ns = 0;
        if (fixed_small + table <= ctx->smem_optin && ntiles * nch <= WALK_WARPS) {
            ns = (int)std::min<size_t>((ctx->smem_optin - fixed_small) / table, (size_t)(WALK_WARPS / (ntiles * nch)));
            if (ntiles * nch > 1) ns = std::min(ns, 15);  // named barriers 1..15
        }
        wp.slot_chunks = nch;
        wp.slot_tiles = ntiles;
        wp.nslots = ns;
        if (ns > 0) {
            const size_t smem = fixed_small + table * ns;
            SB_CUDA(cudaMemsetAsync(counter, 0, 8, ctx->stream));
            const int grid = ctx->sm_count * (smem * 2 <= ctx->smem_optin ? 2 : 1);
#define SB_WALKS(SH, JJ)                                                        \
    do {                                                                        \
        allow_smem(k3_walk_small<SH, JJ>, smem);                                \
        SB_LAUNCH(ctx, (k3_walk_small<SH, JJ>), grid, WALK_THREADS, smem, wp);  \
    } while (0)
#define SB_WALKS_J(SH)                  \
    do {                                \
        if (J == 1) SB_WALKS(SH, 1);    \
        else if (J == 2) SB_WALKS(SH, 2); \
        else if (J == 4) SB_WALKS(SH, 4); \
        else SB_WALKS(SH, 8);           \
    } while (0)
            if (shift) SB_WALKS_J(true); else SB_WALKS_J(false);
#undef SB_WALKS_J
#undef SB_WALKS
        } else {
            wp.by_task = 1;  // no room for per-warp tables: the large-unit kernel walks every unit
        }
    }

    // ---- large units: one CTA per unit (32-bit counts up to 65536 entries, 64-bit beyond) ----
    const bool any_large = wp.by_task || max_m > (int64_t)big_total;
    for (int pass = 0; pass < 2 && any_large; ++pass) {
        const size_t acc_size = pass == 0 ? 4 : 8;
        if (pass == 1 && max_m <= 65536) break;
        wp.pass = pass;
        // Labels per warp for this pass: the widest tile (fewest decodes of an entry) whose tiles all fit
        // shared memory at once, so that the stream is walked once; with hundreds of groups the
        // padding of wide tiles decides (201 labels: seven tiles of 32 fit, one tile of 256 does not).
        // If not even the narrow tiles fit together (64-bit counts), they are walked in batches.
        // Chunk groups never exceed the warps left per tile, which bounds the per-chunk count rows.
        int Jp = J;
        auto groups_max = [&](int jj) { return std::max(1, std::min(WALK_MAX_GROUPS, WALK_WARPS / std::min((Gt + 32 * jj - 1) / (32 * jj), WALK_WARPS))); };
        auto all_fit = [&](int jj) {
            const size_t need = (size_t)Gt * 32 * jj * acc_size * ((Gt + 32 * jj - 1) / (32 * jj)) + walk_fixed_for(groups_max(jj));
            return need <= ctx->smem_optin;
        };
        while (Jp > 1 && !all_fit(Jp)) Jp >>= 1;
        const int hist_groups = groups_max(Jp);
        wp.hist_groups = hist_groups;
        const size_t walk_fixed = walk_fixed_for(hist_groups);
        if (walk_fixed > ctx->smem_optin) throw Error("too many groups inside one block for the shared-memory count table");
        const size_t budget = ctx->smem_optin - walk_fixed;
        const int Wp = 32 * Jp, ntiles_p = (Gt + Wp - 1) / Wp;
        const size_t per_tile = (size_t)Gt * Wp * acc_size;
        if (per_tile > budget) throw Error("too many groups inside one block for the shared-memory count table");
        int tpb, groups;
        int copies = 1;
        wp.shared_table = 0;
        if (per_tile * ntiles_p <= budget) {
            tpb = std::min(ntiles_p, WALK_WARPS);
            if (pass == 0) {  // 32-bit cells take atomic adds: every warp walks a position chunk into one shared table
                wp.shared_table = 1;
                groups = std::max(1, std::min(hist_groups, WALK_WARPS / std::max(1, tpb)));
            } else {
                groups = (int)std::min<size_t>(budget / (per_tile * ntiles_p), (size_t)hist_groups);
                groups = std::max(1, std::min(groups, WALK_WARPS / std::max(1, tpb)));
                copies = groups;
            }
        } else {
            tpb = (int)std::max<size_t>(1, std::min<size_t>(budget / per_tile, (size_t)ntiles_p));
            tpb = std::min(tpb, WALK_WARPS);
            groups = 1;
        }
        wp.tiles_per_batch = tpb;
        wp.groups = groups;
        if (pass == 0) {
            wp.groups0 = std::max(1, std::min(groups, WALK_WARPS / std::max(1, tpb)));
            wp.shared0 = wp.shared_table;
        }
        const size_t smem = walk_fixed + per_tile * tpb * copies;
        SB_CUDA(cudaMemsetAsync(counter, 0, 8, ctx->stream));
        const int grid = (int)std::min<int64_t>(ntasks, smem * 2 <= ctx->smem_optin ? walk_grid_max : ctx->sm_count);
#define SB_WALK(ACC, SH, JJ)                                              \
    do {                                                                  \
        allow_smem(k3_walk<ACC, SH, JJ>, smem);                           \
        SB_LAUNCH(ctx, (k3_walk<ACC, SH, JJ>), grid, WALK_THREADS, smem, wp); \
    } while (0)
#define SB_WALK_J(ACC, SH)                  \
    do {                                    \
        if (Jp == 1) SB_WALK(ACC, SH, 1);    \
        else if (Jp == 2) SB_WALK(ACC, SH, 2); \
        else if (Jp == 4) SB_WALK(ACC, SH, 4); \
        else SB_WALK(ACC, SH, 8);           \
    } while (0)
        if (pass == 0) {
            if (shift) SB_WALK_J(uint32_t, true); else SB_WALK_J(uint32_t, false);
        } else {
            if (shift) SB_WALK_J(unsigned long long, true); else SB_WALK_J(unsigned long long, false);
        }
#undef SB_WALK_J
#undef SB_WALK
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
        tp.maxG = L.maxG;
        const size_t tie_smem = (size_t)(L.maxG + 2 + 12) * 4 + (size_t)TIE_EVENTS * 2;
        if (shift) {
            allow_smem(k3_ties<true>, tie_smem);
            SB_LAUNCH(ctx, k3_ties<true>, (unsigned)ntasks, 256, tie_smem, tp);
        } else {
            allow_smem(k3_ties<false>, tie_smem);
            SB_LAUNCH(ctx, k3_ties<false>, (unsigned)ntasks, 256, tie_smem, tp);
        }
    }
    phase_mark(ctx, PH_COUNT);
}

}  // namespace sb
