// Development probe (not part of the library): cycle counts of the shared-memory radix sort phases.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define SORT_PROBE 1
#include "../scran_b200/csrc/common.cuh"
#include "../scran_b200/csrc/sort.cuh"
using namespace sb;

__global__ void __launch_bounds__(SORT_THREADS) probe(const uint64_t* keys, int n, int reps, long long* out, int top_bit) {
    extern __shared__ __align__(16) unsigned char raw[];
    const int cap = 6784;
    uint64_t* ka = (uint64_t*)raw; uint64_t* kb = ka + cap;
    uint32_t* pa = (uint32_t*)(kb + cap); uint32_t* pb = pa + cap;
    uint32_t* whist = pb + cap;
    long long t_total = 0;
    for (int r = 0; r < reps; ++r) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) { ka[i] = keys[(size_t)blockIdx.x * n + i]; pa[i] = i; }
        __syncthreads();
        long long t0 = clock64();
        int cur = block_sort_keys<uint32_t>(ka, pa, kb, pb, n, whist, top_bit);
        long long t1 = clock64();
        t_total += t1 - t0;
        if (r == 0 && threadIdx.x == 0) {
            const uint64_t* k = cur ? kb : ka; int bad = 0;
            for (int i = 1; i < n; ++i) bad += k[i-1] > k[i];
            out[2 + blockIdx.x % 4] = bad;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t_total / reps; }
}

int main() {
    const int n = 6000, grid = 148;
    std::vector<uint64_t> h((size_t)grid * n);
    for (auto& v : h) { double x = 0.5 + 7.5 * (rand() / (double)RAND_MAX); v = key_of(x); }
    uint64_t* d; long long* out;
    cudaMalloc(&d, h.size() * 8); cudaMalloc(&out, 64); cudaMemset(out, 0, 64);
    cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
    size_t smem = 6784 * 24 + SORT_HIST_WORDS * 4 + 64;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int top : {64, 52, 44}) {
        for (int g : {1, 148}) {
            probe<<<g, SORT_THREADS, smem>>>(d, n, 20, out, top);
            cudaError_t e = cudaDeviceSynchronize();
            long long ho[8]; cudaMemcpy(ho, out, 64, cudaMemcpyDeviceToHost);
            long long pr[8]; cudaMemcpyFromSymbol(pr, g_probe, 64); long long z[8]={0}; cudaMemcpyToSymbol(g_probe, z, 64);
            if (pr[5]) printf("   per pass (thread 0 of all CTAs, avg): zero+hist %lld  wait %lld  scan %lld  reorder %lld  tailwait %lld  (passes %lld)\n", pr[0]/pr[5], pr[1]/pr[5], pr[2]/pr[5], pr[3]/pr[5], pr[4]/pr[5], pr[5]);
            printf("top_bit %d grid %3d: %lld cycles per sort of %d keys (%.1f cyc/key), unsorted pairs %lld  [%s]\n", top, g, ho[0], n,
                   ho[0] / (double)n, ho[2], cudaGetErrorString(e));
        }
    }
    return 0;
}
