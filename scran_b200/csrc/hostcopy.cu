// Host -> device copies of the matrix arrays. Pinned (page-locked) sources go to the copy engine directly
// at PCIe speed. An R process hands over ordinary pageable vectors, for which cudaMemcpy stages every byte
// through one driver thread (about 12 GB/s measured here, a quarter of the link). Pageable sources are
// therefore staged by the library itself: a few host threads each copy their slices of the source into
// their own pair of pinned buffers and queue the DMA on their own stream, so the memcpy of one slice
// overlaps the DMA of the previous one and the threads' memcpys add up.
#include <algorithm>
#include <thread>

#include "common.cuh"

namespace sb {

namespace {

constexpr size_t STAGE_BYTES = 8u << 20;  // per staging buffer; two per worker
constexpr int MAX_WORKERS = 8;

bool is_pinned(const void* p) {
    cudaPointerAttributes attr;
    const cudaError_t e = cudaPointerGetAttributes(&attr, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged;
}

}  // namespace

void stage_release(scranb200_ctx* ctx) {
    if (ctx->stage_pinned) cudaFreeHost(ctx->stage_pinned);
    ctx->stage_pinned = nullptr;
    for (int k = 0; k < ctx->stage_workers; ++k) {
        if (ctx->stage_streams[k]) cudaStreamDestroy(ctx->stage_streams[k]);
        ctx->stage_streams[k] = nullptr;
    }
    ctx->stage_workers = 0;
}

// Synchronous: the data is on the device (and the source free again) when this returns. Work queued on the
// context's stream before the call is NOT waited for; the destination must not be in use.
void h2d_copy(scranb200_ctx* ctx, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return;
    if (bytes < 4 * STAGE_BYTES || is_pinned(src)) {
        SB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
        SB_CUDA(cudaStreamSynchronize(ctx->stream));
        return;
    }
    if (ctx->stage_workers == 0) {
        int w = (int)std::thread::hardware_concurrency() / 2;
        w = std::max(2, std::min(w, MAX_WORKERS));
        void* p = nullptr;
        SB_CUDA(cudaHostAlloc(&p, (size_t)w * 2 * STAGE_BYTES, cudaHostAllocDefault));
        ctx->stage_pinned = p;
        for (int k = 0; k < w; ++k) SB_CUDA(cudaStreamCreateWithFlags(&ctx->stage_streams[k], cudaStreamNonBlocking));
        ctx->stage_workers = w;
    }
    const int W = ctx->stage_workers;
    const size_t nslices = (bytes + STAGE_BYTES - 1) / STAGE_BYTES;
    std::vector<cudaError_t> errs(W, cudaSuccess);
    auto work = [&](int t) {
        cudaError_t e = cudaSetDevice(ctx->device);
        cudaEvent_t done[2] = {nullptr, nullptr};
        for (int b = 0; b < 2 && e == cudaSuccess; ++b) e = cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming);
        unsigned char* mine = static_cast<unsigned char*>(ctx->stage_pinned) + (size_t)t * 2 * STAGE_BYTES;
        int which = 0;
        bool used[2] = {false, false};
        for (size_t s = (size_t)t; s < nslices && e == cudaSuccess; s += (size_t)W) {
            const size_t off = s * STAGE_BYTES, n = std::min(STAGE_BYTES, bytes - off);
            unsigned char* buf = mine + (size_t)which * STAGE_BYTES;
            if (used[which]) e = cudaEventSynchronize(done[which]);  // the DMA that last read this buffer
            if (e != cudaSuccess) break;
            memcpy(buf, static_cast<const unsigned char*>(src) + off, n);
            e = cudaMemcpyAsync(static_cast<unsigned char*>(dst) + off, buf, n, cudaMemcpyHostToDevice, ctx->stage_streams[t]);
            if (e == cudaSuccess) e = cudaEventRecord(done[which], ctx->stage_streams[t]);
            used[which] = true;
            which ^= 1;
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stage_streams[t]);
        for (int b = 0; b < 2; ++b) {
            if (done[b]) cudaEventDestroy(done[b]);
        }
        errs[t] = e;
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < W; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
    for (int t = 0; t < W; ++t) SB_CUDA(errs[t]);
}

}  // namespace sb
