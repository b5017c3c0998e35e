// noalloc.cu - b200pt_intersect_closest_noalloc: the production closest-hit traversal
// (traverse_persist.cu) for callers that may not allocate (code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  The kernel's scratch - the ray-fetch counter and the list of
// rays the wide walk hands to the exact binary walk - lives in a file-scope __device__ array of a fixed
// size instead of the context's grow-only cudaMalloc'ed buffers.  No counterpart in the reference.
//
// Every call on a device shares the one array: calls that overlap on the GPU (two streams, two contexts on
// the same device) race on it.  The array keeps the last call's contents; the two words that must start at
// zero (fetch counter, list length) are cleared by cudaMemsetAsync on the call's stream before each launch
// (LaunchTraversePersist does that for any scratch it is given).
#include <cuda_runtime.h>

#include <mutex>

#include "../../include/b200pt.h"
#include "api_internal.h"
#include "kernels.h"

namespace {

// [0] ray-fetch counter; [kListAt] length of the exact-walk list; [kListAt + 1 ...] the list (one int per ray)
constexpr int kListAt = 64;
__device__ __align__(256) int g_noallocScratch[kListAt + 1 + B200PT_NOALLOC_MAX_RAYS];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
int *g_scratchOnDevice[kMaxDevices];  // device address of g_noallocScratch per device, looked up once

// `g_noallocScratch` in host code is the host-side stand-in of the symbol, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress with the device current.
cudaError_t ScratchAddress(int device, int **out) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratchOnDevice[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noallocScratch);
        if (e != cudaSuccess) return e;
        g_scratchOnDevice[device] = static_cast<int *>(p);
    }
    *out = g_scratchOnDevice[device];
    return cudaSuccess;
}

}  // namespace

extern "C" int b200pt_intersect_closest_noalloc(b200pt_scene *s, const float *rays, int64_t n, b200pt_hit *hits) {
    if (!s || n < 0 || (n > 0 && (!rays || !hits))) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = s->ctx;
    if (n > B200PT_NOALLOC_MAX_RAYS) {
        ctx->err = "b200pt_intersect_closest_noalloc: n = " + std::to_string(n) + " rays exceeds B200PT_NOALLOC_MAX_RAYS = " +
                   std::to_string(B200PT_NOALLOC_MAX_RAYS) + " (the static device scratch); split the batch or use b200pt_intersect_closest";
        return B200PT_ERR_INVALID;
    }
    if (ctx->device < 0 || ctx->device >= kMaxDevices) {
        ctx->err = "b200pt_intersect_closest_noalloc: device index beyond the per-device scratch table";
        return B200PT_ERR_INVALID;
    }
    if (n == 0) return B200PT_OK;
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    int *scratch = nullptr;
    B2_CUDA(ctx, ScratchAddress(ctx->device, &scratch));
    B2_CUDA(ctx, b200pt::LaunchTraversePersist(s->dev, rays, n, hits, nullptr, false, scratch, nullptr,
                                               ctx->smCount, int(s->dev.maxDepth) + 1, ctx->refillMin, ctx->nodeMin,
                                               ctx->stream, false, nullptr, ctx->wideKernel ? scratch + kListAt : nullptr));
    return B200PT_OK;
}
