// api.cpp - the extern "C" boundary declared in include/b200pt.h.
// Host code only; kernels are reached through kernels.h.  No CPU fallback
// exists anywhere below: if there is no device, context creation fails.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/b200pt.h"
#include "bvh2.h"
#include "device_scene.h"
#include "kernels.h"
#include "scene_host.h"
#include "api_internal.h"

using namespace b200pt;

std::string g_createErr;

int EnsureExactScratch(b200pt_ctx *ctx, size_t n, int **out) {
    *out = nullptr;
    if (!ctx->wideKernel) return B200PT_OK;
    if (n + 1 > ctx->dExactCap) {
        if (ctx->dExact) {
            B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            cudaFree(ctx->dExact);
        }
        ctx->dExact = nullptr;
        ctx->dExactCap = 0;
        B2_CUDA(ctx, cudaMalloc((void **)&ctx->dExact, (n + 1) * sizeof(int)));
        ctx->dExactCap = n + 1;
    }
    *out = ctx->dExact;
    return B200PT_OK;
}


extern "C" {

int b200pt_version(void) { return B200PT_VERSION; }

int b200pt_device_count(int *count) {
    if (!count) return B200PT_ERR_INVALID;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        n = 0;
    }
    *count = n;
    return B200PT_OK;
}

int b200pt_context_create(int device, b200pt_ctx **out) {
    if (!out) return B200PT_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        (void)cudaGetLastError();
        g_createErr = "no CUDA device visible; b200pt has no CPU path";
        return B200PT_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= n) {
        g_createErr = "device index out of range";
        return B200PT_ERR_INVALID;
    }
    b200pt_ctx *c = new (std::nothrow) b200pt_ctx;
    if (!c) return B200PT_ERR_NOMEM;
    c->device = device;
    cudaDeviceProp prop;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        g_createErr = std::string("cudaSetDevice: ") + cudaGetErrorString(cudaGetLastError());
        delete c;
        return B200PT_ERR_CUDA;
    }
    if (prop.major < 10) {
        g_createErr = "device is not sm_100 class; b200pt ships sm_100a code only";
        delete c;
        return B200PT_ERR_NO_DEVICE;
    }
    c->smCount = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess ||
        cudaMalloc((void **)&c->dCounters, 8 * sizeof(unsigned long long)) != cudaSuccess) {
        g_createErr = std::string("context setup: ") + cudaGetErrorString(cudaGetLastError());
        delete c;
        return B200PT_ERR_CUDA;
    }
    c->ownStream = true;
    const char *mode = getenv("B200PT_TRAVERSAL");
    c->simpleKernel = mode && std::string(mode) == "simple";
    // default: the exact binary walk - measured faster than the 8-wide walk on every BASELINE config (profiles/r02_*);
    // B200PT_TRAVERSAL=wide selects the 8-wide walk
    c->wideKernel = mode && std::string(mode) == "wide";
    if (const char *v = getenv("B200PT_REFILL_MIN")) c->refillMin = std::max(1, std::min(32, atoi(v)));
    if (const char *v = getenv("B200PT_NODE_MIN")) c->nodeMin = std::max(0, std::min(32, atoi(v)));
    *out = c;
    return B200PT_OK;
}

int b200pt_context_set_stream(b200pt_ctx *ctx, void *cuda_stream) {
    if (!ctx) return B200PT_ERR_INVALID;
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->ownStream) {
        cudaStreamDestroy(ctx->stream);
        ctx->ownStream = false;
    }
    ctx->stream = (cudaStream_t)cuda_stream;  // NULL = the legacy default stream
    return B200PT_OK;
}

int b200pt_context_sync(b200pt_ctx *ctx) {
    if (!ctx) return B200PT_ERR_INVALID;
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200PT_OK;
}

int b200pt_context_destroy(b200pt_ctx *ctx) {
    if (!ctx) return B200PT_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->dIn) cudaFree(ctx->dIn);
    if (ctx->dOut) cudaFree(ctx->dOut);
    if (ctx->dCounters) cudaFree(ctx->dCounters);
    if (ctx->dExact) cudaFree(ctx->dExact);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->ownStream) cudaStreamDestroy(ctx->stream);
    for (cudaEvent_t e : ctx->evIn) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->evK) cudaEventDestroy(e);
    if (ctx->evStart) cudaEventDestroy(ctx->evStart);
    if (ctx->sIn) cudaStreamDestroy(ctx->sIn);
    if (ctx->sOut) cudaStreamDestroy(ctx->sOut);
    delete ctx;
    return B200PT_OK;
}

const char *b200pt_last_error(const b200pt_ctx *ctx) { return ctx ? ctx->err.c_str() : g_createErr.c_str(); }

int b200pt_scene_create(b200pt_ctx *ctx, const b200pt_scene_desc *desc, b200pt_scene **out) {
    if (!ctx || !out) return B200PT_ERR_INVALID;
    *out = nullptr;
    b200pt_scene *s = new (std::nothrow) b200pt_scene;
    if (!s) return B200PT_ERR_NOMEM;
    s->ctx = ctx;
    int rc = AssembleScene(desc, s->host, ctx->err);
    if (rc != B200PT_OK) {
        delete s;
        return rc;
    }
    auto t0 = std::chrono::steady_clock::now();
    HostScene &h = s->host;
    size_t nodeBytes = h.nodes.size() * sizeof(Node32), primBytes = h.prims.size() * sizeof(PrimSlot);
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e == cudaSuccess) e = cudaMalloc(&s->dNodes, nodeBytes);
    if (e == cudaSuccess) e = cudaMalloc(&s->dPrims, primBytes);
    size_t pairBytes = std::max<size_t>(h.pairs.size(), 1) * sizeof(PairNode);
    if (e == cudaSuccess) e = cudaMalloc(&s->dPairs, pairBytes);
    if (e == cudaSuccess && !h.pairs.empty())
        e = cudaMemcpyAsync(s->dPairs, h.pairs.data(), h.pairs.size() * sizeof(PairNode), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(s->dNodes, h.nodes.data(), nodeBytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(s->dPrims, h.prims.data(), primBytes, cudaMemcpyHostToDevice, ctx->stream);
    size_t wideBytes = h.wide.size() * sizeof(Wide8);
    if (e == cudaSuccess && wideBytes) e = cudaMalloc(&s->dWide, wideBytes);
    if (e == cudaSuccess && wideBytes)
        e = cudaMemcpyAsync(s->dWide, h.wide.data(), wideBytes, cudaMemcpyHostToDevice, ctx->stream);
    size_t leafBytes = h.leafBoxes.size() * sizeof(LeafBox);
    if (e == cudaSuccess && leafBytes) e = cudaMalloc(&s->dLeafBoxes, leafBytes);
    if (e == cudaSuccess && leafBytes)
        e = cudaMemcpyAsync(s->dLeafBoxes, h.leafBoxes.data(), leafBytes, cudaMemcpyHostToDevice, ctx->stream);
    size_t instBytes = h.instances.size() * sizeof(InstanceDev);
    if (e == cudaSuccess && instBytes) e = cudaMalloc(&s->dInst, instBytes);
    if (e == cudaSuccess && instBytes)
        e = cudaMemcpyAsync(s->dInst, h.instances.data(), instBytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) {
        ctx->err = std::string("scene upload: ") + cudaGetErrorString(e);
        b200pt_scene_destroy(s);
        return e == cudaErrorMemoryAllocation ? B200PT_ERR_NOMEM : B200PT_ERR_CUDA;
    }
    s->dev.nodes = s->dNodes;
    s->dev.prims = s->dPrims;
    s->dev.pairs = s->dPairs;
    std::memcpy(s->dev.rootBox, h.bounds, 24);
    s->dev.rootRef = h.rootRef;
    s->dev.wide = s->dWide;  // null when every tree is a single leaf: the binary kernel handles that
    s->dev.rootWide = h.rootWide;
    s->dev.leafBoxes = s->dLeafBoxes;
    s->dev.maxStackWide = h.maxStackWide;
    s->dev.instances = (const InstanceDev *)s->dInst;
    s->dev.rootNode = h.rootNode;
    s->dev.nNodes = uint32_t(h.nodes.size());
    s->dev.nPrims = uint32_t(h.prims.size());
    s->dev.nInstances = uint32_t(h.instances.size());
    s->dev.maxDepth = h.maxDepth;
    s->dev.hasPatches = h.hasPatches ? 1u : 0u;
    s->stats.n_triangles = h.nTriangles;
    s->stats.n_instances = h.instances.size();
    s->stats.n_bvh2_nodes = h.nodes.size();
    s->stats.n_wide_nodes = h.wide.size();
    s->stats.device_bytes = nodeBytes + primBytes + pairBytes + instBytes + wideBytes + leafBytes;
    s->stats.build_seconds = h.buildSeconds;
    s->stats.upload_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    *out = s;
    return B200PT_OK;
}

int b200pt_scene_destroy(b200pt_scene *s) {
    if (!s) return B200PT_OK;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    if (s->dNodes) cudaFree(s->dNodes);
    if (s->dPrims) cudaFree(s->dPrims);
    if (s->dPairs) cudaFree(s->dPairs);
    if (s->dWide) cudaFree(s->dWide);
    if (s->dLeafBoxes) cudaFree(s->dLeafBoxes);
    if (s->dInst) cudaFree(s->dInst);
    delete s;
    return B200PT_OK;
}

int b200pt_scene_bounds(const b200pt_scene *s, float b[6]) {
    if (!s || !b) return B200PT_ERR_INVALID;
    std::memcpy(b, s->host.bounds, 24);
    return B200PT_OK;
}

int b200pt_scene_get_stats(const b200pt_scene *s, b200pt_scene_stats *out) {
    if (!s || !out) return B200PT_ERR_INVALID;
    *out = s->stats;
    return B200PT_OK;
}

static int EnsureStaging(b200pt_ctx *ctx, size_t inBytes, size_t outBytes) {
    if (inBytes > ctx->dInCap) {
        if (ctx->dIn) cudaFree(ctx->dIn);
        ctx->dIn = nullptr;
        ctx->dInCap = 0;
        B2_CUDA(ctx, cudaMalloc(&ctx->dIn, inBytes));
        ctx->dInCap = inBytes;
    }
    if (outBytes > ctx->dOutCap) {
        if (ctx->dOut) cudaFree(ctx->dOut);
        ctx->dOut = nullptr;
        ctx->dOutCap = 0;
        B2_CUDA(ctx, cudaMalloc(&ctx->dOut, outBytes));
        ctx->dOutCap = outBytes;
    }
    return B200PT_OK;
}

// Host-buffer batches (B200PT_MEM_HOST) go through in chunks on three streams: while chunk k is traversed, chunk k + 1
// is on its way in and chunk k - 1 on its way out (PCIe is full duplex), so the call costs about the longest of the
// three legs instead of their sum.  Chunks are independent ray ranges: results are those of one launch.
static int IntersectHostPipelinedImpl(b200pt_scene *s, const float *rays, int64_t n, void *out, bool any);
static int IntersectHostPipelined(b200pt_scene *s, const float *rays, int64_t n, void *out, bool any) {
    int rc = IntersectHostPipelinedImpl(s, rays, n, out, any);
    if (rc != B200PT_OK) {
        // copies against the caller's buffers may still be in flight on the three streams: drain them before the
        // caller gets its buffers back (and may free them)
        b200pt_ctx *ctx = s->ctx;
        if (ctx->sIn) cudaStreamSynchronize(ctx->sIn);
        cudaStreamSynchronize(ctx->stream);
        if (ctx->sOut) cudaStreamSynchronize(ctx->sOut);
        (void)cudaGetLastError();
    }
    return rc;
}
static int IntersectHostPipelinedImpl(b200pt_scene *s, const float *rays, int64_t n, void *out, bool any) {
    b200pt_ctx *ctx = s->ctx;
    const size_t outRec = any ? 1 : sizeof(b200pt_hit);
    int64_t chunk = 1 << 21;
    if (const char *v = getenv("B200PT_HOST_CHUNK")) chunk = std::max<int64_t>(1 << 16, atoll(v));
    const int64_t nChunks = (n + chunk - 1) / chunk;
    int *exact = nullptr;
    if (int rc = EnsureExactScratch(ctx, size_t(std::min(chunk, n)), &exact)) return rc;
    if (!ctx->sIn) {
        B2_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->sIn, cudaStreamNonBlocking));
        B2_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->sOut, cudaStreamNonBlocking));
        B2_CUDA(ctx, cudaEventCreateWithFlags(&ctx->evStart, cudaEventDisableTiming));
    }
    while (int64_t(ctx->evIn.size()) < nChunks) {
        cudaEvent_t a, b;
        B2_CUDA(ctx, cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        ctx->evIn.push_back(a);
        B2_CUDA(ctx, cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        ctx->evK.push_back(b);
    }
    // whatever the caller queued on the context's stream before this call comes first
    B2_CUDA(ctx, cudaEventRecord(ctx->evStart, ctx->stream));
    B2_CUDA(ctx, cudaStreamWaitEvent(ctx->sIn, ctx->evStart, 0));
    B2_CUDA(ctx, cudaStreamWaitEvent(ctx->sOut, ctx->evStart, 0));
    B2_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    for (int64_t k = 0; k < nChunks; ++k) {
        const int64_t first = k * chunk, cnt = std::min(chunk, n - first);
        const float *dR = (const float *)ctx->dIn + 7 * first;
        char *dO = (char *)ctx->dOut + size_t(first) * outRec;
        B2_CUDA(ctx, cudaMemcpyAsync((void *)dR, rays + 7 * first, size_t(cnt) * 28, cudaMemcpyHostToDevice, ctx->sIn));
        B2_CUDA(ctx, cudaEventRecord(ctx->evIn[k], ctx->sIn));
        B2_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->evIn[k], 0));
        B2_CUDA(ctx, LaunchTraversePersist(s->dev, dR, cnt, any ? nullptr : (b200pt_hit *)dO, any ? (unsigned char *)dO : nullptr, any,
                                           (int *)(ctx->dCounters + 7), nullptr, ctx->smCount, int(s->dev.maxDepth) + 1,
                                           ctx->refillMin, ctx->nodeMin, ctx->stream, false, nullptr, exact));
        B2_CUDA(ctx, cudaEventRecord(ctx->evK[k], ctx->stream));
        B2_CUDA(ctx, cudaStreamWaitEvent(ctx->sOut, ctx->evK[k], 0));
        B2_CUDA(ctx, cudaMemcpyAsync((char *)out + size_t(first) * outRec, dO, size_t(cnt) * outRec, cudaMemcpyDeviceToHost, ctx->sOut));
    }
    B2_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->sOut));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200PT_OK;
}

static int Intersect(b200pt_scene *s, const float *rays, int64_t n, void *out, uint32_t mem, bool any) {
    if (!s || n < 0 || (n > 0 && (!rays || !out))) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = s->ctx;
    const bool device = (mem & 0xFF) == B200PT_MEM_DEVICE;
    const bool count = (mem & B200PT_TRACE_COUNTERS) != 0;
    if ((mem & 0xFF) > B200PT_MEM_DEVICE) {
        ctx->err = "unknown memory kind";
        return B200PT_ERR_INVALID;
    }
    s->trace = b200pt_trace_stats{};
    s->trace.rays = uint64_t(n);
    if (n == 0) return B200PT_OK;
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t outRec = any ? 1 : sizeof(b200pt_hit);
    const float *dRays = rays;
    void *dOut = out;
    if (!device) {
        int rc = EnsureStaging(ctx, size_t(n) * 28, size_t(n) * outRec);
        if (rc != B200PT_OK) return rc;
        if (!count && !ctx->simpleKernel && !getenv("B200PT_HOST_SERIAL")) return IntersectHostPipelined(s, rays, n, out, any);
        B2_CUDA(ctx, cudaMemcpyAsync(ctx->dIn, rays, size_t(n) * 28, cudaMemcpyHostToDevice, ctx->stream));
        dRays = (const float *)ctx->dIn;
        dOut = ctx->dOut;
    }
    if (count) B2_CUDA(ctx, cudaMemsetAsync(ctx->dCounters, 0, 8 * sizeof(unsigned long long), ctx->stream));
    B2_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    if (ctx->simpleKernel && s->dev.nInstances) {
        ctx->err = "B200PT_TRAVERSAL=simple (the parity kernel) does not handle object instances";
        return B200PT_ERR_INVALID;
    }
    int *exact = nullptr;
    if (!ctx->simpleKernel)
        if (int rc = EnsureExactScratch(ctx, size_t(n), &exact)) return rc;
    if (ctx->simpleKernel)
        B2_CUDA(ctx, LaunchTraverseBVH2(s->dev, dRays, n, any ? nullptr : (b200pt_hit *)dOut,
                                        any ? (unsigned char *)dOut : nullptr, any,
                                        count ? ctx->dCounters : nullptr, ctx->smCount, ctx->stream));
    else
        B2_CUDA(ctx, LaunchTraversePersist(s->dev, dRays, n, any ? nullptr : (b200pt_hit *)dOut,
                                           any ? (unsigned char *)dOut : nullptr, any,
                                           (int *)(ctx->dCounters + 7), count ? ctx->dCounters : nullptr,
                                           ctx->smCount, int(s->dev.maxDepth) + 1, ctx->refillMin, ctx->nodeMin,
                                           ctx->stream, false, nullptr, exact));
    B2_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    if (!device) {
        B2_CUDA(ctx, cudaMemcpyAsync(out, dOut, size_t(n) * outRec, cudaMemcpyDeviceToHost, ctx->stream));
        B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    if (count) {
        unsigned long long c[8];
        B2_CUDA(ctx, cudaMemcpyAsync(c, ctx->dCounters, sizeof c, cudaMemcpyDeviceToHost, ctx->stream));
        B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        s->trace.nodes_visited = c[0];
        s->trace.leaves_visited = c[1];
        s->trace.prims_tested = c[2];
        s->trace.wide_nodes_visited = c[3];
        s->trace.leaf_boxes_tested = c[4];
        if (exact) {
            int nExact = 0;
            B2_CUDA(ctx, cudaMemcpyAsync(&nExact, exact, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
            B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            s->trace.exact_rays = uint64_t(nExact);
        }
    }
    return B200PT_OK;
}

int b200pt_intersect_closest(b200pt_scene *s, const float *rays, int64_t n, b200pt_hit *hits, uint32_t mem) {
    return Intersect(s, rays, n, hits, mem, false);
}

int b200pt_intersect_any(b200pt_scene *s, const float *rays, int64_t n, uint8_t *occluded, uint32_t mem) {
    return Intersect(s, rays, n, occluded, mem, true);
}

int b200pt_get_trace_stats(const b200pt_scene *s, b200pt_trace_stats *out) {
    if (!s || !out) return B200PT_ERR_INVALID;
    b200pt_scene *ms = const_cast<b200pt_scene *>(s);
    // kernel time of the last call (waits for it)
    if (s->trace.rays && cudaEventSynchronize(s->ctx->ev1) == cudaSuccess) {
        float ms_ = 0;
        if (cudaEventElapsedTime(&ms_, s->ctx->ev0, s->ctx->ev1) == cudaSuccess)
            ms->trace.kernel_ms = ms_;
    }
    (void)cudaGetLastError();
    *out = s->trace;
    return B200PT_OK;
}

int b200pt_bvh2_build(const float *bounds, uint32_t n, int max_prims_in_node, void *nodes_out,
                      uint32_t *n_nodes_out, uint32_t *order_out) {
    if (!bounds || !n || !nodes_out || !n_nodes_out || !order_out || max_prims_in_node < 1)
        return B200PT_ERR_INVALID;
    std::vector<Node32> nodes;
    std::vector<uint32_t> order;
    BuildBVH2(reinterpret_cast<const Box *>(bounds), n, max_prims_in_node, nodes, order);
    std::memcpy(nodes_out, nodes.data(), nodes.size() * sizeof(Node32));
    std::memcpy(order_out, order.data(), order.size() * sizeof(uint32_t));
    *n_nodes_out = uint32_t(nodes.size());
    return B200PT_OK;
}

}  // extern "C"
