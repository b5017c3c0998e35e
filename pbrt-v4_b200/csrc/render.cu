// render.cu - the wavefront kernels around the traversal kernel: camera-ray generation, hit
// classification into material-sorted queues, the per-vertex shading step (emission, BSDF, light
// sampling with its shadow ray, Russian roulette), shadow-ray resolution, sample finish (sensor RGB)
// and film accumulation.  All path logic is in render_core.h (shared with the CPU check in
// tests/native/rendercheck.cpp); these kernels move a path between its structure-of-arrays slot and
// registers and feed the queues.
//
// Stage order per pass (render_api.cpp), mirroring WavefrontPathIntegrator::Render
// (wavefront/integrator.cpp:290-493):
//   generate -> for each depth { closest-hit ; classify ; shade<material> ... ; any-hit ; resolve }
//            -> finish -> film.
#include <cuda_runtime.h>

#include <algorithm>

#include "render_core.h"
#include "render_kernels.h"

namespace b200pt {

constexpr int kShadeBlock = 128;

static size_t Al(size_t b) { return (b + 255) & ~size_t(255); }

size_t PassBufferBytes(long long n, int depthSlots, bool vol) {
    size_t t = 0;
    if (vol)
        t += 5 * Al(sizeof(float) * kNLambda * n) + Al(sizeof(uint32_t) * kShadowAuxWords * n) +
             Al(sizeof(uint32_t) * 6 * n) + Al(sizeof(int));
    t += 2 * Al(sizeof(float) * kNLambda * n);         // beta, L
    t += Al(sizeof(float) * kPathScalars * n);
    t += Al(sizeof(int) * kPathInts * n);
    t += Al(sizeof(float) * 4 * n);                    // rgbw
    t += 2 * Al(sizeof(float) * 7 * n) + 2 * Al(sizeof(int) * n);  // rays, rayPath
    t += Al(sizeof(uint32_t) * 6 * n);                 // hits
    t += kMatQueues * Al(sizeof(int) * n);
    t += Al(sizeof(float) * 7 * n) + Al(sizeof(int) * n) + Al(sizeof(float) * kNLambda * n) + Al(size_t(n));
    t += Al(sizeof(int) * kCountStride * depthSlots);
    return t;
}

void CarvePassBuffers(void *base, long long n, int depthSlots, bool vol, PassBuffers &pb) {
    char *p = static_cast<char *>(base);
    auto take = [&](size_t b) {
        char *r = p;
        p += Al(b);
        return r;
    };
    pb.n = (int)n;
    pb.beta = (float *)take(sizeof(float) * kNLambda * n);
    pb.L = (float *)take(sizeof(float) * kNLambda * n);
    pb.scalars = (float *)take(sizeof(float) * kPathScalars * n);
    pb.ints = (int *)take(sizeof(int) * kPathInts * n);
    pb.rgbw = (float *)take(sizeof(float) * 4 * n);
    for (int i = 0; i < 2; ++i) pb.rays[i] = (float *)take(sizeof(float) * 7 * n);
    for (int i = 0; i < 2; ++i) pb.rayPath[i] = (int *)take(sizeof(int) * n);
    pb.hits = (uint32_t *)take(sizeof(uint32_t) * 6 * n);
    for (int m = 0; m < kMatQueues; ++m) pb.matQ[m] = (int *)take(sizeof(int) * n);
    pb.shRays = (float *)take(sizeof(float) * 7 * n);
    pb.shPath = (int *)take(sizeof(int) * n);
    pb.Ld = (float *)take(sizeof(float) * kNLambda * n);
    pb.occluded = (unsigned char *)take(size_t(n));
    pb.counts = (int *)take(sizeof(int) * kCountStride * depthSlots);
    pb.maxDepthSlots = depthSlots;
    pb.ru = pb.rl = pb.shT = pb.shRl = pb.shRu = nullptr;
    pb.shAux = pb.shHits = nullptr;
    pb.shContinue = nullptr;
    if (vol) {
        pb.ru = (float *)take(sizeof(float) * kNLambda * n);
        pb.rl = (float *)take(sizeof(float) * kNLambda * n);
        pb.shT = (float *)take(sizeof(float) * kNLambda * n);
        pb.shRl = (float *)take(sizeof(float) * kNLambda * n);
        pb.shRu = (float *)take(sizeof(float) * kNLambda * n);
        pb.shAux = (uint32_t *)take(sizeof(uint32_t) * kShadowAuxWords * n);
        pb.shHits = (uint32_t *)take(sizeof(uint32_t) * 6 * n);
        pb.shContinue = (int *)take(sizeof(int));
    }
}

// ---- queue push: one atomic per group of converged lanes
__device__ __forceinline__ int QueuePush(int *counter) {
    unsigned m = __activemask();
    int lane = threadIdx.x & 31, leader = __ffs(m) - 1, base = 0;
    if (lane == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(m, base, leader);
    return base + __popc(m & ((1u << lane) - 1));
}

__device__ __forceinline__ void StoreRay(float *rays, int i, V3 o, V3 d, float tMax) {
    float *q = rays + 7ll * i;
    q[0] = o.x; q[1] = o.y; q[2] = o.z;
    q[3] = d.x; q[4] = d.y; q[5] = d.z;
    q[6] = tMax;
}

// ---- path <-> SoA (beta is handled by the kernels: it is not needed by every path)
__device__ __forceinline__ void StoreHeader(const PassBuffers &pb, int i, const PathRegs &p) {
    const long long n = pb.n;
    float *s = pb.scalars + i;
    s[0 * n] = p.lambdaU;
    s[1 * n] = p.etaScale;
    s[2 * n] = p.pB;
    s[3 * n] = p.prev.pi.lo.x; s[4 * n] = p.prev.pi.lo.y; s[5 * n] = p.prev.pi.lo.z;
    s[6 * n] = p.prev.pi.hi.x; s[7 * n] = p.prev.pi.hi.y; s[8 * n] = p.prev.pi.hi.z;
    s[9 * n] = p.prev.n.x; s[10 * n] = p.prev.n.y; s[11 * n] = p.prev.n.z;
    s[12 * n] = p.prev.ns.x; s[13 * n] = p.prev.ns.y; s[14 * n] = p.prev.ns.z;
    int *q = pb.ints + i;
    q[0 * n] = p.sampleIndex;
    q[1 * n] = p.dimension;
    q[2 * n] = p.depth;
    q[3 * n] = (int)p.flags;
    q[4 * n] = p.medium;
}

__device__ __forceinline__ void LoadHeader(const PassBuffers &pb, int i, PathRegs &p) {
    const long long n = pb.n;
    const float *s = pb.scalars + i;
    p.lambdaU = s[0 * n];
    p.etaScale = s[1 * n];
    p.pB = s[2 * n];
    p.prev.pi.lo = mk(s[3 * n], s[4 * n], s[5 * n]);
    p.prev.pi.hi = mk(s[6 * n], s[7 * n], s[8 * n]);
    p.prev.n = mk(s[9 * n], s[10 * n], s[11 * n]);
    p.prev.ns = mk(s[12 * n], s[13 * n], s[14 * n]);
    const int *q = pb.ints + i;
    p.sampleIndex = q[0 * n];
    p.dimension = q[1 * n];
    p.depth = q[2 * n];
    p.flags = (uint32_t)q[3 * n];
    p.medium = q[4 * n];
    p.filterWeight = 1.f;
}

// IO of ShadeVertex on the device: radiance goes straight to the path's L, a light sample goes to a
// fresh slot of the shadow queue of this depth.
struct DevIO {
    const PassBuffers &pb;
    int path;
    int *nShadow;
    int slot;
    float *beta;  // this thread's column of the block's [31][kShadeBlock] shared array
    __device__ DevIO(const PassBuffers &b, int pathId, int *shadowCount, float *sbeta)
        : pb(b), path(pathId), nShadow(shadowCount), slot(-1), beta(sbeta) {}
    __device__ __forceinline__ float Beta(int k) const { return beta[k * kShadeBlock]; }
    __device__ __forceinline__ void SetBeta(int k, float v) { beta[k * kShadeBlock] = v; }
    __device__ __forceinline__ void AddL(int k, float v) { pb.L[(long long)k * pb.n + path] += v; }
    __device__ __forceinline__ void ShadowBegin() { slot = QueuePush(nShadow); }
    __device__ __forceinline__ void SetLd(int k, float v) { pb.Ld[(long long)k * pb.n + slot] = v; }
    __device__ __forceinline__ void ShadowEnd(bool valid, V3 o, V3 d, float tMax) {
        StoreRay(pb.shRays, slot, o, d, valid ? tMax : -1.f);
        pb.shPath[slot] = valid ? path : -1;
    }
    // volpath: r_u / r_l live in global memory ([k][path]); a light sample also records what its walk needs
    __device__ __forceinline__ float Ru(int k) const { return pb.ru[(long long)k * pb.n + path]; }
    __device__ __forceinline__ void SetRu(int k, float v) { pb.ru[(long long)k * pb.n + path] = v; }
    __device__ __forceinline__ float Rl(int k) const { return pb.rl[(long long)k * pb.n + path]; }
    __device__ __forceinline__ void SetRl(int k, float v) { pb.rl[(long long)k * pb.n + path] = v; }
    __device__ void ShadowEndVol(bool valid, V3 o, V3 d, float tMax, const ShadowVolInfo &info) {
        ShadowEnd(valid, o, d, tMax);
        if (!valid) return;
        uint32_t *a = pb.shAux + (long long)kShadowAuxWords * slot;
        a[0] = __float_as_uint(info.p_l);
        a[1] = __float_as_uint(info.scatterPDF);
        a[2] = (uint32_t)info.medium;
        a[3] = 0;  // the walk has not started: T_ray = r_l = r_u = 1, rng seeded from the ray
        const float lp[9] = {info.lightPi.lo.x, info.lightPi.lo.y, info.lightPi.lo.z, info.lightPi.hi.x, info.lightPi.hi.y,
                             info.lightPi.hi.z, info.lightN.x, info.lightN.y, info.lightN.z};
        for (int i = 0; i < 9; ++i) a[8 + i] = __float_as_uint(lp[i]);
    }
};

// IO of the medium kernel and of escaped rays: everything in global memory, indexed by path
struct GlobalIO {
    const PassBuffers &pb;
    int path;
    int *nShadow;
    int slot;
    __device__ GlobalIO(const PassBuffers &b, int pathId, int *shadowCount) : pb(b), path(pathId), nShadow(shadowCount), slot(-1) {}
    __device__ __forceinline__ float Beta(int k) const { return pb.beta[(long long)k * pb.n + path]; }
    __device__ __forceinline__ void SetBeta(int k, float v) { pb.beta[(long long)k * pb.n + path] = v; }
    __device__ __forceinline__ float Ru(int k) const { return pb.ru[(long long)k * pb.n + path]; }
    __device__ __forceinline__ void SetRu(int k, float v) { pb.ru[(long long)k * pb.n + path] = v; }
    __device__ __forceinline__ float Rl(int k) const { return pb.rl[(long long)k * pb.n + path]; }
    __device__ __forceinline__ void SetRl(int k, float v) { pb.rl[(long long)k * pb.n + path] = v; }
    __device__ __forceinline__ void AddL(int k, float v) { pb.L[(long long)k * pb.n + path] += v; }
    __device__ __forceinline__ void ShadowBegin() { slot = QueuePush(nShadow); }
    __device__ __forceinline__ void SetLd(int k, float v) { pb.Ld[(long long)k * pb.n + slot] = v; }
    __device__ void ShadowEndVol(bool valid, V3 o, V3 d, float tMax, const ShadowVolInfo &info) {
        DevIO tmp(pb, path, nShadow, nullptr);
        tmp.slot = slot;
        tmp.ShadowEndVol(valid, o, d, tMax, info);
    }
};

// IO of the medium kernel: beta, r_u, r_l staged in shared memory ([k][thread]) for the walk, which updates all
// three at every tentative collision; light samples go to the shadow queue like DevIO's
struct MediumIO {
    const PassBuffers &pb;
    int path;
    int *nShadow;
    int slot;
    float *sb;  // this thread's column of the block's [3][31][kShadeBlock] shared array
    __device__ MediumIO(const PassBuffers &b, int pathId, int *shadowCount, float *s)
        : pb(b), path(pathId), nShadow(shadowCount), slot(-1), sb(s) {}
    __device__ __forceinline__ float Beta(int k) const { return sb[k * kShadeBlock]; }
    __device__ __forceinline__ void SetBeta(int k, float v) { sb[k * kShadeBlock] = v; }
    __device__ __forceinline__ float Ru(int k) const { return sb[(kNLambda + k) * kShadeBlock]; }
    __device__ __forceinline__ void SetRu(int k, float v) { sb[(kNLambda + k) * kShadeBlock] = v; }
    __device__ __forceinline__ float Rl(int k) const { return sb[(2 * kNLambda + k) * kShadeBlock]; }
    __device__ __forceinline__ void SetRl(int k, float v) { sb[(2 * kNLambda + k) * kShadeBlock] = v; }
    __device__ __forceinline__ void AddL(int k, float v) { pb.L[(long long)k * pb.n + path] += v; }
    __device__ __forceinline__ void ShadowBegin() { slot = QueuePush(nShadow); }
    __device__ __forceinline__ void SetLd(int k, float v) { pb.Ld[(long long)k * pb.n + slot] = v; }
    __device__ void ShadowEndVol(bool valid, V3 o, V3 d, float tMax, const ShadowVolInfo &info) {
        DevIO tmp(pb, path, nShadow, nullptr);
        tmp.slot = slot;
        tmp.ShadowEndVol(valid, o, d, tMax, info);
    }
};

// ---- camera rays: RayIntegrator::EvaluatePixelSample up to GenerateRay (cpu/integrators.cpp:228-262)
__global__ void __launch_bounds__(kShadeBlock)
GenerateKernel(RenderSceneDev rs, PassBuffers pb, int nPixels, int width, int nActive, int firstSample,
               int sampleStride) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) pb.counts[0] = nActive;  // ray queue of depth 0: one camera ray per path, slot = path
    if (i >= nActive) return;
    int pix = i % nPixels, s = i / nPixels;
    PathRegs p;
    GenerateCameraPath(rs, p, rs.film.x0 + pix % width, rs.film.y0 + pix / width, firstSample + s * sampleStride);
    const long long n = pb.n;
#pragma unroll
    for (int k = 0; k < kNLambda; ++k) {
        pb.beta[k * n + i] = 1.f;
        pb.L[k * n + i] = 0.f;
    }
    if (pb.ru) {
#pragma unroll
        for (int k = 0; k < kNLambda; ++k) {
            pb.ru[k * n + i] = 1.f;
            pb.rl[k * n + i] = 1.f;
        }
    }
    StoreHeader(pb, i, p);
    pb.scalars[kCameraWeightScalar * n + i] = p.cameraWeight;
    // a sample without a camera ray (vignetted by the lens system) keeps its queue slot with a ray that
    // hits nothing (tMax < 0); the classify kernel drops it
    StoreRay(pb.rays[0], i, p.rayO, p.rayD, p.rayTMax);
    pb.rayPath[0][i] = i;
}

// ---- hits -> material-sorted queues (the role of EnqueueWorkAfterIntersection, wavefront/intersect.h:67-170)
__global__ void __launch_bounds__(256)
ClassifyKernel(RenderSceneDev rs, PassBuffers pb, int depth) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    int *counts = pb.counts + kCountStride * depth;
    if (t >= counts[0]) return;
    uint32_t slot = pb.hits[6ll * t];
    if (slot == kConsumedTag) return;  // absorbed or scattered in a medium (VolMediumKernel)
    if (slot == kMissTag) {  // escaped: infinite lights add their radiance, the path ends
        if (rs.nInfiniteLights && rs.integrator.type == 1 && pb.rays[depth & 1][7ll * t + 6] >= 0) {
            GlobalIO io(pb, pb.rayPath[depth & 1][t], nullptr);
            EscapedRadianceVol(rs, pb.scalars[io.path], io);
        } else if (rs.nInfiniteLights && pb.rays[depth & 1][7ll * t + 6] >= 0) {
            struct {
                const PassBuffers &pb;
                int path;
                __device__ float Beta(int k) const { return pb.beta[(long long)k * pb.n + path]; }
                __device__ void AddL(int k, float v) { pb.L[(long long)k * pb.n + path] += v; }
            } io{pb, pb.rayPath[depth & 1][t]};
            const long long n = pb.n;
            uint32_t flags = (uint32_t)pb.ints[3 * n + io.path];
            EscapedRadiance(rs, pb.scalars[io.path], depth == 0 || (flags & PF_SPECULAR), pb.scalars[2 * n + io.path], io);
        }
        return;
    }
    const PrimSlot *ps = reinterpret_cast<const PrimSlot *>(rs.prims) + slot;
    int material = rs.meshes[ps->mesh].material;
    uint32_t type = material >= 0 ? rs.materials[material].type : 0u;  // null material: any queue ends the path
    if (type >= (uint32_t)kMatQueues) type = 0;
    // per-type push: lanes of one type share an atomic
    for (uint32_t m = 0; m < (uint32_t)kMatQueues; ++m)
        if (type == m) pb.matQ[m][QueuePush(counts + 2 + m)] = t;
}

#ifndef B2_SHADE_MIN_BLOCKS
#define B2_SHADE_MIN_BLOCKS 4  // 128 registers: the layered walks are latency bound, four blocks per SM beat three
#endif
// ---- one path vertex: PathIntegrator::Li loop body (cpu/integrators.cpp:640-760)
template <int MAT, int LBVH, int VOL, int EXT>
__global__ void __launch_bounds__(kShadeBlock, B2_SHADE_MIN_BLOCKS)
ShadeKernel(RenderSceneDev rs, PassBuffers pb, int depth, int nPixels, int width) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    int *counts = pb.counts + kCountStride * depth;
    if (t >= counts[2 + MAT]) return;
    const int cur = depth & 1, nxt = cur ^ 1;
    int q = pb.matQ[MAT][t];
    int path = pb.rayPath[cur][q];
    const long long n = pb.n;
    // throughput in shared memory, [k][thread]: indexable by k without costing registers or local memory
    __shared__ float sbeta[kNLambda * kShadeBlock];
    float *myBeta = sbeta + threadIdx.x;
    PathRegs p;
    LoadHeader(pb, path, p);
#pragma unroll 8
    for (int k = 0; k < kNLambda; ++k) myBeta[k * kShadeBlock] = pb.beta[k * n + path];
    int pix = path % nPixels;
    p.px = rs.film.x0 + pix % width;
    p.py = rs.film.y0 + pix / width;
    const float *r = pb.rays[cur] + 7ll * q;
    p.rayO = mk(r[0], r[1], r[2]);
    p.rayD = mk(r[3], r[4], r[5]);
    p.rayTMax = r[6];
    uint32_t hit[6];
    const uint2 *h = reinterpret_cast<const uint2 *>(pb.hits + 6ll * q);
    uint2 h0 = h[0], h1 = h[1], h2 = h[2];
    hit[0] = h0.x; hit[1] = h0.y; hit[2] = h1.x; hit[3] = h1.y; hit[4] = h2.x; hit[5] = h2.y;
    DevIO io(pb, path, counts + 1, myBeta);
    ShadeVertex<DevIO, MAT, LBVH, VOL, EXT>(rs, p, hit, io);
    if (p.flags & PF_ALIVE) {
        int *next = pb.counts + kCountStride * (depth + 1);
        int slot = QueuePush(next);
        StoreRay(pb.rays[nxt], slot, p.rayO, p.rayD, p.rayTMax);
        pb.rayPath[nxt][slot] = path;
#pragma unroll 8
        for (int k = 0; k < kNLambda; ++k) pb.beta[k * n + path] = myBeta[k * kShadeBlock];
        StoreHeader(pb, path, p);
    } else {
        pb.ints[3 * n + path] = (int)p.flags;  // the finish kernel needs PF_TERMINATED_SECONDARY
    }
}

// ---- volpath: the rays of this iteration that travel in a medium (cpu/integrators.cpp:968-1075).  Absorbed and
// scattered paths are finished here (their hit slot is marked consumed); survivors go on to classify / shade
// with beta, r_u, r_l rescaled.
// Work lists by kind of medium, so that a warp walks one kind (a homogeneous medium is one segment and a few
// collisions, a density grid a DDA with many): mode 0 = the iteration's rays -> matQ[0] (homogeneous) / matQ[1]
// (grid), lengths in counts[6] / counts[7]; mode 1 = the live light-sample rays -> matQ[0] (no medium) / matQ[1]
// (homogeneous) / matQ[2] (grid), lengths in counts[2..4] (the material queues of this iteration are spent by then).
__global__ void __launch_bounds__(256)
VolSortKernel(RenderSceneDev rs, PassBuffers pb, int depth, int mode) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    int *counts = pb.counts + kCountStride * depth;
    const long long n = pb.n;
    if (mode == 0) {
        if (t >= counts[0]) return;
        if (pb.rays[depth & 1][7ll * t + 6] < 0) return;  // a sample without a camera ray
        int medium = pb.ints[4 * n + pb.rayPath[depth & 1][t]];
        if (medium < 0) return;
        int kind = rs.media[medium].type == 0 ? 0 : 1;
        for (int k = 0; k < 2; ++k)
            if (kind == k) pb.matQ[k][QueuePush(counts + 6 + k)] = t;
    } else {
        if (t >= counts[1]) return;
        if (pb.shPath[t] < 0) return;
        int medium = (int)pb.shAux[(long long)kShadowAuxWords * t + 2];
        int kind = medium < 0 ? 0 : (rs.media[medium].type == 0 ? 1 : 2);
        for (int k = 0; k < 3; ++k)
            if (kind == k) pb.matQ[k][QueuePush(counts + 2 + k)] = t;
    }
}

#ifndef B2_VOL_MIN_BLOCKS
#define B2_VOL_MIN_BLOCKS 3
#endif
template <int LBVH, int EXT>
__global__ void __launch_bounds__(kShadeBlock, B2_VOL_MIN_BLOCKS)
VolMediumKernel(RenderSceneDev rs, PassBuffers pb, int depth, int nPixels, int width, int kind) {
    int qi = blockIdx.x * blockDim.x + threadIdx.x;
    int *counts = pb.counts + kCountStride * depth;
    if (qi >= counts[6 + kind]) return;
    const int t = pb.matQ[kind][qi];
    const int cur = depth & 1, nxt = cur ^ 1;
    const long long n = pb.n;
    int path = pb.rayPath[cur][t];
    const float *r = pb.rays[cur] + 7ll * t;
    PathRegs p;
    LoadHeader(pb, path, p);
    int pix = path % nPixels;
    p.px = rs.film.x0 + pix % width;
    p.py = rs.film.y0 + pix / width;
    p.rayO = mk(r[0], r[1], r[2]);
    p.rayD = mk(r[3], r[4], r[5]);
    p.rayTMax = r[6];
    uint32_t slot = pb.hits[6ll * t];
    float tHit = slot == kMissTag ? INFINITY : __uint_as_float(pb.hits[6ll * t + 1]);
    __shared__ float sstate[3 * kNLambda * kShadeBlock];
    float *mine = sstate + threadIdx.x;
#pragma unroll 8
    for (int k = 0; k < kNLambda; ++k) {
        mine[k * kShadeBlock] = pb.beta[k * n + path];
        mine[(kNLambda + k) * kShadeBlock] = pb.ru[k * n + path];
        mine[(2 * kNLambda + k) * kShadeBlock] = pb.rl[k * n + path];
    }
    MediumIO io(pb, path, counts + 1, mine);
    int res = VolMediumStep<MediumIO, LBVH, EXT>(rs, p, tHit, io);
    if (res != VOL_TERMINATED) {
#pragma unroll 8
        for (int k = 0; k < kNLambda; ++k) {
            pb.beta[k * n + path] = mine[k * kShadeBlock];
            pb.ru[k * n + path] = mine[(kNLambda + k) * kShadeBlock];
            pb.rl[k * n + path] = mine[(2 * kNLambda + k) * kShadeBlock];
        }
    }
    if (res == VOL_SURVIVE) {
        StoreHeader(pb, path, p);
        return;
    }
    pb.hits[6ll * t] = kConsumedTag;
    if (res == VOL_TERMINATED) {
        pb.ints[3 * n + path] = (int)(p.flags & ~PF_ALIVE);
        return;
    }
    int *next = pb.counts + kCountStride * (depth + 1);
    int ns = QueuePush(next);
    StoreRay(pb.rays[nxt], ns, p.rayO, p.rayD, p.rayTMax);
    pb.rayPath[nxt][ns] = path;
    StoreHeader(pb, path, p);
}

// ---- volpath: one segment of every light sample's transmittance walk (cpu/integrators.cpp:1334-1395).  pb.shHits
// holds the closest hits of the shadow rays.  A walk that reaches the light adds its estimate to the path's
// radiance; one that meets an interface leaves its next ray in the same slot and counts in pb.shContinue.
template <int EXT>
__global__ void __launch_bounds__(kShadeBlock, B2_VOL_MIN_BLOCKS)
ShadowTrKernel(RenderSceneDev rs, PassBuffers pb, int depth, int kind) {
    int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= pb.counts[kCountStride * depth + 2 + kind]) return;
    const int t = pb.matQ[kind][qi];
    int path = pb.shPath[t];
    const long long n = pb.n;
    float *ray = pb.shRays + 7ll * t;
    uint32_t *a = pb.shAux + (long long)kShadowAuxWords * t;
    ShadowRegs sh;
    sh.o = mk(ray[0], ray[1], ray[2]);
    sh.d = mk(ray[3], ray[4], ray[5]);
    sh.medium = (int)a[2];
    sh.lightPi.lo = mk(__uint_as_float(a[8]), __uint_as_float(a[9]), __uint_as_float(a[10]));
    sh.lightPi.hi = mk(__uint_as_float(a[11]), __uint_as_float(a[12]), __uint_as_float(a[13]));
    sh.lightN = mk(__uint_as_float(a[14]), __uint_as_float(a[15]), __uint_as_float(a[16]));
    float T[kNLambda], rl[kNLambda], ru[kNLambda];
    const bool started = a[3] != 0;
    if (started) {
        sh.rng.state = (uint64_t)a[4] | ((uint64_t)a[5] << 32);
        sh.rng.inc = (uint64_t)a[6] | ((uint64_t)a[7] << 32);
        for (int k = 0; k < kNLambda; ++k) {
            T[k] = pb.shT[k * n + t];
            rl[k] = pb.shRl[k * n + t];
            ru[k] = pb.shRu[k * n + t];
        }
    } else {
        sh.rng = PcgSeed(Hash3f(sh.o.x, sh.o.y, sh.o.z), Hash3f(sh.d.x, sh.d.y, sh.d.z));
        for (int k = 0; k < kNLambda; ++k) T[k] = rl[k] = ru[k] = 1.f;
    }
    uint32_t hit[6];
    for (int i = 0; i < 6; ++i) hit[i] = pb.shHits[6ll * t + i];
    int res = ShadowTrSegment<EXT>(rs, pb.scalars[path], sh, hit, T, rl, ru);
    if (res == SHADOW_CONTINUE) {
        StoreRay(pb.shRays, t, sh.o, sh.d, ray[6]);
        a[2] = (uint32_t)sh.medium;
        a[3] = 1;
        a[4] = (uint32_t)sh.rng.state; a[5] = (uint32_t)(sh.rng.state >> 32);
        a[6] = (uint32_t)sh.rng.inc;   a[7] = (uint32_t)(sh.rng.inc >> 32);
        for (int k = 0; k < kNLambda; ++k) {
            pb.shT[k * n + t] = T[k];
            pb.shRl[k * n + t] = rl[k];
            pb.shRu[k * n + t] = ru[k];
        }
        atomicAdd(pb.shContinue, 1);
        return;
    }
    if (res == SHADOW_REACHED)
        ShadowTrFinish(T, rl, ru, __uint_as_float(a[0]), __uint_as_float(a[1]), [&](int k) { return pb.Ld[k * n + t]; },
                       [&](int k) { return pb.ru[k * n + path]; }, [&](int k, float v) { pb.L[k * n + path] += v; });
    pb.shPath[t] = -1;
    ray[6] = -1.f;
}

}  // namespace b200pt
#include "render_volume_warp.cuh"
namespace b200pt {

// ---- unoccluded light samples join the path's radiance (RecordShadowRayResult, wavefront/intersect.h:176-190)
__global__ void __launch_bounds__(256)
ResolveShadowKernel(PassBuffers pb, int depth) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= pb.counts[kCountStride * depth + 1]) return;
    int path = pb.shPath[t];
    if (path < 0 || pb.occluded[t]) return;
    const long long n = pb.n;
#pragma unroll
    for (int k = 0; k < kNLambda; ++k) pb.L[k * n + path] += pb.Ld[k * n + t];
}

// ---- finished sample -> weighted sensor RGB (RGBFilm::AddSample, film.h:241-258)
__global__ void __launch_bounds__(kShadeBlock)
FinishKernel(RenderSceneDev rs, PassBuffers pb, int nActive) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nActive) return;
    const long long n = pb.n;
    const float *L = pb.L + i;
    float rgbw[4];
    const float cw = pb.scalars[kCameraWeightScalar * n + i];
    FinishPath(rs, pb.scalars[i], (uint32_t)pb.ints[3 * n + i], 1.f, [&](int k) { return cw * L[k * n]; }, rgbw);
    reinterpret_cast<float4 *>(pb.rgbw)[i] = make_float4(rgbw[0], rgbw[1], rgbw[2], rgbw[3]);
}

// ---- film: one thread per pixel adds the pass's samples in sample order, as the CPU film does for
// a pixel (each pixel is owned by one thread there too), so the double sums are the reference's.
__global__ void FilmAddKernel(PassBuffers pb, int nPixels, int nSamplesInPass, double *film) {
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= nPixels) return;
    double2 *dst = reinterpret_cast<double2 *>(film) + 2ll * pix;
    double2 a = dst[0], b = dst[1];
    for (int s = 0; s < nSamplesInPass; ++s) {
        float4 v = reinterpret_cast<const float4 *>(pb.rgbw)[(long long)s * nPixels + pix];
        a.x += v.x;
        a.y += v.y;
        b.x += v.z;
        b.y += v.w;
    }
    dst[0] = a;
    dst[1] = b;
}

// Spectral film: after the RGB sums, each sample's 31 wavelengths go to their buckets (planes of n pixels,
// so neighbouring pixels' accesses to one bucket share sectors); same order as the CPU film: samples of a
// pixel in index order, wavelengths in order.
__global__ void FilmAddSpectralKernel(FilmDev fd, PassBuffers pb, int nPixels, int nSamplesInPass, double *film) {
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= nPixels) return;
    double *bucketSums = film + 4ll * nPixels + pix;
    double *weightSums = bucketSums + (long long)fd.nBuckets * nPixels;
    const long long n = pb.n;
    for (int s = 0; s < nSamplesInPass; ++s) {
        long long path = (long long)s * nPixels + pix;
        const float *L = pb.L + path;
        const float cw = pb.scalars[kCameraWeightScalar * n + path];
        SpectralFilmAdd(fd, pb.scalars[path], 1.f, [&](int k) { return cw * L[k * n]; },
                        [&](int b, float v, float w) {
                            bucketSums[(long long)b * nPixels] += v;
                            weightSums[(long long)b * nPixels] += w;
                        });
    }
}

// SpectralFilm::GetImage's spectral channels (film.cpp:1262-1272): bucketSums / weightSums in double, to float
__global__ void FilmSpectralImageKernel(int nBuckets, const double *acc, int nPixels, float *out) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= (long long)nPixels * nBuckets) return;
    int pix = (int)(i / nBuckets), b = (int)(i % nBuckets);
    double sum = acc[4ll * nPixels + (long long)b * nPixels + pix];
    double w = acc[4ll * nPixels + (long long)(nBuckets + b) * nPixels + pix];
    out[i] = w > 0 ? (float)(sum / w) : 0.f;
}

// RGBFilm::GetPixelRGB (film.h:261-277)
__global__ void FilmImageKernel(FilmDev film, const double *acc, int nPixels, float *rgb) {
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= nPixels) return;
    const double *a = acc + 4ll * pix;
    float r = (float)a[0], g = (float)a[1], b = (float)a[2], w = (float)a[3];
    if (w != 0) {
        r /= w;
        g /= w;
        b /= w;
    }
    const float *m = film.outputRGBFromSensorRGB;
    rgb[3ll * pix + 0] = m[0] * r + m[1] * g + m[2] * b;
    rgb[3ll * pix + 1] = m[3] * r + m[4] * g + m[5] * b;
    rgb[3ll * pix + 2] = m[6] * r + m[7] * g + m[8] * b;
}

__global__ void SumCountsKernel(PassBuffers pb, int depths, unsigned long long *totals) {
    unsigned long long c = 0, s = 0;
    for (int d = 0; d < depths; ++d) {
        c += (unsigned long long)pb.counts[kCountStride * d];
        s += (unsigned long long)pb.counts[kCountStride * d + 1];
    }
    totals[0] += c;
    totals[1] += s;
}

static inline int Blocks(long long n, int b) { return (int)((n + b - 1) / b); }

cudaError_t LaunchGenerate(const RenderSceneDev &rs, const PassBuffers &pb, int nPixels, int width, int nActive,
                           int firstSample, int sampleStride, cudaStream_t stream) {
    GenerateKernel<<<Blocks(nActive, kShadeBlock), kShadeBlock, 0, stream>>>(rs, pb, nPixels, width, nActive,
                                                                             firstSample, sampleStride);
    return cudaGetLastError();
}
cudaError_t LaunchClassify(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nActive,
                           cudaStream_t stream) {
    ClassifyKernel<<<Blocks(nActive, 256), 256, 0, stream>>>(rs, pb, depth);
    return cudaGetLastError();
}
cudaError_t LaunchShade(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nPixels, int width,
                        int nActive, uint32_t matMask, cudaStream_t stream, int *launches) {
    int grid = Blocks(nActive, kShadeBlock);
    const bool lbvh = rs.integrator.lightSampler == 2, vol = rs.integrator.type == 1;
    const bool ext = rs.geomExt != 0;  // patches / per-vertex normals or uv somewhere in the scene
#define B2_SHADE_EXT(MAT, L, V)                                                                         \
    do {                                                                                                \
        if (ext)                                                                                        \
            ShadeKernel<MAT, L, V, 1><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width); \
        else                                                                                            \
            ShadeKernel<MAT, L, V, 0><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width); \
    } while (0)
#define B2_SHADE(bit, MAT)                                                                              \
    if (matMask & (bit)) {                                                                              \
        if (vol && lbvh)                                                                                \
            B2_SHADE_EXT(MAT, 1, 1);                                                                    \
        else if (vol)                                                                                   \
            B2_SHADE_EXT(MAT, 0, 1);                                                                    \
        else if (lbvh)                                                                                  \
            B2_SHADE_EXT(MAT, 1, 0);                                                                    \
        else                                                                                            \
            B2_SHADE_EXT(MAT, 0, 0);                                                                    \
        ++*launches;                                                                                    \
    }
    B2_SHADE(1u, MAT_DIFFUSE)
    B2_SHADE(2u, MAT_CONDUCTOR)
    B2_SHADE(4u, MAT_DIELECTRIC)
    B2_SHADE(8u, MAT_COATED_DIFFUSE)
#undef B2_SHADE
#undef B2_SHADE_EXT
    return cudaGetLastError();
}
constexpr int kVolWarpGrid = 148 * 16;  // persistent-style grid of the warp-per-ray kernels (stride loop over the list)

cudaError_t LaunchVolMedium(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nPixels, int width, int nActive,
                            cudaStream_t stream, bool warpPerRay) {
    int grid = Blocks(nActive, kShadeBlock);
    VolSortKernel<<<Blocks(nActive, 256), 256, 0, stream>>>(rs, pb, depth, 0);
    for (int kind = 0; kind < 2 && warpPerRay; ++kind) {
        int g = (int)std::min<long long>(kVolWarpGrid, Blocks(nActive, kVolWarpBlock / 32));
        if (rs.integrator.lightSampler == 2)
            VolMediumWarpKernel<1><<<g, kVolWarpBlock, 0, stream>>>(rs, pb, depth, nPixels, width, kind);
        else
            VolMediumWarpKernel<0><<<g, kVolWarpBlock, 0, stream>>>(rs, pb, depth, nPixels, width, kind);
    }
    for (int kind = 0; kind < 2 && !warpPerRay; ++kind) {
        const bool lb = rs.integrator.lightSampler == 2;
        if (rs.geomExt) {
            if (lb)
                VolMediumKernel<1, 1><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width, kind);
            else
                VolMediumKernel<0, 1><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width, kind);
        } else {
            if (lb)
                VolMediumKernel<1, 0><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width, kind);
            else
                VolMediumKernel<0, 0><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width, kind);
        }
    }
    return cudaGetLastError();
}
cudaError_t LaunchShadowTr(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nActive, cudaStream_t stream,
                           bool warpPerRay) {
    // counts[2..4] of this iteration: the material queues are spent, the slots hold the walk queues' lengths
    cudaError_t e = cudaMemsetAsync(pb.counts + kCountStride * depth + 2, 0, 3 * sizeof(int), stream);
    if (e != cudaSuccess) return e;
    VolSortKernel<<<Blocks(nActive, 256), 256, 0, stream>>>(rs, pb, depth, 1);
    for (int kind = 0; kind < 3; ++kind) {
        // walks outside any medium (kind 0) only tell occluders from interfaces: a thread per ray is the better fit
        if (warpPerRay && kind > 0)
            ShadowTrWarpKernel<<<(int)std::min<long long>(kVolWarpGrid, Blocks(nActive, kVolWarpBlock / 32)), kVolWarpBlock, 0,
                                 stream>>>(rs, pb, depth, kind);
        else if (rs.geomExt)
            ShadowTrKernel<1><<<Blocks(nActive, kShadeBlock), kShadeBlock, 0, stream>>>(rs, pb, depth, kind);
        else
            ShadowTrKernel<0><<<Blocks(nActive, kShadeBlock), kShadeBlock, 0, stream>>>(rs, pb, depth, kind);
    }
    return cudaGetLastError();
}
cudaError_t LaunchResolveShadows(const PassBuffers &pb, int depth, int nActive, cudaStream_t stream) {
    ResolveShadowKernel<<<Blocks(nActive, 256), 256, 0, stream>>>(pb, depth);
    return cudaGetLastError();
}
cudaError_t LaunchFinish(const RenderSceneDev &rs, const PassBuffers &pb, int nActive, cudaStream_t stream) {
    FinishKernel<<<Blocks(nActive, kShadeBlock), kShadeBlock, 0, stream>>>(rs, pb, nActive);
    return cudaGetLastError();
}
cudaError_t LaunchFilmAdd(const PassBuffers &pb, int nPixels, int nSamplesInPass, double *film, cudaStream_t stream) {
    FilmAddKernel<<<Blocks(nPixels, 128), 128, 0, stream>>>(pb, nPixels, nSamplesInPass, film);
    return cudaGetLastError();
}
cudaError_t LaunchFilmAddSpectral(const FilmDev &film, const PassBuffers &pb, int nPixels, int nSamplesInPass,
                                  double *acc, cudaStream_t stream) {
    FilmAddSpectralKernel<<<Blocks(nPixels, 128), 128, 0, stream>>>(film, pb, nPixels, nSamplesInPass, acc);
    return cudaGetLastError();
}
cudaError_t LaunchFilmSpectralImage(const FilmDev &film, const double *acc, int nPixels, float *out,
                                    cudaStream_t stream) {
    long long n = (long long)nPixels * film.nBuckets;
    FilmSpectralImageKernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(film.nBuckets, acc, nPixels, out);
    return cudaGetLastError();
}
cudaError_t LaunchFilmImage(const FilmDev &film, const double *acc, int nPixels, float *rgb, cudaStream_t stream) {
    FilmImageKernel<<<Blocks(nPixels, 128), 128, 0, stream>>>(film, acc, nPixels, rgb);
    return cudaGetLastError();
}
cudaError_t LaunchSumCounts(const PassBuffers &pb, int depths, unsigned long long *totals, cudaStream_t stream) {
    SumCountsKernel<<<1, 1, 0, stream>>>(pb, depths, totals);
    return cudaGetLastError();
}

}  // namespace b200pt
