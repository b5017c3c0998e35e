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

#include "../../include/b200pt.h"
#include "render_core.h"
#include "render_kernels.h"

namespace b200pt {

constexpr int kShadeBlock = 128;

static size_t Al(size_t b) { return (b + 255) & ~size_t(255); }

// The translation unit is compiled once per part (-DB2_RENDER_PART=k, pbrt-v4_b200/_build.py) so that the heavy template
// kernels build in parallel: part 0 = everything but the kernels below + the diffuse shade kernels, parts 1-3 = the
// conductor / dielectric / coated-diffuse shade kernels, part 4 = the volume-walk kernels.
#ifndef B2_RENDER_PART
#error "compile with -DB2_RENDER_PART=0..4"
#endif

#if B2_RENDER_PART == 0
size_t PassBufferBytes(long long n, int depthSlots, bool vol, bool aov) {
    size_t t = 0;
    if (aov) t += Al(sizeof(float) * kAovWords * n);
    if (vol)
        t += 5 * Al(sizeof(float) * kNLambda * n) + Al(sizeof(uint32_t) * kShadowAuxWords * n) +
             Al(sizeof(uint32_t) * 6 * n) + Al(2 * sizeof(int));
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

void CarvePassBuffers(void *base, long long n, int depthSlots, bool vol, PassBuffers &pb, bool aov) {
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
    pb.aov = aov ? (float *)take(sizeof(float) * kAovWords * n) : nullptr;
    if (vol) {
        pb.ru = (float *)take(sizeof(float) * kNLambda * n);
        pb.rl = (float *)take(sizeof(float) * kNLambda * n);
        pb.shT = (float *)take(sizeof(float) * kNLambda * n);
        pb.shRl = (float *)take(sizeof(float) * kNLambda * n);
        pb.shRu = (float *)take(sizeof(float) * kNLambda * n);
        pb.shAux = (uint32_t *)take(sizeof(uint32_t) * kShadowAuxWords * n);
        pb.shHits = (uint32_t *)take(sizeof(uint32_t) * 6 * n);
        pb.shContinue = (int *)take(2 * sizeof(int));  // [0] walks that go on, [1] work cursor of the persistent walk kernels
    }
}

#endif  // part 0

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
    float *refl;  // the same for the vertex's cached reflectance (diffuse / coated diffuse kernels; null elsewhere)
    __device__ DevIO(const PassBuffers &b, int pathId, int *shadowCount, float *sbeta, float *srefl = nullptr)
        : pb(b), path(pathId), nShadow(shadowCount), slot(-1), beta(sbeta), refl(srefl) {}
    __device__ __forceinline__ float R(int k) const { return refl[k * kShadeBlock]; }
    __device__ __forceinline__ void SetR(int k, float v) { refl[k * kShadeBlock] = v; }
    __device__ __forceinline__ float Beta(int k) const { return beta[k * kShadeBlock]; }
    __device__ __forceinline__ void SetBeta(int k, float v) { beta[k * kShadeBlock] = v; }
    __device__ __forceinline__ void AddL(int k, float v) { pb.L[(long long)k * pb.n + path] += v; }
    __device__ __forceinline__ void VisibleSurface(V3 p, V3 n, V3 ns, int material, int instance) {
        if (!pb.aov) return;
        float *a = pb.aov + path;
        const long long N = pb.n;
        a[0] = 1.f;
        a[1 * N] = p.x; a[2 * N] = p.y; a[3 * N] = p.z;
        a[4 * N] = n.x; a[5 * N] = n.y; a[6 * N] = n.z;
        a[7 * N] = ns.x; a[8 * N] = ns.y; a[9 * N] = ns.z;
        a[10 * N] = (float)material;
        a[11 * N] = (float)instance;
    }
    // the surface's albedo as XYZ, under the wavelength pdf of a path that keeps all its wavelengths [0..2] and of one
    // that ends with the secondary ones terminated [3..5]: the film converts with the lambda the path ENDS with
    __device__ __forceinline__ bool WantsAlbedo() const { return pb.aov != nullptr; }
    __device__ __forceinline__ void VisibleSurfaceAlbedo(const float xyz[6]) {
        float *a = pb.aov + path;
        const long long N = pb.n;
        for (int i = 0; i < 6; ++i) a[(12 + i) * N] = xyz[i];
    }
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
    __device__ __forceinline__ void VisibleSurface(V3, V3, V3, int, int) {}
    __device__ __forceinline__ bool WantsAlbedo() const { return false; }
    __device__ __forceinline__ void VisibleSurfaceAlbedo(const float *) {}
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

#if B2_RENDER_PART == 0
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
    if (pb.aov) pb.aov[i] = 0.f;  // no VisibleSurface yet
    StoreHeader(pb, i, p);
    pb.scalars[kCameraWeightScalar * n + i] = p.cameraWeight;
    pb.scalars[kFilterWeightScalar * n + i] = p.filterWeight;
    // a sample without a camera ray (vignetted by the lens system) keeps its queue slot with a ray that
    // hits nothing (tMax < 0); the classify kernel drops it
    StoreRay(pb.rays[0], i, p.rayO, p.rayD, p.rayTMax);
    pb.rayPath[0][i] = i;
}

// ---- hits -> material-sorted queues (the role of EnqueueWorkAfterIntersection, wavefront/intersect.h:67-170)
__global__ void __launch_bounds__(256, 3)  // the escaped-ray branch (image infinite light) must not set the register count
ClassifyKernel(RenderSceneDev rs, PassBuffers pb, int depth) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    int *counts = pb.counts + kCountStride * depth;
    if (t >= counts[0]) return;
    uint32_t slot = pb.hits[6ll * t];
    if (slot == kConsumedTag) return;  // absorbed or scattered in a medium (VolMediumKernel)
    if (slot == kMissTag) {  // escaped: infinite lights add their radiance, the path ends
        const float *ray = pb.rays[depth & 1] + 7ll * t;
        auto escapeCtx = [&](int path) {  // the ray and the previous vertex' light-sample context (StoreHeader's layout)
            const long long n = pb.n;
            const float *s = pb.scalars + path;
            EscapeCtx ec;
            ec.unweighted = depth == 0 || ((uint32_t)pb.ints[3 * n + path] & PF_SPECULAR) != 0;
            ec.d = mk(ray[3], ray[4], ray[5]);
            ec.prevP = (mk(s[3 * n], s[4 * n], s[5 * n]) + mk(s[6 * n], s[7 * n], s[8 * n])) / 2;
            ec.prevNs = mk(s[12 * n], s[13 * n], s[14 * n]);
            return ec;
        };
        const bool lightBvh = rs.integrator.lightSampler == 2;
        if (rs.nInfiniteLights && rs.integrator.type == 1 && pb.rays[depth & 1][7ll * t + 6] >= 0) {
            GlobalIO io(pb, pb.rayPath[depth & 1][t], nullptr);
            EscapedRadianceVol(rs, pb.scalars[io.path], io, escapeCtx(io.path), lightBvh);
        } else if (rs.nInfiniteLights && pb.rays[depth & 1][7ll * t + 6] >= 0) {
            struct {
                const PassBuffers &pb;
                int path;
                __device__ float Beta(int k) const { return pb.beta[(long long)k * pb.n + path]; }
                __device__ void AddL(int k, float v) { pb.L[(long long)k * pb.n + path] += v; }
            } io{pb, pb.rayPath[depth & 1][t]};
            const long long n = pb.n;
            EscapedRadiance(rs, pb.scalars[io.path], pb.scalars[2 * n + io.path], io, escapeCtx(io.path), lightBvh);
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

#endif  // part 0

#ifndef B2_SHADE_MIN_BLOCKS
#define B2_SHADE_MIN_BLOCKS 4  // 128 registers: the layered walks are latency bound, four blocks per SM beat three
#endif
// ---- one path vertex: PathIntegrator::Li loop body (cpu/integrators.cpp:640-760)
template <int MAT, int LBVH, int VOL, int EXT>
#ifndef B2_COATED_MIN_BLOCKS
#define B2_COATED_MIN_BLOCKS 6  // the layered walks issue in 9 % of the slots: more warps beat fewer spills (config 3: 99.5 -> 101.8 Msamples/s)
#endif
__global__ void __launch_bounds__(kShadeBlock, MAT == (int)MAT_COATED_DIFFUSE ? B2_COATED_MIN_BLOCKS : B2_SHADE_MIN_BLOCKS)
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
    constexpr bool kCachesR = MAT == (int)MAT_DIFFUSE || MAT == (int)MAT_COATED_DIFFUSE;
    __shared__ float srefl[kCachesR ? kNLambda * kShadeBlock : 1];
    DevIO io(pb, path, counts + 1, myBeta, kCachesR ? srefl + threadIdx.x : nullptr);
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

#if B2_RENDER_PART == 4
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
// The two walk kernels are PERSISTENT with lane refill: a thread takes rays off the work lists until they are used up,
// and the warp executes one common body per round - one step of the delta / ratio-tracking walk (TMajWalk::Step) for
// every lane that is walking.  Walk lengths differ by orders of magnitude between rays (vacuum, thin fog, the dense
// grid): with one ray per thread a warp waited for its longest walk at ~6 active lanes (profiles/r01_config5_*).
// Starting a ray (header, spectra, DDA set-up) and finishing one (light / phase sampling, stores) are long divergent
// blocks, so they are batched too: lanes wait until kVolBatch of them are ready (or nobody is walking).
constexpr int kVolBatch = 8;
enum : int { W_IDLE = 0, W_WALK = 1, W_END = 2, W_OUT = 3, W_COLLIDE = 4 };
constexpr int kVolCellsPerRound = 4;  // majorant segments a walking lane may cross per round

// takes `nIdle` consecutive items off the cursor for the idle lanes; -1 = the lists are used up
__device__ __forceinline__ int VolFetch(int *cursor, unsigned idleMask, bool idle, int total) {
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == __ffs(idleMask) - 1) base = atomicAdd(cursor, __popc(idleMask));
    base = __shfl_sync(0xffffffffu, base, __ffs(idleMask) - 1);
    if (!idle) return -1;
    int qi = base + __popc(idleMask & ((1u << lane) - 1));
    return qi < total ? qi : -1;
}

template <int LBVH, int EXT>
__global__ void __launch_bounds__(kShadeBlock, B2_VOL_MIN_BLOCKS)
VolMediumKernel(RenderSceneDev rs, PassBuffers pb, int depth, int nPixels, int width, int *cursor) {
    int *counts = pb.counts + kCountStride * depth;
    const int n0 = counts[6], total = n0 + counts[7];  // homogeneous media first, then grids (VolSortKernel)
    const int cur = depth & 1, nxt = cur ^ 1;
    const long long n = pb.n;
    __shared__ float sstate[3 * kNLambda * kShadeBlock];
    float *mine = sstate + threadIdx.x;
    const unsigned FULL = 0xffffffffu;
    int state = W_IDLE, t = 0, path = 0;
    bool exhausted = false;
    PathRegs p;
    VolMediumState st;
    TMajFastWalk fw;
    bool tail = false;
    while (true) {
        const unsigned mAdv = __ballot_sync(FULL, state == W_WALK), mColl = __ballot_sync(FULL, state == W_COLLIDE);
        const unsigned mWalk = mAdv | mColl;
        unsigned mEnd = __ballot_sync(FULL, state == W_END);
        if (mEnd && (__popc(mEnd) >= kVolBatch || !mWalk)) {
            if (state == W_END) {
                MediumIO io(pb, path, counts + 1, mine);
                if (tail) fw.TMaj(st.sigT, st.T_maj);
                tail = false;
                int res = VolMediumEnd<MediumIO, LBVH, EXT>(rs, p, st, io);
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
                } else {
                    pb.hits[6ll * t] = kConsumedTag;
                    if (res == VOL_TERMINATED) {
                        pb.ints[3 * n + path] = (int)(p.flags & ~PF_ALIVE);
                    } else {
                        int *next = pb.counts + kCountStride * (depth + 1);
                        int ns = QueuePush(next);
                        StoreRay(pb.rays[nxt], ns, p.rayO, p.rayD, p.rayTMax);
                        pb.rayPath[nxt][ns] = path;
                        StoreHeader(pb, path, p);
                    }
                }
                state = W_IDLE;
            }
        }
        unsigned mIdle = __ballot_sync(FULL, state == W_IDLE);
        if (mIdle && !exhausted && (__popc(mIdle) >= kVolBatch || !mWalk)) {
            int qi = VolFetch(cursor, mIdle, state == W_IDLE, total);
            if (state == W_IDLE) {
                if (qi < 0) {
                    state = W_OUT;
                } else {
                    t = qi < n0 ? pb.matQ[0][qi] : pb.matQ[1][qi - n0];
                    path = pb.rayPath[cur][t];
                    const float *r = pb.rays[cur] + 7ll * t;
                    LoadHeader(pb, path, p);
                    int pix = path % nPixels;
                    p.px = rs.film.x0 + pix % width;
                    p.py = rs.film.y0 + pix / width;
                    p.rayO = mk(r[0], r[1], r[2]);
                    p.rayD = mk(r[3], r[4], r[5]);
                    p.rayTMax = r[6];
                    uint32_t slot = pb.hits[6ll * t];
                    float tHit = slot == kMissTag ? INFINITY : __uint_as_float(pb.hits[6ll * t + 1]);
#pragma unroll 8
                    for (int k = 0; k < kNLambda; ++k) {
                        mine[k * kShadeBlock] = pb.beta[k * n + path];
                        mine[(kNLambda + k) * kShadeBlock] = pb.ru[k * n + path];
                        mine[(2 * kNLambda + k) * kShadeBlock] = pb.rl[k * n + path];
                    }
                    VolMediumBeginFast(rs, p, tHit, st, fw);
                    state = W_WALK;
                }
            }
            if (__any_sync(FULL, state == W_OUT)) exhausted = true;
        }
        if (!__any_sync(FULL, state == W_WALK || state == W_COLLIDE || state == W_END || (state == W_IDLE && !exhausted))) break;
        // the pending tentative collisions once enough lanes have one (or nobody is left to advance), else the scalar walk
        if (mColl && (__popc(mColl) >= kVolBatch || !mAdv)) {
            if (state == W_COLLIDE) {
                MediumIO io(pb, path, counts + 1, mine);
                state = VolMediumCollide(rs, p, st, fw, io) ? W_WALK : W_END;
            }
        } else if (state == W_WALK) {
            int r = fw.Advance(rs.media[p.medium], st.sigT[0], st.rng, kVolCellsPerRound);
            if (r == TMAJ_COLLISION) {
                state = W_COLLIDE;
            } else if (r == TMAJ_RAN_OUT) {
                tail = true;  // T_maj of the stretch after the last collision: formed with the batch in W_END
                state = W_END;
            }
        }
    }
}

// ---- volpath: one segment of every light sample's transmittance walk (cpu/integrators.cpp:1334-1395).  pb.shHits
// holds the closest hits of the shadow rays.  A walk that reaches the light adds its estimate to the path's
// radiance; one that meets an interface leaves its next ray in the same slot and counts in pb.shContinue.
// Work lists: walks outside any medium, in a homogeneous medium, in a grid (counts[2..4], VolSortKernel).
template <int EXT>
__global__ void __launch_bounds__(kShadeBlock, B2_VOL_MIN_BLOCKS)
ShadowTrKernel(RenderSceneDev rs, PassBuffers pb, int depth, int *cursor) {
    const int *counts = pb.counts + kCountStride * depth;
    const int n0 = counts[2], n1 = n0 + counts[3], total = n1 + counts[4];
    const long long n = pb.n;
    const unsigned FULL = 0xffffffffu;
    int state = W_IDLE, t = 0, path = 0, res = -1;
    bool exhausted = false;
    ShadowRegs sh;
    ShadowTrState st;
    TMajFastWalk fw;
    bool tail = false;
    float T[kNLambda], rl[kNLambda], ru[kNLambda];
    while (true) {
        const unsigned mAdv = __ballot_sync(FULL, state == W_WALK), mColl = __ballot_sync(FULL, state == W_COLLIDE);
        const unsigned mWalk = mAdv | mColl;
        unsigned mEnd = __ballot_sync(FULL, state == W_END);
        if (mEnd && (__popc(mEnd) >= kVolBatch || !mWalk)) {
            if (state == W_END) {
                float *ray = pb.shRays + 7ll * t;
                uint32_t *a = pb.shAux + (long long)kShadowAuxWords * t;
                if (tail) fw.TMaj(st.sigT, st.T_maj);
                tail = false;
                if (res < 0) res = ShadowTrEnd(rs, sh, st, T, rl, ru);
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
                } else {
                    if (res == SHADOW_REACHED)
                        ShadowTrFinish(T, rl, ru, __uint_as_float(a[0]), __uint_as_float(a[1]),
                                       [&](int k) { return pb.Ld[k * n + t]; }, [&](int k) { return pb.ru[k * n + path]; },
                                       [&](int k, float v) { pb.L[k * n + path] += v; });
                    pb.shPath[t] = -1;
                    ray[6] = -1.f;
                }
                state = W_IDLE;
            }
        }
        unsigned mIdle = __ballot_sync(FULL, state == W_IDLE);
        if (mIdle && !exhausted && (__popc(mIdle) >= kVolBatch || !mWalk)) {
            int qi = VolFetch(cursor, mIdle, state == W_IDLE, total);
            if (state == W_IDLE) {
                if (qi < 0) {
                    state = W_OUT;
                } else {
                    t = qi < n0 ? pb.matQ[0][qi] : (qi < n1 ? pb.matQ[1][qi - n0] : pb.matQ[2][qi - n1]);
                    path = pb.shPath[t];
                    const float *ray = pb.shRays + 7ll * t;
                    const uint32_t *a = pb.shAux + (long long)kShadowAuxWords * t;
                    sh.o = mk(ray[0], ray[1], ray[2]);
                    sh.d = mk(ray[3], ray[4], ray[5]);
                    sh.medium = (int)a[2];
                    sh.lightPi.lo = mk(__uint_as_float(a[8]), __uint_as_float(a[9]), __uint_as_float(a[10]));
                    sh.lightPi.hi = mk(__uint_as_float(a[11]), __uint_as_float(a[12]), __uint_as_float(a[13]));
                    sh.lightN = mk(__uint_as_float(a[14]), __uint_as_float(a[15]), __uint_as_float(a[16]));
                    if (a[3] != 0) {  // the walk goes on from an earlier segment
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
                    res = ShadowTrBeginFast<EXT>(rs, pb.scalars[path], sh, hit, st, fw);
                    state = (res < 0 && st.walking) ? W_WALK : W_END;
                }
            }
            if (__any_sync(FULL, state == W_OUT)) exhausted = true;
        }
        if (!__any_sync(FULL, state == W_WALK || state == W_COLLIDE || state == W_END || (state == W_IDLE && !exhausted))) break;
        if (mColl && (__popc(mColl) >= kVolBatch || !mAdv)) {
            if (state == W_COLLIDE) state = ShadowTrCollide(rs, sh, st, fw, T, rl, ru) ? W_WALK : W_END;
        } else if (state == W_WALK) {
            int r = fw.Advance(rs.media[sh.medium], st.sigT[0], sh.rng, kVolCellsPerRound);
            if (r == TMAJ_COLLISION) {
                state = W_COLLIDE;
            } else if (r == TMAJ_RAN_OUT) {
                tail = true;
                state = W_END;
            }
        }
    }
}


#endif  // part 4

#if B2_RENDER_PART == 0
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
    FinishPath(rs, pb.scalars[i], (uint32_t)pb.ints[3 * n + i], pb.scalars[kFilterWeightScalar * n + i],
               [&](int k) { return cw * L[k * n]; }, rgbw);
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
        SpectralFilmAdd(fd, pb.scalars[path], pb.scalars[kFilterWeightScalar * n + path], [&](int k) { return cw * L[k * n]; },
                        [&](int b, float v, float w) {
                            bucketSums[(long long)b * nPixels] += v;
                            if (fd.type == 1) weightSums[(long long)b * nPixels] += w;  // (GBufferFilm never reads them)
                        });
    }
}

// GBufferFilm::AddSample (film.cpp:597-666) beyond the RGB sums and the 31 radiance buckets (which are the spectral
// film's: FilmAddSpectralKernel): the VisibleSurface AOVs.  One thread per pixel, samples in index order; the sums are
// kept in double (the reference's are float, its variance a float Welford update: equal within the image tolerance);
// materialId / instanceId are the last sample's, as in the reference, and live outside the summed film (`ids`).
__global__ void FilmAddGBufferKernel(FilmDev fd, PassBuffers pb, int nPixels, int nSamplesInPass, double *film, float *ids,
                                     const int32_t *materialIds) {
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= nPixels) return;
    double *g = film + (4ll + fd.nBuckets) * nPixels + pix;  // plane q at g[q * nPixels] (no weightSums planes in this film)
    const long long n = pb.n, P = nPixels;
    const float *M = fd.outM, *I = fd.outMInv;
    double acc[kGBufferPlanes];
    for (int q = 0; q < kGBufferPlanes; ++q) acc[q] = g[q * P];
    float idM = ids[pix], idI = ids[P + pix];
    for (int s = 0; s < nSamplesInPass; ++s) {
        const long long path = (long long)s * nPixels + pix;
        const float4 v = reinterpret_cast<const float4 *>(pb.rgbw)[path];
        const float *a = pb.aov + path;
        if (a[0] != 0.f && v.w != 0.f) {
            const float w = v.w;
            const float px = a[1 * n], py = a[2 * n], pz = a[3 * n];
            // Transform::operator()(Point3f) (util/transform.h:270-281) and (Normal3f) = transpose(mInv) n (:300-306)
            float xp = M[0] * px + M[1] * py + M[2] * pz + M[3], yp = M[4] * px + M[5] * py + M[6] * pz + M[7];
            float zp = M[8] * px + M[9] * py + M[10] * pz + M[11], wp = M[12] * px + M[13] * py + M[14] * pz + M[15];
            if (wp != 1) { xp /= wp; yp /= wp; zp /= wp; }
            acc[0] += w;
            acc[1] += w * xp; acc[2] += w * yp; acc[3] += w * zp;
            for (int k = 0; k < 2; ++k) {
                const float nx = a[(4 + 3 * k) * n], ny = a[(5 + 3 * k) * n], nz = a[(6 + 3 * k) * n];
                acc[4 + 3 * k] += w * (I[0] * nx + I[4] * ny + I[8] * nz);
                acc[5 + 3 * k] += w * (I[1] * nx + I[5] * ny + I[9] * nz);
                acc[6 + 3 * k] += w * (I[2] * nx + I[6] * ny + I[10] * nz);
            }
            const float rgb[3] = {v.x / w, v.y / w, v.z / w};
            for (int c = 0; c < 3; ++c) {
                acc[10 + c] += rgb[c];
                acc[13 + c] += double(rgb[c]) * rgb[c];
            }
            acc[16] += 1;
            if (fd.gbufferAlbedo) {  // rgbAlbedoSum += weight * (albedo * illuminant).ToRGB(lambda, colorSpace) (film.cpp:640-644)
                const float *x = a + (((uint32_t)pb.ints[3 * n + path] & PF_TERMINATED_SECONDARY) ? 15 : 12) * n;
                const float X = x[0], Y = x[n], Z = x[2 * n];
                const float *m = fd.rgbFromXYZ;
                acc[17] += w * (m[0] * X + m[1] * Y + m[2] * Z);
                acc[18] += w * (m[3] * X + m[4] * Y + m[5] * Z);
                acc[19] += w * (m[6] * X + m[7] * Y + m[8] * Z);
            }
        }
        // p.materialId = visibleSurface->materialId: every sample writes it.  A sample without a VisibleSurface (no camera
        // ray through the lens, an escaped ray) hands the film a default-constructed one whose ids are never initialised
        // (film.h:155-156): in the CPU build that is the stack slot of the previous sample's, so the ids of the last
        // sample that HAD a surface stay - which is what is kept here (within the pixel; 0 before the first)
        if (a[0] != 0.f) {
            const int mi = (int)a[10 * n];
            idM = mi < 0 ? 0.f : (materialIds ? (float)materialIds[mi] : (float)(mi + 1));
            idI = a[11 * n];
        }
    }
    for (int q = 0; q < kGBufferPlanes; ++q) g[q * P] = acc[q];
    ids[pix] = idM;
    ids[P + pix] = idI;
}

// GBufferFilm::GetImage (film.cpp:743-1044), channels-last float image.  Radiance.Cxx = L[b] / rgbWeightSum - the
// bucket sum over the sum of ALL sample weights, not the bucket's own: `pixel.LSplat` is never allocated, so the
// per-bucket normalisation of film.cpp:889-894 does not run.
__global__ void FilmGBufferImageKernel(FilmDev film, const double *acc, const float *ids, int nPixels, uint32_t channels,
                                       int nChannels, float *out) {
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= nPixels) return;
    const long long P = nPixels;
    const double *a = acc + 4ll * pix;
    float *o = out + (long long)nChannels * pix;
    float r = (float)a[0], g = (float)a[1], b = (float)a[2], w = (float)a[3];
    if (w != 0) { r /= w; g /= w; b /= w; }
    const float *m = film.outputRGBFromSensorRGB;
    int c = 0;
    o[c++] = m[0] * r + m[1] * g + m[2] * b;
    o[c++] = m[3] * r + m[4] * g + m[5] * b;
    o[c++] = m[6] * r + m[7] * g + m[8] * b;
    if (channels & B200PT_GBUFFER_RADIANCE)
        for (int k = 0; k < film.nBuckets; ++k) {
            float L = (float)acc[(4ll + k) * P + pix];
            o[c++] = w != 0 ? L / w : L;
        }
    const double *gq = acc + (4ll + film.nBuckets) * P + pix;
    const float gw = (float)gq[0];
    if (channels & B200PT_GBUFFER_ALBEDO)  // albedoRgb / rgbWeightSum, as it is (film.cpp:846-858, 914-918)
        for (int k = 0; k < 3; ++k) {
            float v = (float)gq[(17 + k) * P];
            o[c++] = w != 0 ? v / w : v;
        }
    if (channels & B200PT_GBUFFER_POSITION)
        for (int k = 0; k < 3; ++k) {
            float v = (float)gq[(1 + k) * P];
            o[c++] = gw != 0 ? v / gw : v;
        }
    if (channels & B200PT_GBUFFER_MATERIAL) o[c++] = ids[pix];
    if (channels & B200PT_GBUFFER_INSTANCE) o[c++] = ids[P + pix];
    for (int which = 1; which >= 0; --which) {  // Ns before N (film.cpp:812-825)
        if (!(channels & (which ? B200PT_GBUFFER_NS : B200PT_GBUFFER_NORMAL))) continue;
        float nx = (float)gq[(4 + 3 * which) * P], ny = (float)gq[(5 + 3 * which) * P], nz = (float)gq[(6 + 3 * which) * P];
        float l2 = nx * nx + ny * ny + nz * nz;
        if (l2 > 0) {
            float l = sqrtf(l2);
            nx /= l; ny /= l; nz /= l;
        } else
            nx = ny = nz = 0;
        o[c++] = nx; o[c++] = ny; o[c++] = nz;
    }
    const double cnt = gq[16 * P];
    float var[3], mean[3];
    for (int k = 0; k < 3; ++k) {  // VarianceEstimator (util/math.h): S / (n - 1), relative = variance / mean
        const double s1 = gq[(10 + k) * P], s2 = gq[(13 + k) * P];
        mean[k] = cnt > 0 ? (float)(s1 / cnt) : 0.f;
        var[k] = cnt > 1 ? (float)(fmax(s2 - s1 * s1 / cnt, 0.0) / (cnt - 1)) : 0.f;
    }
    if (channels & B200PT_GBUFFER_VARIANCE)
        for (int k = 0; k < 3; ++k) o[c++] = var[k];
    if (channels & B200PT_GBUFFER_RELATIVE_VARIANCE)
        for (int k = 0; k < 3; ++k) o[c++] = (cnt < 1 || mean[k] == 0) ? 0.f : var[k] / mean[k];
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

#endif  // part 0

static inline int Blocks(long long n, int b) { return (int)((n + b - 1) / b); }

#if B2_RENDER_PART == 0
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
#endif  // part 0

// one material's shade kernels per part
#define B2_SHADE_MAT_FN(NAME, MAT)                                                                       \
    cudaError_t NAME(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nPixels, int width,  \
                     int nActive, cudaStream_t stream) {                                                  \
        const int grid = Blocks(nActive, kShadeBlock);                                                    \
        const bool lbvh = rs.integrator.lightSampler == 2, vol = rs.integrator.type == 1;                 \
        const bool ext = rs.geomExt != 0; /* patches / per-vertex normals or uv somewhere in the scene */ \
        if (vol && lbvh) {                                                                                \
            if (ext) ShadeKernel<MAT, 1, 1, 1><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width); \
            else ShadeKernel<MAT, 1, 1, 0><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width);     \
        } else if (vol) {                                                                                 \
            if (ext) ShadeKernel<MAT, 0, 1, 1><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width); \
            else ShadeKernel<MAT, 0, 1, 0><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width);     \
        } else if (lbvh) {                                                                                \
            if (ext) ShadeKernel<MAT, 1, 0, 1><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width); \
            else ShadeKernel<MAT, 1, 0, 0><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width);     \
        } else {                                                                                          \
            if (ext) ShadeKernel<MAT, 0, 0, 1><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width); \
            else ShadeKernel<MAT, 0, 0, 0><<<grid, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width);     \
        }                                                                                                 \
        return cudaGetLastError();                                                                        \
    }
cudaError_t LaunchShadeDiffuse(const RenderSceneDev &, const PassBuffers &, int, int, int, int, cudaStream_t);
cudaError_t LaunchShadeConductor(const RenderSceneDev &, const PassBuffers &, int, int, int, int, cudaStream_t);
cudaError_t LaunchShadeDielectric(const RenderSceneDev &, const PassBuffers &, int, int, int, int, cudaStream_t);
cudaError_t LaunchShadeCoated(const RenderSceneDev &, const PassBuffers &, int, int, int, int, cudaStream_t);
#if B2_RENDER_PART == 0
B2_SHADE_MAT_FN(LaunchShadeDiffuse, MAT_DIFFUSE)
cudaError_t LaunchShade(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nPixels, int width,
                        int nActive, uint32_t matMask, cudaStream_t stream, int *launches) {
    cudaError_t e = cudaSuccess;
    if ((matMask & 1u) && e == cudaSuccess) e = (++*launches, LaunchShadeDiffuse(rs, pb, depth, nPixels, width, nActive, stream));
    if ((matMask & 2u) && e == cudaSuccess) e = (++*launches, LaunchShadeConductor(rs, pb, depth, nPixels, width, nActive, stream));
    if ((matMask & 4u) && e == cudaSuccess) e = (++*launches, LaunchShadeDielectric(rs, pb, depth, nPixels, width, nActive, stream));
    if ((matMask & 8u) && e == cudaSuccess) e = (++*launches, LaunchShadeCoated(rs, pb, depth, nPixels, width, nActive, stream));
    return e;
}
#elif B2_RENDER_PART == 1
B2_SHADE_MAT_FN(LaunchShadeConductor, MAT_CONDUCTOR)
#elif B2_RENDER_PART == 2
B2_SHADE_MAT_FN(LaunchShadeDielectric, MAT_DIELECTRIC)
#elif B2_RENDER_PART == 3
B2_SHADE_MAT_FN(LaunchShadeCoated, MAT_COATED_DIFFUSE)
#endif
#undef B2_SHADE_MAT_FN

#if B2_RENDER_PART == 4
// resident blocks of the persistent walk kernels: SMs x B2_VOL_MIN_BLOCKS
static int VolPersistGrid() {
    static int g = 0;
    if (!g) {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        g = sms * B2_VOL_MIN_BLOCKS;
    }
    return g;
}  // persistent-style grid of the warp-per-ray kernels (stride loop over the list)

cudaError_t LaunchVolMedium(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nPixels, int width, int nActive,
                            cudaStream_t stream) {
    VolSortKernel<<<Blocks(nActive, 256), 256, 0, stream>>>(rs, pb, depth, 0);
    // persistent: the resident set of blocks works the two lists off a cursor
    int *cursor = pb.shContinue + 1;
    cudaError_t e = cudaMemsetAsync(cursor, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    const int g = (int)std::min<long long>(Blocks(nActive, kShadeBlock), (long long)VolPersistGrid());
    const bool lb = rs.integrator.lightSampler == 2;
    if (rs.geomExt) {
        if (lb)
            VolMediumKernel<1, 1><<<g, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width, cursor);
        else
            VolMediumKernel<0, 1><<<g, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width, cursor);
    } else {
        if (lb)
            VolMediumKernel<1, 0><<<g, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width, cursor);
        else
            VolMediumKernel<0, 0><<<g, kShadeBlock, 0, stream>>>(rs, pb, depth, nPixels, width, cursor);
    }
    return cudaGetLastError();
}
cudaError_t LaunchShadowTr(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nActive, cudaStream_t stream) {
    // counts[2..4] of this iteration: the material queues are spent, the slots hold the walk queues' lengths
    cudaError_t e = cudaMemsetAsync(pb.counts + kCountStride * depth + 2, 0, 3 * sizeof(int), stream);
    if (e != cudaSuccess) return e;
    VolSortKernel<<<Blocks(nActive, 256), 256, 0, stream>>>(rs, pb, depth, 1);
    int *cursor = pb.shContinue + 1;
    e = cudaMemsetAsync(cursor, 0, sizeof(int), stream);
    if (e != cudaSuccess) return e;
    const int g = (int)std::min<long long>(Blocks(nActive, kShadeBlock), (long long)VolPersistGrid());
    if (rs.geomExt)
        ShadowTrKernel<1><<<g, kShadeBlock, 0, stream>>>(rs, pb, depth, cursor);
    else
        ShadowTrKernel<0><<<g, kShadeBlock, 0, stream>>>(rs, pb, depth, cursor);
    return cudaGetLastError();
}
#endif  // part 4

#if B2_RENDER_PART == 0
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
cudaError_t LaunchFilmAddGBuffer(const FilmDev &film, const PassBuffers &pb, int nPixels, int nSamplesInPass, double *acc,
                                 float *ids, const int32_t *materialIds, cudaStream_t stream) {
    FilmAddGBufferKernel<<<Blocks(nPixels, 128), 128, 0, stream>>>(film, pb, nPixels, nSamplesInPass, acc, ids, materialIds);
    return cudaGetLastError();
}
cudaError_t LaunchFilmGBufferImage(const FilmDev &film, const double *acc, const float *ids, int nPixels, uint32_t channels,
                                   int nChannels, float *out, cudaStream_t stream) {
    FilmGBufferImageKernel<<<Blocks(nPixels, 128), 128, 0, stream>>>(film, acc, ids, nPixels, channels, nChannels, out);
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

#endif  // part 0

}  // namespace b200pt
