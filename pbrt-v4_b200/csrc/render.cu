// render.cu - the wavefront kernels around the traversal kernel: camera-ray generation, the
// per-vertex shading step (emission, BSDF, light sampling with its shadow ray, Russian roulette),
// sample finish (sensor RGB) and film accumulation.  All path logic is in render_core.h (shared with
// the CPU check in tests/native/rendercheck.cpp); these kernels only move a path between its
// structure-of-arrays slot and registers.
//
// Stage order per pass (api.cpp RenderRun), mirroring WavefrontPathIntegrator::Render
// (wavefront/integrator.cpp:290-493): generate -> for each depth { closest-hit ; any-hit of the shadow
// rays of the previous vertex ; shade } -> finish -> film.
#include <cuda_runtime.h>

#include "render_core.h"
#include "render_kernels.h"

namespace b200pt {

constexpr int kShadeBlock = 128;

size_t PathBufferBytes(long long n) {
    auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
    size_t t = 0;
    t += al(sizeof(float) * 3 * kNLambda * n);
    t += al(sizeof(float) * kPathScalars * n);
    t += al(sizeof(int) * kPathInts * n);
    t += 2 * al(sizeof(float) * 7 * n);
    t += al(sizeof(uint32_t) * 6 * n);
    t += al(size_t(n));
    t += al(sizeof(float) * 4 * n);
    return t;
}

void CarvePathBuffers(void *base, long long n, PathBuffers &pb) {
    char *p = static_cast<char *>(base);
    auto take = [&](size_t b) {
        char *r = p;
        p += (b + 255) & ~size_t(255);
        return r;
    };
    pb.n = (int)n;
    pb.spec = (float *)take(sizeof(float) * 3 * kNLambda * n);
    pb.scalars = (float *)take(sizeof(float) * kPathScalars * n);
    pb.ints = (int *)take(sizeof(int) * kPathInts * n);
    pb.rays = (float *)take(sizeof(float) * 7 * n);
    pb.shadowRays = (float *)take(sizeof(float) * 7 * n);
    pb.hits = (uint32_t *)take(sizeof(uint32_t) * 6 * n);
    pb.occluded = (unsigned char *)take(size_t(n));
    pb.rgbw = (float *)take(sizeof(float) * 4 * n);
}

// ---- path <-> SoA
__device__ __forceinline__ void StoreRay(float *rays, int i, V3 o, V3 d, float tMax) {
    float *q = rays + 7ll * i;
    q[0] = o.x; q[1] = o.y; q[2] = o.z;
    q[3] = d.x; q[4] = d.y; q[5] = d.z;
    q[6] = tMax;
}

__device__ __forceinline__ void StorePath(const PathBuffers &pb, int i, const PathRegs &p, bool storeLd) {
    const long long n = pb.n;
    float *beta = pb.spec + i, *L = beta + kNLambda * n, *Ld = L + kNLambda * n;
#pragma unroll
    for (int k = 0; k < kNLambda; ++k) {
        beta[k * n] = p.beta[k];
        L[k * n] = p.L[k];
    }
    if (storeLd) {
#pragma unroll
        for (int k = 0; k < kNLambda; ++k) Ld[k * n] = p.Ld[k];
    }
    float *s = pb.scalars + i;
    s[0 * n] = p.lambdaU;
    s[1 * n] = p.etaScale;
    s[2 * n] = p.pB;
    s[3 * n] = p.filterWeight;
    s[4 * n] = p.prev.pi.lo.x; s[5 * n] = p.prev.pi.lo.y; s[6 * n] = p.prev.pi.lo.z;
    s[7 * n] = p.prev.pi.hi.x; s[8 * n] = p.prev.pi.hi.y; s[9 * n] = p.prev.pi.hi.z;
    s[10 * n] = p.prev.n.x; s[11 * n] = p.prev.n.y; s[12 * n] = p.prev.n.z;
    s[13 * n] = p.prev.ns.x; s[14 * n] = p.prev.ns.y; s[15 * n] = p.prev.ns.z;
    int *q = pb.ints + i;
    q[0 * n] = p.sampleIndex;
    q[1 * n] = p.dimension;
    q[2 * n] = p.depth;
    q[3 * n] = (int)p.flags;
    StoreRay(pb.rays, i, p.rayO, p.rayD, p.rayTMax);
    StoreRay(pb.shadowRays, i, p.shO, p.shD, p.shTMax);
}

__device__ __forceinline__ void LoadPath(const PathBuffers &pb, int i, PathRegs &p, bool loadLd) {
    const long long n = pb.n;
    const float *beta = pb.spec + i, *L = beta + kNLambda * n, *Ld = L + kNLambda * n;
#pragma unroll
    for (int k = 0; k < kNLambda; ++k) {
        p.beta[k] = beta[k * n];
        p.L[k] = L[k * n];
    }
#pragma unroll
    for (int k = 0; k < kNLambda; ++k) p.Ld[k] = loadLd ? Ld[k * n] : 0.f;
    const float *s = pb.scalars + i;
    p.lambdaU = s[0 * n];
    p.etaScale = s[1 * n];
    p.pB = s[2 * n];
    p.filterWeight = s[3 * n];
    p.prev.pi.lo = mk(s[4 * n], s[5 * n], s[6 * n]);
    p.prev.pi.hi = mk(s[7 * n], s[8 * n], s[9 * n]);
    p.prev.n = mk(s[10 * n], s[11 * n], s[12 * n]);
    p.prev.ns = mk(s[13 * n], s[14 * n], s[15 * n]);
    const int *q = pb.ints + i;
    p.sampleIndex = q[0 * n];
    p.dimension = q[1 * n];
    p.depth = q[2 * n];
    p.flags = (uint32_t)q[3 * n];
    const float *r = pb.rays + 7ll * i;
    p.rayO = mk(r[0], r[1], r[2]);
    p.rayD = mk(r[3], r[4], r[5]);
    p.rayTMax = r[6];
    p.shTMax = -1.f;
    p.shO = mk(0, 0, 0);
    p.shD = mk(0, 0, 0);
}

__device__ __forceinline__ void CountRays(bool closest, bool shadow, unsigned long long *counters) {
    unsigned c = __ballot_sync(0xffffffffu, closest), s = __ballot_sync(0xffffffffu, shadow);
    if ((threadIdx.x & 31) == 0) {
        if (c) atomicAdd(counters + 0, (unsigned long long)__popc(c));
        if (s) atomicAdd(counters + 1, (unsigned long long)__popc(s));
    }
}

// ---- camera rays: RayIntegrator::EvaluatePixelSample up to GenerateRay (cpu/integrators.cpp:228-262)
__global__ void __launch_bounds__(kShadeBlock)
GenerateKernel(RenderSceneDev rs, PathBuffers pb, int nPixels, int width, int nActive, int firstSample,
               int sampleStride, unsigned long long *counters) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool live = i < nActive;
    if (live) {
        int pix = i % nPixels, s = i / nPixels;
        PathRegs p;
        p.prev.pi.lo = p.prev.pi.hi = p.prev.n = p.prev.ns = mk(0, 0, 0);
        p.shO = p.shD = mk(0, 0, 0);
        GenerateCameraPath(rs, p, rs.film.x0 + pix % width, rs.film.y0 + pix / width,
                           firstSample + s * sampleStride);
        StorePath(pb, i, p, false);
        pb.occluded[i] = 0;
    }
    CountRays(live, false, counters);
}

// ---- one path vertex: PathIntegrator::Li loop body (cpu/integrators.cpp:640-760)
__global__ void __launch_bounds__(kShadeBlock)
ShadeKernel(RenderSceneDev rs, PathBuffers pb, int nPixels, int width, int nActive, unsigned long long *counters) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool nextRay = false, shadowRay = false;
    if (i < nActive) {
        uint32_t flags = (uint32_t)pb.ints[3ll * pb.n + i];
        if (flags & (PF_ALIVE | PF_SHADOW_PENDING)) {
            PathRegs p;
            LoadPath(pb, i, p, (flags & PF_SHADOW_PENDING) != 0);
            int pix = i % nPixels;
            p.px = rs.film.x0 + pix % width;
            p.py = rs.film.y0 + pix / width;
            uint32_t hit[6];
            const uint2 *h = reinterpret_cast<const uint2 *>(pb.hits + 6ll * i);
            uint2 h0 = h[0], h1 = h[1], h2 = h[2];
            hit[0] = h0.x; hit[1] = h0.y; hit[2] = h1.x; hit[3] = h1.y; hit[4] = h2.x; hit[5] = h2.y;
            ShadeVertex(rs, p, hit, pb.occluded[i] != 0);
            StorePath(pb, i, p, (p.flags & PF_SHADOW_PENDING) != 0);
            nextRay = (p.flags & PF_ALIVE) != 0;
            shadowRay = (p.flags & PF_SHADOW_PENDING) != 0;
        }
    }
    CountRays(nextRay, shadowRay, counters);
}

// ---- finished sample -> weighted sensor RGB (RGBFilm::AddSample, film.h:241-258)
__global__ void __launch_bounds__(kShadeBlock)
FinishKernel(RenderSceneDev rs, PathBuffers pb, int nPixels, int nActive) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nActive) return;
    const long long n = pb.n;
    PathRegs p;
    const float *L = pb.spec + kNLambda * n + i, *Ld = L + kNLambda * n;
    p.flags = (uint32_t)pb.ints[3 * n + i];
    bool pending = (p.flags & PF_SHADOW_PENDING) != 0;
#pragma unroll
    for (int k = 0; k < kNLambda; ++k) {
        p.L[k] = L[k * n];
        p.Ld[k] = pending ? Ld[k * n] : 0.f;
    }
    p.lambdaU = pb.scalars[i];
    p.filterWeight = pb.scalars[3 * n + i];
    float rgbw[4];
    FinishPath(rs, p, pb.occluded[i] != 0, rgbw);
    reinterpret_cast<float4 *>(pb.rgbw)[i] = make_float4(rgbw[0], rgbw[1], rgbw[2], rgbw[3]);
}

// ---- film: one thread per pixel adds the pass's samples in sample order, as the CPU film does for
// a pixel (each pixel is owned by one thread there too), so the double sums are the reference's.
__global__ void FilmAddKernel(PathBuffers pb, int nPixels, int nSamplesInPass, double *film) {
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= nPixels) return;
    double4 *dst = reinterpret_cast<double4 *>(film) + pix;
    double2 a = reinterpret_cast<double2 *>(dst)[0], b = reinterpret_cast<double2 *>(dst)[1];
    for (int s = 0; s < nSamplesInPass; ++s) {
        float4 v = reinterpret_cast<const float4 *>(pb.rgbw)[(long long)s * nPixels + pix];
        a.x += v.x;
        a.y += v.y;
        b.x += v.z;
        b.y += v.w;
    }
    reinterpret_cast<double2 *>(dst)[0] = a;
    reinterpret_cast<double2 *>(dst)[1] = b;
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

static inline int Blocks(long long n, int b) { return (int)((n + b - 1) / b); }

cudaError_t LaunchGenerate(const RenderSceneDev &rs, const PathBuffers &pb, int nPixels, int width, int nActive,
                           int firstSample, int sampleStride, unsigned long long *counters, cudaStream_t stream) {
    GenerateKernel<<<Blocks(nActive, kShadeBlock), kShadeBlock, 0, stream>>>(rs, pb, nPixels, width, nActive,
                                                                             firstSample, sampleStride, counters);
    return cudaGetLastError();
}
cudaError_t LaunchShade(const RenderSceneDev &rs, const PathBuffers &pb, int nPixels, int width, int nActive,
                        unsigned long long *counters, cudaStream_t stream) {
    ShadeKernel<<<Blocks(nActive, kShadeBlock), kShadeBlock, 0, stream>>>(rs, pb, nPixels, width, nActive, counters);
    return cudaGetLastError();
}
cudaError_t LaunchFinish(const RenderSceneDev &rs, const PathBuffers &pb, int nPixels, int nActive,
                         cudaStream_t stream) {
    FinishKernel<<<Blocks(nActive, kShadeBlock), kShadeBlock, 0, stream>>>(rs, pb, nPixels, nActive);
    return cudaGetLastError();
}
cudaError_t LaunchFilmAdd(const PathBuffers &pb, int nPixels, int nSamplesInPass, double *film, cudaStream_t stream) {
    FilmAddKernel<<<Blocks(nPixels, 128), 128, 0, stream>>>(pb, nPixels, nSamplesInPass, film);
    return cudaGetLastError();
}
cudaError_t LaunchFilmImage(const FilmDev &film, const double *acc, int nPixels, float *rgb, cudaStream_t stream) {
    FilmImageKernel<<<Blocks(nPixels, 128), 128, 0, stream>>>(film, acc, nPixels, rgb);
    return cudaGetLastError();
}

}  // namespace b200pt
