// render_kernels.h - launchers of the render-path kernels (render.cu) and the per-pass path buffers.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "render_types.h"

namespace b200pt {

// One pass = nPixels x nSamplesInPass paths; path i is pixel i % nPixels, sample slot i / nPixels.
// Spectral arrays are [k][path] and scalar state [field][path], so a warp's accesses coalesce.
struct PathBuffers {
    int n;                 // capacity in paths
    float *spec;           // 3 x 31 x n : beta, L, Ld
    float *scalars;        // kPathScalars x n
    int *ints;             // kPathInts x n
    float *rays;           // n x 7  {o, d, tMax}; tMax < 0 = no ray
    float *shadowRays;     // n x 7
    uint32_t *hits;        // n x 6  internal record {slot, t, b0, b1, b2, 0}
    unsigned char *occluded;  // n
    float *rgbw;           // n x 4  finished sample: weight * sensor rgb, weight
};
constexpr int kPathScalars = 16;  // lambdaU, etaScale, pB, filterWeight, prev{pi.lo, pi.hi, n, ns}
constexpr int kPathInts = 4;      // sampleIndex, dimension, depth, flags

size_t PathBufferBytes(long long n);
void CarvePathBuffers(void *base, long long n, PathBuffers &pb);

// counters: [0] closest rays issued, [1] shadow rays issued
cudaError_t LaunchGenerate(const RenderSceneDev &rs, const PathBuffers &pb, int nPixels, int width, int nActive,
                           int firstSample, int sampleStride, unsigned long long *counters, cudaStream_t stream);
cudaError_t LaunchShade(const RenderSceneDev &rs, const PathBuffers &pb, int nPixels, int width, int nActive,
                        unsigned long long *counters, cudaStream_t stream);
cudaError_t LaunchFinish(const RenderSceneDev &rs, const PathBuffers &pb, int nPixels, int nActive,
                         cudaStream_t stream);
cudaError_t LaunchFilmAdd(const PathBuffers &pb, int nPixels, int nSamplesInPass, double *film, cudaStream_t stream);
cudaError_t LaunchFilmImage(const FilmDev &film, const double *acc, int nPixels, float *rgb, cudaStream_t stream);

}  // namespace b200pt
