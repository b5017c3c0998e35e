// render_kernels.h - launchers of the render-path kernels (render.cu) and the per-pass buffers.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "render_types.h"

namespace b200pt {

constexpr int kPathScalars = 17;  // lambdaU, etaScale, pB, prev{pi.lo, pi.hi, n, ns}, cameraWeight, filterWeight
constexpr int kFilterWeightScalar = 16;  // FilterSample::weight of the camera sample (1 for the box filter)
constexpr int kCameraWeightScalar = 15;  // written by the generate kernel, read when the sample is finished
constexpr int kPathInts = 5;      // sampleIndex, dimension, depth, flags, medium
constexpr int kShadowAuxWords = 20;  // volpath light sample: p_l, scatterPDF, medium, started, rng[4], lightPi[6], lightN[3]
constexpr uint32_t kConsumedTag = 0xFFFFFFFDu;  // hit slot of a ray the medium kernel already dealt with
constexpr int kVolMaxIterations = 96;  // wavefront iterations of a volpath pass (interfaces do not count as depth)
constexpr int kMatQueues = 4;     // one shading queue per MaterialType
constexpr int kCountStride = 8;   // ints per depth in PassBuffers::counts
constexpr int kAovWords = 18;     // PassBuffers::aov planes: {set, p, n, ns, material, instance}, then the albedo's XYZ twice (below)
// GBufferFilm accumulators behind the rgbw and the 2 x 31 bucket planes, one plane of nPixels doubles each:
// gBufferWeightSum, pSum[3], nSum[3], nsSum[3], sum rgb[3], sum rgb^2[3], n (VarianceEstimator as sums)
constexpr int kGBufferPlanes = 20;  // ..., 17-19: rgbAlbedoSum

// One pass = nPixels x nSamplesInPass paths; path i is pixel i % nPixels, sample slot i / nPixels.
// Path state is indexed by path id and laid out [field][path] so a warp's accesses coalesce; the
// queues (rays, hits, shadow rays, material queues) are dense and indexed by queue slot.  Queue
// lengths live on the device (counts): the surface path never reads them on the host inside a pass, launches are
// sized for the worst case and threads past the length exit (as the reference's wavefront does,
// wavefront/integrator.cpp:403-406 "maxQueueSize").  The volumetric path reads two lengths per iteration and one per
// walk segment (render_api.cpp: interface crossings make the iteration count data dependent).
struct PassBuffers {
    int n;                 // capacity in paths
    // ---- per path
    float *beta;           // 31 x n
    float *L;              // 31 x n   radiance so far
    float *scalars;        // kPathScalars x n
    int *ints;             // kPathInts x n
    float *rgbw;           // n x 4    finished sample: weight * sensor rgb, weight
    float *aov;            // kAovWords x n, GBufferFilm only (else null): the sample's VisibleSurface {set, p, n, ns, material, instance}
    // ---- per queue slot
    float *rays[2];        // n x 7 {o, d, tMax}, current / next depth
    int *rayPath[2];       // n     path id of the ray in that slot
    uint32_t *hits;        // n x 6 internal record {slot, t, b0, b1, b2, 0}
    int *matQ[kMatQueues]; // n     ray slots whose hit has material type m
    float *shRays;         // n x 7 shadow rays
    int *shPath;           // n     path id, -1 = void slot
    float *Ld;             // 31 x n  contribution to add when unoccluded (already x beta)
    unsigned char *occluded;  // n
    // ---- volpath only (null otherwise): rescaled path probabilities per path, the transmittance walk's state per
    // shadow-queue slot, closest hits of the shadow rays, number of walks that go on to another segment
    float *ru, *rl;        // 31 x n
    float *shT, *shRl, *shRu;  // 31 x n
    uint32_t *shAux;       // n x kShadowAuxWords
    uint32_t *shHits;      // n x 6
    int *shContinue;       // 1
    // ---- queue lengths: counts[kCountStride * depth + {0 rays, 1 shadow, 2.. material queues}]
    int *counts;
    int maxDepthSlots;
};

size_t PassBufferBytes(long long n, int depthSlots, bool vol, bool aov = false);
void CarvePassBuffers(void *base, long long n, int depthSlots, bool vol, PassBuffers &pb, bool aov = false);

cudaError_t LaunchGenerate(const RenderSceneDev &rs, const PassBuffers &pb, int nPixels, int width, int nActive,
                           int firstSample, int sampleStride, cudaStream_t stream);
cudaError_t LaunchClassify(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nActive,
                           cudaStream_t stream);
// matMask: bit m set = material type m occurs in the scene (its queue is shaded)
cudaError_t LaunchShade(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nPixels, int width,
                        int nActive, uint32_t matMask, cudaStream_t stream, int *launches);
// volpath: delta tracking of the rays that travel in a medium (before classify), one segment of every light
// sample's transmittance walk (after a closest-hit pass over the shadow rays into pb.shHits)
cudaError_t LaunchVolMedium(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nPixels, int width, int nActive,
                            cudaStream_t stream);
cudaError_t LaunchShadowTr(const RenderSceneDev &rs, const PassBuffers &pb, int depth, int nActive, cudaStream_t stream);
cudaError_t LaunchResolveShadows(const PassBuffers &pb, int depth, int nActive, cudaStream_t stream);
cudaError_t LaunchFinish(const RenderSceneDev &rs, const PassBuffers &pb, int nActive, cudaStream_t stream);
cudaError_t LaunchFilmAdd(const PassBuffers &pb, int nPixels, int nSamplesInPass, double *film, cudaStream_t stream);
// spectral film: the bucket planes behind the 4 * nPixels rgbw doubles of `film`
// GBufferFilm (film_type 2): the AOV planes behind the bucket planes, and the ids of the last sample of each pixel
cudaError_t LaunchFilmAddGBuffer(const FilmDev &film, const PassBuffers &pb, int nPixels, int nSamplesInPass, double *acc,
                                 float *ids, const int32_t *materialIds, cudaStream_t stream);
// GBufferFilm::GetImage (film.cpp:743-1044): `channels` (B200PT_GBUFFER_*) in the reference's channel order
cudaError_t LaunchFilmGBufferImage(const FilmDev &film, const double *acc, const float *ids, int nPixels, uint32_t channels,
                                   int nChannels, float *out, cudaStream_t stream);
// exr_pack.cu: channels-last float image -> scanline chunks of an uncompressed EXR (order[c] = image channel of file channel c)
cudaError_t LaunchExrPack(const float *img, int width, int height, int nChannels, const int *order, bool half, int y0,
                          unsigned char *out, cudaStream_t stream);
cudaError_t LaunchFilmAddSpectral(const FilmDev &film, const PassBuffers &pb, int nPixels, int nSamplesInPass,
                                  double *acc, cudaStream_t stream);
cudaError_t LaunchFilmSpectralImage(const FilmDev &film, const double *acc, int nPixels, float *out,
                                    cudaStream_t stream);
cudaError_t LaunchFilmImage(const FilmDev &film, const double *acc, int nPixels, float *rgb, cudaStream_t stream);
// sums counts over depths into totals[0] (closest rays) and totals[1] (shadow rays)
cudaError_t LaunchSumCounts(const PassBuffers &pb, int depths, unsigned long long *totals, cudaStream_t stream);

}  // namespace b200pt
