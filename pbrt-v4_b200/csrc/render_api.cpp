// render_api.cpp - the render entry points of include/b200pt.h: table upload, the wavefront pass
// loop that strings the kernels together (the role of WavefrontPathIntegrator::Render,
// wavefront/integrator.cpp:290-493), film read-out.  Host code only; no CPU rendering path exists.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <mutex>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "api_internal.h"
#include "kernels.h"
#include "render_host.h"
#include "render_kernels.h"

using namespace b200pt;

struct b200pt_render {
    b200pt_scene *scene = nullptr;
    RenderHost host;
    RenderSceneDev dev{};
    void *dSpectra = nullptr, *dMaterials = nullptr, *dLights = nullptr, *dMeshes = nullptr;
    void *dLightBvh = nullptr, *dBitTrails = nullptr, *dInfinite = nullptr, *dLightGeom = nullptr, *dFilterTable = nullptr, *dEnvImage = nullptr, *dEnvDist = nullptr;
    void *dLens = nullptr, *dPupil = nullptr, *dCornerAttr = nullptr, *dMaterialIds = nullptr;
    float *dFilmIds = nullptr;  // GBufferFilm: materialId / instanceId of the last sample per pixel (2 x nPixels)
    void *dMedia = nullptr;
    std::vector<void *> dGrids;  // density and majorant grids of the media
    double *dFilm = nullptr;
    float *dImage = nullptr;
    void *dPaths = nullptr;
    PassBuffers pb{};
    long long pathCap = 0;
    uint32_t matMask = 0;                     // material types present in the scene
    unsigned long long *dCounters = nullptr;  // [0] closest rays, [1] shadow rays
    int *dRayCounter = nullptr;               // work counter of the persistent traversal kernel
    int nPixels = 0;
    size_t filmDoubles = 0;  // 4 per pixel, + 2 * nBuckets per pixel for a spectral film
    // paths per pass: up to 32 Mi when 40 % of the free device memory holds their buffers (fewer, longer launches; a
    // 1920 x 1080 frame takes 16 samples per pixel per pass), never below 4 Mi; B200PT_PASS_PATHS overrides
    long long maxPathsPerPass = 4ll << 20;
    b200pt_render_stats stats{};
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
};

namespace {
template <class T>
int Upload(b200pt_ctx *ctx, const std::vector<T> &v, void **d) {
    size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
    B2_CUDA(ctx, cudaMalloc(d, bytes));
    if (!v.empty()) B2_CUDA(ctx, cudaMemcpyAsync(*d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    return B200PT_OK;
}
}  // namespace

extern "C" {

int b200pt_render_destroy(b200pt_render *r) {
    if (!r) return B200PT_OK;
    if (r->scene && r->scene->ctx) {
        cudaSetDevice(r->scene->ctx->device);
        cudaStreamSynchronize(r->scene->ctx->stream);
    }
    cudaFree(r->dSpectra);
    cudaFree(r->dMaterials);
    cudaFree(r->dMaterialIds);
    cudaFree(r->dFilmIds);
    cudaFree(r->dLights);
    cudaFree(r->dMeshes);
    cudaFree(r->dLightBvh);
    cudaFree(r->dBitTrails);
    cudaFree(r->dInfinite);
    cudaFree(r->dLightGeom);
    cudaFree(r->dFilterTable);
    cudaFree(r->dEnvImage);
    cudaFree(r->dEnvDist);
    cudaFree(r->dLens);
    cudaFree(r->dMedia);
    for (void *g : r->dGrids) cudaFree(g);
    cudaFree(r->dPupil);
    cudaFree(r->dCornerAttr);
    cudaFree(r->dFilm);
    cudaFree(r->dImage);
    cudaFree(r->dPaths);
    cudaFree(r->dCounters);
    if (r->ev0) cudaEventDestroy(r->ev0);
    if (r->ev1) cudaEventDestroy(r->ev1);
    delete r;
    return B200PT_OK;
}

int b200pt_render_create(b200pt_scene *scene, const b200pt_render_desc *desc, b200pt_render **out) {
    if (!scene || !desc || !out) return B200PT_ERR_INVALID;
    *out = nullptr;
    b200pt_ctx *ctx = scene->ctx;
    b200pt_render *r = new (std::nothrow) b200pt_render;
    if (!r) return B200PT_ERR_NOMEM;
    r->scene = scene;
    int rc = BuildRenderTables(desc, scene->host, scene->host.nMeshes, r->host, ctx->err);
    if (rc != B200PT_OK) {
        delete r;
        return rc;
    }
    auto fail = [&](int code) {
        b200pt_render_destroy(r);
        return code;
    };
    if (cudaSetDevice(ctx->device) != cudaSuccess) {
        ctx->err = "cudaSetDevice failed";
        return fail(B200PT_ERR_CUDA);
    }
    if ((rc = Upload(ctx, r->host.spectra, &r->dSpectra)) || (rc = Upload(ctx, r->host.materials, &r->dMaterials)) ||
        (rc = Upload(ctx, r->host.lights, &r->dLights)) || (rc = Upload(ctx, r->host.meshes, &r->dMeshes)) ||
        (rc = Upload(ctx, r->host.lightBvhNodes, &r->dLightBvh)) ||
        (rc = Upload(ctx, r->host.lightBitTrails, &r->dBitTrails)) ||
        (rc = Upload(ctx, r->host.infiniteLights, &r->dInfinite)) || (rc = Upload(ctx, r->host.lens, &r->dLens)) ||
        (rc = Upload(ctx, r->host.pupilBounds, &r->dPupil)) || (rc = Upload(ctx, r->host.cornerAttr, &r->dCornerAttr)) ||
        (rc = Upload(ctx, r->host.materialIds, &r->dMaterialIds)) || (rc = Upload(ctx, r->host.lightGeom, &r->dLightGeom)) ||
        (rc = Upload(ctx, r->host.filterTable, &r->dFilterTable)) || (rc = Upload(ctx, r->host.envImage, &r->dEnvImage)) ||
        (rc = Upload(ctx, r->host.envDist, &r->dEnvDist)))
        return fail(rc);
    for (size_t i = 0; i < r->host.media.size(); ++i) {  // grids first: the medium table points at them
        if (r->host.media[i].type != 1) continue;
        void *dd = nullptr, *dm = nullptr;
        if ((rc = Upload(ctx, r->host.mediumGrids[i].first, &dd))) return fail(rc);
        r->dGrids.push_back(dd);
        if ((rc = Upload(ctx, r->host.mediumGrids[i].second, &dm))) return fail(rc);
        r->dGrids.push_back(dm);
        r->host.media[i].density = (const float *)dd;
        r->host.media[i].majorants = (const float *)dm;
        if (!r->host.mediumLeGrids[i].empty()) {
            void *dl = nullptr;
            if ((rc = Upload(ctx, r->host.mediumLeGrids[i], &dl))) return fail(rc);
            r->dGrids.push_back(dl);
            r->host.media[i].leScale = (const float *)dl;
        }
        if (!r->host.mediumTempGrids[i].empty()) {
            void *dt = nullptr;
            if ((rc = Upload(ctx, r->host.mediumTempGrids[i], &dt))) return fail(rc);
            r->dGrids.push_back(dt);
            r->host.media[i].temperature = (const float *)dt;
        }
    }
    if ((rc = Upload(ctx, r->host.media, &r->dMedia))) return fail(rc);
    r->nPixels = r->host.width * r->host.height;
    r->dev = r->host.rs;
    r->dev.media = (const MediumDev *)r->dMedia;
    r->dev.spectra = (const SpectrumTable *)r->dSpectra;
    r->dev.materials = (const MaterialDev *)r->dMaterials;
    r->dev.materialIds = r->host.materialIds.empty() ? nullptr : (const int32_t *)r->dMaterialIds;
    r->dev.lights = (const LightDev *)r->dLights;
    r->dev.meshes = (const MeshAttrDev *)r->dMeshes;
    r->dev.lightBvhNodes = r->dLightBvh;
    r->dev.lightBitTrails = (const uint32_t *)r->dBitTrails;
    r->dev.infiniteLights = (const int32_t *)r->dInfinite;
    r->dev.lightGeom = r->host.lightGeom.empty() ? nullptr : (const float *)r->dLightGeom;
    r->dev.film.filterTable = r->host.filterTable.empty() ? nullptr : (const float *)r->dFilterTable;
    r->dev.envImage = (const float *)r->dEnvImage;
    r->dev.envDist = (const float *)r->dEnvDist;
    r->dev.camera.lens = (const float *)r->dLens;
    r->dev.camera.pupilBounds = (const float *)r->dPupil;
    r->dev.cornerAttr = (const float *)r->dCornerAttr;
    r->dev.prims = scene->dev.prims;
    r->dev.instances = scene->dev.instances;
    for (const MaterialDev &m : r->host.materials) r->matMask |= 1u << m.type;
    for (const MeshAttrDev &m : r->host.meshes)
        if (m.material < 0) r->matMask |= 1u;  // null materials go through the diffuse queue, which ends the path
    if (r->matMask == 0) r->matMask = 1;
    {
        size_t freeB = 0, totalB = 0;
        if (cudaMemGetInfo(&freeB, &totalB) == cudaSuccess) {
            const bool volPath = r->dev.integrator.type == 1;
            const size_t perPath = PassBufferBytes(1 << 20, volPath ? kVolMaxIterations + 2 : r->dev.integrator.maxDepth + 2, volPath, r->dev.film.type == 2) >> 20;
            const long long fit = (long long)(0.4 * double(freeB) / double(std::max<size_t>(perPath, 1)));
            r->maxPathsPerPass = std::max(4ll << 20, std::min(32ll << 20, fit));
        } else {
            (void)cudaGetLastError();
        }
    }
    if (const char *v = getenv("B200PT_PASS_PATHS")) r->maxPathsPerPass = std::max(1ll, atoll(v));
    // RGBFilm: {rgbSum[3], weightSum}; SpectralFilm: + bucketSums and weightSums planes; GBufferFilm: + the 31 radiance
    // planes (its image divides by the RGB weight sum, film.cpp:846-850: no per-bucket weights) + the AOV planes
    r->filmDoubles = size_t(r->nPixels) * (r->dev.film.type == 2 ? 4 + size_t(r->dev.film.nBuckets) + kGBufferPlanes
                                                                 : 4 + 2 * size_t(r->dev.film.nBuckets));
    if (r->dev.film.type == 2 &&
        (cudaMalloc((void **)&r->dFilmIds, sizeof(float) * 2 * size_t(r->nPixels)) != cudaSuccess ||
         cudaMemsetAsync(r->dFilmIds, 0, sizeof(float) * 2 * size_t(r->nPixels), ctx->stream) != cudaSuccess)) {
        ctx->err = "render setup: gbuffer id planes";
        return fail(B200PT_ERR_CUDA);
    }
    if (cudaMalloc((void **)&r->dFilm, sizeof(double) * r->filmDoubles) != cudaSuccess ||
        cudaMalloc((void **)&r->dImage, sizeof(float) * std::max(3, r->dev.film.nBuckets + 32) * size_t(r->nPixels)) != cudaSuccess ||
        cudaMalloc((void **)&r->dCounters, 4 * sizeof(unsigned long long)) != cudaSuccess ||
        cudaEventCreate(&r->ev0) != cudaSuccess || cudaEventCreate(&r->ev1) != cudaSuccess ||
        cudaMemsetAsync(r->dFilm, 0, sizeof(double) * r->filmDoubles, ctx->stream) != cudaSuccess ||
        cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
        ctx->err = std::string("render setup: ") + cudaGetErrorString(cudaGetLastError());
        return fail(B200PT_ERR_CUDA);
    }
    r->dRayCounter = (int *)(r->dCounters + 3);
    *out = r;
    return B200PT_OK;
}

int b200pt_film_clear(b200pt_render *r) {
    if (!r) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = r->scene->ctx;
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    B2_CUDA(ctx, cudaMemsetAsync(r->dFilm, 0, sizeof(double) * r->filmDoubles, ctx->stream));
    if (r->dFilmIds) B2_CUDA(ctx, cudaMemsetAsync(r->dFilmIds, 0, sizeof(float) * 2 * size_t(r->nPixels), ctx->stream));
    return B200PT_OK;
}

int b200pt_render_run(b200pt_render *r, int32_t first_sample, int32_t n_samples, int32_t sample_stride) {
    if (!r) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = r->scene->ctx;
    if (first_sample < 0 || n_samples < 0 || sample_stride < 1) {
        ctx->err = "render_run: bad sample range";
        return B200PT_ERR_INVALID;
    }
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    const int nPix = r->nPixels, width = r->host.width;
    // samples of every pixel per pass, bounded by the path-buffer budget
    int perPass = (int)std::max<long long>(1, std::min<long long>(n_samples, r->maxPathsPerPass / nPix));
    long long need = (long long)perPass * nPix;
    if (need > 0x7fffffffll) {
        ctx->err = "render_run: film too large for one pass";
        return B200PT_ERR_INVALID;
    }
    const DeviceScene &sc = r->scene->dev;
    const int stackDepth = int(sc.maxDepth) + 1;
    const int maxDepth = r->dev.integrator.maxDepth;
    const bool vol = r->dev.integrator.type == 1;
    const int depthSlots = vol ? kVolMaxIterations + 2 : maxDepth + 2;
    if (need > r->pathCap) {
        B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(r->dPaths);
        r->dPaths = nullptr;
        r->pathCap = 0;
        if (cudaMalloc(&r->dPaths, PassBufferBytes(need, depthSlots, vol, r->dev.film.type == 2)) != cudaSuccess) {
            (void)cudaGetLastError();
            ctx->err = "render_run: out of device memory for the path buffers";
            return B200PT_ERR_NOMEM;
        }
        CarvePassBuffers(r->dPaths, need, depthSlots, vol, r->pb, r->dev.film.type == 2);
        r->pathCap = need;
    }
    const PassBuffers &pb = r->pb;
    int *exact = nullptr;
    if (int rc = EnsureExactScratch(ctx, size_t(need), &exact)) return rc;
    int launches = 0;
    uint64_t volPathsCut = 0, volWalksCut = 0;
    B2_CUDA(ctx, cudaMemsetAsync(r->dCounters, 0, 4 * sizeof(unsigned long long), ctx->stream));
    B2_CUDA(ctx, cudaEventRecord(r->ev0, ctx->stream));
    for (int done = 0; done < n_samples; done += perPass) {
        int ns = std::min(perPass, n_samples - done);
        int nActive = ns * nPix;
        B2_CUDA(ctx, cudaMemsetAsync(pb.counts, 0, sizeof(int) * kCountStride * depthSlots, ctx->stream));
        B2_CUDA(ctx, LaunchGenerate(r->dev, pb, nPix, width, nActive, first_sample + done * sample_stride,
                                    sample_stride, ctx->stream));
        ++launches;
        if (vol) {
            // VolPathIntegrator: iteration j = one pass of Li's while loop for every live path (a medium event, a
            // surface vertex or an interface crossing), until no ray is left.  Queue lengths are read back once per
            // iteration (interface crossings do not count as depth, so the iteration count is not known ahead).
            int usedIters = 0;
            for (int j = 0; j < kVolMaxIterations; ++j) {
                const int *nRays = pb.counts + kCountStride * j, *nShadow = nRays + 1;
                B2_CUDA(ctx, LaunchTraversePersist(sc, pb.rays[j & 1], nActive, (b200pt_hit *)pb.hits, nullptr, false,
                                                   r->dRayCounter, nullptr, ctx->smCount, stackDepth, ctx->refillMin,
                                                   ctx->nodeMin, ctx->stream, true, nRays, exact));
                ++launches;
                if (r->dev.nMedia) {
                    B2_CUDA(ctx, LaunchVolMedium(r->dev, pb, j, nPix, width, nActive, ctx->stream));
                    launches += 2;
                }
                B2_CUDA(ctx, LaunchClassify(r->dev, pb, j, nActive, ctx->stream));
                ++launches;
                B2_CUDA(ctx, LaunchShade(r->dev, pb, j, nPix, width, nActive, r->matMask, ctx->stream, &launches));
                int hc[2] = {0, 0};  // light samples of this iteration, rays of the next
                B2_CUDA(ctx, cudaMemcpyAsync(&hc[0], nShadow, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
                B2_CUDA(ctx, cudaMemcpyAsync(&hc[1], pb.counts + kCountStride * (j + 1), sizeof(int), cudaMemcpyDeviceToHost,
                                             ctx->stream));
                B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                usedIters = j + 1;
                // the transmittance walks: closest hit of every light-sample ray, one segment, until none goes on
                int live = hc[0];
                for (int seg = 0; live > 0 && seg < 64; ++seg) {
                    B2_CUDA(ctx, cudaMemsetAsync(pb.shContinue, 0, sizeof(int), ctx->stream));
                    B2_CUDA(ctx, LaunchTraversePersist(sc, pb.shRays, hc[0], (b200pt_hit *)pb.shHits, nullptr, false,
                                                       r->dRayCounter, nullptr, ctx->smCount, stackDepth, ctx->refillMin,
                                                       ctx->nodeMin, ctx->stream, true, nShadow, exact));
                    B2_CUDA(ctx, LaunchShadowTr(r->dev, pb, j, hc[0], ctx->stream));
                    launches += 3;
                    B2_CUDA(ctx, cudaMemcpyAsync(&live, pb.shContinue, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
                    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                }
                volWalksCut += uint64_t(live);  // walks still crossing interfaces after 64 segments are dropped
                if (hc[1] == 0) break;
                if (j == kVolMaxIterations - 1) volPathsCut += uint64_t(hc[1]);  // paths alive at the iteration cap
            }
            B2_CUDA(ctx, LaunchFinish(r->dev, pb, nActive, ctx->stream));
            B2_CUDA(ctx, LaunchFilmAdd(pb, nPix, ns, r->dFilm, ctx->stream));
            if (r->dev.film.type >= 1) {
                B2_CUDA(ctx, LaunchFilmAddSpectral(r->dev.film, pb, nPix, ns, r->dFilm, ctx->stream));
                ++launches;
            }
            if (r->dev.film.type == 2) {
                B2_CUDA(ctx, LaunchFilmAddGBuffer(r->dev.film, pb, nPix, ns, r->dFilm, r->dFilmIds, r->dev.materialIds, ctx->stream));
                ++launches;
            }
            B2_CUDA(ctx, LaunchSumCounts(pb, usedIters, r->dCounters, ctx->stream));
            launches += 3;
            continue;
        }
        // vertex j: closest hit of the rays in queue j -> material queues -> shading (next rays into queue
        // j+1, light samples into the shadow queue) -> visibility of the light samples -> radiance.
        // Vertex maxDepth only collects emission (cpu/integrators.cpp:690-692).
        for (int j = 0; j <= maxDepth; ++j) {
            const int *nRays = pb.counts + kCountStride * j, *nShadow = nRays + 1;
            B2_CUDA(ctx, LaunchTraversePersist(sc, pb.rays[j & 1], nActive, (b200pt_hit *)pb.hits, nullptr, false,
                                               r->dRayCounter, nullptr, ctx->smCount, stackDepth, ctx->refillMin,
                                               ctx->nodeMin, ctx->stream, true, nRays, exact));
            B2_CUDA(ctx, LaunchClassify(r->dev, pb, j, nActive, ctx->stream));
            launches += 2;
            B2_CUDA(ctx, LaunchShade(r->dev, pb, j, nPix, width, nActive, r->matMask, ctx->stream, &launches));
            if (j < maxDepth) {
                B2_CUDA(ctx, LaunchTraversePersist(sc, pb.shRays, nActive, nullptr, pb.occluded, true, r->dRayCounter,
                                                   nullptr, ctx->smCount, stackDepth, ctx->refillMin, ctx->nodeMin,
                                                   ctx->stream, false, nShadow, exact));
                B2_CUDA(ctx, LaunchResolveShadows(pb, j, nActive, ctx->stream));
                launches += 2;
            }
        }
        B2_CUDA(ctx, LaunchFinish(r->dev, pb, nActive, ctx->stream));
        B2_CUDA(ctx, LaunchFilmAdd(pb, nPix, ns, r->dFilm, ctx->stream));
        if (r->dev.film.type >= 1) {
            B2_CUDA(ctx, LaunchFilmAddSpectral(r->dev.film, pb, nPix, ns, r->dFilm, ctx->stream));
            ++launches;
        }
        if (r->dev.film.type == 2) {
            B2_CUDA(ctx, LaunchFilmAddGBuffer(r->dev.film, pb, nPix, ns, r->dFilm, r->dFilmIds, r->dev.materialIds, ctx->stream));
            ++launches;
        }
        B2_CUDA(ctx, LaunchSumCounts(pb, maxDepth + 1, r->dCounters, ctx->stream));
        launches += 3;
    }
    B2_CUDA(ctx, cudaEventRecord(r->ev1, ctx->stream));
    r->timed = true;
    r->stats.samples = (uint64_t)n_samples * nPix;
    r->stats.kernel_launches = (uint32_t)launches;
    r->stats.vol_paths_cut = volPathsCut;
    r->stats.vol_walks_cut = volWalksCut;
    return B200PT_OK;
}

int b200pt_film_device_ptr(b200pt_render *r, void **ptr, uint64_t *n_doubles) {
    if (!r || !ptr) return B200PT_ERR_INVALID;
    *ptr = r->dFilm;
    if (n_doubles) *n_doubles = r->filmDoubles;
    return B200PT_OK;
}

// ncclReduce out of the NCCL the process already has (the host made `comm` with it), else out of libnccl.so.2: the
// library itself does not link NCCL, a single-GPU host never needs it.
int b200pt_film_reduce(b200pt_render *r, void *nccl_comm, int32_t root, uint64_t n_doubles) {
    if (!r || !nccl_comm || root < 0) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = r->scene->ctx;
    if (n_doubles == 0 || n_doubles > r->filmDoubles) n_doubles = r->filmDoubles;
    using ReduceFn = int (*)(const void *, void *, size_t, int, int, int, void *, cudaStream_t);
    static ReduceFn reduce = nullptr;
    static std::mutex mu;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!reduce) {
            void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
            if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
            if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
            reduce = h ? reinterpret_cast<ReduceFn>(dlsym(h, "ncclReduce")) : nullptr;
            if (!reduce) {
                ctx->err = "b200pt_film_reduce: no NCCL in this process and libnccl.so.2 cannot be loaded";
                return B200PT_ERR_INVALID;
            }
        }
    }
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    // ncclReduce(send, recv, count, ncclFloat64 = 8, ncclSum = 0, root, comm, stream), in place on the film buffer
    int rc = reduce(r->dFilm, r->dFilm, size_t(n_doubles), 8, 0, root, nccl_comm, ctx->stream);
    if (rc != 0) {
        ctx->err = "b200pt_film_reduce: ncclReduce returned " + std::to_string(rc);
        return B200PT_ERR_CUDA;
    }
    return B200PT_OK;
}

int b200pt_film_read(b200pt_render *r, double *rgb_weight) {
    if (!r || !rgb_weight) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = r->scene->ctx;
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    B2_CUDA(ctx, cudaMemcpyAsync(rgb_weight, r->dFilm, sizeof(double) * r->filmDoubles, cudaMemcpyDeviceToHost,
                                 ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200PT_OK;
}

int b200pt_film_image(b200pt_render *r, float *rgb) {
    if (!r || !rgb) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = r->scene->ctx;
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    B2_CUDA(ctx, LaunchFilmImage(r->dev.film, r->dFilm, r->nPixels, r->dImage, ctx->stream));
    B2_CUDA(ctx, cudaMemcpyAsync(rgb, r->dImage, sizeof(float) * 3 * r->nPixels, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200PT_OK;
}

int b200pt_film_gbuffer_image(b200pt_render *r, uint32_t channels, float *out, int32_t *n_channels) {
    if (!r || !n_channels) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = r->scene->ctx;
    if (r->dev.film.type != 2) {
        ctx->err = "b200pt_film_gbuffer_image: the film is not a gbuffer film";
        return B200PT_ERR_INVALID;
    }
    int nc = 3;
    if (channels & B200PT_GBUFFER_RADIANCE) nc += r->dev.film.nBuckets;
    if (channels & B200PT_GBUFFER_ALBEDO) nc += 3;
    if (channels & B200PT_GBUFFER_POSITION) nc += 3;
    if (channels & B200PT_GBUFFER_MATERIAL) nc += 1;
    if (channels & B200PT_GBUFFER_INSTANCE) nc += 1;
    if (channels & B200PT_GBUFFER_NS) nc += 3;
    if (channels & B200PT_GBUFFER_NORMAL) nc += 3;
    if (channels & B200PT_GBUFFER_VARIANCE) nc += 3;
    if (channels & B200PT_GBUFFER_RELATIVE_VARIANCE) nc += 3;
    *n_channels = nc;
    if (!out) return B200PT_OK;
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    B2_CUDA(ctx, LaunchFilmGBufferImage(r->dev.film, r->dFilm, r->dFilmIds, r->nPixels, channels, nc, r->dImage, ctx->stream));
    B2_CUDA(ctx, cudaMemcpyAsync(out, r->dImage, sizeof(float) * size_t(nc) * r->nPixels, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200PT_OK;
}

int b200pt_film_write_exr(b200pt_render *r, uint32_t channels, const char *path, int32_t fp16) {
    if (!r || !path) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = r->scene->ctx;
    const FilmDev &film = r->dev.film;
    const int W = r->host.width, H = r->host.height, nB = film.nBuckets;
    // ---- channel names in the order the image kernels write them
    std::vector<std::string> names = {"R", "G", "B"};
    auto numbered = [&](const char *prefix, int n) {
        for (int i = 1; i <= n; ++i) {
            char b[32];
            snprintf(b, sizeof b, "%s%02d", prefix, i);
            names.push_back(b);
        }
    };
    if (film.type == 2) {
        if (channels & B200PT_GBUFFER_RADIANCE) numbered("Radiance.C", nB);
        if (channels & B200PT_GBUFFER_ALBEDO) for (const char *c : {"Albedo.R", "Albedo.G", "Albedo.B"}) names.push_back(c);
        if (channels & B200PT_GBUFFER_POSITION) for (const char *c : {"P.X", "P.Y", "P.Z"}) names.push_back(c);
        if (channels & B200PT_GBUFFER_MATERIAL) names.push_back("MaterialId");
        if (channels & B200PT_GBUFFER_INSTANCE) names.push_back("InstanceId");
        if (channels & B200PT_GBUFFER_NS) for (const char *c : {"Ns.X", "Ns.Y", "Nx.Z"}) names.push_back(c);  // (sic, film.cpp:765)
        if (channels & B200PT_GBUFFER_NORMAL) for (const char *c : {"N.X", "N.Y", "N.Z"}) names.push_back(c);
        if (channels & B200PT_GBUFFER_VARIANCE) for (const char *c : {"Variance.R", "Variance.G", "Variance.B"}) names.push_back(c);
        if (channels & B200PT_GBUFFER_RELATIVE_VARIANCE)
            for (const char *c : {"RelativeVariance.R", "RelativeVariance.G", "RelativeVariance.B"}) names.push_back(c);
    } else if (film.type == 1) {
        for (int i = 0; i < nB; ++i) {  // SpectralFilm::GetImage (film.cpp:1235-1243): bucket centre, '.' written as ','
            const float t = (i + 0.5f) / nB;
            float lambda = (1 - t) * 395.f + t * 705.f;  // Lerp(t, lambdaMin, lambdaMax)
            char b[48];
            snprintf(b, sizeof b, "S0.%.3fnm", lambda);
            std::string s = b;
            std::replace(s.begin() + 3, s.end(), '.', ',');
            names.push_back(s);
        }
    }
    const int nc = int(names.size());
    // ---- the float image on the device
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    float *dImg = nullptr;
    unsigned char *dOut = nullptr, *hOut = nullptr;
    int *dOrder = nullptr;
    const size_t px = size_t(W) * H, bytes = fp16 ? 2 : 4;
    const size_t payload = size_t(H) * (8 + size_t(nc) * W * bytes);
    auto cleanup = [&]() {
        cudaFree(dImg);
        cudaFree(dOut);
        cudaFree(dOrder);
        cudaFreeHost(hOut);
    };
#define B2_TRY(call)                                                            \
    do {                                                                        \
        cudaError_t e_ = (call);                                                \
        if (e_ != cudaSuccess) {                                                \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);      \
            cleanup();                                                          \
            return B200PT_ERR_CUDA;                                             \
        }                                                                       \
    } while (0)
    B2_TRY(cudaMalloc((void **)&dImg, sizeof(float) * px * nc));
    if (film.type == 2) {
        B2_TRY(LaunchFilmGBufferImage(film, r->dFilm, r->dFilmIds, r->nPixels, channels, nc, dImg, ctx->stream));
    } else {
        // R G B (channels-last, 3 wide) and the buckets (nB wide) side by side into the nc-wide image
        B2_TRY(LaunchFilmImage(film, r->dFilm, r->nPixels, r->dImage, ctx->stream));
        B2_TRY(cudaMemcpy2DAsync(dImg, sizeof(float) * nc, r->dImage, sizeof(float) * 3, sizeof(float) * 3, px,
                                 cudaMemcpyDeviceToDevice, ctx->stream));
        if (film.type == 1) {
            B2_TRY(LaunchFilmSpectralImage(film, r->dFilm, r->nPixels, r->dImage, ctx->stream));
            B2_TRY(cudaMemcpy2DAsync(dImg + 3, sizeof(float) * nc, r->dImage, sizeof(float) * nB, sizeof(float) * nB, px,
                                     cudaMemcpyDeviceToDevice, ctx->stream));
        }
    }
    // ---- file order = alphabetical (OpenEXR stores the channel list sorted by name)
    std::vector<int> order(nc);
    for (int i = 0; i < nc; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return names[a] < names[b]; });
    B2_TRY(cudaMalloc((void **)&dOrder, sizeof(int) * nc));
    B2_TRY(cudaMemcpyAsync(dOrder, order.data(), sizeof(int) * nc, cudaMemcpyHostToDevice, ctx->stream));
    B2_TRY(cudaMalloc((void **)&dOut, payload));
    B2_TRY(cudaMallocHost((void **)&hOut, payload));
    B2_TRY(LaunchExrPack(dImg, W, H, nc, dOrder, fp16 != 0, film.y0, dOut, ctx->stream));
    B2_TRY(cudaMemcpyAsync(hOut, dOut, payload, cudaMemcpyDeviceToHost, ctx->stream));
    B2_TRY(cudaStreamSynchronize(ctx->stream));
#undef B2_TRY
    // ---- header (OpenEXR file layout: magic, version, attributes, offset table, chunks)
    std::string h;
    auto put32 = [&](int32_t v) { h.append(reinterpret_cast<const char *>(&v), 4); };
    auto putf = [&](float v) { h.append(reinterpret_cast<const char *>(&v), 4); };
    auto attr = [&](const char *name, const char *type, const std::string &value) {
        h += name;
        h.push_back('\0');
        h += type;
        h.push_back('\0');
        put32(int32_t(value.size()));
        h += value;
    };
    put32(20000630);
    put32(2);
    {
        std::string v;
        for (int c = 0; c < nc; ++c) {
            v += names[order[c]];
            v.push_back('\0');
            int32_t type = fp16 ? 1 : 2, one = 1;
            v.append(reinterpret_cast<const char *>(&type), 4);
            v.append(4, '\0');  // pLinear + reserved
            v.append(reinterpret_cast<const char *>(&one), 4);
            v.append(reinterpret_cast<const char *>(&one), 4);
        }
        v.push_back('\0');
        attr("channels", "chlist", v);
    }
    attr("compression", "compression", std::string(1, '\0'));
    auto box = [&](int x0, int y0, int x1, int y1) {
        std::string v;
        for (int32_t q : {x0, y0, x1, y1}) v.append(reinterpret_cast<const char *>(&q), 4);
        return v;
    };
    attr("dataWindow", "box2i", box(film.x0, film.y0, film.x1 - 1, film.y1 - 1));
    attr("displayWindow", "box2i", box(0, 0, film.fullResX - 1, film.fullResY - 1));
    attr("lineOrder", "lineOrder", std::string(1, '\0'));
    {
        std::string s = h;
        h.clear();
        putf(1.f);
        std::string one = h;
        h.clear();
        putf(0.f);
        putf(0.f);
        std::string centre = h;
        h = s;
        attr("pixelAspectRatio", "float", one);
        attr("screenWindowCenter", "v2f", centre);
        attr("screenWindowWidth", "float", one);
    }
    h.push_back('\0');
    const uint64_t first = h.size() + 8ull * H, lineChunk = 8 + uint64_t(nc) * W * bytes;
    for (int y = 0; y < H; ++y) {
        uint64_t off = first + lineChunk * y;
        h.append(reinterpret_cast<const char *>(&off), 8);
    }
    FILE *f = fopen(path, "wb");
    bool ok = f && fwrite(h.data(), 1, h.size(), f) == h.size() && fwrite(hOut, 1, payload, f) == payload;
    if (f) ok = (fclose(f) == 0) && ok;
    cleanup();
    if (!ok) {
        ctx->err = std::string("b200pt_film_write_exr: cannot write ") + path;
        return B200PT_ERR_INVALID;
    }
    return B200PT_OK;
}

int b200pt_film_spectral_image(b200pt_render *r, float *buckets) {
    if (!r || !buckets) return B200PT_ERR_INVALID;
    b200pt_ctx *ctx = r->scene->ctx;
    if (r->dev.film.type != 1) {  // (a gbuffer film's buckets come out through b200pt_film_gbuffer_image)
        ctx->err = "b200pt_film_spectral_image: the film is not a spectral film";
        return B200PT_ERR_INVALID;
    }
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    B2_CUDA(ctx, LaunchFilmSpectralImage(r->dev.film, r->dFilm, r->nPixels, r->dImage, ctx->stream));
    B2_CUDA(ctx, cudaMemcpyAsync(buckets, r->dImage, sizeof(float) * size_t(r->dev.film.nBuckets) * r->nPixels,
                                 cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200PT_OK;
}

int b200pt_render_get_stats(const b200pt_render *cr, b200pt_render_stats *out) {
    if (!cr || !out) return B200PT_ERR_INVALID;
    b200pt_render *r = const_cast<b200pt_render *>(cr);
    b200pt_ctx *ctx = r->scene->ctx;
    if (r->timed) {
        B2_CUDA(ctx, cudaSetDevice(ctx->device));
        B2_CUDA(ctx, cudaEventSynchronize(r->ev1));
        float ms = 0;
        B2_CUDA(ctx, cudaEventElapsedTime(&ms, r->ev0, r->ev1));
        unsigned long long c[2] = {0, 0};
        B2_CUDA(ctx, cudaMemcpy(c, r->dCounters, sizeof(c), cudaMemcpyDeviceToHost));
        r->stats.gpu_ms = ms;
        r->stats.closest_rays = c[0];
        r->stats.shadow_rays = c[1];
    }
    *out = r->stats;
    return B200PT_OK;
}

}  // extern "C"
