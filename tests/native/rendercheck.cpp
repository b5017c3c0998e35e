// tests/native/rendercheck.cpp - TEST-ONLY shared library that runs the render path's host+device
// functions (pbrt-v4_b200/csrc/render_*.h) on the CPU, one path at a time, with the PairNode walk of
// traverse_persist.cu for the rays.  Lets the whole pipeline be compared with pbrt_ref in the
// GPU-less container.  Not part of libb200pt.so; the product never calls it.
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../pbrt-v4_b200/csrc/render_core.h"
#include "../../pbrt-v4_b200/csrc/render_host.h"
#include "../../pbrt-v4_b200/csrc/scene_host.h"
#include "trace_host.h"

using namespace b200pt;

namespace {

// closest hit -> internal record {slot, t, b0, b1, b2, instance + 1}; any = occlusion only
bool Trace(const HostScene &hs, V3 o, V3 d, float tMax, bool any, uint32_t hit[6]) {
    float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    Best best;
    bool h = TracePairs(hs, hs.rootRef, hs.bounds, oo, dd, tMax, any, best, -1, false);
    if (any) return h;
    if (best.slot < 0) { hit[0] = kMissTag; return false; }
    hit[0] = uint32_t(best.slot);
    memcpy(&hit[1], &tMax, 4); memcpy(&hit[2], &best.b0, 4); memcpy(&hit[3], &best.b1, 4); memcpy(&hit[4], &best.b2, 4);
    hit[5] = uint32_t(best.inst + 1);
    return true;
}

// the IO of ShadeVertex on the host: plain arrays, shadow ray resolved right after the vertex
struct HostIO {
    float L[kNLambda], Ld[kNLambda], beta[kNLambda];
    float ru[kNLambda], rl[kNLambda];  // volpath: rescaled path probabilities
    ShadowVolInfo info;
    float Ru(int k) const { return ru[k]; }
    void SetRu(int k, float v) { ru[k] = v; }
    float Rl(int k) const { return rl[k]; }
    void SetRl(int k, float v) { rl[k] = v; }
    void ShadowEndVol(bool valid, V3 o, V3 d, float tMax, const ShadowVolInfo &i) {
        ShadowEnd(valid, o, d, tMax);
        info = i;
    }
    bool shadow;
    V3 shO, shD;
    float shTMax;
    float Beta(int k) const { return beta[k]; }
    void SetBeta(int k, float v) { beta[k] = v; }
    float refl[kNLambda];  // the vertex's cached reflectance
    float R(int k) const { return refl[k]; }
    void SetR(int k, float v) { refl[k] = v; }
    void VisibleSurface(V3, V3, V3, int, int) {}
    bool WantsAlbedo() const { return false; }
    void VisibleSurfaceAlbedo(const float *) {}
    void AddL(int k, float v) { L[k] += v; }
    void ShadowBegin() {}
    void SetLd(int k, float v) { Ld[k] = v; }
    void ShadowEnd(bool valid, V3 o, V3 d, float tMax) {
        shadow = valid;
        shO = o;
        shD = d;
        shTMax = tMax;
    }
};

// One pixel sample through the host+device path functions: the depth loop of the kernels, one path at a time.
void RunPath(const RenderSceneDev &rs, const HostScene &hs, PathRegs &p, HostIO &io) {
    const bool vol = rs.integrator.type == 1;
    for (int i = 0; i < kNLambda; ++i) io.L[i] = 0.f, io.beta[i] = io.ru[i] = io.rl[i] = 1.f;
    while (p.flags & PF_ALIVE) {
        uint32_t hit[6] = {kMissTag, 0, 0, 0, 0, 0};
        Trace(hs, p.rayO, p.rayD, p.rayTMax, false, hit);
        io.shadow = false;
        int r = VOL_SURVIVE;
        if (vol && p.medium >= 0) {
            float tHit = INFINITY;
            if (hit[0] != kMissTag) memcpy(&tHit, &hit[1], 4);
            r = VolMediumStep(rs, p, tHit, io);
            if (r == VOL_TERMINATED) KillPath(p);
        }
        if (r == VOL_SURVIVE) ShadeVertex(rs, p, hit, io);
        if (!io.shadow) continue;
        if (!vol) {
            if (!Trace(hs, io.shO, io.shD, io.shTMax, true, nullptr))
                for (int i = 0; i < kNLambda; ++i) io.L[i] += io.Ld[i];
            continue;
        }
        // SampleLd's transmittance walk (cpu/integrators.cpp:1329-1395)
        ShadowRegs sh;
        sh.o = io.shO;
        sh.d = io.shD;
        sh.medium = io.info.medium;
        sh.rng = PcgSeed(Hash3f(sh.o.x, sh.o.y, sh.o.z), Hash3f(sh.d.x, sh.d.y, sh.d.z));
        sh.lightPi = io.info.lightPi;
        sh.lightN = io.info.lightN;
        float T[kNLambda], rl[kNLambda], ru[kNLambda];
        for (int i = 0; i < kNLambda; ++i) T[i] = rl[i] = ru[i] = 1.f;
        int res;
        do {
            uint32_t sh_hit[6] = {kMissTag, 0, 0, 0, 0, 0};
            Trace(hs, sh.o, sh.d, io.shTMax, false, sh_hit);
            res = ShadowTrSegment(rs, p.lambdaU, sh, sh_hit, T, rl, ru);
        } while (res == SHADOW_CONTINUE);
        if (res == SHADOW_REACHED)
            ShadowTrFinish(T, rl, ru, io.info.p_l, io.info.scatterPDF, [&](int k) { return io.Ld[k]; },
                           [&](int k) { return io.ru[k]; }, [&](int k, float v) { io.L[k] += v; });
    }
}

}  // namespace

// rgbw: 4 doubles per pixel (rgbSum, weightSum), row major; a spectral film appends its bucket planes
// (layout of b200pt_film_read).  Returns 0 or an error code; err text in msg.
extern "C" int rendercheck_run(const b200pt_scene_desc *sd, const b200pt_render_desc *rd, int firstSample,
                               int nSamples, int stride, double *rgbw, char *msg, int msgLen, int nThreads) {
    HostScene hs;
    RenderHost rh;
    std::string err;
    int rc = AssembleScene(sd, hs, err);
    if (rc == 0) rc = BuildRenderTables(rd, hs, sd->n_meshes, rh, err);
    if (rc != 0) {
        snprintf(msg, msgLen, "%s", err.c_str());
        return rc;
    }
    RenderSceneDev rs = rh.rs;
    rs.spectra = rh.spectra.data();
    rs.materials = rh.materials.data();
    rs.lights = rh.lights.data();
    rs.meshes = rh.meshes.data();
    rs.lightBvhNodes = rh.lightBvhNodes.data();
    rs.lightBitTrails = rh.lightBitTrails.data();
    rs.infiniteLights = rh.infiniteLights.data();
    rs.lightGeom = rh.lightGeom.empty() ? nullptr : rh.lightGeom.data();
    rs.film.filterTable = rh.filterTable.empty() ? nullptr : rh.filterTable.data();
    rs.envImage = rh.envImage.data();
    rs.envDist = rh.envDist.data();
    rs.prims = hs.prims.data();
    rs.instances = hs.instances.data();
    rs.camera.lens = rh.lens.data();
    rs.camera.pupilBounds = rh.pupilBounds.data();
    rs.cornerAttr = rh.cornerAttr.data();
    for (size_t i = 0; i < rh.media.size(); ++i) {
        rh.media[i].density = rh.mediumGrids[i].first.data();
        rh.media[i].majorants = rh.mediumGrids[i].second.data();
        rh.media[i].leScale = rh.mediumLeGrids[i].empty() ? nullptr : rh.mediumLeGrids[i].data();
        rh.media[i].temperature = rh.mediumTempGrids[i].empty() ? nullptr : rh.mediumTempGrids[i].data();
    }
    rs.media = rh.media.data();
    const int W = rh.width, H = rh.height;
    if (nThreads < 1) nThreads = 1;
    std::vector<std::thread> pool;
    for (int t = 0; t < nThreads; ++t)
        pool.emplace_back([&, t]() {
            for (int y = t; y < H; y += nThreads)
                for (int x = 0; x < W; ++x)
                    for (int k = 0; k < nSamples; ++k) {
                        PathRegs p;
                        GenerateCameraPath(rs, p, rs.film.x0 + x, rs.film.y0 + y, firstSample + k * stride);
                        HostIO io;
                        RunPath(rs, hs, p, io);
                        float out[4];
                        const float cw = p.cameraWeight;
                        FinishPath(rs, p.lambdaU, p.flags, p.filterWeight, [&](int i) { return cw * io.L[i]; }, out);
                        double *px = rgbw + 4 * (size_t(y) * W + x);
                        for (int c = 0; c < 4; ++c) px[c] += out[c];
                        if (rs.film.type == 1) {  // spectral film: planes behind the rgbw block, as on the device
                            size_t nPix = size_t(W) * H, pix = size_t(y) * W + x;
                            double *bs = rgbw + 4 * nPix + pix, *ws = bs + size_t(rs.film.nBuckets) * nPix;
                            SpectralFilmAdd(rs.film, p.lambdaU, p.filterWeight, [&](int i) { return cw * io.L[i]; },
                                            [&](int b, float v, float wt) {
                                                bs[size_t(b) * nPix] += v;
                                                ws[size_t(b) * nPix] += wt;
                                            });
                        }
                    }
        });
    for (auto &th : pool) th.join();
    return 0;
}

// Debug twin of `dumpscene_ref scene.pbrt --path x y s`: radiance of one pixel sample with maxDepth cut
// at 0, 1, 2, ... and the vertices visited.
extern "C" int rendercheck_path(const b200pt_scene_desc *sd, const b200pt_render_desc *rd, int x, int y, int si) {
    HostScene hs;
    RenderHost rh;
    std::string err;
    int rc = AssembleScene(sd, hs, err);
    if (rc == 0) rc = BuildRenderTables(rd, hs, sd->n_meshes, rh, err);
    if (rc != 0) return rc;
    RenderSceneDev rs = rh.rs;
    rs.spectra = rh.spectra.data();
    rs.materials = rh.materials.data();
    rs.lights = rh.lights.data();
    rs.meshes = rh.meshes.data();
    rs.lightBvhNodes = rh.lightBvhNodes.data();
    rs.lightBitTrails = rh.lightBitTrails.data();
    rs.infiniteLights = rh.infiniteLights.data();
    rs.lightGeom = rh.lightGeom.empty() ? nullptr : rh.lightGeom.data();
    rs.film.filterTable = rh.filterTable.empty() ? nullptr : rh.filterTable.data();
    rs.envImage = rh.envImage.data();
    rs.envDist = rh.envDist.data();
    rs.prims = hs.prims.data();
    rs.instances = hs.instances.data();
    rs.camera.lens = rh.lens.data();
    rs.camera.pupilBounds = rh.pupilBounds.data();
    rs.cornerAttr = rh.cornerAttr.data();
    for (size_t i = 0; i < rh.media.size(); ++i) {
        rh.media[i].density = rh.mediumGrids[i].first.data();
        rh.media[i].majorants = rh.mediumGrids[i].second.data();
        rh.media[i].leScale = rh.mediumLeGrids[i].empty() ? nullptr : rh.mediumLeGrids[i].data();
        rh.media[i].temperature = rh.mediumTempGrids[i].empty() ? nullptr : rh.mediumTempGrids[i].data();
    }
    rs.media = rh.media.data();
    int full = rs.integrator.maxDepth;
    for (int d = 0; d <= full; ++d) {
        rs.integrator.maxDepth = d;
        PathRegs p;
        GenerateCameraPath(rs, p, x, y, si);
        HostIO io;
        for (int i = 0; i < kNLambda; ++i) io.L[i] = 0.f, io.beta[i] = 1.f;
        V3 o = p.rayO, dd = p.rayD;
        while (p.flags & PF_ALIVE) {
            uint32_t hit[6] = {kMissTag, 0, 0, 0, 0, 0};
            Trace(hs, p.rayO, p.rayD, p.rayTMax, false, hit);
            if (d == full) {
                float t;
                memcpy(&t, &hit[1], 4);
                const PrimSlot *ps = hit[0] == kMissTag ? nullptr : &hs.prims[hit[0]];
                printf("   vertex depth %d: o %.9g %.9g %.9g d %.9g %.9g %.9g -> mesh %d tri %d t %.9g  beta0 %.9g dim %d\n",
                       p.depth, p.rayO.x, p.rayO.y, p.rayO.z, p.rayD.x, p.rayD.y, p.rayD.z, ps ? int(ps->mesh) : -1,
                       ps ? int(ps->tri) : -1, t, io.beta[0], p.dimension); printf("      etaScale %.9g beta0*etaScale %.9g\n", p.etaScale, io.beta[0]*p.etaScale);
            }
            io.shadow = false;
            ShadeVertex(rs, p, hit, io);
            bool occ = io.shadow && Trace(hs, io.shO, io.shD, io.shTMax, true, nullptr);
            if (d == full && io.shadow) printf("      shadow ray occluded %d Ld0 %.9g\n", int(occ), io.Ld[0]);
            if (io.shadow && !occ)
                for (int i = 0; i < kNLambda; ++i) io.L[i] += io.Ld[i];
        }
        printf("maxdepth %d  lambda0 %.9g  o %.9g %.9g %.9g  d %.9g %.9g %.9g  L0 %.9g L1 %.9g L15 %.9g L30 %.9g\n", d,
               LambdaIter(p.lambdaU).lam, o.x, o.y, o.z, dd.x, dd.y, dd.z, io.L[0], io.L[1], io.L[15], io.L[30]);
    }
    fflush(stdout);
    return 0;
}

// ---- small probes for tests/test_render_core.py
// BVHLightSampler::Sample at n reference points {p, ns, u}; PMF of light query[i] at the same point.
extern "C" int rendercheck_lightbvh(const b200pt_scene_desc *sd, const b200pt_render_desc *rd, const float *in,
                                    const int *query, int n, int *light, float *pmf) {
    HostScene hs;
    RenderHost rh;
    std::string err;
    int rc = AssembleScene(sd, hs, err);
    if (rc == 0) rc = BuildRenderTables(rd, hs, sd->n_meshes, rh, err);
    if (rc != 0) return rc;
    RenderSceneDev rs = rh.rs;
    rs.lights = rh.lights.data();
    rs.lightBvhNodes = rh.lightBvhNodes.data();
    rs.lightBitTrails = rh.lightBitTrails.data();
    rs.infiniteLights = rh.infiniteLights.data();
    for (int i = 0; i < n; ++i) {
        const float *c = in + 7 * i;
        float pm = 0.f;
        light[i] = LightBvhSample(rs, mk(c[0], c[1], c[2]), mk(c[3], c[4], c[5]), c[6], &pm);
        pmf[2 * i] = light[i] >= 0 ? pm : 0.f;
        pmf[2 * i + 1] = LightBvhPMF(rs, mk(c[0], c[1], c[2]), mk(c[3], c[4], c[5]), query[i]);
    }
    return 0;
}

extern "C" uint32_t rendercheck_sobol_column(int dim, int j) { return b200pt::SobolColumn(dim, j); }
extern "C" uint32_t rendercheck_sobol_dim1(uint64_t a) { return b200pt::SobolDim1(a); }
extern "C" int rendercheck_permute(int p, int digit) { return b200pt::ZSobolPermute(p, digit); }
extern "C" float rendercheck_spectrum_at(const b200pt_spectrum *s, float lambda) {
    return b200pt::SpectrumAt(*reinterpret_cast<const b200pt::SpectrumTable *>(s), lambda);
}
