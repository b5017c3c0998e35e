// oracle/flatten_ref.h - the reference's scene objects (BasicScene after parsing, built in RenderCPU's order)
// flattened into the named tables a host adaptor hands to b200pt_scene_create / b200pt_render_create.
// Shared by oracle/dumpscene_ref.cpp (tables -> .b2rs file) and oracle/pbrt_b200.cpp (tables -> C ABI).
// TEST INFRASTRUCTURE ONLY: includes the reference's headers with private members opened in this TU.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#define private public
#define protected public
#include <pbrt/pbrt.h>
#include <pbrt/cameras.h>
#include <pbrt/cpu/aggregates.h>
#include <pbrt/cpu/integrators.h>
#include <pbrt/cpu/primitive.h>
#include <pbrt/film.h>
#include <pbrt/filters.h>
#include <pbrt/lights.h>
#include <pbrt/lightsamplers.h>
#include <pbrt/materials.h>
#include <pbrt/media.h>
#include <pbrt/options.h>
#include <pbrt/parser.h>
#include <pbrt/samplers.h>
#include <pbrt/scene.h>
#include <pbrt/shapes.h>
#include <pbrt/textures.h>
#include <pbrt/util/mesh.h>
#include <pbrt/util/sampling.h>
#include <pbrt/util/spectrum.h>
#undef private
#undef protected

namespace pbrt {
struct alignas(32) LinearBVHNode {  // layout of cpu/aggregates.cpp:129-137
    Bounds3f bounds;
    union {
        int primitivesOffset;
        int secondChildOffset;
    };
    uint16_t nPrimitives;
    uint8_t axis;
};
}  // namespace pbrt
using namespace pbrt;

static void die(const std::string &msg) {
    fprintf(stderr, "flatten_ref: %s\n", msg.c_str());
    exit(2);
}

static std::vector<float> Mat(const SquareMatrix<4> &m) {
    std::vector<float> v(16);
    for (int i = 0; i < 16; ++i) v[i] = m[i / 4][i % 4];
    return v;
}

// spectra pool: 1 nm tables 395..705
struct SpectraPool {
    std::vector<float> values;
    std::vector<uint32_t> interp;
    std::map<std::vector<uint32_t>, int> seen;  // tables equal bit for bit are stored once
    int add(Spectrum s, uint32_t mode) {
        std::vector<uint32_t> key(312);
        float t[311];
        for (int l = 395; l <= 705; ++l) t[l - 395] = s(Float(l));
        memcpy(key.data(), t, sizeof(t));
        key[311] = mode;
        auto it = seen.find(key);
        if (it != seen.end()) return it->second;
        values.insert(values.end(), t, t + 311);
        interp.push_back(mode);
        return seen[key] = int(interp.size()) - 1;
    }
};

static int countNodes(const LinearBVHNode *nodes) {
    int maxIdx = 0;
    std::vector<int> stack{0};
    while (!stack.empty()) {
        int i = stack.back();
        stack.pop_back();
        maxIdx = std::max(maxIdx, i);
        if (nodes[i].nPrimitives == 0) {
            stack.push_back(i + 1);
            stack.push_back(nodes[i].secondChildOffset);
        }
    }
    return maxIdx + 1;
}


// The objects RenderCPU creates (cpu/render.cpp:25-55), kept alive for the flattening and for the caller.
struct RefObjects {
    std::map<std::string, Medium> media;
    NamedTextures textures;
    std::map<int, pstd::vector<Light> *> shapeIndexToAreaLights;
    std::vector<Light> lights;
    std::map<std::string, pbrt::Material> namedMaterials;
    std::vector<pbrt::Material> materials;
    std::string integratorName;
    ParameterDictionary integratorParams;
    Primitive accel;
    Camera camera;
    Film film;
    Sampler sampler;
};

inline void CreateRefObjects(BasicScene &scene, RefObjects &o) {
    o.media = scene.CreateMedia();
    o.textures = scene.CreateTextures();
    o.lights = scene.CreateLights(o.textures, &o.shapeIndexToAreaLights);
    scene.CreateMaterials(o.textures, &o.namedMaterials, &o.materials);
    o.integratorName = scene.integrator.name;
    o.integratorParams = scene.integrator.parameters;
    o.accel = scene.CreateAggregate(o.textures, o.shapeIndexToAreaLights, o.media, o.namedMaterials, o.materials);
    o.camera = scene.GetCamera();
    o.film = o.camera.GetFilm();
    o.sampler = scene.GetSampler();
    // Integrator constructor (cpu/integrators.h:66-73): infinite lights learn the scene's bounding sphere
    Bounds3f sceneBounds = o.accel ? o.accel.Bounds() : Bounds3f();
    for (auto &light : o.lights) light.Preprocess(sceneBounds);
}

// W: anything with f32 / i32 / u32 (name, dims, values) - the .b2rs Writer or an in-memory table set
template <class W>
void FlattenScene(RefObjects &ro, W &w, bool dumpBvh, size_t *nMeshOut = nullptr, size_t *nTriOut = nullptr,
                  size_t *nMaterialOut = nullptr, size_t *nSpectraOut = nullptr) {
    std::map<std::string, Medium> &media = ro.media;
    std::vector<Light> &lights = ro.lights;
    std::string &integratorName = ro.integratorName;
    ParameterDictionary &integratorParams = ro.integratorParams;
    Primitive &accel = ro.accel;
    Camera &camera = ro.camera;
    Film &film = ro.film;
    Sampler &sampler = ro.sampler;
    SpectraPool pool;

    // ---- camera: perspective, or the RealisticCamera's lens table and exit-pupil bounds (cameras.h:599-603)
    if (camera.Is<PerspectiveCamera>()) {
        const PerspectiveCamera *pc = camera.Cast<PerspectiveCamera>();
        w.u32("camera_type", {1}, {0});
        w.f32("camera_from_raster", {16}, Mat(pc->cameraFromRaster.GetMatrix()));
        w.f32("render_from_camera", {16}, Mat(pc->cameraTransform.renderFromCamera.startTransform.GetMatrix()));
        w.f32("camera_params", {2}, {pc->lensRadius, pc->focalDistance});
    } else if (camera.Is<RealisticCamera>()) {
        const RealisticCamera *rc = camera.Cast<RealisticCamera>();
        if (rc->apertureImage) die("aperture images are not dumped");
        w.u32("camera_type", {1}, {1});
        w.f32("render_from_camera", {16}, Mat(rc->cameraTransform.renderFromCamera.startTransform.GetMatrix()));
        std::vector<float> lens, pupil;
        for (const auto &e : rc->elementInterfaces)
            lens.insert(lens.end(), {e.curvatureRadius, e.thickness, e.eta, e.apertureRadius});
        for (const Bounds2f &b : rc->exitPupilBounds) pupil.insert(pupil.end(), {b.pMin.x, b.pMin.y, b.pMax.x, b.pMax.y});
        w.f32("camera_lens", {uint32_t(lens.size() / 4), 4}, lens);
        w.f32("camera_exit_pupil", {uint32_t(pupil.size() / 4), 4}, pupil);
        const Bounds2f &pe = rc->physicalExtent;
        w.f32("camera_extent", {5}, {pe.pMin.x, pe.pMin.y, pe.pMax.x, pe.pMax.y, film.Diagonal()});
    } else if (camera.Is<OmniCamera>()) {
        // ISET OmniCamera without a microlens array (cameras.h:856-1040): the RealisticCamera tables plus one
        // rigid transform per element interface and the chromatic-aberration switch
        const OmniCamera *oc = camera.Cast<OmniCamera>();
        if (oc->apertureImage) die("aperture images are not dumped");
        if (oc->microlens.elementInterfaces.size() > 0) {
            // the microlens array (cameras.h:940-960): interfaces + transforms as for the main lens, the grid, the
            // per-lens centre offsets, and the two scalars TraceFullLensSystemFromFilm / SampleMicrolensPupil read
            // (microlens.simulationRadius is assigned from a not yet initialised member in the constructor,
            // cameras.cpp:2530 - the value the object really holds is what is dumped)
            std::vector<float> ml, mxf, off;
            for (const auto &e : oc->microlensElementInterfaces) {
                if (e.asphericCoefficients.size() > 0) die("aspheric elements are not dumped");
                ml.insert(ml.end(), {e.curvatureRadius.x, e.thickness, e.eta, e.apertureRadius.x});
                std::vector<float> m = Mat(e.transform.GetMatrix()), mi = Mat(e.transform.GetInverseMatrix());
                mxf.insert(mxf.end(), m.begin(), m.end());
                mxf.insert(mxf.end(), mi.begin(), mi.end());
            }
            Vector2i dims = oc->microlensDimensions;
            if (oc->microlens.dimensions != dims) die("microlens dimensions disagree");
            if (oc->microlensOffsets.size() != size_t(dims.x) * dims.y) die("microlens offsets missing");
            for (const Vector2f &o : oc->microlensOffsets) off.insert(off.end(), {o.x, o.y});
            int R = oc->microlens.simulationRadius;
            if (R < 0 || R > 8) die("microlens.simulationRadius holds an indeterminate value in this run");
            w.f32("microlens_lens", {uint32_t(ml.size() / 4), 4}, ml);
            w.f32("microlens_transforms", {uint32_t(mxf.size() / 32), 32}, mxf);
            w.f32("microlens_offsets", {uint32_t(off.size() / 2), 2}, off);
            w.i32("microlens_ints", {3}, {dims.x, dims.y, R});
            w.f32("microlens_params", {1}, {oc->microlens.offsetFromSensor});
            fprintf(stderr, "dumpscene_ref: microlens %d x %d, simulation radius %d (asked: %d)\n", dims.x, dims.y, R,
                    oc->microlensSimulationRadius);
        }
        w.u32("camera_type", {1}, {2});
        w.f32("render_from_camera", {16}, Mat(oc->cameraTransform.renderFromCamera.startTransform.GetMatrix()));
        std::vector<float> lens, pupil, xf;
        for (const auto &e : oc->elementInterfaces) {
            if (e.asphericCoefficients.size() > 0) die("aspheric elements are not dumped (the reference skips them too)");
            lens.insert(lens.end(), {e.curvatureRadius.x, e.thickness, e.eta, e.apertureRadius.x});
            std::vector<float> m = Mat(e.transform.GetMatrix()), mi = Mat(e.transform.GetInverseMatrix());
            xf.insert(xf.end(), m.begin(), m.end());
            xf.insert(xf.end(), mi.begin(), mi.end());
        }
        for (const Bounds2f &b : oc->exitPupilBounds) pupil.insert(pupil.end(), {b.pMin.x, b.pMin.y, b.pMax.x, b.pMax.y});
        w.f32("camera_lens", {uint32_t(lens.size() / 4), 4}, lens);
        w.f32("camera_lens_transforms", {uint32_t(xf.size() / 32), 32}, xf);
        w.f32("camera_exit_pupil", {uint32_t(pupil.size() / 4), 4}, pupil);
        const Bounds2f &pe = oc->physicalExtent;
        w.f32("camera_extent", {5}, {pe.pMin.x, pe.pMin.y, pe.pMax.x, pe.pMax.y, film.Diagonal()});
        w.u32("camera_flags", {1}, {oc->caFlag ? 1u : 0u});
        if (getenv("DUMPSCENE_CAMDEBUG")) {
            fprintf(stderr, "omni: back thickness %g rearZ %g pupil[0] %g %g %g %g\n", oc->elementInterfaces.back().thickness,
                    oc->LensRearZ(), pupil[0], pupil[1], pupil[2], pupil[3]);
            CameraSample cs;
            cs.pFilm = Point2f(film.FullResolution().x / 2 + 0.5f, film.FullResolution().y / 2 + 0.5f);
            cs.pLens = Point2f(0.5f, 0.5f);
            SampledWavelengths lam = SampledWavelengths::SampleUniform(0.5f);
            auto cr = oc->GenerateRay(cs, lam);
            fprintf(stderr, "omni: centre ray %s\n", cr ? "ok" : "vignetted");
        }
    } else
        die("only the perspective, realistic and omni cameras are dumped so far");

    // ---- film (RGBFilm) and its sensor
    auto dumpFilm = [&](const auto *rf) {
        Bounds2i pb = rf->pixelBounds;
        w.i32("film_ints", {6}, {pb.pMin.x, pb.pMin.y, pb.pMax.x, pb.pMax.y, rf->fullResolution.x, rf->fullResolution.y});
        // filter_type as include/b200pt.h numbers them (4 = triangle)
        const uint32_t ft = rf->filter.template Is<BoxFilter>() ? 0u : rf->filter.template Is<GaussianFilter>() ? 1u
                          : rf->filter.template Is<MitchellFilter>() ? 2u : rf->filter.template Is<LanczosSincFilter>() ? 3u : 4u;
        w.u32("filter_type", {1}, {ft});
        // the filters sampled through a FilterSampler: its tabulated function (filters.cpp:92-118), [ny][nx]
        auto dumpTable = [&](const FilterSampler &fs) {
            std::vector<float> tab;
            for (int y = 0; y < fs.f.YSize(); ++y)
                for (int x = 0; x < fs.f.XSize(); ++x) tab.push_back(fs.f(x, y));
            w.f32("filter_table", {uint32_t(fs.f.YSize()), uint32_t(fs.f.XSize())}, tab);
        };
        if (ft == 1) dumpTable(rf->filter.template Cast<GaussianFilter>()->sampler);
        if (ft == 2) dumpTable(rf->filter.template Cast<MitchellFilter>()->sampler);
        if (ft == 3) dumpTable(rf->filter.template Cast<LanczosSincFilter>()->sampler);
        Vector2f fr = rf->filter.Radius();
        w.f32("film_params", {4}, {fr.x, fr.y, rf->sensor->imagingRatio, rf->maxComponentValue});
        std::vector<float> om(9);
        for (int i = 0; i < 9; ++i) om[i] = rf->outputRGBFromSensorRGB[i / 3][i % 3];
        w.f32("film_matrix", {9}, om);
        int sr = pool.add(&rf->sensor->r_bar, 0), sg = pool.add(&rf->sensor->g_bar, 0), sb = pool.add(&rf->sensor->b_bar, 0);
        w.i32("film_sensor", {3}, {sr, sg, sb});
    };
    if (film.Is<RGBFilm>()) {
        w.u32("film_type", {1}, {0});
        dumpFilm(film.Cast<RGBFilm>());
    } else if (film.Is<SpectralFilm>()) {  // film.h:430-560
        const SpectralFilm *sf = film.Cast<SpectralFilm>();
        w.u32("film_type", {1}, {1});
        dumpFilm(sf);
        w.f32("film_spectral", {3}, {float(sf->nBuckets), sf->lambdaMin, sf->lambdaMax});
    } else if (film.Is<GBufferFilm>()) {  // film.h:322-408, the ISET spectral film
        const GBufferFilm *gf = film.Cast<GBufferFilm>();
        w.u32("film_type", {1}, {2});
        dumpFilm(gf);
        w.f32("film_spectral", {3}, {31.f, 395.f, 705.f});  // LambdaToBucket, film.h:383-389
        // the AOVs' space: ApplyInverse(outputFromRender) for "camera" (film.cpp:1088-1090), outputFromRender for "world"
        Transform ofr = gf->outputFromRender.startTransform;
        Transform eff = gf->applyInverse ? Inverse(ofr) : ofr;
        std::vector<float> m(16), mi(16);
        for (int i = 0; i < 16; ++i) {
            m[i] = eff.GetMatrix()[i / 4][i % 4];
            mi[i] = eff.GetInverseMatrix()[i / 4][i % 4];
        }
        w.f32("gbuffer_from_render", {16}, m);
        w.f32("gbuffer_from_render_inv", {16}, mi);
        if (gf->writeAlbedo) {  // Albedo.*: what (albedo * illuminant).ToRGB(lambda, colorSpace) needs (film.cpp:640-644)
            w.i32("gbuffer_albedo", {4}, {pool.add(&Spectra::X(), 0), pool.add(&Spectra::Y(), 0), pool.add(&Spectra::Z(), 0),
                                          pool.add(&gf->colorSpace->illuminant, 0)});
            std::vector<float> rx(9);
            for (int i = 0; i < 9; ++i) rx[i] = gf->colorSpace->RGBFromXYZ[i / 3][i % 3];
            w.f32("gbuffer_rgb_from_xyz", {9}, rx);
        }
    } else
        die("only the rgb, spectral and gbuffer films are dumped so far");

    // ---- sampler
    if (sampler.Is<ZSobolSampler>()) {
        const ZSobolSampler *zs = sampler.Cast<ZSobolSampler>();
        w.i32("sampler", {6}, {0, 1 << zs->log2SamplesPerPixel, zs->seed, int(zs->randomize),
                               zs->log2SamplesPerPixel, zs->nBase4Digits});
    } else if (sampler.Is<IndependentSampler>()) {
        const IndependentSampler *is = sampler.Cast<IndependentSampler>();
        w.i32("sampler", {6}, {1, is->samplesPerPixel, is->seed, 0, 0, 0});
    } else
        die("only the zsobol and independent samplers are dumped so far");

    // ---- integrator parameters (cpu/integrators.cpp PathIntegrator::Create)
    int maxDepth = integratorParams.GetOneInt("maxdepth", 5);
    std::string ls = integratorParams.GetOneString("lightsampler", "bvh");
    bool regularize = integratorParams.GetOneBool("regularize", false);
    int lsId = ls == "uniform" ? 0 : (ls == "power" ? 1 : 2);
    w.i32("integrator", {3}, {maxDepth, regularize ? 1 : 0, lsId});
    if (integratorName != "path" && integratorName != "volpath") die("only the path and volpath integrators are dumped");
    const bool volpath = integratorName == "volpath";
    w.u32("integrator_type", {1}, {volpath ? 1u : 0u});
    // media in the order of the scene's name -> Medium map; index -1 = no medium
    std::map<const void *, int> mediumId;
    {
        int k = 0;
        for (auto &kv : media) mediumId[kv.second.ptr()] = k++;
    }
    auto mediumIndex = [&](Medium m) -> int {
        if (!m) return -1;
        auto it = mediumId.find(m.ptr());
        if (it == mediumId.end()) die("a medium that is not in the scene's medium table");
        return it->second;
    };

    // ---- geometry: walk the top-level BVH's ordered primitives (triangles only)
    if (!accel.Is<BVHAggregate>()) die("top-level aggregate is not a BVH");
    const BVHAggregate *bvh = accel.Cast<BVHAggregate>();
    // mesh ids: the triangle meshes first, the bilinear patch meshes after them
    const auto &allMeshes = *Triangle::allMeshes;
    const size_t nTriMeshes = allMeshes.size();
    const size_t nPatchMeshes = BilinearPatch::allMeshes ? BilinearPatch::allMeshes->size() : 0;
    auto patchMesh = [&](size_t m) { return (*BilinearPatch::allMeshes)[m - nTriMeshes]; };
    size_t nMesh = nTriMeshes + nPatchMeshes;
    std::vector<int32_t> meshAttr(nMesh * 3, -1);
    std::vector<const void *> meshMaterial(nMesh, nullptr);
    std::map<const void *, int> materialId;
    std::vector<pbrt::Material> materialList;
    std::map<const void *, int> lightIndex;
    for (size_t i = 0; i < lights.size(); ++i) lightIndex[lights[i].ptr()] = int(i);
    std::vector<uint32_t> bvhPrims;
    std::vector<int32_t> lightMeshTri(lights.size() * 2, -1);
    std::vector<int32_t> meshObject(nMesh, -1);
    // one triangle primitive of the world or of an object definition -> its mesh / material / light records
    std::map<int, std::pair<int, int>> meshMediaOf;  // reference mesh index -> {inside, outside}
    std::map<int, float> meshAlphaOf;                // reference mesh index -> constant alpha
    auto visitTriangle = [&](const Primitive &p, int object, std::vector<uint32_t> &primsOut) {
        Shape shape;
        pbrt::Material mtl;
        Light area;
        if (const SimplePrimitive *sp = p.CastOrNullptr<SimplePrimitive>()) {
            shape = sp->shape;
            mtl = sp->material;
        } else if (const GeometricPrimitive *gp = p.CastOrNullptr<GeometricPrimitive>()) {
            shape = gp->shape;
            mtl = gp->material;
            area = gp->areaLight;
            if (gp->alpha) {  // constant alpha only: one value per mesh (the cut-out test runs inside traversal)
                const FloatConstantTexture *ct = gp->alpha.CastOrNullptr<FloatConstantTexture>();
                if (!ct || !shape.Is<Triangle>()) die("only constant alpha textures on triangle meshes are dumped");
                int am = shape.Cast<Triangle>()->meshIndex;
                if (meshAlphaOf.count(am) && meshAlphaOf[am] != ct->Evaluate(TextureEvalContext())) die("two alpha values on one mesh");
                meshAlphaOf[am] = ct->Evaluate(TextureEvalContext());
            }
            if (volpath && gp->mediumInterface.IsMediumTransition()) {
                int mm = shape.Is<Triangle>() ? shape.Cast<Triangle>()->meshIndex
                                              : int(nTriMeshes) + shape.Cast<BilinearPatch>()->meshIndex;
                meshMediaOf[mm] = {mediumIndex(gp->mediumInterface.inside), mediumIndex(gp->mediumInterface.outside)};
            }
        } else
            die("only triangle and bilinear patch primitives are dumped so far");
        if (const BilinearPatch *bp = shape.CastOrNullptr<BilinearPatch>()) {
            int m = int(nTriMeshes) + bp->meshIndex;
            primsOut.insert(primsOut.end(), {2u, uint32_t(m), uint32_t(bp->blpIndex)});
            if (mtl && !materialId.count(mtl.ptr())) {
                materialId[mtl.ptr()] = int(materialList.size());
                materialList.push_back(mtl);
            }
            if (meshObject[m] != -1 && meshObject[m] != object) die("mesh used by two object definitions");
            meshObject[m] = object;
            meshAttr[3 * m + 0] = mtl ? materialId[mtl.ptr()] : -1;
            meshAttr[3 * m + 2] = (patchMesh(m)->reverseOrientation ^ patchMesh(m)->transformSwapsHandedness) ? 1 : 0;
            if (area) {
                int li = lightIndex.at(area.ptr());
                lightMeshTri[2 * li] = m;
                lightMeshTri[2 * li + 1] = bp->blpIndex;
                int first = li - bp->blpIndex;
                if (meshAttr[3 * m + 1] != -1 && meshAttr[3 * m + 1] != first) die("area lights of a mesh are not consecutive");
                meshAttr[3 * m + 1] = first;
            }
            return;
        }
        if (!shape.Is<Triangle>()) die("only triangles and bilinear patches are dumped so far");
        const Triangle *t = shape.Cast<Triangle>();
        int m = t->meshIndex;
        primsOut.insert(primsOut.end(), {0u, uint32_t(m), uint32_t(t->triIndex)});
        if (mtl && !materialId.count(mtl.ptr())) {
            materialId[mtl.ptr()] = int(materialList.size());
            materialList.push_back(mtl);
        }
        if (meshObject[m] != -1 && meshObject[m] != object) die("mesh used by two object definitions");
        meshObject[m] = object;
        meshAttr[3 * m + 0] = mtl ? materialId[mtl.ptr()] : -1;
        meshAttr[3 * m + 2] = (allMeshes[m]->reverseOrientation ^ allMeshes[m]->transformSwapsHandedness) ? 1 : 0;
        if (area) {
            int li = lightIndex.at(area.ptr());
            lightMeshTri[2 * li] = m;
            lightMeshTri[2 * li + 1] = t->triIndex;
            int first = li - t->triIndex;
            if (meshAttr[3 * m + 1] != -1 && meshAttr[3 * m + 1] != first) die("area lights of a mesh are not consecutive");
            meshAttr[3 * m + 1] = first;
        }
    };
    // object definitions met through TransformedPrimitives (scene.cpp:1523-1575)
    std::map<const void *, int> objectId;
    std::vector<std::vector<uint32_t>> objNodes, objPrims;
    std::vector<int32_t> instObject;
    std::vector<float> instM, instMInv;
    std::map<int, std::pair<int, const Transform *>> instById;
    for (const Primitive &p : bvh->primitives) {
        const TransformedPrimitive *tp = p.CastOrNullptr<TransformedPrimitive>();
        if (!tp) {
            visitTriangle(p, -2, bvhPrims);  // -2 = world (kept apart from "unseen" = -1 until the remap below)
            continue;
        }
        const void *key = tp->primitive.ptr();
        if (!objectId.count(key)) {
            int id = int(objectId.size());
            objectId[key] = id;
            objNodes.emplace_back();
            objPrims.emplace_back();
            if (const BVHAggregate *ob = tp->primitive.CastOrNullptr<BVHAggregate>()) {
                for (const Primitive &q : ob->primitives) visitTriangle(q, id, objPrims[id]);
                int nn = countNodes(ob->nodes);
                objNodes[id].resize(size_t(nn) * 8);
                memcpy(objNodes[id].data(), ob->nodes, size_t(nn) * 32);
            } else {
                // a one-primitive object has no BVH (scene.cpp:1551): hand over a one-leaf tree of its bounds
                visitTriangle(tp->primitive, id, objPrims[id]);
                Bounds3f bb = tp->primitive.Bounds();
                LinearBVHNode leaf;
                memset(&leaf, 0, sizeof leaf);
                leaf.bounds = bb;
                leaf.primitivesOffset = 0;
                leaf.nPrimitives = 1;
                objNodes[id].resize(8);
                memcpy(objNodes[id].data(), &leaf, 32);
            }
        }
        bvhPrims.insert(bvhPrims.end(), {1u, tp->instanceId, 0u});
        instById[int(tp->instanceId)] = {objectId[key], tp->renderFromPrimitive};
    }
    for (size_t k = 0; k < instById.size(); ++k) {
        auto it = instById.find(int(k) + 1);
        if (it == instById.end()) die("instance ids are not 1..n (empty instances are not handled)");
        instObject.push_back(it->second.first);
        std::vector<float> a = Mat(it->second.second->GetMatrix()), b = Mat(it->second.second->GetInverseMatrix());
        instM.insert(instM.end(), a.begin(), a.end());
        instMInv.insert(instMInv.end(), b.begin(), b.end());
    }
    // Meshes that no primitive of the aggregate refers to (the private shape copies the area
    // lights keep) are left out; the rest are renumbered densely.
    std::vector<int> meshRemap(nMesh, -1);
    {
        std::vector<char> used(nMesh, 0);
        for (size_t m = 0; m < nMesh; ++m) used[m] = meshObject[m] != -1;
        int next = 0;
        for (size_t m = 0; m < nMesh; ++m)
            if (used[m]) meshRemap[m] = next++;
        std::vector<int32_t> attr2(size_t(next) * 3);
        for (size_t m = 0; m < nMesh; ++m)
            if (used[m]) memcpy(&attr2[3 * meshRemap[m]], &meshAttr[3 * m], 12);
        meshAttr.swap(attr2);
        for (size_t i = 0; i < bvhPrims.size(); i += 3)
            if (bvhPrims[i] != 1) bvhPrims[i + 1] = meshRemap[bvhPrims[i + 1]];
        for (auto &op : objPrims)
            for (size_t i = 0; i < op.size(); i += 3) op[i + 1] = meshRemap[op[i + 1]];
        std::vector<int32_t> mo2(next);
        for (size_t m = 0; m < nMesh; ++m)
            if (used[m]) mo2[meshRemap[m]] = meshObject[m] == -2 ? -1 : meshObject[m];
        meshObject.swap(mo2);
        for (size_t i = 0; i < lightMeshTri.size(); i += 2)
            if (lightMeshTri[i] >= 0) lightMeshTri[i] = meshRemap[lightMeshTri[i]];
    }
    size_t nMeshAll = nMesh;
    nMesh = meshAttr.size() / 3;
    std::vector<uint32_t> nv, nt;
    std::vector<float> P;
    std::vector<int32_t> idx, patchIdx;
    std::vector<uint32_t> meshKind, meshHasN, meshHasUV, meshHasS;
    std::vector<float> cornerN, cornerUV, cornerS;
    for (size_t m = 0; m < nMeshAll; ++m) {
        if (meshRemap[m] < 0) continue;
        if (m >= nTriMeshes) {
            const BilinearPatchMesh *mesh = patchMesh(m);
            if (mesh->imageDistribution) die("image-distributed patch emission is not dumped");
            nv.push_back(mesh->nVertices);
            nt.push_back(mesh->nPatches);
            meshKind.push_back(1);
            meshHasN.push_back(mesh->n ? 1 : 0);  // four corner normals per patch
            meshHasUV.push_back(mesh->uv ? 1 : 0);  // four corner (s, t) per patch
            meshHasS.push_back(0);
            for (int i = 0; mesh->n && i < 4 * mesh->nPatches; ++i) {
                const Normal3f &vn = mesh->n[mesh->vertexIndices[i]];
                cornerN.insert(cornerN.end(), {vn.x, vn.y, vn.z});
            }
            for (int i = 0; mesh->uv && i < 4 * mesh->nPatches; ++i) {
                const Point2f &st = mesh->uv[mesh->vertexIndices[i]];
                cornerUV.insert(cornerUV.end(), {st.x, st.y});
            }
            for (int i = 0; i < mesh->nVertices; ++i) P.insert(P.end(), {mesh->p[i].x, mesh->p[i].y, mesh->p[i].z});
            patchIdx.insert(patchIdx.end(), mesh->vertexIndices, mesh->vertexIndices + 4 * mesh->nPatches);
            continue;
        }
        meshKind.push_back(0);
        const TriangleMesh *mesh = allMeshes[m];
        // per-vertex normals / uv leave as per-corner arrays (3 per triangle), so the device needs no index buffer
        meshHasN.push_back(mesh->n ? 1 : 0);
        meshHasUV.push_back(mesh->uv ? 1 : 0);
        meshHasS.push_back(mesh->s ? 1 : 0);
        for (int i = 0; i < 3 * mesh->nTriangles; ++i) {
            int v = mesh->vertexIndices[i];
            if (mesh->n) cornerN.insert(cornerN.end(), {mesh->n[v].x, mesh->n[v].y, mesh->n[v].z});
            if (mesh->uv) cornerUV.insert(cornerUV.end(), {mesh->uv[v].x, mesh->uv[v].y});
            if (mesh->s) cornerS.insert(cornerS.end(), {mesh->s[v].x, mesh->s[v].y, mesh->s[v].z});
        }
        nv.push_back(mesh->nVertices);
        nt.push_back(mesh->nTriangles);
        for (int i = 0; i < mesh->nVertices; ++i) P.insert(P.end(), {mesh->p[i].x, mesh->p[i].y, mesh->p[i].z});
        idx.insert(idx.end(), mesh->vertexIndices, mesh->vertexIndices + 3 * mesh->nTriangles);
    }
    w.u32("mesh_nverts", {uint32_t(nMesh)}, nv);
    w.u32("mesh_ntris", {uint32_t(nMesh)}, nt);
    w.f32("P", {uint32_t(P.size() / 3), 3}, P);
    w.i32("indices", {uint32_t(idx.size() / 3), 3}, idx);
    if (!cornerS.empty()) {  // per-vertex tangents (TriangleMesh::s), per corner like the normals
        w.u32("mesh_has_s", {uint32_t(nMesh)}, meshHasS);
        w.f32("corner_s", {uint32_t(cornerS.size() / 3), 3}, cornerS);
    }
    if (!cornerN.empty() || !cornerUV.empty()) {
        w.u32("mesh_has_n", {uint32_t(nMesh)}, meshHasN);
        w.u32("mesh_has_uv", {uint32_t(nMesh)}, meshHasUV);
        w.f32("corner_n", {uint32_t(cornerN.size() / 3), 3}, cornerN);
        w.f32("corner_uv", {uint32_t(cornerUV.size() / 2), 2}, cornerUV);
    }
    if (!patchIdx.empty()) {
        // mesh_kind 1 = bilinear patch mesh: its mesh_ntris patches take 4 indices each (p00 p10 p01 p11) from patch_indices
        w.u32("mesh_kind", {uint32_t(nMesh)}, meshKind);
        w.i32("patch_indices", {uint32_t(patchIdx.size() / 4), 4}, patchIdx);
    }
    w.i32("mesh_attr", {uint32_t(nMesh), 3}, meshAttr);
    w.i32("mesh_object", {uint32_t(nMesh)}, meshObject);
    if (!meshAlphaOf.empty()) {  // -1 = no alpha texture
        std::vector<float> ma(nMesh, -1.f);
        for (auto &kv : meshAlphaOf)
            if (meshRemap[kv.first] >= 0) ma[meshRemap[kv.first]] = kv.second;
        w.f32("mesh_alpha", {uint32_t(nMesh)}, ma);
    }
    if (volpath) {
        std::vector<int32_t> mm(nMesh * 2, -1);
        for (auto &kv : meshMediaOf)
            if (meshRemap[kv.first] >= 0) {
                mm[2 * meshRemap[kv.first]] = kv.second.first;
                mm[2 * meshRemap[kv.first] + 1] = kv.second.second;
            }
        w.i32("mesh_media", {uint32_t(nMesh), 2}, mm);
        Medium camMedium = camera.Is<PerspectiveCamera>() ? camera.Cast<PerspectiveCamera>()->medium
                           : camera.Is<OmniCamera>()      ? camera.Cast<OmniCamera>()->medium
                                                          : camera.Cast<RealisticCamera>()->medium;
        w.i32("camera_medium", {1}, {mediumIndex(camMedium)});
        // media_i: {type, sigma_a, sigma_s, nx, ny, nz, mres[3]}; media_f: {g, bounds[6], mediumFromRender[16],
        // renderFromMedium[16]} (include/b200pt.h b200pt_medium)
        std::vector<int32_t> mi, mle;
        std::vector<float> mf;
        int k = 0;
        for (auto &kv : media) {
            Medium m = kv.second;
            // media_le: {Le_spec or -1, LeScale grid nx ny nz} per medium; medium_lescale_<k>: the grid of an emissive grid medium
            int32_t le[4] = {-1, 0, 0, 0};
            float rec[39] = {0};
            if (const HomogeneousMedium *hm = m.CastOrNullptr<HomogeneousMedium>()) {
                mi.insert(mi.end(), {0, pool.add(&hm->sigma_a_spec, 0), pool.add(&hm->sigma_s_spec, 0), 0, 0, 0, 0, 0, 0});
                rec[0] = hm->phase.g;
                if (m.IsEmissive()) le[0] = pool.add(&hm->Le_spec, 0);  // (already times Lescale / its photometric norm)
            } else if (const GridMedium *gm = m.CastOrNullptr<GridMedium>()) {
                const SampledGrid<Float> &dg = gm->densityGrid;
                const MajorantGrid &mg = gm->majorantGrid;
                mi.insert(mi.end(), {1, pool.add(&gm->sigma_a_spec, 0), pool.add(&gm->sigma_s_spec, 0), dg.nx, dg.ny, dg.nz,
                                     mg.res.x, mg.res.y, mg.res.z});
                rec[0] = gm->phase.g;
                const Bounds3f &b = gm->bounds;
                float bb[6] = {b.pMin.x, b.pMin.y, b.pMin.z, b.pMax.x, b.pMax.y, b.pMax.z};
                memcpy(rec + 1, bb, 24);
                std::vector<float> a = Mat(gm->renderFromMedium.GetInverseMatrix()), c = Mat(gm->renderFromMedium.GetMatrix());
                memcpy(rec + 7, a.data(), 64);
                memcpy(rec + 23, c.data(), 64);
                std::vector<float> dens(dg.values.begin(), dg.values.end()), maj(mg.voxels.begin(), mg.voxels.end());
                w.f32("medium_density_" + std::to_string(k), {uint32_t(dg.nz), uint32_t(dg.ny), uint32_t(dg.nx)}, dens);
                w.f32("medium_majorants_" + std::to_string(k), {uint32_t(mg.res.z), uint32_t(mg.res.y), uint32_t(mg.res.x)}, maj);
                if (m.IsEmissive()) {
                    if (gm->temperatureGrid) {  // blackbody emission: the grid and its scale / offset
                        const SampledGrid<Float> &tg = *gm->temperatureGrid;
                        std::vector<float> tv(tg.values.begin(), tg.values.end());
                        w.f32("medium_temperature_" + std::to_string(k), {uint32_t(tg.nz), uint32_t(tg.ny), uint32_t(tg.nx)}, tv);
                        w.f32("medium_temperature_params_" + std::to_string(k), {2}, {gm->temperatureScale, gm->temperatureOffset});
                    }
                    const SampledGrid<Float> &lg = gm->LeScale;
                    le[0] = pool.add(&gm->Le_spec, 0);
                    le[1] = lg.nx;
                    le[2] = lg.ny;
                    le[3] = lg.nz;
                    std::vector<float> lv(lg.values.begin(), lg.values.end());
                    w.f32("medium_lescale_" + std::to_string(k), {uint32_t(lg.nz), uint32_t(lg.ny), uint32_t(lg.nx)}, lv);
                }
            } else
                die("only homogeneous and uniformgrid media are dumped so far");
            mf.insert(mf.end(), rec, rec + 39);
            mle.insert(mle.end(), le, le + 4);
            ++k;
        }
        w.i32("media_i", {uint32_t(k), 9}, mi);
        w.f32("media_f", {uint32_t(k), 39}, mf);
        w.i32("media_le", {uint32_t(k), 4}, mle);
    }
    if (!instObject.empty()) {
        w.i32("instance_object", {uint32_t(instObject.size())}, instObject);
        w.f32("instance_m", {uint32_t(instObject.size()), 16}, instM);
        w.f32("instance_minv", {uint32_t(instObject.size()), 16}, instMInv);
        for (size_t k = 0; dumpBvh && k < objNodes.size(); ++k) {
            w.u32("obj_bvh_nodes_" + std::to_string(k), {uint32_t(objNodes[k].size() / 8), 8}, objNodes[k]);
            w.u32("obj_bvh_prims_" + std::to_string(k), {uint32_t(objPrims[k].size() / 3), 3}, objPrims[k]);
        }
    }
    int nNodes = countNodes(bvh->nodes);
    std::vector<uint32_t> nodeWords(size_t(nNodes) * 8);
    memcpy(nodeWords.data(), bvh->nodes, size_t(nNodes) * 32);
    if (dumpBvh) {
        w.u32("bvh_nodes", {uint32_t(nNodes), 8}, nodeWords);
        w.u32("bvh_prims", {uint32_t(bvhPrims.size() / 3), 3}, bvhPrims);
    }

    // ---- materials
    std::vector<int32_t> matI;
    std::vector<float> matF;
    auto constSpec = [&](SpectrumTexture t) -> int {
        if (!t) return -1;
        const SpectrumConstantTexture *ct = t.CastOrNullptr<SpectrumConstantTexture>();
        if (!ct) die("only constant spectrum textures are dumped so far");
        return pool.add(ct->value, 1);
    };
    auto constFloat = [&](FloatTexture t) -> float {
        const FloatConstantTexture *ct = t.CastOrNullptr<FloatConstantTexture>();
        if (!ct) die("only constant float textures are dumped so far");
        return ct->value;
    };
    // materials_i: {type, spectrum[4]}; materials_f: params[12] (layout: include/b200pt.h b200pt_material)
    for (pbrt::Material mtl : materialList) {
        int32_t rec[5] = {-1, -1, -1, -1, -1};
        float prm[12] = {0};
        if (!mtl) {
            rec[0] = -1;
        } else if (const DiffuseMaterial *dm = mtl.CastOrNullptr<DiffuseMaterial>()) {
            if (dm->displacement || dm->normalMap) die("displacement / normal maps are not dumped");
            rec[0] = 0;
            rec[1] = constSpec(dm->reflectance);
        } else if (const DiffuseTransmissionMaterial *dt = mtl.CastOrNullptr<DiffuseTransmissionMaterial>()) {
            if (dt->displacement || dt->normalMap) die("displacement / normal maps are not dumped");
            rec[0] = 0;  // the diffuse queue, params[4] = transmission
            rec[1] = constSpec(dt->reflectance);
            rec[2] = constSpec(dt->transmittance);
            prm[4] = 1.f;
            prm[5] = dt->scale;
        } else if (const ConductorMaterial *cm = mtl.CastOrNullptr<ConductorMaterial>()) {
            if (cm->displacement || cm->normalMap) die("displacement / normal maps are not dumped");
            rec[0] = 1;
            rec[1] = constSpec(cm->eta);
            rec[2] = constSpec(cm->k);
            rec[3] = constSpec(cm->reflectance);
            prm[0] = constFloat(cm->uRoughness);
            prm[1] = constFloat(cm->vRoughness);
            prm[2] = cm->remapRoughness ? 1.f : 0.f;
        } else if (const DielectricMaterial *em = mtl.CastOrNullptr<DielectricMaterial>()) {
            if (em->displacement || em->normalMap) die("displacement / normal maps are not dumped");
            rec[0] = 2;
            rec[1] = pool.add(em->eta, 1);
            prm[0] = constFloat(em->uRoughness);
            prm[1] = constFloat(em->vRoughness);
            prm[2] = em->remapRoughness ? 1.f : 0.f;
            prm[3] = em->eta.Is<ConstantSpectrum>() ? 1.f : 0.f;
        } else if (const ThinDielectricMaterial *tm = mtl.CastOrNullptr<ThinDielectricMaterial>()) {
            if (tm->displacement || tm->normalMap) die("displacement / normal maps are not dumped");
            rec[0] = 2;  // the dielectric queue, params[4] = thin
            rec[1] = pool.add(tm->eta, 1);
            prm[3] = tm->eta.Is<ConstantSpectrum>() ? 1.f : 0.f;
            prm[4] = 1.f;
        } else if (const CoatedDiffuseMaterial *cd = mtl.CastOrNullptr<CoatedDiffuseMaterial>()) {
            if (cd->displacement || cd->normalMap) die("displacement / normal maps are not dumped");
            rec[0] = 3;
            rec[1] = constSpec(cd->reflectance);
            rec[2] = constSpec(cd->albedo);
            rec[3] = pool.add(cd->eta, 1);
            prm[0] = constFloat(cd->uRoughness);
            prm[1] = constFloat(cd->vRoughness);
            prm[2] = cd->remapRoughness ? 1.f : 0.f;
            prm[3] = cd->eta.Is<ConstantSpectrum>() ? 1.f : 0.f;
            prm[4] = constFloat(cd->thickness);
            prm[5] = constFloat(cd->g);
            prm[6] = float(cd->maxDepth);
            prm[7] = float(cd->nSamples);
        } else if (const CoatedConductorMaterial *cc = mtl.CastOrNullptr<CoatedConductorMaterial>()) {
            if (cc->displacement || cc->normalMap) die("displacement / normal maps are not dumped");
            const SpectrumConstantTexture *alb = cc->albedo.CastOrNullptr<SpectrumConstantTexture>();
            if (!alb) die("only constant spectrum textures are dumped so far");
            for (int l = 395; l <= 705; ++l)
                if (alb->value(Float(l)) != 0) die("coatedconductor with albedo != 0 (a medium between the layers) is not dumped");
            rec[0] = 3;  // the coated queue, params[8] = conductor base
            rec[1] = constSpec(cc->conductorEta);
            rec[2] = constSpec(cc->k);
            rec[3] = pool.add(cc->interfaceEta, 1);
            rec[4] = constSpec(cc->reflectance);
            prm[0] = constFloat(cc->interfaceURoughness);
            prm[1] = constFloat(cc->interfaceVRoughness);
            prm[2] = cc->remapRoughness ? 1.f : 0.f;
            prm[3] = cc->interfaceEta.Is<ConstantSpectrum>() ? 1.f : 0.f;
            prm[4] = constFloat(cc->thickness);
            prm[5] = constFloat(cc->g);
            prm[6] = float(cc->maxDepth);
            prm[7] = float(cc->nSamples);
            prm[8] = 1.f;
            prm[9] = constFloat(cc->conductorURoughness);
            prm[10] = constFloat(cc->conductorVRoughness);
        } else
            die("only diffuse, diffusetransmission, coatedconductor, conductor, dielectric, thindielectric and coateddiffuse materials are dumped so far");
        matI.insert(matI.end(), rec, rec + 5);
        matF.insert(matF.end(), prm, prm + 12);
    }
    {  // Material::materialId (materials.cpp:738), GBufferFilm's MaterialId channel
        std::vector<int32_t> ids;
        for (pbrt::Material mtl : materialList) ids.push_back(mtl ? int32_t(mtl.materialId) : 0);
        w.i32("material_ids", {uint32_t(ids.size())}, ids);
    }
    w.i32("materials_i", {uint32_t(materialList.size()), 5}, matI);
    w.f32("materials_f", {uint32_t(materialList.size()), 12}, matF);

    // ---- lights + the power sampler's alias table (lightsamplers.cpp:76-96, util/sampling.cpp:562-617)
    PowerLightSampler pls(pstd::span<const Light>(lights.data(), lights.size()), Allocator());
    std::vector<int32_t> lightI;
    std::vector<float> lightF, lightGeom;
    std::map<const void *, int> lemitId;
    for (size_t i = 0; i < lights.size(); ++i) {
        const auto &bin = pls.aliasTable.bins[i];
        if (const UniformInfiniteLight *ul = lights[i].CastOrNullptr<UniformInfiniteLight>()) {
            if (!lemitId.count(ul->Lemit)) lemitId[ul->Lemit] = pool.add(ul->Lemit, 0);
            lightI.insert(lightI.end(), {-1, -1, lemitId[ul->Lemit], 0, bin.alias, -1, 1});
            lightF.insert(lightF.end(), {ul->scale, 0.f, 0.f, 0.f, bin.q, bin.p});
            continue;
        }
        // the delta lights: placement in lights_geom (include/b200pt.h b200pt_render_desc::light_geometry)
        auto deltaLight = [&](int type, const DenselySampledSpectrum *spec, float scale, const float (&g)[24]) {
            if (!lemitId.count(spec)) lemitId[spec] = pool.add(spec, 0);
            lightI.insert(lightI.end(), {-1, -1, lemitId[spec], 0, bin.alias, -1, type});
            lightF.insert(lightF.end(), {scale, 0.f, 0.f, 0.f, bin.q, bin.p});
            if (lightGeom.empty()) lightGeom.assign(24 * lights.size(), 0.f);
            std::copy(g, g + 24, lightGeom.begin() + 24 * i);
        };
        if (const PointLight *pl = lights[i].CastOrNullptr<PointLight>()) {
            Point3f p = pl->renderFromLight(Point3f(0, 0, 0));
            float g[24] = {p.x, p.y, p.z};
            deltaLight(2, pl->I, pl->scale, g);
            continue;
        }
        if (const SpotLight *sl = lights[i].CastOrNullptr<SpotLight>()) {
            Point3f p = sl->renderFromLight(Point3f(0, 0, 0));
            const SquareMatrix<4> &mi = sl->renderFromLight.GetInverseMatrix();
            float g[24] = {p.x, p.y, p.z, 0.f, mi[0][0], mi[0][1], mi[0][2], mi[1][0], mi[1][1], mi[1][2],
                           mi[2][0], mi[2][1], mi[2][2], sl->cosFalloffStart, sl->cosFalloffEnd, 0.f};
            deltaLight(3, sl->Iemit, sl->scale, g);
            continue;
        }
        if (const DistantLight *dl = lights[i].CastOrNullptr<DistantLight>()) {
            Vector3f wi = Normalize(dl->renderFromLight(Vector3f(0, 0, 1)));
            float g[24] = {wi.x, wi.y, wi.z, dl->sceneRadius};
            deltaLight(4, dl->Lemit, dl->scale, g);
            continue;
        }
        if (const ImageInfiniteLight *il = lights[i].CastOrNullptr<ImageInfiniteLight>()) {
            // image and sampling distribution: include/b200pt.h b200pt_render_desc::env_image / env_distribution
            const SquareMatrix<4> &m = il->renderFromLight.GetMatrix(), &mi = il->renderFromLight.GetInverseMatrix();
            float g[24] = {m[0][0], m[0][1], m[0][2], m[1][0], m[1][1], m[1][2], m[2][0], m[2][1], m[2][2],
                           mi[0][0], mi[0][1], mi[0][2], mi[1][0], mi[1][1], mi[1][2], mi[2][0], mi[2][1], mi[2][2],
                           il->sceneRadius};
            // `spectrum` = the colour space's illuminant, as RGBIlluminantSpectrum samples it (DenselySampledSpectrum)
            const DenselySampledSpectrum *illum = &il->imageColorSpace->illuminant;
            deltaLight(5, illum, il->scale, g);
            const int res = il->image.Resolution().x;
            std::vector<float> px(size_t(res) * res * 4), dist(size_t(res) * res);
            for (int y = 0; y < res; ++y)
                for (int x = 0; x < res; ++x) {
                    RGB rgb;
                    for (int c = 0; c < 3; ++c) rgb[c] = il->image.GetChannel({x, y}, c);
                    RGBIlluminantSpectrum spec(*il->imageColorSpace, ClampZero(rgb));
                    float *o = &px[(size_t(y) * res + x) * 4];
                    o[0] = spec.rsp.c0;
                    o[1] = spec.rsp.c1;
                    o[2] = spec.rsp.c2;
                    o[3] = spec.scale;
                    dist[size_t(y) * res + x] = il->compensatedDistribution.pConditionalV[y].func[x];
                }
            w.f32("env_image", {uint32_t(res), uint32_t(res), 4}, px);
            w.f32("env_distribution", {uint32_t(res), uint32_t(res)}, dist);
            continue;
        }
        const DiffuseAreaLight *al = lights[i].CastOrNullptr<DiffuseAreaLight>();
        if (!al) die("only diffuse area, uniform infinite, point, spot and distant lights are dumped so far");
        if (al->image) die("image area lights are not dumped");
        if (!lemitId.count(al->Lemit)) lemitId[al->Lemit] = pool.add(al->Lemit, 0);
        int m = lightMeshTri[2 * i], tri = lightMeshTri[2 * i + 1];
        if (m < 0) die("area light without a triangle in the aggregate");
        lightI.insert(lightI.end(), {m, tri, lemitId[al->Lemit], al->twoSided ? 1 : 0, bin.alias, meshAttr[3 * m + 2], 0});
        lightF.insert(lightF.end(), {al->scale, al->cosFalloffEnd, al->tanFalloffEnd, al->normalize_falloffEnd, bin.q, bin.p});
    }
    w.i32("lights_i", {uint32_t(lights.size()), 7}, lightI);
    w.f32("lights_f", {uint32_t(lights.size()), 6}, lightF);
    if (!lightGeom.empty()) w.f32("lights_geom", {uint32_t(lights.size()), 24}, lightGeom);

    // ---- the light BVH (lightsamplers.cpp:109-232), node fields as include/b200pt.h lists them
    if (lsId == 2) {
        BVHLightSampler bls(pstd::span<const Light>(lights.data(), lights.size()), Allocator());
        std::vector<uint32_t> nodes;
        for (const LightBVHNode &nd : bls.nodes) {
            const CompactLightBounds &cb = nd.lightBounds;
            uint32_t phiBits;
            memcpy(&phiBits, &cb.phi, 4);
            nodes.insert(nodes.end(),
                         {uint32_t(cb.w.x) | uint32_t(cb.w.y) << 16, phiBits,
                          uint32_t(cb.qCosTheta_o) | uint32_t(cb.qCosTheta_e) << 15 | uint32_t(cb.twoSided) << 30,
                          uint32_t(cb.qb[0][0]) | uint32_t(cb.qb[0][1]) << 16,
                          uint32_t(cb.qb[0][2]) | uint32_t(cb.qb[1][0]) << 16,
                          uint32_t(cb.qb[1][1]) | uint32_t(cb.qb[1][2]) << 16,
                          uint32_t(nd.childOrLightIndex) | uint32_t(nd.isLeaf) << 31, 0u});
        }
        std::vector<uint32_t> trails(lights.size(), 0xFFFFFFFFu);
        for (size_t i = 0; i < lights.size(); ++i)
            if (bls.lightToBitTrail.HasKey(lights[i])) trails[i] = bls.lightToBitTrail[lights[i]];
        const Bounds3f &ab = bls.allLightBounds;
        w.u32("lightbvh_nodes", {uint32_t(bls.nodes.size()), 8}, nodes);
        w.u32("lightbvh_trails", {uint32_t(lights.size())}, trails);
        w.f32("lightbvh_bounds", {6}, {ab.pMin.x, ab.pMin.y, ab.pMin.z, ab.pMax.x, ab.pMax.y, ab.pMax.z});
        // known answers of BVHLightSampler::Sample / PMF at random reference points (p, ns, u); ns = 0 for a
        // third of them (points in media).  Checked bit for bit by tests/test_render_core.py.
        const int nProbe = 768;
        RNG rng(7, 11);
        Bounds3f sb = accel.Bounds();
        std::vector<float> pin, pout;
        std::vector<int32_t> pidx;
        for (int i = 0; i < nProbe; ++i) {
            Point3f p = sb.Lerp(Point3f(rng.Uniform<Float>() * 1.2f - .1f, rng.Uniform<Float>() * 1.2f - .1f,
                                        rng.Uniform<Float>() * 1.2f - .1f));
            Normal3f n(0, 0, 0);
            if (i % 3) n = Normal3f(SampleUniformSphere(Point2f(rng.Uniform<Float>(), rng.Uniform<Float>())));
            Float u = rng.Uniform<Float>();
            LightSampleContext ctx(Point3fi(p), n, n);
            pstd::optional<SampledLight> sl = bls.Sample(ctx, u);
            int li = sl ? lightIndex[sl->light.ptr()] : -1;
            // PMF of the sampled light, or of light (i mod nLights) when nothing was sampled
            int lq = li >= 0 ? li : int(i % lights.size());
            pin.insert(pin.end(), {p.x, p.y, p.z, n.x, n.y, n.z, u});
            pidx.insert(pidx.end(), {li, lq});
            pout.insert(pout.end(), {sl ? sl->p : 0.f, bls.PMF(ctx, lights[lq])});
        }
        w.f32("lightbvh_probe_in", {uint32_t(nProbe), 7}, pin);
        w.i32("lightbvh_probe_light", {uint32_t(nProbe), 2}, pidx);
        w.f32("lightbvh_probe_pmf", {uint32_t(nProbe), 2}, pout);
    }

    w.f32("spectra_values", {uint32_t(pool.interp.size()), 311}, pool.values);
    w.u32("spectra_interp", {uint32_t(pool.interp.size())}, pool.interp);
    if (nMeshOut) *nMeshOut = nMesh;
    if (nTriOut) *nTriOut = idx.size() / 3;
    if (nMaterialOut) *nMaterialOut = materialList.size();
    if (nSpectraOut) *nSpectraOut = pool.interp.size();
}
