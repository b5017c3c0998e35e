// render_path.h - one path in registers and the steps the kernels apply to it (included by
// render_shade.h).  GenerateCameraPath = RayIntegrator::EvaluatePixelSample up to the camera ray
// (cpu/integrators.cpp:228-262); ShadeVertex = one iteration of PathIntegrator::Li's loop
// (cpu/integrators.cpp:640-760) with SampleLd (:764-805) split at the shadow ray: the contribution
// goes to a shadow-queue slot and is added to L when the any-hit kernel finds the ray unoccluded
// (the reference's wavefront does the same: wavefront/intersect.h RecordShadowRayResult);
// FinishPath = the film's AddSample.
//
// Spectral work is written as streams over k = 0..30 with the wavelength carried along a
// recurrence (LambdaIter), never as arrays of 31 that must stay live across the whole step: only
// beta, L and the pending Ld are reached through the IO object (beta in shared memory and L / Ld in
// global memory on the device, plain arrays in tests/native/rendercheck.cpp).  The loops are kept
// rolled (B2_SPECTRAL_LOOP): fully unrolled, the 31 copies of every loop body made the shading
// kernel ~20 K instructions and instruction-fetch bound (profiles/r01_render_v2_*).
#pragma once
#include "render_bsdf.h"
#include "render_camera.h"
#include "render_layered.h"
#include "render_lightbvh.h"

namespace b200pt {

enum : uint32_t {
    PF_ALIVE = 1u, PF_SPECULAR = 2u, PF_ANY_NONSPECULAR = 4u,
    PF_TERMINATED_SECONDARY = 16u  // SampledWavelengths::TerminateSecondary happened (util/spectrum.h:346-353)
};

// the 31-trip spectral loops stay (mostly) rolled on the device: unrolled by B2_SPECTRAL_UNROLL (2: the shade kernels
// are register bound; the volume-walk kernels, whose spectral state lives in local memory, take 8 so that the loads of
// several wavelengths are in flight together - set per translation unit by the build)
#ifndef B2_SPECTRAL_UNROLL
#define B2_SPECTRAL_UNROLL 2
#endif
#define B2_PRAGMA_(x) _Pragma(#x)
#define B2_PRAGMA(x) B2_PRAGMA_(x)
#if defined(__CUDA_ARCH__)
#define B2_SPECTRAL_LOOP B2_PRAGMA(unroll B2_SPECTRAL_UNROLL)
#else
#define B2_SPECTRAL_LOOP
#endif

struct PathRegs {
    float lambdaU;
    int px, py, sampleIndex, dimension, depth;
    uint32_t flags;
    float etaScale, pB, filterWeight;
    int medium;          // Ray::medium of p.ray, index into the scene's media, -1 = none (volpath only)
    float cameraWeight;  // CameraRay::weight: L = weight * Li (cpu/integrators.cpp:247-252); 0 = no ray
    LightCtx prev;
    V3 rayO, rayD;
    float rayTMax;
};

template <int EXT = 1>
B2_HD PathSampler SamplerFor(const RenderSceneDev &rs, const PathRegs &p) {
    PathSampler s;
    PathSamplerStart<EXT>(s, rs.sampler, p.px, p.py, p.sampleIndex, p.dimension);
    return s;
}

// RayIntegrator::EvaluatePixelSample (cpu/integrators.cpp:228-262) + GetCameraSample (samplers.h:797-815)
B2_HD void GenerateCameraPath(const RenderSceneDev &rs, PathRegs &p, int px, int py, int sampleIndex) {
    p.px = px;
    p.py = py;
    p.sampleIndex = sampleIndex;
    p.dimension = 0;
    PathSampler s = SamplerFor<1>(rs, p);
    p.lambdaU = SamplerGet1D(s);
    float f0, f1;
    SamplerGet2D(s, f0, f1);  // GetPixel2D -> filter.Sample
    float fx, fy;
    SampleFilter(rs.film, f0, f1, fx, fy, p.filterWeight);
    float pFilmX = px + fx + 0.5f, pFilmY = py + fy + 0.5f;
    (void)SamplerGet1D(s);  // time
    float l0, l1;
    SamplerGet2D(s, l0, l1);
    p.dimension = s.z.dimension;
    p.cameraWeight = 1.f;
    if (rs.camera.type >= 1)  // realistic, omni
        p.cameraWeight = GenerateRealisticRay(rs.camera, rs.film, pFilmX, pFilmY, l0, l1, p.rayO, p.rayD);
    else
        GeneratePerspectiveRay(rs.camera, pFilmX, pFilmY, l0, l1, p.rayO, p.rayD);
    p.rayTMax = INFINITY;
    p.depth = 0;
    p.flags = PF_ALIVE;
    if (p.cameraWeight == 0) {  // no camera ray: the sample is added to the film with L = 0
        p.flags = 0;
        p.rayO = p.rayD = mk(0, 0, 0);
        p.rayTMax = -1.f;
    }
    p.etaScale = 1.f;
    p.pB = 0.f;
    p.medium = rs.integrator.type == 1 ? rs.camera.medium : -1;
    p.prev.pi.lo = p.prev.pi.hi = p.prev.n = p.prev.ns = mk(0, 0, 0);
}

B2_HD void KillPath(PathRegs &p) {
    p.flags &= ~PF_ALIVE;
    p.rayTMax = -1.f;
}

// Material::GetBxDF for the surface (materials.h: Diffuse 430-476, Conductor 491-509, Dielectric 183-203)
// MAT >= 0: the caller knows the material type (a material-sorted queue); the constant lets the
// compiler drop the other models from that kernel.
// EXT = 0: a scene without the less common features (RenderSceneDev::geomExt) - here the material variants
// (ThinDielectric, DiffuseTransmission) - whose branches the plain kernels then do not carry.
template <int MAT, int EXT = 1>
B2_HD Bsdf MakeBsdf(const RenderSceneDev &rs, const MaterialDev &mat, float lambda0, PathRegs &p) {
    Bsdf b;
    b.type = MAT >= 0 ? uint32_t(MAT) : mat.type;
    b.tr = TRDist{0, 0};
    b.eta = 1;
    b.thin = 0;
    b.trans = 0;
    b.dscale = 1;
    b.tMax = 0;
    b.specA = mat.spectrum[0];
    b.specB = -1;
    b.thickness = 0;
    b.rMax = 0;
    b.maxDepth = b.nSamples = 0;
    b.layerSeed = 0;  // Options::seed (options.h), 0 unless the host was started with --seed
    b.coatCond = 0;
    b.condTr = TRDist{0, 0};
    if (b.type == MAT_CONDUCTOR || b.type == MAT_DIELECTRIC || b.type == MAT_COATED_DIFFUSE) {
        float ur = mat.params[0], vr = mat.params[1];
        if (mat.params[2] != 0) {  // remaproughness: RoughnessToAlpha = sqrt
            ur = sqrtf(ur);
            vr = sqrtf(vr);
        }
        b.tr = MakeTR(ur, vr);
    }
    if (b.type == MAT_CONDUCTOR) {
        if (mat.spectrum[0] >= 0) {
            b.specA = mat.spectrum[0];
            b.specB = mat.spectrum[1];
        } else {
            b.specA = mat.spectrum[2];  // reflectance; k derived per wavelength
            b.specB = -1;
        }
    }
    if (b.type == MAT_DIELECTRIC) {
        float e = SpectrumAt(rs.spectra[mat.spectrum[0]], lambda0);
        if (mat.params[3] == 0) p.flags |= PF_TERMINATED_SECONDARY;  // eta is not a ConstantSpectrum
        if (e == 0) e = 1;
        b.eta = e;
        if (EXT && mat.params[4] != 0) {  // ThinDielectricMaterial (materials.h:215-262): no roughness
            b.thin = 1;
            b.tr = TRDist{0, 0};
        }
    }
    if (EXT && b.type == MAT_DIFFUSE && mat.params[4] != 0) {  // DiffuseTransmissionMaterial::GetBxDF (materials.h:891-897)
        b.trans = 1;
        b.specB = mat.spectrum[1];
        b.dscale = mat.params[5];
        float rm = 0, tm = 0;
        LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k, li.Next()) {
            rm = fmaxf(rm, Clampf(b.dscale * SpectrumAt(rs.spectra[b.specA], li.lam), 0, 1));
            tm = fmaxf(tm, Clampf(b.dscale * SpectrumAt(rs.spectra[b.specB], li.lam), 0, 1));
        }
        b.rMax = rm;
        b.tMax = tm;
    }
    if (b.type == MAT_COATED_DIFFUSE) {  // CoatedDiffuseMaterial::GetBxDF (materials.cpp:301-329)
        float e = SpectrumAt(rs.spectra[mat.spectrum[2]], lambda0);
        if (mat.params[3] == 0) p.flags |= PF_TERMINATED_SECONDARY;
        if (e == 0) e = 1;
        b.eta = e;
        b.thickness = fmaxf(mat.params[4], 1.17549435e-38f);
        b.maxDepth = (int)mat.params[6];
        b.nSamples = (int)mat.params[7];
        if (EXT && mat.params[8] != 0) {  // CoatedConductorMaterial::GetBxDF (materials.cpp:392-438)
            b.coatCond = 1;
            float cu = mat.params[9], cv = mat.params[10];
            if (mat.params[2] != 0) {
                cu = sqrtf(cu);
                cv = sqrtf(cv);
            }
            b.condTr = MakeTR(cu, cv);
            if (mat.spectrum[0] >= 0) {
                b.specA = mat.spectrum[0];
                b.specB = mat.spectrum[1];
            } else {
                b.specA = mat.spectrum[3];  // reflectance; k derived per wavelength
                b.specB = -1;
            }
            return b;
        }
        float rm = 0;
        LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k, li.Next()) rm = fmaxf(rm, Clampf(SpectrumAt(rs.spectra[b.specA], li.lam), 0, 1));
        b.rMax = rm;
    }
    return b;
}

// the four BxDF entry points with the layered model added (it needs its own header)
// (rs, lambdaU: the coated conductor's walks need the base's spectra at the path's wavelengths)
// spec: kNLambda floats for its per-wavelength value (BsdfEval kind 8)
B2_HD BsdfEval EvalF(const RenderSceneDev &rs, float lambdaU, const Bsdf &b, V3 wo, V3 wi, float *spec) {
    if (b.type == MAT_COATED_DIFFUSE && b.coatCond) return CoatedCondF(rs, b, lambdaU, wo, wi, b.layerSeed, spec);
    return b.type == MAT_COATED_DIFFUSE ? CoatedF(b, wo, wi, b.layerSeed) : BsdfF(b, wo, wi);
}
B2_HD float EvalPDF(const Bsdf &b, V3 wo, V3 wi) {
    if (b.type == MAT_COATED_DIFFUSE && b.coatCond) return CoatedCondPDF(b, wo, wi, b.layerSeed);
    return b.type == MAT_COATED_DIFFUSE ? CoatedPDF(b, wo, wi, b.layerSeed) : BsdfPDF(b, wo, wi);
}
B2_HD BsdfSample EvalSample(const RenderSceneDev &rs, float lambdaU, const Bsdf &b, V3 wo, float uc, float u0, float u1,
                            BsdfEval &ev, float *spec) {
    if (b.type == MAT_COATED_DIFFUSE && b.coatCond) return CoatedCondSample(rs, b, lambdaU, wo, uc, u0, u1, b.layerSeed, ev, spec);
    return b.type == MAT_COATED_DIFFUSE ? CoatedSample(b, wo, uc, u0, u1, b.layerSeed, ev)
                                        : BsdfSampleF(b, wo, uc, u0, u1, ev);
}

// f_k: the per-wavelength factor applied to the scalar evaluation `e` (see BsdfEval)
// Rk: the vertex's cached Clamp(R(lambda_k), 0, 1) of a diffuse / coated diffuse surface (ShadeVertex fills IO::SetR
// once per vertex: the flags, the light sample and the BSDF sample all need the same 31 look-ups)
template <int EXT = 1>
B2_HD float BsdfSpectral(const RenderSceneDev &rs, const Bsdf &b, const BsdfEval &e, float lam, float Rk, int k, const float *spec) {
    if (EXT && e.kind == 8) return spec[k];  // coated conductor: the walk's value at wavelength k (EXT = 1 kernels only)
    switch (e.kind) {
    case 1: return e.scale;
    case 2: return Rk * e.scale;  // DiffuseMaterial clamps R
    case 6: return Clampf(b.dscale * SpectrumAt(rs.spectra[b.specA], lam), 0, 1) * e.scale;  // DiffuseTransmission: R
    case 7: return Clampf(b.dscale * SpectrumAt(rs.spectra[b.specB], lam), 0, 1) * e.scale;  //   T (materials.h:894-895)
    case 5: {  // coated diffuse: polynomial in R_k (render_layered.h)
        float R = Rk;
        float v = e.poly[kLayerPoly - 1];
        for (int a = kLayerPoly - 2; a >= 0; --a) v = v * R + e.poly[a];
        return v;
    }
    case 3:
    case 4: {
        float etaK, kK;
        if (b.specB >= 0) {
            etaK = SpectrumAt(rs.spectra[b.specA], lam);
            kK = SpectrumAt(rs.spectra[b.specB], lam);
        } else {  // materials.h:503-507
            float r = Clampf(SpectrumAt(rs.spectra[b.specA], lam), 0, .9999f);
            etaK = 1.f;
            kK = 2 * sqrtf(r) / sqrtf(fmaxf(0.f, 1 - r));
        }
        float F = FrComplex(e.cosF, etaK, kK);
        return e.kind == 3 ? (e.D * F) * e.G / e.denom : F / e.denom;
    }
    }
    return 0.f;
}

// What a light sample needs beyond its ray when media are present (SampleLd's state before the walk,
// cpu/integrators.cpp:1296-1332)
struct ShadowVolInfo {
    float p_l, scatterPDF;
    int medium;   // medium of the ray's first segment, -1 = none
    P3i lightPi;  // ls->pLight: the walk re-aims at it after every interface (SpawnRayTo)
    V3 lightN;
};

// Sum of r_u over the wavelengths / N (SampledSpectrum::Average), and of r_u + r_l
template <class IO>
B2_HD float AvgRu(const IO &io) {
    float sum = io.Ru(0);
    for (int k = 1; k < kNLambda; ++k) sum += io.Ru(k);
    return sum / kNLambda;
}
template <class IO>
B2_HD float AvgRuRl(const IO &io) {
    float sum = io.Ru(0) + io.Rl(0);
    for (int k = 1; k < kNLambda; ++k) sum += io.Ru(k) + io.Rl(k);
    return sum / kNLambda;
}

// What an escaped ray needs for the MIS weight of an infinite light that light sampling can also pick (an image
// infinite light; a uniform one has PDF_Li = 0 under allowIncompletePDF): the ray's direction and the previous
// vertex' light-sample context.
struct EscapeCtx {
    bool unweighted;  // the ray left the camera or a specular bounce
    V3 d, prevP, prevNs;
};
// LightSampler::PMF(ctx, light) for the scene's sampler (lightsamplers.h:48-53, 78-82, 320-358)
B2_HD float LightPMF(const RenderSceneDev &rs, bool lightBvh, V3 p, V3 ns, int light) {
    if (lightBvh) return LightBvhPMF(rs, p, ns, light);
    if (rs.integrator.lightSampler == 0) return 1.f / rs.nLights;
    return rs.lights[light].aliasP;
}

// VolPathIntegrator's escaped ray (cpu/integrators.cpp:1081-1099): beta * Le / r_u.Average() from the camera or a
// specular bounce; else r_l is first scaled by the light's p_l (0 for a UniformInfiniteLight under allowIncompletePDF)
// and the estimate is beta * Le / (r_u + r_l).Average().
template <class IO>
B2_HD void EscapedRadianceVol(const RenderSceneDev &rs, float lambdaU, IO &io, const EscapeCtx &ec, bool lightBvh) {
    for (uint32_t i = 0; i < rs.nInfiniteLights; ++i) {
        const int lidx = rs.infiniteLights[i];
        const LightDev &light = rs.lights[lidx];
        if (light.type == LIGHT_IMAGE_INFINITE) {
            const float *g = rs.lightGeom + size_t(kLightGeomFloats) * lidx;
            const int pixel = EnvPixelOfRay(rs, g, ec.d);
            float avg;
            if (ec.unweighted)
                avg = AvgRu(io);
            else {
                float p_l = LightPMF(rs, lightBvh, ec.prevP, ec.prevNs, lidx) * EnvPDF(rs, g, ec.d);
                for (int k = 0; k < kNLambda; ++k) io.SetRl(k, io.Rl(k) * p_l);
                avg = AvgRuRl(io);
            }
            LambdaIter li(lambdaU);
B2_SPECTRAL_LOOP
            for (int k = 0; k < kNLambda; ++k, li.Next()) {
                float Le = EnvLe(rs, light, pixel, li.lam);
                if (Le != 0) io.AddL(k, io.Beta(k) * Le / avg);
            }
            continue;
        }
        if (light.type != LIGHT_UNIFORM_INFINITE) continue;  // a DistantLight: in the sampler's list, LightType::DeltaDirection
        const float avg = AvgRu(io);
        const SpectrumTable &sp = rs.spectra[light.spectrum];
        LambdaIter li(lambdaU);
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k, li.Next()) {
            float Le = light.scale * SpectrumAt(sp, li.lam);
            if (Le != 0) io.AddL(k, io.Beta(k) * Le / avg);
        }
    }
}

// Escaped ray: every infinite light adds beta * Le, MIS-weighted against light sampling unless the ray
// left the camera or a specular bounce (cpu/integrators.cpp:645-662).  UniformInfiniteLight::PDF_Li is 0
// under allowIncompletePDF (lights.cpp:998-1003), so its weight is PowerHeuristic(p_b, 0).
template <class IO>
B2_HD void EscapedRadiance(const RenderSceneDev &rs, float lambdaU, float pB, IO &io, const EscapeCtx &ec, bool lightBvh) {
    for (uint32_t i = 0; i < rs.nInfiniteLights; ++i) {
        const int lidx = rs.infiniteLights[i];
        const LightDev &light = rs.lights[lidx];
        if (light.type != LIGHT_UNIFORM_INFINITE && light.type != LIGHT_IMAGE_INFINITE) continue;
        const bool env = light.type == LIGHT_IMAGE_INFINITE;
        const float *g = rs.lightGeom + size_t(kLightGeomFloats) * lidx;
        const int pixel = env ? EnvPixelOfRay(rs, g, ec.d) : 0;
        float p_l = 0;
        if (env && !ec.unweighted) p_l = LightPMF(rs, lightBvh, ec.prevP, ec.prevNs, lidx) * EnvPDF(rs, g, ec.d);
        const SpectrumTable &sp = rs.spectra[light.spectrum];
        float w = ec.unweighted ? 1.f : PowerHeuristic(pB, p_l);
        LambdaIter li(lambdaU);
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k, li.Next()) {
            float Le = env ? EnvLe(rs, light, pixel, li.lam) : light.scale * SpectrumAt(sp, li.lam);
            io.AddL(k, ec.unweighted ? io.Beta(k) * Le : io.Beta(k) * w * Le);
        }
    }
}

// One path vertex.  `hit` is the internal record {slot, t, b0, b1, b2} of the closest-hit kernel for
// p.ray.  IO: AddL(k, v) adds to the path's radiance; ShadowBegin() / SetLd(k, v) / ShadowEnd(valid,
// o, d, tMax) record a light sample whose visibility is still unknown; Beta(k) / SetBeta(k, v) are the
// path throughput.
// MAT / LBVH: material type and "the light sampler is the BVH one" fixed at compile time (the kernels);
// -1 = read from the scene (the host check).
// VOL: 1 = the surface half of VolPathIntegrator::Li (cpu/integrators.cpp:1100-1268): MIS through the rescaled path
// probabilities r_u / r_l (IO::Ru / Rl), interface surfaces are crossed, light samples carry what the
// transmittance walk needs (IO::ShadowEndVol), the ray's medium follows the surface's MediumInterface.
template <class IO, int MAT = -1, int LBVH = -1, int VOL = -1, int EXT = 1>
B2_HD void ShadeVertex(const RenderSceneDev &rs, PathRegs &p, const uint32_t hit[6], IO &io) {
    const bool lightBvh = LBVH < 0 ? rs.integrator.lightSampler == 2 : LBVH == 1;
    const bool vol = VOL < 0 ? rs.integrator.type == 1 : VOL == 1;
    if (hit[0] == kMissTag) {  // escaped ray (cpu/integrators.cpp:645-662); the kernels do this in ClassifyKernel
        if (rs.nInfiniteLights) {
            EscapeCtx ec{p.depth == 0 || (p.flags & PF_SPECULAR) != 0, p.rayD, Mid(p.prev.pi), p.prev.ns};
            if (vol)
                EscapedRadianceVol(rs, p.lambdaU, io, ec, lightBvh);
            else
                EscapedRadiance(rs, p.lambdaU, p.pB, io, ec, lightBvh);
        }
        KillPath(p);
        return;
    }
    uint32_t slot = hit[0];
    float b0, b1, b2;
    memcpy(&b0, &hit[2], 4);
    memcpy(&b1, &hit[3], 4);
    memcpy(&b2, &hit[4], 4);
    TriVerts tv = LoadTri(rs.prims, slot);
    const MeshAttrDev mesh = rs.meshes[tv.mesh];
    V3 wo = -p.rayD;
    SurfHit si;
    if (hit[5] != 0) {
        // hit inside an object instance: the interaction is built in the object's space from the
        // instance ray (cpu/primitive.cpp:116-125) and transformed back
        const InstanceDev &in = reinterpret_cast<const InstanceDev *>(rs.instances)[hit[5] - 1];
        const float *mi = in.mInv;
        V3 dInst = mk(mi[0] * p.rayD.x + mi[1] * p.rayD.y + mi[2] * p.rayD.z,
                      mi[4] * p.rayD.x + mi[5] * p.rayD.y + mi[6] * p.rayD.z,
                      mi[8] * p.rayD.x + mi[9] * p.rayD.y + mi[10] * p.rayD.z);
        si = SlotInteraction<EXT>(rs.prims, slot, tv, b0, b1, b2, -dInst, mesh.flipNormal != 0, CornersOf(rs, mesh, tv.tri, tv.kind));
        TransformSurfHit(in, si);
    } else {
        si = SlotInteraction<EXT>(rs.prims, slot, tv, b0, b1, b2, wo, mesh.flipNormal != 0, CornersOf(rs, mesh, tv.tri, tv.kind));
    }

    // ---- emitted light at the hit (cpu/integrators.cpp:664-679)
    if (mesh.lightFirst >= 0) {
        const LightDev al = rs.lights[mesh.lightFirst + tv.tri];
        if (vol && AreaLightEmits(al, si.n, wo)) {  // cpu/integrators.cpp:1102-1115
            const SpectrumTable &sp = rs.spectra[al.spectrum];
            bool anyLe = false;
            {
                LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
                for (int k = 0; k < kNLambda; ++k, li.Next()) anyLe |= al.scale * SpectrumAt(sp, li.lam) != 0;
            }
            if (anyLe) {
                float avg;
                if (p.depth == 0 || (p.flags & PF_SPECULAR))
                    avg = AvgRu(io);
                else {
                    float pmf = al.aliasP;
                    if (rs.integrator.lightSampler == 0) pmf = 1.f / rs.nLights;
                    if (lightBvh) pmf = LightBvhPMF(rs, Mid(p.prev.pi), p.prev.ns, mesh.lightFirst + (int)tv.tri);
                    float p_l = pmf * ShapePDF<EXT>(rs.prims, slot, tv, p.prev, p.rayD, al.flipNormal != 0);
                    for (int k = 0; k < kNLambda; ++k) io.SetRl(k, io.Rl(k) * p_l);
                    avg = AvgRuRl(io);
                }
                LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
                for (int k = 0; k < kNLambda; ++k, li.Next()) {
                    float Le = al.scale * SpectrumAt(sp, li.lam);
                    io.AddL(k, io.Beta(k) * Le / avg);
                }
            }
        } else if (AreaLightEmits(al, si.n, wo)) {
            float w = 1.f;
            if (!(p.depth == 0 || (p.flags & PF_SPECULAR))) {
                float pmf = al.aliasP;  // PowerLightSampler::PMF (lightsamplers.h:78-82)
                if (rs.integrator.lightSampler == 0) pmf = 1.f / rs.nLights;
                if (lightBvh)
                    pmf = LightBvhPMF(rs, Mid(p.prev.pi), p.prev.ns, mesh.lightFirst + (int)tv.tri);
                float p_l = pmf * ShapePDF<EXT>(rs.prims, slot, tv, p.prev, p.rayD, al.flipNormal != 0);
                w = PowerHeuristic(p.pB, p_l);
            }
            const SpectrumTable &sp = rs.spectra[al.spectrum];
            LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
            for (int k = 0; k < kNLambda; ++k, li.Next()) {
                float Le = al.scale * SpectrumAt(sp, li.lam);
                if (Le != 0) io.AddL(k, io.Beta(k) * w * Le);
            }
        }
    }

    // ---- BSDF
    if (mesh.material < 0) {  // interface material: isect.SkipIntersection(&ray, tHit) and go on (:1118-1122)
        if (!vol) {
            KillPath(p);
            return;
        }
        p.rayO = OffsetRayOrigin(si.pi, si.n, p.rayD);  // SpawnRay(ray->d)
        p.rayTMax = INFINITY;
        p.medium = mesh.mediumInside == mesh.mediumOutside ? p.medium
                                                           : (DotN(si.n, p.rayD) > 0 ? mesh.mediumOutside : mesh.mediumInside);
        return;
    }
    const MaterialDev mat = rs.materials[mesh.material];
    Bsdf bsdf = MakeBsdf<MAT, EXT>(rs, mat, LambdaIter(p.lambdaU).lam, p);
    // regularize first: it can turn a specular lobe glossy, which changes Flags() (cpu/integrators.cpp:714-718)
    if (rs.integrator.regularize && (p.flags & PF_ANY_NONSPECULAR)) BsdfRegularize(bsdf);
    // the film's VisibleSurface (film.cpp:197-213) at the first vertex with a BSDF (cpu/integrators.cpp:721-732): p, the
    // geometric and the shading normal turned to wo, the material and the instance (GBufferFilm's AOVs)
    if (p.depth == 0)
        io.VisibleSurface(Mid(si.pi), DotN(si.n, si.wo) < 0 ? -si.n : si.n, DotN(si.ns, si.wo) < 0 ? -si.ns : si.ns,
                          mesh.material, (int)hit[5]);
    uint32_t bflags = BsdfFlags(bsdf);
    float spec8[EXT ? kNLambda : 1];  // a coated conductor's value per wavelength (BsdfEval kind 8)
    const bool cachesR = (bsdf.type == MAT_DIFFUSE && !bsdf.trans) || (bsdf.type == MAT_COATED_DIFFUSE && !bsdf.coatCond);
    if (cachesR) {  // R at the path's wavelengths, once (DiffuseMaterial / CoatedDiffuseMaterial::GetBxDF clamp it)
        bool anyR = false;
        LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k, li.Next()) {
            float R = Clampf(SpectrumAt(rs.spectra[bsdf.specA], li.lam), 0, 1);
            io.SetR(k, R);
            anyR |= R != 0;
        }
        // DiffuseBxDF::Flags: Unset when R is black (bxdfs.h:33-35)
        if (bsdf.type == MAT_DIFFUSE && !anyR) bflags = 0;
    }
    // BSDF shading frame: Frame::FromXZ(Normalize(dpdus), ns) (bsdf.h:24-25)
    V3 fx = Normalize(si.dpdus), fz = si.ns, fy = Cross(fz, fx);
    V3 woL = mk(Dot(wo, fx), Dot(wo, fy), Dot(wo, fz));
    // GBufferFilm's albedo of the first visible surface (cpu/integrators.cpp:690-711): bsdf.rho(isect.wo, ucRho, uRho),
    // 16 BSDF samples at fixed points; kept as the XYZ of albedo * illuminant (SampledSpectrum::ToXYZ, util/spectrum.cpp)
    // under both wavelength pdfs the path can end with (the film converts with the final lambda, film.cpp:640-642)
    if (EXT && p.depth == 0 && rs.film.gbufferAlbedo && io.WantsAlbedo()) {
        const float ucRho[16] = {0.75741637f, 0.37870818f, 0.7083487f, 0.18935409f, 0.9149363f, 0.35417435f, 0.5990858f, 0.09467703f,
                                 0.8578725f, 0.45746812f, 0.686759f, 0.17708716f, 0.9674518f, 0.2995429f, 0.5083201f, 0.047338516f};
        const float uRho[16][2] = {{0.855985f, 0.570367f}, {0.381823f, 0.851844f}, {0.285328f, 0.764262f}, {0.733380f, 0.114073f},
                                   {0.542663f, 0.344465f}, {0.127274f, 0.414848f}, {0.964700f, 0.947162f}, {0.594089f, 0.643463f},
                                   {0.095109f, 0.170369f}, {0.825444f, 0.263359f}, {0.429467f, 0.454469f}, {0.244460f, 0.816459f},
                                   {0.756135f, 0.731258f}, {0.516165f, 0.152852f}, {0.180888f, 0.214174f}, {0.898579f, 0.503897f}};
        float alb[kNLambda];
        for (int k = 0; k < kNLambda; ++k) alb[k] = 0;
        const V3 woRho = mk(Dot(si.wo, fx), Dot(si.wo, fy), Dot(si.wo, fz));  // RenderToLocal(isect.wo)
        for (int i = 0; i < 16 && woRho.z != 0; ++i) {  // BxDF::rho (bxdfs.cpp:1135-1148)
            BsdfEval ev;
            BsdfSample bs = EvalSample(rs, p.lambdaU, bsdf, woRho, ucRho[i], uRho[i][0], uRho[i][1], ev, spec8);
            if (!bs.ok || !(bs.pdf > 0)) continue;
            const float c = AbsCosTheta(bs.wi) / bs.pdf;
            LambdaIter li(p.lambdaU);
            for (int k = 0; k < kNLambda; ++k, li.Next())
                alb[k] += BsdfSpectral<EXT>(rs, bsdf, ev, li.lam, cachesR ? io.R(k) : 0.f, k, spec8) * c;
        }
        float xyz[6] = {0, 0, 0, 0, 0, 0};
        const float pdfU = 1 / (kLambdaMax - kLambdaMin);
        LambdaIter li(p.lambdaU);
        for (int k = 0; k < kNLambda; ++k, li.Next()) {
            const float a = (alb[k] / 16) * SpectrumAt(rs.spectra[rs.film.albedoIlluminant], li.lam);
            for (int c = 0; c < 3; ++c) {
                const float v = SpectrumAt(rs.spectra[rs.film.xyzSpectrum[c]], li.lam) * a;
                xyz[c] += v / pdfU;
                if (k == 0) xyz[3 + c] = v / (pdfU / kNLambda);  // after TerminateSecondary: pdf[0] / N, the rest 0
            }
        }
        for (int c = 0; c < 6; ++c) xyz[c] = xyz[c] / kNLambda / kCIE_Y_integral;
        io.VisibleSurfaceAlbedo(xyz);
    }

    if (p.depth++ == rs.integrator.maxDepth) {
        KillPath(p);
        return;
    }
    PathSampler s = SamplerFor<EXT>(rs, p);

    // ---- direct lighting (SampleLd, cpu/integrators.cpp:764-805)
    if (bflags & (BX_DIFFUSE | BX_GLOSSY)) {
        LightCtx ctx;
        ctx.pi = si.pi;
        ctx.n = si.n;
        ctx.ns = si.ns;
        bool refl = (bflags & BX_REFLECTION) != 0, trans = (bflags & BX_TRANSMISSION) != 0;
        if (refl != trans) {  // nudge to the side light can arrive from
            V3 nudged = OffsetRayOrigin(si.pi, si.n, refl ? si.wo : -si.wo);
            ctx.pi.lo = nudged;
            ctx.pi.hi = nudged;
        }
        // SampleLd evaluates the BSDF with intr.wo (normalised); Sample_f below gets -ray.d as it is
        V3 woLd = mk(Dot(si.wo, fx), Dot(si.wo, fy), Dot(si.wo, fz));
        float uL = SamplerGet1D(s);
        float uLight0, uLight1;
        SamplerGet2D(s, uLight0, uLight1);
        if (rs.nLights > 0) {
            // AliasTable::Sample (util/sampling.cpp:619-647) / UniformLightSampler (lightsamplers.h:33-53)
            int nl = (int)rs.nLights, lidx;
            float pmf;
            if (lightBvh) {
                lidx = LightBvhSample(rs, Mid(ctx.pi), ctx.ns, uL, &pmf);
            } else if (rs.integrator.lightSampler == 0) {
                lidx = (int)fminf(uL * nl, (float)(nl - 1));
                pmf = 1.f / nl;
            } else if (false) {
                lidx = LightBvhSample(rs, Mid(ctx.pi), ctx.ns, uL, &pmf);
            } else {
                int offset = (int)fminf(uL * nl, (float)(nl - 1));
                float up = fminf(uL * nl - offset, kOneMinusEps);
                lidx = (up < rs.lights[offset].aliasQ) ? offset : rs.lights[offset].alias;
                pmf = rs.lights[lidx].aliasP;
            }
            // no light chosen, or a UniformInfiniteLight: its SampleLi gives nothing under
            // allowIncompletePDF (lights.cpp:986-990); it is found by BSDF sampling alone
            const LightDev light = rs.lights[lidx < 0 ? 0 : lidx];
            ShapeSampleOut ss;
            ss.ok = false;
            if (lidx >= 0 && light.type == 0) {
                TriVerts lt = LoadTri(rs.prims, light.slot);
                ss = SampleShapeSolidAngle<EXT>(rs.prims, light.slot, lt, ctx, uLight0, uLight1, light.flipNormal != 0, CornersOf(rs, rs.meshes[lt.mesh], lt.tri, lt.kind).n);
            }
            V3 cp = Mid(ctx.pi);
            // the delta lights (point, spot, distant): pdf 1, no MIS (cpu/integrators.cpp:798-799, 1386-1387)
            const bool delta = EXT && lidx >= 0 && light.type >= LIGHT_POINT && light.type <= LIGHT_DISTANT;
            V3 wiDelta = mk(0, 0, 0);
            float attenDelta = 1.f;
            if (delta) SampleDeltaLight(light, rs.lightGeom + size_t(kLightGeomFloats) * lidx, cp, ss, wiDelta, attenDelta);
            // an image infinite light: direction from its compensated distribution, MIS as for an area light
            const bool env = EXT && lidx >= 0 && light.type == LIGHT_IMAGE_INFINITE;
            int envPixel = 0;
            if (env) SampleEnvLight(rs, rs.lightGeom + size_t(kLightGeomFloats) * lidx, cp, uLight0, uLight1, ss, wiDelta, envPixel);
            if (ss.ok && ss.pdf != 0 && LengthSquared(Mid(ss.pi) - cp) != 0) {
                V3 wi = (delta || env) ? wiDelta : Normalize(Mid(ss.pi) - cp);
                if (delta || env || AreaLightEmits(light, ss.n, -wi)) {
                    float atten = attenDelta;
                    if (!delta && !env && light.cosFalloffEnd > 0) {  // ISET spread angle (lights.cpp:762-770)
                        float cos_a = -DotN(ss.n, wi);
                        float sin_a = SafeSqrt(1 - SqrF(cos_a));
                        float tan_a = sin_a / cos_a;
                        atten = fmaxf((1.0f - (light.tanFalloffEnd * tan_a)) * light.normalizeFalloffEnd, 0.0f);
                    }
                    V3 wiL = mk(Dot(wi, fx), Dot(wi, fy), Dot(wi, fz));
                    BsdfEval ev = (woLd.z == 0) ? ZeroEval() : EvalF(rs, p.lambdaU, bsdf, woLd, wiL, spec8);  // BSDF::f (bsdf.h:37-43)
                    if (ev.kind != 0) {
                        float absCos = AbsDotN(si.ns, wi);
                        float p_l = pmf * ss.pdf;
                        // a delta light: w_l = PowerHeuristic(p_l, 0) = 1 and the volpath estimate divides by r_l alone
                        float p_b = (woLd.z == 0 || delta) ? 0.f : EvalPDF(bsdf, woLd, wiL);
                        float w_l = PowerHeuristic(p_l, p_b);
                        const SpectrumTable &sp = rs.spectra[light.spectrum];
                        bool anyL = false, anyF = false;
                        io.ShadowBegin();
                        LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
                        for (int k = 0; k < kNLambda; ++k, li.Next()) {
                            float Le = env ? EnvLe(rs, light, envPixel, li.lam) : light.scale * SpectrumAt(sp, li.lam) * atten;
                            float f = BsdfSpectral<EXT>(rs, bsdf, ev, li.lam, cachesR ? io.R(k) : 0.f, k, spec8) * absCos;
                            // volpath keeps beta * f_hat * L; the transmittance walk supplies T_ray and the MIS weight
                            io.SetLd(k, vol ? io.Beta(k) * f * Le : io.Beta(k) * (w_l * Le * f / p_l));
                            anyL |= Le != 0;
                            anyF |= f != 0;
                        }
                        // Unoccluded(intr, ls->pLight): intr.SpawnRayTo(pLight) (ray.h:106-111), tMax 1 - ShadowEpsilon
                        V3 pf = OffsetRayOrigin(si.pi, si.n, Mid(ss.pi) - Mid(si.pi));
                        V3 pt = OffsetRayOrigin(ss.pi, ss.n, pf - Mid(ss.pi));
                        if (vol) {
                            ShadowVolInfo info;
                            info.p_l = p_l;
                            info.scatterPDF = p_b;
                            info.medium = mesh.mediumInside == mesh.mediumOutside
                                              ? p.medium
                                              : (DotN(si.n, pt - pf) > 0 ? mesh.mediumOutside : mesh.mediumInside);
                            info.lightPi = ss.pi;
                            info.lightN = ss.n;
                            io.ShadowEndVol(anyL && anyF, pf, pt - pf, 1 - kShadowEpsilon, info);
                        } else
                            io.ShadowEnd(anyL && anyF, pf, pt - pf, 1 - kShadowEpsilon);
                    }
                }
            }
        }
    }

    // ---- sample the BSDF for the next direction (cpu/integrators.cpp:728-745, BSDF::Sample_f bsdf.h:58-80)
    float uc = SamplerGet1D(s);
    float u0, u1;
    SamplerGet2D(s, u0, u1);
    p.dimension = s.z.dimension;
    if (woL.z == 0 || !(bflags & (BX_REFLECTION | BX_TRANSMISSION))) {
        KillPath(p);
        return;
    }
    BsdfEval ev;
    BsdfSample bs = EvalSample(rs, p.lambdaU, bsdf, woL, uc, u0, u1, ev, spec8);
    if (!bs.ok || bs.pdf == 0 || bs.wi.z == 0) {
        KillPath(p);
        return;
    }
    V3 wi = bs.wi.x * fx + bs.wi.y * fy + bs.wi.z * fz;  // Frame::FromLocal
    float absCos = AbsDotN(si.ns, wi);
#ifdef B2_PATH_DEBUG
    printf("      sample: f0 %.9g absdot %.9g pdf %.9g eta %.9g woL %.9g %.9g %.9g\n", BsdfSpectral<EXT>(rs, bsdf, ev, 500.f, 0.f, 0, spec8),
           absCos, bs.pdf, bs.eta, woL.x, woL.y, woL.z);
#endif
    bool anyF = false;
    {
        LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k, li.Next()) {
            float f = BsdfSpectral<EXT>(rs, bsdf, ev, li.lam, cachesR ? io.R(k) : 0.f, k, spec8);
            anyF |= f != 0;
            io.SetBeta(k, io.Beta(k) * (f * absCos / bs.pdf));
        }
    }
    if (!anyF) {  // !bs->f
        KillPath(p);
        return;
    }
    p.pB = bs.pdfIsProportional ? EvalPDF(bsdf, woL, bs.wi) : bs.pdf;
    if (vol) {  // r_l = r_u / pdf (cpu/integrators.cpp:1190-1193); the ray's medium follows the surface's interface
        for (int k = 0; k < kNLambda; ++k) io.SetRl(k, io.Ru(k) / p.pB);
        p.medium = mesh.mediumInside == mesh.mediumOutside ? p.medium
                                                           : (DotN(si.n, wi) > 0 ? mesh.mediumOutside : mesh.mediumInside);
    }
    bool specular = (bs.flags & BX_SPECULAR) != 0;
    p.flags = specular ? (p.flags | PF_SPECULAR) : ((p.flags & ~PF_SPECULAR) | PF_ANY_NONSPECULAR);
    if (bs.flags & BX_TRANSMISSION) p.etaScale *= SqrF(bs.eta);
    p.prev.pi = si.pi;
    p.prev.n = si.n;
    p.prev.ns = si.ns;
    p.rayO = OffsetRayOrigin(si.pi, si.n, wi);  // isect.SpawnRay(wi)
    p.rayD = wi;
    p.rayTMax = INFINITY;

    // ---- Russian roulette (cpu/integrators.cpp:748-756; volpath :1255-1266: rrBeta = beta * etaScale /
    // r_u.Average(), and uRR is drawn whether or not it is used)
    if (vol) {
        const float avg = AvgRu(io);
        float mc = io.Beta(0) * p.etaScale / avg;
        for (int k = 1; k < kNLambda; ++k) mc = fmaxf(mc, io.Beta(k) * p.etaScale / avg);
        s = SamplerFor<EXT>(rs, p);
        float uRR = SamplerGet1D(s);
        p.dimension = s.z.dimension;
        if (mc < 1 && p.depth > 1) {
            float q = fmaxf(0.f, 1 - mc);
            if (uRR < q) {
                KillPath(p);
                return;
            }
            for (int k = 0; k < kNLambda; ++k) io.SetBeta(k, io.Beta(k) / (1 - q));
        }
        return;
    }
    float maxComp = io.Beta(0) * p.etaScale;
B2_SPECTRAL_LOOP
    for (int k = 1; k < kNLambda; ++k) maxComp = fmaxf(maxComp, io.Beta(k) * p.etaScale);
#ifdef B2_PATH_DEBUG
    printf("      rr maxComp %.9g\n", maxComp);
#endif
    if (maxComp < 1 && p.depth > 1) {
        float q = fmaxf(0.f, 1 - maxComp);
        s = SamplerFor<EXT>(rs, p);
        float ur = SamplerGet1D(s);
        p.dimension = s.z.dimension;
        if (ur < q) {
            KillPath(p);
            return;
        }
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) io.SetBeta(k, io.Beta(k) / (1 - q));
    }
}

// RGBFilm::AddSample (film.h:241-258) with PixelSensor::ToSensorRGB (film.h:95-100) on
// L = SafeDiv(L, lambda.PDF()) (cpu/integrators.cpp:264-268).  L(k) reads the path's radiance.
template <class LFn>
B2_HD void FinishPath(const RenderSceneDev &rs, float lambdaU, uint32_t flags, float filterWeight, LFn L,
                      float rgbw[4]) {
    const float pdfU = 1 / (kLambdaMax - kLambdaMin);
    bool term = (flags & PF_TERMINATED_SECONDARY) != 0;
    bool bad = false;
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k) {
        float v = L(k);
        bad |= std::isnan(v) | std::isinf(v);
    }
    const SpectrumTable &rb = rs.spectra[rs.film.sensorSpectrum[0]];
    const SpectrumTable &gb = rs.spectra[rs.film.sensorSpectrum[1]];
    const SpectrumTable &bb = rs.spectra[rs.film.sensorSpectrum[2]];
    float sum[3] = {0, 0, 0};
    LambdaIter li(lambdaU);
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k, li.Next()) {
        // after TerminateSecondary: pdf[0] = pdfU / 31, pdf[k > 0] = 0 (SafeDiv gives 0)
        float pdf = term ? (k == 0 ? pdfU / kNLambda : 0.f) : pdfU;
        float Lk = (bad || pdf == 0) ? 0.f : L(k) / pdf;
        sum[0] += SpectrumAt(rb, li.lam) * Lk;
        sum[1] += SpectrumAt(gb, li.lam) * Lk;
        sum[2] += SpectrumAt(bb, li.lam) * Lk;
    }
    float rgb[3];
    for (int c = 0; c < 3; ++c) rgb[c] = rs.film.imagingRatio * (sum[c] / kNLambda);
    float m = fmaxf(rgb[0], fmaxf(rgb[1], rgb[2]));
    if (m > rs.film.maxComponentValue) {
        float sc = rs.film.maxComponentValue / m;
        rgb[0] *= sc;
        rgb[1] *= sc;
        rgb[2] *= sc;
    }
    rgbw[0] = filterWeight * rgb[0];
    rgbw[1] = filterWeight * rgb[1];
    rgbw[2] = filterWeight * rgb[2];
    rgbw[3] = filterWeight;
}

// SpectralFilm::AddSample, the spectral half (film.h:466-488): L clamped as a whole to maxComponentValue,
// times weight * CIE_Y_integral, each wavelength added to its bucket.  L is the sample's radiance as the
// integrator returned it (after the NaN / Inf guard of cpu/integrators.cpp:270-279), NOT divided by the
// wavelength pdf; after TerminateSecondary the secondary wavelengths are added too, as the reference does.
// acc(bucket, value, weight) performs `bucketSums[b] += value; weightSums[b] += weight` in doubles.
B2_HD int LambdaToBucket(float lambda, int nBuckets) {  // film.h:534-539
    int bucket = (int)(nBuckets * (lambda - kLambdaMin) / (kLambdaMax - kLambdaMin));
    return bucket < 0 ? 0 : (bucket > nBuckets - 1 ? nBuckets - 1 : bucket);
}
template <class LFn, class Acc>
B2_HD void SpectralFilmAdd(const FilmDev &film, float lambdaU, float weight, LFn L, Acc acc) {
    bool bad = false;
    float lm = L(0);
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k) {
        float v = L(k);
        bad |= std::isnan(v) | std::isinf(v);
        lm = fmaxf(lm, v);
    }
    bool clamp = !bad && lm > film.maxComponentValue;
    float sc = clamp ? film.maxComponentValue / lm : 1.f;
    float wY = weight * kCIE_Y_integral;
    LambdaIter li(lambdaU);
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k, li.Next()) {
        float v = bad ? 0.f : L(k);
        if (clamp) v *= sc;
        v *= wY;
        acc(LambdaToBucket(li.lam, film.nBuckets), v, weight);
    }
}

}  // namespace b200pt

#include "render_volpath.h"
