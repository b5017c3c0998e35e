// render_path.h - one path in registers and the steps the kernels apply to it (included by
// render_shade.h).  GenerateCameraPath = RayIntegrator::EvaluatePixelSample up to the camera ray
// (cpu/integrators.cpp:228-262); ShadeVertex = one iteration of PathIntegrator::Li's loop
// (cpu/integrators.cpp:640-760) with SampleLd (:764-805) split at the shadow ray, whose
// visibility comes back from the any-hit kernel before the next step; FinishPath = film AddSample.
#pragma once

namespace b200pt {

enum : uint32_t { PF_ALIVE = 1u, PF_SPECULAR = 2u, PF_ANY_NONSPECULAR = 4u, PF_SHADOW_PENDING = 8u };

struct PathRegs {
    float beta[kNLambda], L[kNLambda], Ld[kNLambda];
    float lambdaU;
    int px, py, sampleIndex, dimension, depth;
    uint32_t flags;
    float etaScale, pB, filterWeight;
    LightCtx prev;
    V3 rayO, rayD;
    float rayTMax;
    V3 shO, shD;
    float shTMax;
};

B2_HD ZSobol SamplerFor(const RenderSceneDev &rs, const PathRegs &p) {
    ZSobol s;
    ZSobolStart(s, rs.sampler, p.px, p.py, p.sampleIndex, p.dimension);
    return s;
}

// RayIntegrator::EvaluatePixelSample (cpu/integrators.cpp:228-262) + GetCameraSample (samplers.h:797-815)
B2_HD void GenerateCameraPath(const RenderSceneDev &rs, PathRegs &p, int px, int py, int sampleIndex) {
    p.px = px;
    p.py = py;
    p.sampleIndex = sampleIndex;
    p.dimension = 0;
    ZSobol s = SamplerFor(rs, p);
    p.lambdaU = ZSobolGet1D(s);
    float f0, f1;
    ZSobolGet2D(s, f0, f1);  // GetPixel2D -> filter.Sample
    // BoxFilter::Sample (filters.h): p = Lerp(u, -radius, radius), weight 1
    float fx = Lerp(f0, -rs.film.filterRadiusX, rs.film.filterRadiusX);
    float fy = Lerp(f1, -rs.film.filterRadiusY, rs.film.filterRadiusY);
    p.filterWeight = 1.f;
    float pFilmX = px + fx + 0.5f, pFilmY = py + fy + 0.5f;
    (void)ZSobolGet1D(s);  // time
    float l0, l1;
    ZSobolGet2D(s, l0, l1);
    p.dimension = s.dimension;
    GeneratePerspectiveRay(rs.camera, pFilmX, pFilmY, l0, l1, p.rayO, p.rayD);
    p.rayTMax = INFINITY;
    for (int k = 0; k < kNLambda; ++k) {
        p.beta[k] = 1.f;
        p.L[k] = 0.f;
        p.Ld[k] = 0.f;
    }
    p.depth = 0;
    p.flags = PF_ALIVE;
    p.etaScale = 1.f;
    p.pB = 0.f;
    p.shTMax = -1.f;
}

B2_HD void KillPath(PathRegs &p) {
    p.flags &= ~PF_ALIVE;
    p.rayTMax = -1.f;  // a ray with tMax < 0 fails the root box test: the traversal kernel drops it at once
}

// One path vertex.  `hit` is the internal record {slot, t, b0, b1, b2} of the closest-hit kernel for
// p.ray; `prevOccluded` the any-hit answer for the shadow ray issued at the previous vertex.
B2_HD void ShadeVertex(const RenderSceneDev &rs, PathRegs &p, const uint32_t hit[6], bool prevOccluded) {
    // ---- finish SampleLd of the previous vertex: L += beta * Ld if unoccluded (cpu/integrators.cpp:795-804)
    if (p.flags & PF_SHADOW_PENDING) {
        if (!prevOccluded)
            for (int k = 0; k < kNLambda; ++k) p.L[k] += p.Ld[k];
        p.flags &= ~PF_SHADOW_PENDING;
    }
    p.shTMax = -1.f;
    if (!(p.flags & PF_ALIVE)) return;

    if (hit[0] == kMissTag) {  // no infinite lights yet: an escaped ray adds nothing (cpu/integrators.cpp:645-662)
        KillPath(p);
        return;
    }
    float lambda[kNLambda];
    SampleWavelengths(p.lambdaU, lambda);
    uint32_t slot = hit[0];
    float b0, b1, b2;
    memcpy(&b0, &hit[2], 4);
    memcpy(&b1, &hit[3], 4);
    memcpy(&b2, &hit[4], 4);
    TriVerts tv = LoadTri(rs.prims, slot);
    const MeshAttrDev &mesh = rs.meshes[tv.mesh];
    V3 wo = -p.rayD;
    SurfHit si = TriangleInteraction(tv, b0, b1, b2, wo, mesh.flipNormal != 0);

    // ---- emitted light at the hit (cpu/integrators.cpp:664-679)
    if (mesh.lightFirst >= 0) {
        const LightDev &al = rs.lights[mesh.lightFirst + tv.tri];
        if (AreaLightEmits(al, si.n, wo)) {
            float w = 1.f;
            if (!(p.depth == 0 || (p.flags & PF_SPECULAR))) {
                float pmf = al.aliasP;  // PowerLightSampler::PMF (lightsamplers.h:78-82)
                if (rs.integrator.lightSampler == 0) pmf = 1.f / rs.nLights;
                float p_l = pmf * TrianglePDF(tv, p.prev, p.rayD, al.flipNormal != 0);
                w = PowerHeuristic(p.pB, p_l);
            }
            const SpectrumTable &sp = rs.spectra[al.spectrum];
            for (int k = 0; k < kNLambda; ++k) p.L[k] += p.beta[k] * w * (al.scale * SpectrumAt(sp, lambda[k]));
        }
    }

    // ---- BSDF (materials.h:430-476 DiffuseMaterial::GetBxDF)
    if (mesh.material < 0) {  // interface material: skip the surface (cpu/integrators.cpp:683-688); not used yet
        KillPath(p);
        return;
    }
    const MaterialDev &mat = rs.materials[mesh.material];
    float R[kNLambda];
    bool anyR = false;
    {
        const SpectrumTable &sp = rs.spectra[mat.spectrum[0]];
        for (int k = 0; k < kNLambda; ++k) {
            R[k] = Clampf(SpectrumAt(sp, lambda[k]), 0, 1);
            anyR |= R[k] != 0;
        }
    }
    // BSDF shading frame: Frame::FromXZ(Normalize(dpdus), ns) (bsdf.h:24-25)
    V3 fx = Normalize(si.dpdus), fz = si.ns, fy = Cross(fz, fx);
    V3 woL = mk(Dot(wo, fx), Dot(wo, fy), Dot(wo, fz));

    if (p.depth++ == rs.integrator.maxDepth) {
        KillPath(p);
        return;
    }
    ZSobol s = SamplerFor(rs, p);

    // ---- direct lighting (SampleLd, cpu/integrators.cpp:764-805); DiffuseBxDF flags: reflective only
    if (anyR) {
        LightCtx ctx;
        ctx.pi = si.pi;
        ctx.n = si.n;
        ctx.ns = si.ns;
        V3 nudged = OffsetRayOrigin(si.pi, si.n, wo);  // ctx.pi = intr.OffsetRayOrigin(intr.wo)
        ctx.pi.lo = nudged;
        ctx.pi.hi = nudged;
        float uL = ZSobolGet1D(s);
        float uLight0, uLight1;
        ZSobolGet2D(s, uLight0, uLight1);
        if (rs.nLights > 0) {
            // AliasTable::Sample (util/sampling.cpp:619-647) / UniformLightSampler (lightsamplers.h:33-53)
            int nl = (int)rs.nLights, li;
            float pmf;
            if (rs.integrator.lightSampler == 0) {
                li = (int)fminf(uL * nl, (float)(nl - 1));
                pmf = 1.f / nl;
            } else {
                int offset = (int)fminf(uL * nl, (float)(nl - 1));
                float up = fminf(uL * nl - offset, kOneMinusEps);
                li = (up < rs.lights[offset].aliasQ) ? offset : rs.lights[offset].alias;
                pmf = rs.lights[li].aliasP;
            }
            const LightDev &light = rs.lights[li];
            TriVerts lt = LoadTri(rs.prims, light.slot);
            ShapeSampleOut ss = SampleTriangleSolidAngle(lt, ctx, uLight0, uLight1, light.flipNormal != 0);
            V3 cp = Mid(ctx.pi);
            if (ss.ok && ss.pdf != 0 && LengthSquared(Mid(ss.pi) - cp) != 0) {
                V3 wi = Normalize(Mid(ss.pi) - cp);
                if (AreaLightEmits(light, ss.n, -wi)) {
                    float atten = 1.f;
                    if (light.cosFalloffEnd > 0) {  // ISET spread angle (lights.cpp:762-770)
                        float cos_a = -Dot(wi, ss.n);
                        float sin_a = SafeSqrt(1 - SqrF(cos_a));
                        float tan_a = sin_a / cos_a;
                        atten = fmaxf((1.0f - (light.tanFalloffEnd * tan_a)) * light.normalizeFalloffEnd, 0.0f);
                    }
                    V3 wiL = mk(Dot(wi, fx), Dot(wi, fy), Dot(wi, fz));
                    bool sameHemi = wiL.z * woL.z > 0;
                    if (woL.z != 0 && sameHemi) {
                        float absCos = AbsDot(wi, si.ns);
                        float p_l = pmf * ss.pdf;
                        float p_b = fabsf(wiL.z) * kInvPi;  // DiffuseBxDF::PDF
                        float w_l = PowerHeuristic(p_l, p_b);
                        const SpectrumTable &sp = rs.spectra[light.spectrum];
                        bool any = false;
                        for (int k = 0; k < kNLambda; ++k) {
                            float Le = light.scale * SpectrumAt(sp, lambda[k]) * atten;
                            float f = R[k] * kInvPi * absCos;
                            float v = w_l * Le * f / p_l;
                            p.Ld[k] = p.beta[k] * v;
                            any |= (Le != 0) & (f != 0);
                        }
                        if (any) {
                            // Unoccluded(intr, ls->pLight): intr.SpawnRayTo(pLight) (ray.h:106-111), tMax 1 - ShadowEpsilon
                            V3 pf = OffsetRayOrigin(si.pi, si.n, Mid(ss.pi) - Mid(si.pi));
                            V3 pt = OffsetRayOrigin(ss.pi, ss.n, pf - Mid(ss.pi));
                            p.shO = pf;
                            p.shD = pt - pf;
                            p.shTMax = 1 - kShadowEpsilon;
                            p.flags |= PF_SHADOW_PENDING;
                        }
                    }
                }
            }
        }
    }

    // ---- sample the BSDF for the next direction (cpu/integrators.cpp:728-745; DiffuseBxDF::Sample_f bxdfs.h:44-56)
    float uc = ZSobolGet1D(s);
    (void)uc;
    float u0, u1;
    ZSobolGet2D(s, u0, u1);
    p.dimension = s.dimension;
    if (woL.z == 0 || !anyR) {
        KillPath(p);
        return;
    }
    V3 wiL = SampleCosineHemisphere(u0, u1);
    if (woL.z < 0) wiL.z *= -1;
    float pdf = fabsf(wiL.z) * kInvPi;
    if (pdf == 0 || wiL.z == 0) {
        KillPath(p);
        return;
    }
    V3 wi = wiL.x * fx + wiL.y * fy + wiL.z * fz;  // Frame::FromLocal
    float absCos = AbsDot(wi, si.ns);
    for (int k = 0; k < kNLambda; ++k) p.beta[k] *= R[k] * kInvPi * absCos / pdf;
    p.pB = pdf;
    p.flags &= ~PF_SPECULAR;
    p.flags |= PF_ANY_NONSPECULAR;
    p.prev.pi = si.pi;
    p.prev.n = si.n;
    p.prev.ns = si.ns;
    p.rayO = OffsetRayOrigin(si.pi, si.n, wi);  // isect.SpawnRay(wi)
    p.rayD = wi;
    p.rayTMax = INFINITY;

    // ---- Russian roulette (cpu/integrators.cpp:748-756)
    float maxComp = p.beta[0] * p.etaScale;
    for (int k = 1; k < kNLambda; ++k) maxComp = fmaxf(maxComp, p.beta[k] * p.etaScale);
    if (maxComp < 1 && p.depth > 1) {
        float q = fmaxf(0.f, 1 - maxComp);
        s = SamplerFor(rs, p);
        float ur = ZSobolGet1D(s);
        p.dimension = s.dimension;
        if (ur < q) {
            KillPath(p);
            return;
        }
        for (int k = 0; k < kNLambda; ++k) p.beta[k] /= 1 - q;
    }
}

// RGBFilm::AddSample (film.h:241-258) with PixelSensor::ToSensorRGB (film.h:95-100).  Returns the
// weighted sensor rgb and weight to add into the pixel.
B2_HD void FinishPath(const RenderSceneDev &rs, PathRegs &p, bool lastOccluded, float rgbw[4]) {
    if (p.flags & PF_SHADOW_PENDING) {
        if (!lastOccluded)
            for (int k = 0; k < kNLambda; ++k) p.L[k] += p.Ld[k];
        p.flags &= ~PF_SHADOW_PENDING;
    }
    float lambda[kNLambda];
    SampleWavelengths(p.lambdaU, lambda);
    const float pdf = 1 / (kLambdaMax - kLambdaMin);
    float rgb[3] = {0, 0, 0};
    bool bad = false;
    for (int k = 0; k < kNLambda; ++k) bad |= std::isnan(p.L[k]) | std::isinf(p.L[k]);
    for (int c = 0; c < 3; ++c) {
        const SpectrumTable &bar = rs.spectra[rs.film.sensorSpectrum[c]];
        float sum = 0;
        for (int k = 0; k < kNLambda; ++k) {
            float Lk = bad ? 0.f : p.L[k] / pdf;  // SafeDiv(L, lambda.PDF())
            sum += SpectrumAt(bar, lambda[k]) * Lk;
        }
        rgb[c] = rs.film.imagingRatio * (sum / kNLambda);
    }
    float m = fmaxf(rgb[0], fmaxf(rgb[1], rgb[2]));
    if (m > rs.film.maxComponentValue) {
        float sc = rs.film.maxComponentValue / m;
        rgb[0] *= sc;
        rgb[1] *= sc;
        rgb[2] *= sc;
    }
    rgbw[0] = p.filterWeight * rgb[0];
    rgbw[1] = p.filterWeight * rgb[1];
    rgbw[2] = p.filterWeight * rgb[2];
    rgbw[3] = p.filterWeight;
}

}  // namespace b200pt
