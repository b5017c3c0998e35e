// render_volpath.h - the medium half of VolPathIntegrator::Li (cpu/integrators.cpp:953-1075: delta tracking
// along the path ray, light and phase sampling at a real scattering event) and the transmittance walk of its
// SampleLd (:1329-1395: ratio tracking from segment to segment through interface surfaces).  The surface half is
// ShadeVertex (render_path.h) in its volumetric mode.  Included at the end of render_path.h.
#pragma once
#include "render_volume.h"

namespace b200pt {

// SurfaceInteraction of a closest-hit record for a ray with direction rayD (as ShadeVertex builds it)
template <int EXT = 1>
B2_HD SurfHit SurfHitFromRecord(const RenderSceneDev &rs, V3 rayD, const uint32_t hit[6], uint32_t *meshOut) {
    float b0, b1, b2;
    memcpy(&b0, &hit[2], 4);
    memcpy(&b1, &hit[3], 4);
    memcpy(&b2, &hit[4], 4);
    TriVerts tv = LoadTri(rs.prims, hit[0]);
    const MeshAttrDev &mesh = rs.meshes[tv.mesh];
    *meshOut = tv.mesh;
    SurfHit si;
    if (hit[5] != 0) {
        const InstanceDev &in = reinterpret_cast<const InstanceDev *>(rs.instances)[hit[5] - 1];
        const float *mi = in.mInv;
        V3 dInst = mk(mi[0] * rayD.x + mi[1] * rayD.y + mi[2] * rayD.z, mi[4] * rayD.x + mi[5] * rayD.y + mi[6] * rayD.z,
                      mi[8] * rayD.x + mi[9] * rayD.y + mi[10] * rayD.z);
        si = SlotInteraction<EXT>(rs.prims, hit[0], tv, b0, b1, b2, -dInst, mesh.flipNormal != 0, CornersOf(rs, mesh, tv.tri, tv.kind));
        TransformSurfHit(in, si);
    } else {
        si = SlotInteraction<EXT>(rs.prims, hit[0], tv, b0, b1, b2, -rayD, mesh.flipNormal != 0, CornersOf(rs, mesh, tv.tri, tv.kind));
    }
    return si;
}

// Interaction::GetMedium(w) (interaction.h:117-121) for a surface of `mesh` met by a ray in medium `rayMedium`
B2_HD int MediumAfter(const MeshAttrDev &mesh, V3 n, V3 w, int rayMedium) {
    if (mesh.mediumInside == mesh.mediumOutside) return rayMedium;  // no transition: SetIntersectionProperties keeps the ray's
    return DotN(n, w) > 0 ? mesh.mediumOutside : mesh.mediumInside;
}

// LightSampler::Sample + Light::SampleLi for a reference point (the light half of SampleLd, cpu/integrators.cpp:1290-1303)
struct LightPick {
    bool env;        // an image infinite light: its radiance comes from pixel envPixel
    int envPixel;
    bool ok, delta;  // delta: pdf 1 and no MIS against phase-function sampling (cpu/integrators.cpp:1386-1387)
    int lidx;
    float pmf, atten;
    ShapeSampleOut ss;
    V3 wi;
};
template <int LBVH, int EXT = 1>
B2_HD LightPick PickAndSampleLight(const RenderSceneDev &rs, const LightCtx &ctx, float uL, float u0, float u1) {
    const bool lightBvh = LBVH < 0 ? rs.integrator.lightSampler == 2 : LBVH == 1;
    LightPick lp;
    lp.ok = false;
    lp.delta = false;
    lp.env = false;
    lp.envPixel = 0;
    lp.atten = 1.f;
    if (rs.nLights == 0) return lp;
    int nl = (int)rs.nLights;
    if (lightBvh) {
        lp.lidx = LightBvhSample(rs, Mid(ctx.pi), ctx.ns, uL, &lp.pmf);
    } else if (rs.integrator.lightSampler == 0) {
        lp.lidx = (int)fminf(uL * nl, (float)(nl - 1));
        lp.pmf = 1.f / nl;
    } else {
        int offset = (int)fminf(uL * nl, (float)(nl - 1));
        float up = fminf(uL * nl - offset, kOneMinusEps);
        lp.lidx = (up < rs.lights[offset].aliasQ) ? offset : rs.lights[offset].alias;
        lp.pmf = rs.lights[lp.lidx].aliasP;
    }
    if (lp.lidx < 0) return lp;
    const LightDev &light = rs.lights[lp.lidx];
    if (EXT && light.type >= LIGHT_POINT && light.type <= LIGHT_DISTANT) {  // point, spot, distant
        lp.delta = true;
        lp.ok = SampleDeltaLight(light, rs.lightGeom + size_t(kLightGeomFloats) * lp.lidx, Mid(ctx.pi), lp.ss, lp.wi, lp.atten);
        return lp;
    }
    if (EXT && light.type == LIGHT_IMAGE_INFINITE) {
        lp.ok = SampleEnvLight(rs, rs.lightGeom + size_t(kLightGeomFloats) * lp.lidx, Mid(ctx.pi), u0, u1, lp.ss, lp.wi, lp.envPixel);
        lp.env = lp.ok;
        return lp;
    }
    if (light.type != 0) return lp;  // a uniform infinite light gives nothing under allowIncompletePDF
    TriVerts lt = LoadTri(rs.prims, light.slot);
    lp.ss = SampleShapeSolidAngle<EXT>(rs.prims, light.slot, lt, ctx, u0, u1, light.flipNormal != 0, CornersOf(rs, rs.meshes[lt.mesh], lt.tri, lt.kind).n);
    V3 cp = Mid(ctx.pi);
    if (!lp.ss.ok || lp.ss.pdf == 0 || LengthSquared(Mid(lp.ss.pi) - cp) == 0) return lp;
    lp.wi = Normalize(Mid(lp.ss.pi) - cp);
    if (!AreaLightEmits(light, lp.ss.n, -lp.wi)) return lp;
    if (light.cosFalloffEnd > 0) {
        float cos_a = -DotN(lp.ss.n, lp.wi);
        float sin_a = SafeSqrt(1 - SqrF(cos_a));
        float tan_a = sin_a / cos_a;
        lp.atten = fmaxf((1.0f - (light.tanFalloffEnd * tan_a)) * light.normalizeFalloffEnd, 0.0f);
    }
    lp.ok = true;
    return lp;
}

template <class IO>
B2_HD bool AnyOf(const IO &io, int which) {
    bool any = false;
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k) any |= (which == 0 ? io.Beta(k) : io.Ru(k)) != 0;
    return any;
}

enum { VOL_SURVIVE = 0, VOL_TERMINATED = 1, VOL_SCATTERED = 2 };

// The `if (ray.medium)` block of VolPathIntegrator::Li for the ray p.ray with closest hit at tHit (INFINITY = none).
// IO adds Ru/SetRu/Rl/SetRl (rescaled path probabilities) and ShadowEndVol to ShadeVertex's IO.
// In three parts so that the delta-tracking walk in the middle can be stepped by a persistent kernel (render.cu
// VolMediumKernel): VolMediumBegin (sampler dimensions, the medium's spectra, the walk's start), the walk's steps
// (TMajWalk::Step with a PathMediumCB built over the same state), VolMediumEnd (what follows the walk).
struct VolMediumState {
    PCG32 rng;
    float sigA[kNLambda], sigS[kNLambda], sigT[kNLambda], T_maj[kNLambda];
    TMajWalk walk;
    bool terminated, scattered;
    V3 pScatter;
};

B2_HD void VolMediumBegin(const RenderSceneDev &rs, PathRegs &p, float tHit, VolMediumState &st) {
    const MediumDev &m = rs.media[p.medium];
    PathSampler s = SamplerFor<1>(rs, p);
    uint64_t hash0 = Hash1f(SamplerGet1D(s));
    uint64_t hash1 = Hash1f(SamplerGet1D(s));
    st.rng = PcgSeed(hash0, hash1);
    float u = SamplerGet1D(s);
    p.dimension = s.z.dimension;
    MediumSpectra(rs, m, p.lambdaU, st.sigA, st.sigS, st.sigT);
    st.terminated = st.scattered = false;
    st.pScatter = mk(0, 0, 0);
    st.walk.Begin(m, p.rayO, p.rayD, tHit, u, st.T_maj);
}

// one step of the walk; TMAJ_GO_ON = call again
template <class IO>
B2_HD int VolMediumWalkStep(const RenderSceneDev &rs, PathRegs &p, VolMediumState &st, IO &io) {
    PathMediumCB<IO> cb(io, st.rng, st.sigA, st.sigS, st.sigT, p.depth, rs.integrator.maxDepth, rs, rs.media[p.medium], p.lambdaU);
    int r = st.walk.Step(rs.media[p.medium], st.sigT, st.rng, st.T_maj, cb);
    st.terminated |= cb.terminated;
    if (cb.scattered) {
        st.scattered = true;
        st.pScatter = cb.pScatter;
    }
    return r;
}

template <class IO, int LBVH = -1, int EXT = 1>
B2_HD int VolMediumEnd(const RenderSceneDev &rs, PathRegs &p, VolMediumState &st, IO &io);

// the fast split walk of the persistent kernel (TMajFastWalk): st.walk is not used, `fw` is
B2_HD void VolMediumBeginFast(const RenderSceneDev &rs, PathRegs &p, float tHit, VolMediumState &st, TMajFastWalk &fw) {
    const MediumDev &m = rs.media[p.medium];
    PathSampler s = SamplerFor<1>(rs, p);
    uint64_t hash0 = Hash1f(SamplerGet1D(s));
    uint64_t hash1 = Hash1f(SamplerGet1D(s));
    st.rng = PcgSeed(hash0, hash1);
    float u = SamplerGet1D(s);
    p.dimension = s.z.dimension;
    MediumSpectra(rs, m, p.lambdaU, st.sigA, st.sigS, st.sigT);
    st.terminated = st.scattered = false;
    st.pScatter = mk(0, 0, 0);
    fw.Begin(m, p.rayO, p.rayD, tHit, u);
}
// the pending collision; false = the walk is over (st.T_maj all ones, as the reference leaves it)
template <class IO>
B2_HD bool VolMediumCollide(const RenderSceneDev &rs, PathRegs &p, VolMediumState &st, TMajFastWalk &fw, IO &io) {
    PathMediumCB<IO> cb(io, st.rng, st.sigA, st.sigS, st.sigT, p.depth, rs.integrator.maxDepth, rs, rs.media[p.medium], p.lambdaU);
    const bool goOn = fw.Collide(rs.media[p.medium], st.sigT, st.T_maj, cb);
    st.terminated |= cb.terminated;
    if (cb.scattered) {
        st.scattered = true;
        st.pScatter = cb.pScatter;
    }
    if (!goOn)
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) st.T_maj[k] = 1.f;
    return goOn;
}

template <class IO, int LBVH = -1, int EXT = 1>
B2_HD int VolMediumStep(const RenderSceneDev &rs, PathRegs &p, float tHit, IO &io) {
    VolMediumState st;
    VolMediumBegin(rs, p, tHit, st);
    while (VolMediumWalkStep(rs, p, st, io) == TMAJ_GO_ON) {
    }
    return VolMediumEnd<IO, LBVH, EXT>(rs, p, st, io);
}

template <class IO, int LBVH, int EXT>
B2_HD int VolMediumEnd(const RenderSceneDev &rs, PathRegs &p, VolMediumState &st, IO &io) {
    const MediumDev &m = rs.media[p.medium];
    const float *T_maj = st.T_maj;
    struct {
        bool terminated, scattered;
        V3 pScatter;
    } cb = {st.terminated, st.scattered, st.pScatter};
    if (cb.terminated || !AnyOf(io, 0) || !AnyOf(io, 1)) return VOL_TERMINATED;
    if (!cb.scattered) {
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) {
            float f = T_maj[k] / T_maj[0];
            io.SetBeta(k, io.Beta(k) * f);
            io.SetRu(k, io.Ru(k) * f);
            io.SetRl(k, io.Rl(k) * f);
        }
        return VOL_SURVIVE;
    }
    // ---- real scattering at cb.pScatter: SampleLd(intr, nullptr, ...) then the phase function (:1025-1049)
    V3 pS = cb.pScatter;
    V3 woN = Normalize(-p.rayD);  // MediumInteraction::wo
    LightCtx ctx;
    ctx.pi = MakeP3i(pS, mk(0, 0, 0));
    ctx.n = ctx.ns = mk(0, 0, 0);
    PathSampler s = SamplerFor<EXT>(rs, p);
    float uL = SamplerGet1D(s);
    float uLight0, uLight1;
    SamplerGet2D(s, uLight0, uLight1);
    LightPick lp = PickAndSampleLight<LBVH, EXT>(rs, ctx, uL, uLight0, uLight1);
    if (lp.ok) {
        float fHat = HenyeyGreenstein(Dot(woN, lp.wi), m.g);
        if (fHat != 0) {
            const LightDev &light = rs.lights[lp.lidx];
            const SpectrumTable &sp = rs.spectra[light.spectrum];
            bool anyL = false;
            io.ShadowBegin();
            LambdaIter li(p.lambdaU);
B2_SPECTRAL_LOOP
            for (int k = 0; k < kNLambda; ++k, li.Next()) {
                float Le = lp.env ? EnvLe(rs, light, lp.envPixel, li.lam) : light.scale * SpectrumAt(sp, li.lam) * lp.atten;
                io.SetLd(k, io.Beta(k) * fHat * Le);
                anyL |= Le != 0;
            }
            ShadowVolInfo info;
            info.p_l = lp.pmf * lp.ss.pdf;
            info.scatterPDF = lp.delta ? 0.f : fHat;
            info.medium = p.medium;
            info.lightPi = lp.ss.pi;
            info.lightN = lp.ss.n;
            V3 pf = OffsetRayOrigin(ctx.pi, ctx.n, Mid(lp.ss.pi) - pS);
            V3 pt = OffsetRayOrigin(lp.ss.pi, lp.ss.n, pf - Mid(lp.ss.pi));
            io.ShadowEndVol(anyL, pf, pt - pf, 1 - kShadowEpsilon, info);
        }
    }
    float u0, u1;
    SamplerGet2D(s, u0, u1);
    p.dimension = s.z.dimension;
    float pdf;
    V3 wi = SampleHenyeyGreenstein(-p.rayD, m.g, u0, u1, &pdf);
    if (pdf == 0) return VOL_TERMINATED;
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k) {
        io.SetBeta(k, io.Beta(k) * (pdf / pdf));
        io.SetRl(k, io.Ru(k) / pdf);
    }
    p.prev = ctx;
    p.rayO = pS;
    p.rayD = wi;
    p.rayTMax = INFINITY;
    p.flags = (p.flags & ~PF_SPECULAR) | PF_ANY_NONSPECULAR;
    return VOL_SCATTERED;
}

// One light-sample ray segment of SampleLd's transmittance walk (cpu/integrators.cpp:1334-1395).
struct ShadowRegs {
    V3 o, d;
    int medium;
    PCG32 rng;
    P3i lightPi;
    V3 lightN;
};
enum { SHADOW_BLOCKED = 0, SHADOW_REACHED = 1, SHADOW_CONTINUE = 2 };
// T_ray / r_l / r_u: the walk's running products (all ones before the first segment).  hit: closest hit of
// (o, d, 1 - ShadowEpsilon).  SHADOW_CONTINUE leaves the next segment's ray in sh.
// In three parts like VolMediumStep: ShadowTrBegin classifies the hit (SHADOW_BLOCKED at an opaque surface; otherwise
// -1, with st.walking = the segment runs through a medium and its ratio-tracking walk has been started), the walk's
// steps (ShadowTrWalkStep), ShadowTrEnd (the rest of the segment).
struct ShadowTrState {
    float sigA[kNLambda], sigS[kNLambda], sigT[kNLambda], T_maj[kNLambda];
    TMajWalk walk;
    SurfHit si;
    uint32_t meshId;
    bool isHit, walking;
};

template <int EXT = 1>
B2_HD int ShadowTrBegin(const RenderSceneDev &rs, float lambdaU, ShadowRegs &sh, const uint32_t hit[6], ShadowTrState &st) {
    st.isHit = hit[0] != kMissTag;
    st.meshId = 0;
    st.walking = false;
    if (st.isHit) {
        st.si = SurfHitFromRecord<EXT>(rs, sh.d, hit, &st.meshId);
        if (rs.meshes[st.meshId].material >= 0) return SHADOW_BLOCKED;  // an opaque surface
    }
    if (sh.medium >= 0) {
        const MediumDev &m = rs.media[sh.medium];
        float tHit;
        memcpy(&tHit, &hit[1], 4);
        float tMax = st.isHit ? tHit : (1 - kShadowEpsilon);
        float u = PcgFloat(sh.rng);
        MediumSpectra(rs, m, lambdaU, st.sigA, st.sigS, st.sigT);
        st.walk.Begin(m, sh.o, sh.d, tMax, u, st.T_maj);
        st.walking = true;
    }
    return -1;
}

// the fast split walk of the persistent kernel: ShadowTrBegin with `fw` started instead of st.walk
template <int EXT = 1>
B2_HD int ShadowTrBeginFast(const RenderSceneDev &rs, float lambdaU, ShadowRegs &sh, const uint32_t hit[6], ShadowTrState &st,
                            TMajFastWalk &fw) {
    st.isHit = hit[0] != kMissTag;
    st.meshId = 0;
    st.walking = false;
    if (st.isHit) {
        st.si = SurfHitFromRecord<EXT>(rs, sh.d, hit, &st.meshId);
        if (rs.meshes[st.meshId].material >= 0) return SHADOW_BLOCKED;  // an opaque surface
    }
    if (sh.medium >= 0) {
        const MediumDev &m = rs.media[sh.medium];
        float tHit;
        memcpy(&tHit, &hit[1], 4);
        float tMax = st.isHit ? tHit : (1 - kShadowEpsilon);
        float u = PcgFloat(sh.rng);
        MediumSpectra(rs, m, lambdaU, st.sigA, st.sigS, st.sigT);
        fw.Begin(m, sh.o, sh.d, tMax, u);
        st.walking = true;
    }
    return -1;
}
B2_HD bool ShadowTrCollide(const RenderSceneDev &rs, ShadowRegs &sh, ShadowTrState &st, TMajFastWalk &fw, float *T_ray,
                           float *r_l, float *r_u) {
    ShadowMediumCB cb(sh.rng, st.sigA, st.sigS, st.sigT, T_ray, r_l, r_u);
    const bool goOn = fw.Collide(rs.media[sh.medium], st.sigT, st.T_maj, cb);
    if (!goOn)
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) st.T_maj[k] = 1.f;
    return goOn;
}

B2_HD int ShadowTrWalkStep(const RenderSceneDev &rs, ShadowRegs &sh, ShadowTrState &st, float *T_ray, float *r_l, float *r_u) {
    ShadowMediumCB cb(sh.rng, st.sigA, st.sigS, st.sigT, T_ray, r_l, r_u);
    return st.walk.Step(rs.media[sh.medium], st.sigT, sh.rng, st.T_maj, cb);
}

B2_HD int ShadowTrEnd(const RenderSceneDev &rs, ShadowRegs &sh, ShadowTrState &st, float *T_ray, float *r_l, float *r_u) {
    if (st.walking) {
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) {
            float f = st.T_maj[k] / st.T_maj[0];
            T_ray[k] *= f;
            r_l[k] *= f;
            r_u[k] *= f;
        }
    }
    bool any = false;
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k) any |= T_ray[k] != 0;
    if (!any) return SHADOW_BLOCKED;
    if (!st.isHit) return SHADOW_REACHED;
    // lightRay = si->intr.SpawnRayTo(ls->pLight) (interaction.h:110-114, ray.h:106-111)
    V3 pf = OffsetRayOrigin(st.si.pi, st.si.n, Mid(sh.lightPi) - Mid(st.si.pi));
    V3 pt = OffsetRayOrigin(sh.lightPi, sh.lightN, pf - Mid(sh.lightPi));
    sh.o = pf;
    sh.d = pt - pf;
    sh.medium = MediumAfter(rs.meshes[st.meshId], st.si.n, sh.d, sh.medium);
    if (sh.d.x == 0 && sh.d.y == 0 && sh.d.z == 0) return SHADOW_REACHED;  // while (lightRay.d != Vector3f(0, 0, 0))
    return SHADOW_CONTINUE;
}

template <int EXT = 1>
B2_HD int ShadowTrSegment(const RenderSceneDev &rs, float lambdaU, ShadowRegs &sh, const uint32_t hit[6], float *T_ray,
                          float *r_l, float *r_u) {
    ShadowTrState st;
    int r = ShadowTrBegin<EXT>(rs, lambdaU, sh, hit, st);
    if (r >= 0) return r;
    if (st.walking)
        while (ShadowTrWalkStep(rs, sh, st, T_ray, r_l, r_u) == TMAJ_GO_ON) {
        }
    return ShadowTrEnd(rs, sh, st, T_ray, r_l, r_u);
}

// The estimate a walk that reached the light returns (cpu/integrators.cpp:1397-1403, area lights): Ld(k) is
// beta * f_hat * L as stored with the light sample, r_p(k) the path's r_u
template <class LdFn, class RpFn, class AddFn>
B2_HD void ShadowTrFinish(const float *T_ray, const float *r_l, const float *r_u, float p_l, float scatterPDF, LdFn Ld,
                          RpFn r_p, AddFn add) {
    float sum = 0;
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k) {
        float rp = r_p(k);
        float a = r_l[k] * (rp * p_l), b = r_u[k] * (rp * scatterPDF);
        sum = k == 0 ? a + b : sum + (a + b);
    }
    float avg = sum / kNLambda;
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k) add(k, Ld(k) * T_ray[k] / avg);
}

}  // namespace b200pt
