// render_layered.h - CoatedDiffuseBxDF = LayeredBxDF<DielectricBxDF, DiffuseBxDF, twoSided = true>
// (bxdfs.h:565-1041) without a medium between the interfaces (albedo = 0, the default of
// CoatedDiffuseMaterial, materials.cpp:363-376).  f, Sample_f and PDF are the reference's stochastic
// random walks, driven by the reference's own generator (PCG32 seeded from MurmurHash64A of the
// directions, util/rng.h, util/hash.h) so the same walk is taken.
//
// The walk multiplies its throughput only by scalars (dielectric top, transmittance, cosines) and by
// the diffuse base's R_k.  So instead of 31-wide `beta` and `f` it carries beta = S * R^a (a scalar and
// an exponent) and f = sum_a c_a R^a (kLayerPoly coefficients); the 31 values come out of a Horner
// loop at the end (BsdfSpectral kind 5).  Russian roulette inside the walk uses max_k beta_k =
// S * (max_k R_k)^a.
#pragma once
#include "render_bsdf.h"
#include "render_core.h"

namespace b200pt {

// Hash(args...) (util/hash.h:95-103) for the argument packs LayeredBxDF uses: (int seed, Vector3f), (Vector3f),
// (Float, Point2f); Murmur64A and the PCG32 generator are in render_math.h
B2_HD uint64_t HashSeedV3(int32_t seed, V3 v) {
    unsigned char buf[16];
    memcpy(buf, &seed, 4);
    memcpy(buf + 4, &v.x, 4);
    memcpy(buf + 8, &v.y, 4);
    memcpy(buf + 12, &v.z, 4);
    return Murmur64A(buf, 16, 0);
}
B2_HD uint64_t Hash3f(float a, float b, float c) {
    unsigned char buf[12];
    memcpy(buf, &a, 4);
    memcpy(buf + 4, &b, 4);
    memcpy(buf + 8, &c, 4);
    return Murmur64A(buf, 12, 0);
}

// the `r` lambda of LayeredBxDF: min(rng.Uniform<Float>(), OneMinusEpsilon)
B2_HD float PcgFloat(PCG32 &g) { return fminf(kOneMinusEps, PcgNext(g) * 0x1p-32f); }

// FastExp of the CPU build (util/math.h:450-476)
B2_HD float FastExpRef(float x) {
    float xp = x * 1.442695041f;
    float fxp = floorf(xp), f = xp - fxp;
    int i = (int)fxp;
    float twoToF = FMAf(f, FMAf(f, FMAf(f, 0.0781455737f, 0.226173572f), 0.695556856f), 1.f);
    uint32_t bits;
    memcpy(&bits, &twoToF, 4);
    int exponent = (int)(bits >> 23) - 127 + i;
    if (exponent < -126) return 0;
    if (exponent > 127) return INFINITY;
    bits &= 0x807fffffu;
    bits |= uint32_t(exponent + 127) << 23;
    float out;
    memcpy(&out, &bits, 4);
    return out;
}
B2_HD float LayerTr(float dz, V3 w) {  // LayeredBxDF::Tr (bxdfs.h:1016-1021)
    if (fabsf(dz) <= 1.17549435e-38f) return 1;
    return FastExpRef(-fabsf(dz / w.z));
}

// S * R^a with R = rMax
B2_HD float PowI(float r, int a) {
    float v = 1;
    for (int i = 0; i < a; ++i) v *= r;
    return v;
}

// Two random numbers for a Point2f: `Point2f(r(), r())` / `{r(), r()}`.  GCC evaluates constructor
// arguments right to left and braced lists left to right; which form a call site uses is noted there.
#define B2_R2_PAREN(g, u0, u1) \
    do {                       \
        u1 = PcgFloat(g);      \
        u0 = PcgFloat(g);      \
    } while (0)
#define B2_R2_BRACE(g, u0, u1) \
    do {                       \
        u0 = PcgFloat(g);      \
        u1 = PcgFloat(g);      \
    } while (0)

// DiffuseBxDF pieces of the bottom interface (bxdfs.h:37-67), R_k factored out
B2_HD float DiffFScalar(V3 wo, V3 wi) { return SameHemisphere(wo, wi) ? kInvPi : 0.f; }
B2_HD float DiffPDF(V3 wo, V3 wi) { return SameHemisphere(wo, wi) ? AbsCosTheta(wi) * kInvPi : 0.f; }

// LayeredBxDF::f (bxdfs.h:608-790), Radiance mode.  Result: ev.poly (already divided by nSamples).
B2_HD_CALL BsdfEval CoatedF(const Bsdf &b, V3 wo, V3 wi, int32_t seed) {
    BsdfEval ev = ZeroEval();
    for (int i = 0; i < kLayerPoly; ++i) ev.poly[i] = 0;
    if (wo.z < 0) {  // twoSided
        wo = -wo;
        wi = -wi;
    }
    // entered through the top; leaving through the bottom is impossible (the diffuse base does not
    // transmit: every sample of the reference `continue`s), so only the same-hemisphere case has a value
    if (!SameHemisphere(wo, wi)) return ev;
    ev.kind = 5;
    const float thickness = b.thickness;
    const bool topSpecular = TRSmooth(b.tr) || b.eta == 1;
    ev.poly[0] = b.nSamples * DielF(b.eta, b.tr, wo, wi, true);
    PCG32 g = PcgSeed(HashSeedV3(seed, wo), Hash3f(wi.x, wi.y, wi.z));
    for (int s = 0; s < b.nSamples; ++s) {
        float uc = PcgFloat(g), u0, u1, fwos, fwis;
        B2_R2_PAREN(g, u0, u1);
        BsdfSample wos = DielSample(b.eta, b.tr, wo, uc, u0, u1, true, SF_TRANSMISSION, fwos);
        if (!wos.ok || fwos == 0 || wos.pdf == 0 || wos.wi.z == 0) continue;
        uc = PcgFloat(g);
        B2_R2_PAREN(g, u0, u1);
        BsdfSample wis = DielSample(b.eta, b.tr, wi, uc, u0, u1, false, SF_TRANSMISSION, fwis);
        if (!wis.ok || fwis == 0 || wis.pdf == 0 || wis.wi.z == 0) continue;
        float S = fwos * AbsCosTheta(wos.wi) / wos.pdf;
        int a = 0;
        float z = thickness;
        V3 w = wos.wi;
        for (int depth = 0; depth < b.maxDepth; ++depth) {
            float bmax = S * PowI(b.rMax, a);
            if (depth > 3 && bmax < 0.25f) {
                float q = fmaxf(0.f, 1 - bmax);
                if (PcgFloat(g) < q) break;
                S /= 1 - q;
            }
            z = (z == thickness) ? 0 : thickness;
            S *= LayerTr(thickness, w);
            if (z == thickness) {  // exit interface (top): reflect back down
                float fbs;
                uc = PcgFloat(g);
                B2_R2_PAREN(g, u0, u1);
                BsdfSample bs = DielSample(b.eta, b.tr, -w, uc, u0, u1, true, SF_REFLECTION, fbs);
                if (!bs.ok || fbs == 0 || bs.pdf == 0 || bs.wi.z == 0) break;
                S *= fbs * AbsCosTheta(bs.wi) / bs.pdf;
                w = bs.wi;
            } else {  // the diffuse base
                float wt = 1;
                if (!topSpecular) wt = PowerHeuristic(wis.pdf, DiffPDF(-w, -wis.wi));
                if (a + 1 < kLayerPoly)
                    ev.poly[a + 1] += S * DiffFScalar(-w, -wis.wi) * AbsCosTheta(wis.wi) * wt * LayerTr(thickness, wis.wi) *
                                      fwis / wis.pdf;
                uc = PcgFloat(g);
                B2_R2_PAREN(g, u0, u1);
                V3 wo2 = -w;
                V3 bwi = SampleCosineHemisphere(u0, u1);
                if (wo2.z < 0) bwi.z *= -1;
                float bpdf = AbsCosTheta(bwi) * kInvPi;
                if (b.rMax == 0 || bpdf == 0 || bwi.z == 0) break;
                S *= kInvPi * AbsCosTheta(bwi) / bpdf;
                a += 1;
                w = bwi;
                if (!topSpecular) {
                    float fExit = DielF(b.eta, b.tr, -w, wi, true);
                    if (fExit != 0) {
                        float exitPDF = DielPDF(b.eta, b.tr, -w, wi, SF_TRANSMISSION);
                        float wt2 = PowerHeuristic(bpdf, exitPDF);
                        if (a < kLayerPoly) ev.poly[a] += S * LayerTr(thickness, bwi) * fExit * wt2;
                    }
                }
            }
        }
    }
    for (int i = 0; i < kLayerPoly; ++i) ev.poly[i] /= b.nSamples;
    return ev;
}

// LayeredBxDF::Sample_f (bxdfs.h:792-905), Radiance mode, all lobes.
B2_HD_CALL BsdfSample CoatedSample(const Bsdf &b, V3 wo, float uc, float u0, float u1, int32_t seed, BsdfEval &ev) {
    BsdfSample out{false, mk(0, 0, 0), 0, 0, 1, false};
    ev = ZeroEval();
    for (int i = 0; i < kLayerPoly; ++i) ev.poly[i] = 0;
    bool flipWi = false;
    if (wo.z < 0) {
        wo = -wo;
        flipWi = true;
    }
    float fbs;
    BsdfSample bs = DielSample(b.eta, b.tr, wo, uc, u0, u1, true, SF_ALL, fbs);
    if (!bs.ok || fbs == 0 || bs.pdf == 0 || bs.wi.z == 0) return out;
    if (bs.flags & BX_REFLECTION) {
        if (flipWi) bs.wi = -bs.wi;
        bs.pdfIsProportional = true;
        ev.kind = 1;
        ev.scale = fbs;
        return bs;
    }
    V3 w = bs.wi;
    bool specularPath = (bs.flags & BX_SPECULAR) != 0;
    PCG32 g = PcgSeed(HashSeedV3(seed, wo), Hash3f(uc, u0, u1));
    float S = fbs * AbsCosTheta(bs.wi), pdf = bs.pdf;
    int a = 0;
    const float thickness = b.thickness;
    float z = thickness;
    for (int depth = 0; depth < b.maxDepth; ++depth) {
        float rrBeta = S * PowI(b.rMax, a) / pdf;
        if (depth > 3 && rrBeta < 0.25f) {
            float q = fmaxf(0.f, 1 - rrBeta);
            if (PcgFloat(g) < q) return out;
            pdf *= 1 - q;
        }
        if (w.z == 0) return out;
        z = (z == thickness) ? 0 : thickness;
        S *= LayerTr(thickness, w);
        float uc2 = PcgFloat(g), v0, v1;
        B2_R2_PAREN(g, v0, v1);
        V3 nwi;
        if (z == 0) {  // diffuse base
            V3 wo2 = -w;
            nwi = SampleCosineHemisphere(v0, v1);
            if (wo2.z < 0) nwi.z *= -1;
            float bpdf = AbsCosTheta(nwi) * kInvPi;
            if (b.rMax == 0 || bpdf == 0 || nwi.z == 0) return out;
            S *= kInvPi;
            a += 1;
            pdf *= bpdf;
            specularPath = false;
            w = nwi;
        } else {
            float f2;
            BsdfSample b2 = DielSample(b.eta, b.tr, -w, uc2, v0, v1, true, SF_ALL, f2);
            if (!b2.ok || f2 == 0 || b2.pdf == 0 || b2.wi.z == 0) return out;
            S *= f2;
            pdf *= b2.pdf;
            specularPath = specularPath && (b2.flags & BX_SPECULAR) != 0;
            w = b2.wi;
            nwi = b2.wi;
            if (b2.flags & BX_TRANSMISSION) {
                out.flags = (SameHemisphere(wo, w) ? BX_REFLECTION : BX_TRANSMISSION) |
                            (specularPath ? BX_SPECULAR : BX_GLOSSY);
                if (flipWi) w = -w;
                out.ok = true;
                out.wi = w;
                out.pdf = pdf;
                out.eta = 1.f;
                out.pdfIsProportional = true;
                ev.kind = 5;
                if (a < kLayerPoly) ev.poly[a] = S;
                return out;
            }
        }
        S *= AbsCosTheta(nwi);
    }
    return out;
}

// LayeredBxDF::PDF (bxdfs.h:907-1011), all lobes.
B2_HD_CALL float CoatedPDF(const Bsdf &b, V3 wo, V3 wi, int32_t seed) {
    if (wo.z < 0) {
        wo = -wo;
        wi = -wi;
    }
    PCG32 g = PcgSeed(HashSeedV3(seed, wi), Hash3f(wo.x, wo.y, wo.z));
    float pdfSum = 0;
    const bool same = SameHemisphere(wo, wi);
    const bool topSpecular = TRSmooth(b.tr) || b.eta == 1;
    if (same) pdfSum += b.nSamples * DielPDF(b.eta, b.tr, wo, wi, SF_REFLECTION);
    for (int s = 0; s < b.nSamples && same; ++s) {
        // TRT term.  `Sample_f(wo, r(), {r(), r()}, ...)`: arguments right to left, braces left to right
        float uc, u0, u1, fwos, fwis;
        B2_R2_BRACE(g, u0, u1);
        uc = PcgFloat(g);
        BsdfSample wos = DielSample(b.eta, b.tr, wo, uc, u0, u1, true, SF_TRANSMISSION, fwos);
        B2_R2_BRACE(g, u0, u1);
        uc = PcgFloat(g);
        BsdfSample wis = DielSample(b.eta, b.tr, wi, uc, u0, u1, false, SF_TRANSMISSION, fwis);
        if (wos.ok && fwos != 0 && wos.pdf > 0 && wis.ok && fwis != 0 && wis.pdf > 0) {
            if (topSpecular) {
                pdfSum += DiffPDF(-wos.wi, -wis.wi);
            } else {
                B2_R2_BRACE(g, u0, u1);
                (void)PcgFloat(g);
                V3 wo2 = -wos.wi;
                V3 rwi = SampleCosineHemisphere(u0, u1);
                if (wo2.z < 0) rwi.z *= -1;
                float rpdf = AbsCosTheta(rwi) * kInvPi;
                // BSDFSample validity as the reference tests it: rs && rs->f && rs->pdf > 0
                if (b.rMax != 0 && rpdf > 0) {
                    float rPDF = DiffPDF(-wos.wi, -wis.wi);
                    float wt = PowerHeuristic(wis.pdf, rPDF);
                    pdfSum += wt * rPDF;
                    float tPDF = DielPDF(b.eta, b.tr, -rwi, wi, SF_ALL);
                    wt = PowerHeuristic(rpdf, tPDF);
                    pdfSum += wt * tPDF;
                }
            }
        }
    }
    // different hemispheres: the TT term needs a transmission through the diffuse base: always 0
    return Lerp(0.9f, 1 / (4 * kPi), pdfSum / b.nSamples);
}

// ---------------------------------------------------------------------------------------------------------------
// CoatedConductorBxDF = LayeredBxDF<DielectricBxDF, ConductorBxDF, true> (bxdfs.h:1298; CoatedConductorMaterial,
// materials.cpp:392-438), no medium between the interfaces.  Every bounce off the base multiplies the walk's
// throughput by the conductor's Fresnel spectrum at that bounce's angle, so the throughput is a scalar S times a
// per-wavelength product `prod[k]` (no polynomial here), and the value comes back as 31 numbers (BsdfEval kind 8).
// The conductor's index is relative to the coating's: eta_k / ieta, k_k / ieta (materials.cpp:421-422).
struct CondSpectra {
    float eta[kNLambda], k[kNLambda];
};
B2_HD void CoatedCondSpectra(const RenderSceneDev &rs, const Bsdf &b, float lambdaU, CondSpectra &c) {
    LambdaIter li(lambdaU);
    for (int i = 0; i < kNLambda; ++i, li.Next()) {
        float e, kk;
        if (b.specB >= 0) {
            e = SpectrumAt(rs.spectra[b.specA], li.lam);
            kk = SpectrumAt(rs.spectra[b.specB], li.lam);
        } else {  // reflectance given (materials.cpp:416-419)
            float r = Clampf(SpectrumAt(rs.spectra[b.specA], li.lam), 0, .9999f);
            e = 1.f;
            kk = 2 * sqrtf(r) / sqrtf(fmaxf(0.f, 1 - r));
        }
        c.eta[i] = e / b.eta;
        c.k[i] = kk / b.eta;
    }
}
// the base as a Bsdf of its own, and one wavelength of a scalar evaluation of it (BsdfSpectral kinds 3 / 4)
B2_HD Bsdf CoatedCondBase(const Bsdf &b) {
    Bsdf c = b;
    c.type = MAT_CONDUCTOR;
    c.tr = b.condTr;
    return c;
}
B2_HD float CondValue(const BsdfEval &e, float eta, float k) {
    float F = FrComplex(e.cosF, eta, k);
    return e.kind == 3 ? (e.D * F) * e.G / e.denom : F / e.denom;
}
B2_HD float MaxProd(const float *prod) {
    float m = prod[0];
    for (int i = 1; i < kNLambda; ++i) m = fmaxf(m, prod[i]);
    return m;
}

// LayeredBxDF::f (bxdfs.h:608-790), Radiance mode.  Result: spec[k] = f at wavelength k (kNLambda floats).
B2_HD_CALL BsdfEval CoatedCondF(const RenderSceneDev &rs, const Bsdf &b, float lambdaU, V3 wo, V3 wi, int32_t seed, float *spec) {
    BsdfEval ev = ZeroEval();
    for (int i = 0; i < kNLambda; ++i) spec[i] = 0;
    if (wo.z < 0) {  // twoSided
        wo = -wo;
        wi = -wi;
    }
    // through the base there is no way out (a conductor does not transmit: the exit-interface sample fails)
    if (!SameHemisphere(wo, wi)) return ev;
    ev.kind = 8;
    CondSpectra cs;
    CoatedCondSpectra(rs, b, lambdaU, cs);
    const Bsdf base = CoatedCondBase(b);
    const float thickness = b.thickness;
    const bool topSpecular = TRSmooth(b.tr) || b.eta == 1, baseSpecular = TRSmooth(b.condTr);
    const float fTop = b.nSamples * DielF(b.eta, b.tr, wo, wi, true);
    float prod[kNLambda];
    PCG32 g = PcgSeed(HashSeedV3(seed, wo), Hash3f(wi.x, wi.y, wi.z));
    for (int s = 0; s < b.nSamples; ++s) {
        float uc = PcgFloat(g), u0, u1, fwos, fwis;
        B2_R2_PAREN(g, u0, u1);
        BsdfSample wos = DielSample(b.eta, b.tr, wo, uc, u0, u1, true, SF_TRANSMISSION, fwos);
        if (!wos.ok || fwos == 0 || wos.pdf == 0 || wos.wi.z == 0) continue;
        uc = PcgFloat(g);
        B2_R2_PAREN(g, u0, u1);
        BsdfSample wis = DielSample(b.eta, b.tr, wi, uc, u0, u1, false, SF_TRANSMISSION, fwis);
        if (!wis.ok || fwis == 0 || wis.pdf == 0 || wis.wi.z == 0) continue;
        float S = fwos * AbsCosTheta(wos.wi) / wos.pdf;
        for (int i = 0; i < kNLambda; ++i) prod[i] = 1;
        float z = thickness;
        V3 w = wos.wi;
        for (int depth = 0; depth < b.maxDepth; ++depth) {
            float bmax = S * MaxProd(prod);
            if (depth > 3 && bmax < 0.25f) {
                float q = fmaxf(0.f, 1 - bmax);
                if (PcgFloat(g) < q) break;
                S /= 1 - q;
            }
            z = (z == thickness) ? 0 : thickness;
            S *= LayerTr(thickness, w);
            if (z == thickness) {  // exit interface (top): reflect back down
                float fbs;
                uc = PcgFloat(g);
                B2_R2_PAREN(g, u0, u1);
                BsdfSample bs = DielSample(b.eta, b.tr, -w, uc, u0, u1, true, SF_REFLECTION, fbs);
                if (!bs.ok || fbs == 0 || bs.pdf == 0 || bs.wi.z == 0) break;
                S *= fbs * AbsCosTheta(bs.wi) / bs.pdf;
                w = bs.wi;
            } else {  // the conductor
                if (!baseSpecular) {  // light along the presampled direction
                    float wt = 1;
                    if (!topSpecular) wt = PowerHeuristic(wis.pdf, BsdfPDF(base, -w, -wis.wi));
                    BsdfEval ce = BsdfF(base, -w, -wis.wi);
                    if (ce.kind != 0) {
                        float c = S * AbsCosTheta(wis.wi) * wt * LayerTr(thickness, wis.wi) * fwis / wis.pdf;
                        for (int i = 0; i < kNLambda; ++i) spec[i] += c * prod[i] * CondValue(ce, cs.eta[i], cs.k[i]);
                    }
                }
                uc = PcgFloat(g);
                B2_R2_PAREN(g, u0, u1);
                BsdfEval se;
                BsdfSample bs = BsdfSampleF(base, -w, uc, u0, u1, se);
                if (!bs.ok || se.kind == 0 || bs.pdf == 0 || bs.wi.z == 0) break;
                S *= AbsCosTheta(bs.wi) / bs.pdf;
                for (int i = 0; i < kNLambda; ++i) prod[i] *= CondValue(se, cs.eta[i], cs.k[i]);
                w = bs.wi;
                if (!topSpecular) {
                    float fExit = DielF(b.eta, b.tr, -w, wi, true);
                    if (fExit != 0) {
                        float wt = 1;
                        if (!baseSpecular) wt = PowerHeuristic(bs.pdf, DielPDF(b.eta, b.tr, -w, wi, SF_TRANSMISSION));
                        float c = S * LayerTr(thickness, bs.wi) * fExit * wt;
                        for (int i = 0; i < kNLambda; ++i) spec[i] += c * prod[i];
                    }
                }
            }
        }
    }
    for (int i = 0; i < kNLambda; ++i) spec[i] = (spec[i] + fTop) / b.nSamples;
    return ev;
}

// LayeredBxDF::Sample_f (bxdfs.h:792-905), Radiance mode, all lobes.
B2_HD_CALL BsdfSample CoatedCondSample(const RenderSceneDev &rs, const Bsdf &b, float lambdaU, V3 wo, float uc, float u0,
                                       float u1, int32_t seed, BsdfEval &ev, float *spec) {
    BsdfSample out{false, mk(0, 0, 0), 0, 0, 1, false};
    ev = ZeroEval();
    bool flipWi = false;
    if (wo.z < 0) {
        wo = -wo;
        flipWi = true;
    }
    float fbs;
    BsdfSample bs = DielSample(b.eta, b.tr, wo, uc, u0, u1, true, SF_ALL, fbs);
    if (!bs.ok || fbs == 0 || bs.pdf == 0 || bs.wi.z == 0) return out;
    if (bs.flags & BX_REFLECTION) {
        if (flipWi) bs.wi = -bs.wi;
        bs.pdfIsProportional = true;
        ev.kind = 1;
        ev.scale = fbs;
        return bs;
    }
    CondSpectra cs;
    CoatedCondSpectra(rs, b, lambdaU, cs);
    const Bsdf base = CoatedCondBase(b);
    V3 w = bs.wi;
    bool specularPath = (bs.flags & BX_SPECULAR) != 0;
    PCG32 g = PcgSeed(HashSeedV3(seed, wo), Hash3f(uc, u0, u1));
    float S = fbs * AbsCosTheta(bs.wi), pdf = bs.pdf;
    float prod[kNLambda];
    for (int i = 0; i < kNLambda; ++i) prod[i] = 1;
    const float thickness = b.thickness;
    float z = thickness;
    for (int depth = 0; depth < b.maxDepth; ++depth) {
        float rrBeta = S * MaxProd(prod) / pdf;
        if (depth > 3 && rrBeta < 0.25f) {
            float q = fmaxf(0.f, 1 - rrBeta);
            if (PcgFloat(g) < q) return out;
            pdf *= 1 - q;
        }
        if (w.z == 0) return out;
        z = (z == thickness) ? 0 : thickness;
        S *= LayerTr(thickness, w);
        float uc2 = PcgFloat(g), v0, v1;
        B2_R2_PAREN(g, v0, v1);
        V3 nwi;
        if (z == 0) {  // the conductor
            BsdfEval se;
            BsdfSample b2 = BsdfSampleF(base, -w, uc2, v0, v1, se);
            if (!b2.ok || se.kind == 0 || b2.pdf == 0 || b2.wi.z == 0) return out;
            for (int i = 0; i < kNLambda; ++i) prod[i] *= CondValue(se, cs.eta[i], cs.k[i]);
            pdf *= b2.pdf;
            specularPath = specularPath && (b2.flags & BX_SPECULAR) != 0;
            w = b2.wi;
            nwi = b2.wi;
        } else {
            float f2;
            BsdfSample b2 = DielSample(b.eta, b.tr, -w, uc2, v0, v1, true, SF_ALL, f2);
            if (!b2.ok || f2 == 0 || b2.pdf == 0 || b2.wi.z == 0) return out;
            S *= f2;
            pdf *= b2.pdf;
            specularPath = specularPath && (b2.flags & BX_SPECULAR) != 0;
            w = b2.wi;
            nwi = b2.wi;
            if (b2.flags & BX_TRANSMISSION) {
                out.flags = (SameHemisphere(wo, w) ? BX_REFLECTION : BX_TRANSMISSION) |
                            (specularPath ? BX_SPECULAR : BX_GLOSSY);
                if (flipWi) w = -w;
                out.ok = true;
                out.wi = w;
                out.pdf = pdf;
                out.eta = 1.f;
                out.pdfIsProportional = true;
                ev.kind = 8;
                for (int i = 0; i < kNLambda; ++i) spec[i] = S * prod[i];
                return out;
            }
        }
        S *= AbsCosTheta(nwi);
    }
    return out;
}

// LayeredBxDF::PDF (bxdfs.h:907-1011), all lobes.
B2_HD_CALL float CoatedCondPDF(const Bsdf &b, V3 wo, V3 wi, int32_t seed) {
    if (wo.z < 0) {
        wo = -wo;
        wi = -wi;
    }
    PCG32 g = PcgSeed(HashSeedV3(seed, wi), Hash3f(wo.x, wo.y, wo.z));
    const Bsdf base = CoatedCondBase(b);
    float pdfSum = 0;
    const bool same = SameHemisphere(wo, wi);
    const bool topSpecular = TRSmooth(b.tr) || b.eta == 1, baseSpecular = TRSmooth(b.condTr);
    if (same) pdfSum += b.nSamples * DielPDF(b.eta, b.tr, wo, wi, SF_REFLECTION);
    for (int s = 0; s < b.nSamples && same; ++s) {
        // TRT term.  `Sample_f(wo, r(), {r(), r()}, ...)`: arguments right to left, braces left to right
        float uc, u0, u1, fwos, fwis;
        B2_R2_BRACE(g, u0, u1);
        uc = PcgFloat(g);
        BsdfSample wos = DielSample(b.eta, b.tr, wo, uc, u0, u1, true, SF_TRANSMISSION, fwos);
        B2_R2_BRACE(g, u0, u1);
        uc = PcgFloat(g);
        BsdfSample wis = DielSample(b.eta, b.tr, wi, uc, u0, u1, false, SF_TRANSMISSION, fwis);
        if (wos.ok && fwos != 0 && wos.pdf > 0 && wis.ok && fwis != 0 && wis.pdf > 0) {
            if (topSpecular) {
                pdfSum += BsdfPDF(base, -wos.wi, -wis.wi);
            } else {
                B2_R2_BRACE(g, u0, u1);
                uc = PcgFloat(g);
                BsdfEval se;
                BsdfSample rs = BsdfSampleF(base, -wos.wi, uc, u0, u1, se);
                if (rs.ok && se.kind != 0 && rs.pdf > 0) {
                    if (baseSpecular) {
                        pdfSum += DielPDF(b.eta, b.tr, -rs.wi, wi, SF_ALL);
                    } else {
                        float rPDF = BsdfPDF(base, -wos.wi, -wis.wi);
                        float wt = PowerHeuristic(wis.pdf, rPDF);
                        pdfSum += wt * rPDF;
                        float tPDF = DielPDF(b.eta, b.tr, -rs.wi, wi, SF_ALL);
                        wt = PowerHeuristic(rs.pdf, tPDF);
                        pdfSum += wt * tPDF;
                    }
                }
            }
        }
    }
    // different hemispheres: the TT term needs a transmission through the conductor: always 0
    return Lerp(0.9f, 1 / (4 * kPi), pdfSum / b.nSamples);
}

}  // namespace b200pt
