// render_bsdf.h - BxDFs of the shading kernels in local shading space (z = shading normal),
// host+device.  Each model is split into a scalar part, evaluated once per direction pair, and a
// per-wavelength factor, so the 31-wide loops of the kernels stay simple streams:
//     f_k(wo, wi) = Spectral(k) applied to the scalar BsdfEval.
// Arithmetic order follows the reference expressions (bxdfs.h / bxdfs.cpp / util/scattering.h).
#pragma once
#include "render_math.h"
#include "render_types.h"

namespace b200pt {

// ---- trigonometry in the shading frame (util/vecmath.h:1685-1727)
B2_HD float CosTheta(V3 w) { return w.z; }
B2_HD float Cos2Theta(V3 w) { return SqrF(w.z); }
B2_HD float AbsCosTheta(V3 w) { return fabsf(w.z); }
B2_HD float Sin2Theta(V3 w) { return fmaxf(0.f, 1 - Cos2Theta(w)); }
B2_HD float SinTheta(V3 w) { return sqrtf(Sin2Theta(w)); }
B2_HD float Tan2Theta(V3 w) { return Sin2Theta(w) / Cos2Theta(w); }
B2_HD float CosPhi(V3 w) {
    float s = SinTheta(w);
    return (s == 0) ? 1 : Clampf(w.x / s, -1, 1);
}
B2_HD float SinPhi(V3 w) {
    float s = SinTheta(w);
    return (s == 0) ? 0 : Clampf(w.y / s, -1, 1);
}
B2_HD bool SameHemisphere(V3 a, V3 b) { return a.z * b.z > 0; }
B2_HD V3 Reflect(V3 wo, V3 n) { return -wo + (2 * Dot(wo, n)) * n; }  // util/scattering.h:18-20

// Refract (util/scattering.h:22-47)
B2_HD bool Refract(V3 wi, V3 n, float eta, float *etap, V3 *wt) {
    float cosTheta_i = DotN(n, wi);  // n is a Normal3f there
    if (cosTheta_i < 0) {
        eta = 1 / eta;
        cosTheta_i = -cosTheta_i;
        n = -n;
    }
    float sin2Theta_i = fmaxf(0.f, 1 - SqrF(cosTheta_i));
    float sin2Theta_t = sin2Theta_i / SqrF(eta);
    if (sin2Theta_t >= 1) return false;
    float cosTheta_t = sqrtf(1 - sin2Theta_t);
    *wt = -wi / eta + (cosTheta_i / eta - cosTheta_t) * n;
    if (etap) *etap = eta;
    return true;
}

// FrDielectric (util/scattering.h:61-79)
B2_HD float FrDielectric(float cosTheta_i, float eta) {
    cosTheta_i = Clampf(cosTheta_i, -1, 1);
    if (cosTheta_i < 0) {
        eta = 1 / eta;
        cosTheta_i = -cosTheta_i;
    }
    float sin2Theta_i = 1 - SqrF(cosTheta_i);
    float sin2Theta_t = sin2Theta_i / SqrF(eta);
    if (sin2Theta_t >= 1) return 1.f;
    float cosTheta_t = SafeSqrt(1 - sin2Theta_t);
    float r_parl = (eta * cosTheta_i - cosTheta_t) / (eta * cosTheta_i + cosTheta_t);
    float r_perp = (cosTheta_i - eta * cosTheta_t) / (cosTheta_i + eta * cosTheta_t);
    return (SqrF(r_parl) + SqrF(r_perp)) / 2;
}

// pstd::complex<float> as the reference defines it (util/pstd.h:1066-1229)
struct Cx {
    float re, im;
};
B2_HD Cx cx(float re, float im) { return Cx{re, im}; }
B2_HD Cx operator+(Cx a, Cx b) { return cx(a.re + b.re, a.im + b.im); }
B2_HD Cx operator-(Cx a, Cx b) { return cx(a.re - b.re, a.im - b.im); }
B2_HD Cx operator*(Cx a, Cx b) { return cx(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
B2_HD Cx operator/(Cx a, Cx z) {
    float scale = 1 / (z.re * z.re + z.im * z.im);
    return cx(scale * (a.re * z.re + a.im * z.im), scale * (a.im * z.re - a.re * z.im));
}
B2_HD float CxNorm(Cx z) { return z.re * z.re + z.im * z.im; }
B2_HD Cx CxSqrt(Cx z) {
    float n = sqrtf(CxNorm(z)), t1 = sqrtf(.5f * (n + fabsf(z.re))), t2 = .5f * z.im / t1;
    if (n == 0) return cx(0, 0);
    if (z.re >= 0) return cx(t1, t2);
    return cx(fabsf(t2), copysignf(t1, z.im));
}
// FrComplex (util/scattering.h:81-92)
B2_HD float FrComplex(float cosTheta_i, float etaRe, float etaIm) {
    Cx eta = cx(etaRe, etaIm);
    cosTheta_i = Clampf(cosTheta_i, 0, 1);
    float sin2Theta_i = 1 - SqrF(cosTheta_i);
    Cx sin2Theta_t = cx(sin2Theta_i, 0) / (eta * eta);
    Cx cosTheta_t = CxSqrt(cx(1, 0) - sin2Theta_t);
    Cx ci = cx(cosTheta_i, 0);
    Cx r_parl = (eta * ci - cosTheta_t) / (eta * ci + cosTheta_t);
    Cx r_perp = (ci - eta * cosTheta_t) / (ci + eta * cosTheta_t);
    return (CxNorm(r_parl) + CxNorm(r_perp)) / 2;
}

// ---- TrowbridgeReitzDistribution (util/scattering.h:109-217)
struct TRDist {
    float ax, ay;
};
B2_HD bool TRSmooth(const TRDist &d) { return fmaxf(d.ax, d.ay) < 1e-3f; }
B2_HD TRDist MakeTR(float ax, float ay) {
    TRDist d{ax, ay};
    if (!TRSmooth(d)) {
        d.ax = fmaxf(d.ax, 1e-4f);
        d.ay = fmaxf(d.ay, 1e-4f);
    }
    return d;
}
B2_HD void TRRegularize(TRDist &d) {
    if (d.ax < 0.3f) d.ax = Clampf(2 * d.ax, 0.1f, 0.3f);
    if (d.ay < 0.3f) d.ay = Clampf(2 * d.ay, 0.1f, 0.3f);
}
B2_HD float TR_D(const TRDist &d, V3 wm) {
    float tan2Theta = Tan2Theta(wm);
    if (std::isinf(tan2Theta)) return 0;
    float cos4Theta = SqrF(Cos2Theta(wm));
    if (cos4Theta < 1e-16f) return 0;
    float e = tan2Theta * (SqrF(CosPhi(wm) / d.ax) + SqrF(SinPhi(wm) / d.ay));
    return 1 / (kPi * d.ax * d.ay * cos4Theta * SqrF(1 + e));
}
B2_HD float TR_Lambda(const TRDist &d, V3 w) {
    float tan2Theta = Tan2Theta(w);
    if (std::isinf(tan2Theta)) return 0;
    float alpha2 = SqrF(CosPhi(w) * d.ax) + SqrF(SinPhi(w) * d.ay);
    return (sqrtf(1 + alpha2 * tan2Theta) - 1) / 2;
}
B2_HD float TR_G1(const TRDist &d, V3 w) { return 1 / (1 + TR_Lambda(d, w)); }
B2_HD float TR_G(const TRDist &d, V3 wo, V3 wi) { return 1 / (1 + TR_Lambda(d, wo) + TR_Lambda(d, wi)); }
B2_HD float TR_PDF(const TRDist &d, V3 w, V3 wm) {  // D(w, wm)
    return TR_G1(d, w) / AbsCosTheta(w) * TR_D(d, wm) * AbsDot(w, wm);
}
B2_HD V3 TR_Sample_wm(const TRDist &d, V3 w, float u0, float u1) {
    V3 wh = Normalize(mk(d.ax * w.x, d.ay * w.y, w.z));
    if (wh.z < 0) wh = -wh;
    V3 T1 = (wh.z < 0.99999f) ? Normalize(Cross(mk(0, 0, 1), wh)) : mk(1, 0, 0);
    V3 T2 = Cross(wh, T1);
    // SampleUniformDiskPolar (util/sampling.h:311-315)
    float r = sqrtf(u0), theta = 2 * kPi * u1;
    float px = r * cosf(theta), py = r * sinf(theta);
    float h = sqrtf(1 - SqrF(px));
    py = Lerp((1 + wh.z) / 2, h, py);
    float pz = sqrtf(fmaxf(0.f, 1 - (SqrF(px) + SqrF(py))));
    V3 nh = px * T1 + py * T2 + pz * wh;
    return Normalize(mk(d.ax * nh.x, d.ay * nh.y, fmaxf(1e-6f, nh.z)));
}

// ---- BxDF flags (base/bxdf.h:33-48)
enum : uint32_t {
    BX_REFLECTION = 1, BX_TRANSMISSION = 2, BX_DIFFUSE = 4, BX_GLOSSY = 8, BX_SPECULAR = 16
};

// What the integrator needs back from Sample_f besides the per-wavelength f.
struct BsdfSample {
    bool ok;
    V3 wi;            // local
    float pdf;
    uint32_t flags;   // BX_* of the sampled lobe
    float eta;        // etap for transmission, 1 otherwise
    bool pdfIsProportional;
};

// Scalar part of an f(wo, wi) evaluation; the spectral factor is applied per wavelength by
// BsdfSpectral().  kind: 0 zero, 1 constant*`scale` (same for all k), 2 diffuse R_k*scale,
// 3 conductor ((D*F_k)*G)/denom with F_k = FrComplex(cosF, eta_k, k_k), 4 specular conductor F_k/denom.
// 5 coated diffuse: sum_a poly[a] * R_k^a (the random walk of LayeredBxDF only ever multiplies the
// throughput by scalars and by the diffuse base's R_k, so its value is a polynomial in R_k).
// 8 coated conductor: the value at each wavelength is in the caller's array of kNLambda floats (render_layered.h).
// 6 / 7 diffuse transmission: clamp(dscale*R_k)*scale on wo's side, clamp(dscale*T_k)*scale on the other.
constexpr int kLayerPoly = 14;
struct BsdfEval {
    int kind;
    float scale, D, G, denom, cosF;
    float poly[kLayerPoly];
};

// One surface's BxDF, as GetBxDF leaves it (materials.h): the spectra stay in their tables and are
// looked up per wavelength when a spectral factor is needed (keeps 31- and 62-float members out of
// registers).
struct Bsdf {
    uint32_t type;      // MaterialType
    TRDist tr;          // conductor, dielectric
    float eta;          // dielectric: eta(lambda[0])
    uint32_t thin;      // dielectric: ThinDielectricBxDF (bxdfs.h:342-410) instead of DielectricBxDF
    uint32_t trans;     // diffuse: DiffuseTransmissionBxDF (bxdfs.h:218-297): specA = reflectance, specB = transmittance,
    float dscale, tMax; //   both times dscale and clamped to [0, 1]; rMax / tMax = their maxima at the path's wavelengths
    int32_t specA;      // diffuse, coated diffuse: reflectance; conductor: eta, or reflectance when specB < 0
    int32_t specB;      // conductor: k
    // coated diffuse (LayeredBxDF<DielectricBxDF, DiffuseBxDF, true>, bxdfs.h:565-1041)
    float thickness, rMax;  // rMax = max_k R_k at this path's wavelengths
    int maxDepth, nSamples;
    int32_t layerSeed;      // Options::seed, hashed into the random walks' generators
    // coated conductor (LayeredBxDF<DielectricBxDF, ConductorBxDF, true>): the base's microfacet distribution; its eta /
    // k (or reflectance) are specA / specB as for a conductor, the coating's index is `eta`
    uint32_t coatCond;
    TRDist condTr;
};

// ---- DielectricBxDF with explicit TransportMode and BxDFReflTransFlags (bxdfs.cpp:77-255): the coated
// materials call their top interface with Importance mode and with one of the two lobes masked out.
enum : uint32_t { SF_REFLECTION = 1, SF_TRANSMISSION = 2, SF_ALL = 3 };

B2_HD_CALL float DielF(float eta, TRDist tr, V3 wo, V3 wi, bool radiance) {
    if (eta == 1 || TRSmooth(tr)) return 0;
    float cosTheta_o = CosTheta(wo), cosTheta_i = CosTheta(wi);
    bool reflect = cosTheta_i * cosTheta_o > 0;
    float etap = 1;
    if (!reflect) etap = cosTheta_o > 0 ? eta : (1 / eta);
    V3 wm = wi * etap + wo;
    if (cosTheta_i == 0 || cosTheta_o == 0 || LengthSquared(wm) == 0) return 0;
    wm = FaceForward(Normalize(wm), mk(0, 0, 1));
    if (Dot(wm, wi) * cosTheta_i < 0 || Dot(wm, wo) * cosTheta_o < 0) return 0;
    float F = FrDielectric(Dot(wo, wm), eta);
    if (reflect) return TR_D(tr, wm) * TR_G(tr, wo, wi) * F / fabsf(4 * cosTheta_i * cosTheta_o);
    float denom = SqrF(Dot(wi, wm) + Dot(wo, wm) / etap) * cosTheta_i * cosTheta_o;
    float ft = TR_D(tr, wm) * (1 - F) * TR_G(tr, wo, wi) * fabsf(Dot(wi, wm) * Dot(wo, wm) / denom);
    if (radiance) ft /= SqrF(etap);
    return ft;
}

B2_HD_CALL float DielPDF(float eta, TRDist tr, V3 wo, V3 wi, uint32_t sflags) {
    if (eta == 1 || TRSmooth(tr)) return 0;
    float cosTheta_o = CosTheta(wo), cosTheta_i = CosTheta(wi);
    bool reflect = cosTheta_i * cosTheta_o > 0;
    float etap = 1;
    if (!reflect) etap = cosTheta_o > 0 ? eta : (1 / eta);
    V3 wm = wi * etap + wo;
    if (cosTheta_i == 0 || cosTheta_o == 0 || LengthSquared(wm) == 0) return 0;
    wm = FaceForward(Normalize(wm), mk(0, 0, 1));
    if (Dot(wm, wi) * cosTheta_i < 0 || Dot(wm, wo) * cosTheta_o < 0) return 0;
    float R = FrDielectric(Dot(wo, wm), eta), T = 1 - R;
    float pr = (sflags & SF_REFLECTION) ? R : 0, pt = (sflags & SF_TRANSMISSION) ? T : 0;
    if (pr == 0 && pt == 0) return 0;
    if (reflect) return TR_PDF(tr, wo, wm) / (4 * AbsDot(wo, wm)) * pr / (pr + pt);
    float denom = SqrF(Dot(wi, wm) + Dot(wo, wm) / etap);
    float dwm_dwi = AbsDot(wi, wm) / denom;
    return TR_PDF(tr, wo, wm) * dwm_dwi * pt / (pr + pt);
}

// returns the sample; `f` is the (wavelength-independent) BSDF value
B2_HD_CALL BsdfSample DielSample(float eta, TRDist tr, V3 wo, float uc, float u0, float u1, bool radiance,
                            uint32_t sflags, float &f) {
    BsdfSample s{false, mk(0, 0, 0), 0, 0, 1, false};
    f = 0;
    if (eta == 1 || TRSmooth(tr)) {
        float R = FrDielectric(CosTheta(wo), eta), T = 1 - R;
        float pr = (sflags & SF_REFLECTION) ? R : 0, pt = (sflags & SF_TRANSMISSION) ? T : 0;
        if (pr == 0 && pt == 0) return s;
        if (uc < pr / (pr + pt)) {
            V3 wi = mk(-wo.x, -wo.y, wo.z);
            f = R / AbsCosTheta(wi);
            s.wi = wi;
            s.pdf = pr / (pr + pt);
            s.flags = BX_SPECULAR | BX_REFLECTION;
            s.ok = true;
            return s;
        }
        V3 wi;
        float etap;
        if (!Refract(wo, mk(0, 0, 1), eta, &etap, &wi)) return s;
        float ft = T / AbsCosTheta(wi);
        if (radiance) ft /= SqrF(etap);
        f = ft;
        s.wi = wi;
        s.pdf = pt / (pr + pt);
        s.flags = BX_SPECULAR | BX_TRANSMISSION;
        s.eta = etap;
        s.ok = true;
        return s;
    }
    V3 wm = TR_Sample_wm(tr, wo, u0, u1);
    float R = FrDielectric(Dot(wo, wm), eta), T = 1 - R;
    float pr = (sflags & SF_REFLECTION) ? R : 0, pt = (sflags & SF_TRANSMISSION) ? T : 0;
    if (pr == 0 && pt == 0) return s;
    if (uc < pr / (pr + pt)) {
        V3 wi = Reflect(wo, wm);
        if (!SameHemisphere(wo, wi)) return s;
        s.pdf = TR_PDF(tr, wo, wm) / (4 * AbsDot(wo, wm)) * pr / (pr + pt);
        f = TR_D(tr, wm) * TR_G(tr, wo, wi) * R / (4 * CosTheta(wi) * CosTheta(wo));
        s.wi = wi;
        s.flags = BX_GLOSSY | BX_REFLECTION;
        s.ok = true;
        return s;
    }
    float etap;
    V3 wi = mk(0, 0, 0);
    bool tir = !Refract(wo, wm, eta, &etap, &wi);
    if (tir || SameHemisphere(wo, wi) || wi.z == 0) return s;
    float denom = SqrF(Dot(wi, wm) + Dot(wo, wm) / etap);
    float dwm_dwi = AbsDot(wi, wm) / denom;
    s.pdf = TR_PDF(tr, wo, wm) * dwm_dwi * pt / (pr + pt);
    float ft = T * TR_D(tr, wm) * TR_G(tr, wo, wi) *
               fabsf(Dot(wi, wm) * Dot(wo, wm) / (CosTheta(wi) * CosTheta(wo) * denom));
    if (radiance) ft /= SqrF(etap);
    f = ft;
    s.wi = wi;
    s.flags = BX_GLOSSY | BX_TRANSMISSION;
    s.eta = etap;
    s.ok = true;
    return s;
}

B2_HD uint32_t BsdfFlags(const Bsdf &b) {
    switch (b.type) {
    case MAT_DIFFUSE:
        if (b.trans)  // bxdfs.h:288-291
            return (b.rMax != 0 ? (BX_REFLECTION | BX_DIFFUSE) : 0u) | (b.tMax != 0 ? (BX_TRANSMISSION | BX_DIFFUSE) : 0u);
        return BX_REFLECTION | BX_DIFFUSE;  // R is never all-zero-checked here; see ShadeVertex
    case MAT_CONDUCTOR: return BX_REFLECTION | (TRSmooth(b.tr) ? BX_SPECULAR : BX_GLOSSY);
    case MAT_DIELECTRIC: {  // bxdfs.h:300-306; thin: bxdfs.h:404-406
        if (b.thin) return BX_REFLECTION | BX_TRANSMISSION | BX_SPECULAR;
        uint32_t f = (b.eta == 1) ? BX_TRANSMISSION : (BX_REFLECTION | BX_TRANSMISSION);
        return f | (TRSmooth(b.tr) ? BX_SPECULAR : BX_GLOSSY);
    }
    case MAT_COATED_DIFFUSE: {  // LayeredBxDF::Flags (bxdfs.h:589-606): dielectric top, diffuse bottom, no medium
        uint32_t f = BX_REFLECTION;
        bool topSpecular = TRSmooth(b.tr);
        if (topSpecular) f |= BX_SPECULAR;
        if (b.coatCond) return f | ((!topSpecular || !TRSmooth(b.condTr)) ? BX_GLOSSY : 0u);  // conductor bottom
        if (b.rMax != 0) f |= BX_DIFFUSE;
        else if (!topSpecular) f |= BX_GLOSSY;
        return f;
    }
    }
    return 0;
}
B2_HD void BsdfRegularize(Bsdf &b) {
    if (b.type == MAT_DIELECTRIC && b.thin) return;  // ThinDielectricBxDF::Regularize does nothing (bxdfs.h:400-401)
    if (b.type == MAT_CONDUCTOR || b.type == MAT_DIELECTRIC || b.type == MAT_COATED_DIFFUSE) TRRegularize(b.tr);
    if (b.type == MAT_COATED_DIFFUSE && b.coatCond) TRRegularize(b.condTr);  // LayeredBxDF::Regularize: top and bottom
}

B2_HD BsdfEval ZeroEval() {
    BsdfEval e;
    e.kind = 0;
    e.scale = e.D = e.G = e.cosF = 0;
    e.denom = 1;
    return e;
}

// scalar part of BxDF::f(wo, wi, Radiance)
B2_HD BsdfEval BsdfF(const Bsdf &b, V3 wo, V3 wi) {
    BsdfEval e = ZeroEval();
    if (b.type == MAT_DIFFUSE) {  // DiffuseBxDF::f (bxdfs.h:37-42)
        if (b.trans) {  // DiffuseTransmissionBxDF::f (bxdfs.h:226-228)
            e.kind = SameHemisphere(wo, wi) ? 6 : 7;
            e.scale = kInvPi;
            return e;
        }
        if (!SameHemisphere(wo, wi)) return e;
        e.kind = 2;
        e.scale = kInvPi;
        return e;
    }
    if (b.type == MAT_CONDUCTOR) {  // ConductorBxDF::f (bxdfs.h:463-481)
        if (!SameHemisphere(wo, wi)) return e;
        if (TRSmooth(b.tr)) return e;
        float cosTheta_o = AbsCosTheta(wo), cosTheta_i = AbsCosTheta(wi);
        if (cosTheta_i == 0 || cosTheta_o == 0) return e;
        V3 wm = wi + wo;
        if (LengthSquared(wm) == 0) return e;
        wm = Normalize(wm);
        e.kind = 3;
        e.cosF = AbsDot(wo, wm);
        e.D = TR_D(b.tr, wm);
        e.G = TR_G(b.tr, wo, wi);
        e.denom = 4 * cosTheta_i * cosTheta_o;
        return e;
    }
    if (b.type == MAT_DIELECTRIC) {  // DielectricBxDF::f (bxdfs.cpp:167-205)
        float f = DielF(b.eta, b.tr, wo, wi, true);
        if (f != 0) {
            e.kind = 1;
            e.scale = f;
        }
        return e;
    }
    return e;
}

// BxDF::PDF(wo, wi, Radiance, All)
B2_HD float BsdfPDF(const Bsdf &b, V3 wo, V3 wi) {
    if (b.type == MAT_DIFFUSE) {  // bxdfs.h:60-67
        if (b.trans) {  // DiffuseTransmissionBxDF::PDF (bxdfs.h:263-278)
            float pr = b.rMax, pt = b.tMax;
            if (pr == 0 && pt == 0) return 0;
            return (SameHemisphere(wo, wi) ? pr : pt) / (pr + pt) * (AbsCosTheta(wi) * kInvPi);
        }
        if (!SameHemisphere(wo, wi)) return 0;
        return AbsCosTheta(wi) * kInvPi;
    }
    if (b.type == MAT_CONDUCTOR) {  // bxdfs.h:484-499
        if (!SameHemisphere(wo, wi)) return 0;
        if (TRSmooth(b.tr)) return 0;
        V3 wm = wo + wi;
        if (LengthSquared(wm) == 0) return 0;
        wm = FaceForward(Normalize(wm), mk(0, 0, 1));
        return TR_PDF(b.tr, wo, wm) / (4 * AbsDot(wo, wm));
    }
    if (b.type == MAT_DIELECTRIC) return DielPDF(b.eta, b.tr, wo, wi, SF_ALL);  // bxdfs.cpp:207-255
    return 0;
}

// BxDF::Sample_f(wo, uc, u, Radiance, All); the per-wavelength f comes back as `ev`.
B2_HD BsdfSample BsdfSampleF(const Bsdf &b, V3 wo, float uc, float u0, float u1, BsdfEval &ev) {
    BsdfSample s{false, mk(0, 0, 0), 0, 0, 1, false};
    ev = ZeroEval();
    if (b.type == MAT_DIFFUSE && b.trans) {  // DiffuseTransmissionBxDF::Sample_f (bxdfs.h:230-261)
        float pr = b.rMax, pt = b.tMax;
        if (pr == 0 && pt == 0) return s;
        bool reflect = uc < pr / (pr + pt);
        V3 wi = SampleCosineHemisphere(u0, u1);
        if (reflect ? wo.z < 0 : wo.z > 0) wi.z *= -1;
        s.wi = wi;
        s.pdf = AbsCosTheta(wi) * kInvPi * (reflect ? pr : pt) / (pr + pt);
        s.flags = BX_DIFFUSE | (reflect ? BX_REFLECTION : BX_TRANSMISSION);
        ev.kind = reflect ? 6 : 7;
        ev.scale = kInvPi;
        s.ok = true;
        return s;
    }
    if (b.type == MAT_DIFFUSE) {  // bxdfs.h:44-58
        V3 wi = SampleCosineHemisphere(u0, u1);
        if (wo.z < 0) wi.z *= -1;
        s.wi = wi;
        s.pdf = AbsCosTheta(wi) * kInvPi;
        s.flags = BX_DIFFUSE | BX_REFLECTION;
        ev.kind = 2;
        ev.scale = kInvPi;
        s.ok = true;
        return s;
    }
    if (b.type == MAT_CONDUCTOR) {  // bxdfs.h:430-461
        if (TRSmooth(b.tr)) {
            V3 wi = mk(-wo.x, -wo.y, wo.z);
            s.wi = wi;
            s.pdf = 1;
            s.flags = BX_SPECULAR | BX_REFLECTION;
            ev.kind = 4;
            ev.cosF = AbsCosTheta(wi);
            ev.denom = AbsCosTheta(wi);
            s.ok = true;
            return s;
        }
        if (wo.z == 0) return s;
        V3 wm = TR_Sample_wm(b.tr, wo, u0, u1);
        V3 wi = Reflect(wo, wm);
        if (!SameHemisphere(wo, wi)) return s;
        float pdf = TR_PDF(b.tr, wo, wm) / (4 * AbsDot(wo, wm));
        float cosTheta_o = AbsCosTheta(wo), cosTheta_i = AbsCosTheta(wi);
        if (cosTheta_i == 0 || cosTheta_o == 0) return s;
        ev.kind = 3;
        ev.cosF = AbsDot(wo, wm);
        ev.D = TR_D(b.tr, wm);
        ev.G = TR_G(b.tr, wo, wi);
        ev.denom = 4 * cosTheta_i * cosTheta_o;
        s.wi = wi;
        s.pdf = pdf;
        s.flags = BX_GLOSSY | BX_REFLECTION;
        s.ok = true;
        return s;
    }
    if (b.type == MAT_DIELECTRIC && b.thin) {  // ThinDielectricBxDF::Sample_f (bxdfs.h:355-386)
        ev.kind = 1;
        float Rf = FrDielectric(AbsCosTheta(wo), b.eta), T = 1 - Rf;
        if (Rf < 1) {  // scattering between the two interfaces
            Rf += SqrF(T) * Rf / (1 - SqrF(Rf));
            T = 1 - Rf;
        }
        const float pr = Rf, pt = T;
        if (pr == 0 && pt == 0) return s;
        s.ok = true;
        if (uc < pr / (pr + pt)) {
            s.wi = mk(-wo.x, -wo.y, wo.z);
            ev.scale = Rf / AbsCosTheta(s.wi);
            s.pdf = pr / (pr + pt);
            s.flags = BX_SPECULAR | BX_REFLECTION;
        } else {
            s.wi = -wo;
            ev.scale = T / AbsCosTheta(s.wi);
            s.pdf = pt / (pr + pt);
            s.flags = BX_SPECULAR | BX_TRANSMISSION;
        }
        return s;
    }
    if (b.type == MAT_DIELECTRIC) {  // bxdfs.cpp:77-165
        ev.kind = 1;
        return DielSample(b.eta, b.tr, wo, uc, u0, u1, true, SF_ALL, ev.scale);
    }
    return s;
}

}  // namespace b200pt
