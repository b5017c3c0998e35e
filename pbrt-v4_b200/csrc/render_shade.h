// render_shade.h - surface interaction, emission, BSDFs, light sampling and the per-vertex step of
// the path integrator (included by render_core.h).
#pragma once

namespace b200pt {

struct TriVerts {
    V3 p0, p1, p2;
    uint32_t mesh, tri;
};
B2_HD TriVerts LoadTri(const void *prims, uint32_t slot) {
    const PrimSlot *s = reinterpret_cast<const PrimSlot *>(prims) + slot;
    TriVerts t;
    t.p0 = mk(s->p0[0], s->p0[1], s->p0[2]);
    t.p1 = mk(s->p1[0], s->p1[1], s->p1[2]);
    t.p2 = mk(s->p2[0], s->p2[1], s->p2[2]);
    t.mesh = s->mesh;
    t.tri = s->tri;
    return t;
}

// What the integrator needs of SurfaceInteraction (interaction.h:158-270).
struct SurfHit {
    P3i pi;
    V3 n, ns, dpdus, wo;
    float u, v;
};

// Triangle::InteractionFromIntersection (shapes.h:884-1010) for meshes without per-vertex n/s/uv:
// default uv (0,0),(1,0),(1,1); dpdu/dpdv from the uv determinant; n = normalized cross(dp02, dp12).
B2_HD SurfHit TriangleInteraction(const TriVerts &t, float b0, float b1, float b2, V3 wo, bool flipNormal) {
    SurfHit h;
    const float uv[3][2] = {{0, 0}, {1, 0}, {1, 1}};
    float duv02x = uv[0][0] - uv[2][0], duv02y = uv[0][1] - uv[2][1];
    float duv12x = uv[1][0] - uv[2][0], duv12y = uv[1][1] - uv[2][1];
    V3 dp02 = t.p0 - t.p2, dp12 = t.p1 - t.p2;
    float determinant = DiffOfProducts(duv02x, duv12y, duv02y, duv12x);
    V3 dpdu = mk(0, 0, 0), dpdv = mk(0, 0, 0);
    bool degenerateUV = fabsf(determinant) < 1e-9f;
    if (!degenerateUV) {
        float invdet = 1 / determinant;
        dpdu = mk(DiffOfProducts(duv12y, dp02.x, duv02y, dp12.x), DiffOfProducts(duv12y, dp02.y, duv02y, dp12.y),
                  DiffOfProducts(duv12y, dp02.z, duv02y, dp12.z)) * invdet;
        dpdv = mk(DiffOfProducts(duv02x, dp12.x, duv12x, dp02.x), DiffOfProducts(duv02x, dp12.y, duv12x, dp02.y),
                  DiffOfProducts(duv02x, dp12.z, duv12x, dp02.z)) * invdet;
    }
    if (degenerateUV || LengthSquared(Cross(dpdu, dpdv)) == 0) {
        V3 ng = Cross(t.p2 - t.p0, t.p1 - t.p0);
        CoordinateSystem(Normalize(ng), &dpdu, &dpdv);
    }
    V3 pHit = b0 * t.p0 + b1 * t.p1 + b2 * t.p2;
    h.u = b0 * uv[0][0] + b1 * uv[1][0] + b2 * uv[2][0];
    h.v = b0 * uv[0][1] + b1 * uv[1][1] + b2 * uv[2][1];
    V3 pAbsSum = Abs(b0 * t.p0) + Abs(b1 * t.p1) + Abs(b2 * t.p2);
    V3 pError = GammaF(7) * pAbsSum;
    h.pi = MakeP3i(pHit, pError);
    V3 n = Normalize(Cross(dp02, dp12));
    if (flipNormal) n = -n;
    h.n = n;
    h.ns = n;
    h.dpdus = dpdu;
    h.wo = Normalize(wo);  // Interaction constructor (interaction.h:36-41)
    return h;
}

// Transform::operator()(SurfaceInteraction) for an object instance (util/transform.cpp:227-258,
// util/transform.h:133-176, 329-334): the hit found in the object's space goes back to render space.
B2_HD void TransformSurfHit(const InstanceDev &in, SurfHit &h) {
    const float *m = in.m, *mi = in.mInv;
    V3 p = Mid(h.pi), e = Err(h.pi);
    const float g3 = GammaF(3);
    float q[3], qe[3];
    for (int r = 0; r < 3; ++r) {
        const float *row = m + 4 * r;
        q[r] = (row[0] * p.x + row[1] * p.y) + (row[2] * p.z + row[3]);
        qe[r] = (g3 + 1) * (fabsf(row[0]) * e.x + fabsf(row[1]) * e.y + fabsf(row[2]) * e.z) +
                g3 * (fabsf(row[0] * p.x) + fabsf(row[1] * p.y) + fabsf(row[2] * p.z) + fabsf(row[3]));
    }
    h.pi = MakeP3i(mk(q[0], q[1], q[2]), mk(qe[0], qe[1], qe[2]));
    float wp = (m[12] * p.x + m[13] * p.y) + (m[14] * p.z + m[15]);
    if (wp != 1) {  // Point3fi / wp (util/math.h:1028-1036); renderFromInstance is normally exactly affine
        float *lo = &h.pi.lo.x, *hi = &h.pi.hi.x;
        for (int r = 0; r < 3; ++r) {
            float a = lo[r], b = hi[r];
            lo[r] = wp > 0 ? NextDown(a / wp) : NextDown(b / wp);
            hi[r] = wp > 0 ? NextUp(b / wp) : NextUp(a / wp);
        }
    }
    auto nrm = [&](V3 n) {  // normals go through the inverse transpose
        return mk(mi[0] * n.x + mi[4] * n.y + mi[8] * n.z, mi[1] * n.x + mi[5] * n.y + mi[9] * n.z,
                  mi[2] * n.x + mi[6] * n.y + mi[10] * n.z);
    };
    auto vec = [&](V3 v) {
        return mk(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z,
                  m[8] * v.x + m[9] * v.y + m[10] * v.z);
    };
    h.n = Normalize(nrm(h.n));
    h.wo = Normalize(vec(h.wo));
    h.dpdus = vec(h.dpdus);
    h.ns = FaceForward(Normalize(nrm(h.ns)), h.n);
}

// ------------------------------------------------------------------ area lights on triangles
struct LightCtx {  // LightSampleContext (lights.h:83-101)
    P3i pi;
    V3 n, ns;
};

// DiffuseAreaLight::L (lights.h:444-470) without image / alpha.
B2_HD bool AreaLightEmits(const LightDev &l, V3 n, V3 w) {
    if (!l.twoSided && DotN(n, w) < 0) return false;
    if (l.cosFalloffEnd > 0) {
        float cosTheta = AbsDotN(n, w);
        if (cosTheta < l.cosFalloffEnd) return false;
    }
    return true;
}

struct ShapeSampleOut {
    P3i pi;
    V3 n;
    float pdf;
    bool ok;
};

B2_HD float TriArea(const TriVerts &t) { return 0.5f * Length(Cross(t.p1 - t.p0, t.p2 - t.p0)); }  // shapes.h:866-871

// Triangle::Sample(u) by area (shapes.h:1012-1047)
B2_HD ShapeSampleOut SampleTriangleArea(const TriVerts &t, float u0, float u1, bool flipNormal) {
    ShapeSampleOut o;
    float b[3];
    SampleUniformTriangle(u0, u1, b);
    V3 p = b[0] * t.p0 + b[1] * t.p1 + b[2] * t.p2;
    V3 n = Normalize(Cross(t.p1 - t.p0, t.p2 - t.p0));
    if (flipNormal) n = n * -1.f;
    V3 pAbsSum = Abs(b[0] * t.p0) + Abs(b[1] * t.p1) + Abs((1 - b[0] - b[1]) * t.p2);
    V3 pError = GammaF(6) * pAbsSum;
    o.pi = MakeP3i(p, pError);
    o.n = n;
    o.pdf = 1 / TriArea(t);
    o.ok = true;
    return o;
}

// SampleSphericalTriangle (util/sampling.cpp:28-107)
B2_HD bool SampleSphericalTriangle(const TriVerts &t, V3 p, float u0, float u1, float b[3], float *pdf) {
    *pdf = 0;
    V3 a = Normalize(t.p0 - p), bb = Normalize(t.p1 - p), c = Normalize(t.p2 - p);
    V3 n_ab = Cross(a, bb), n_bc = Cross(bb, c), n_ca = Cross(c, a);
    if (LengthSquared(n_ab) == 0 || LengthSquared(n_bc) == 0 || LengthSquared(n_ca) == 0) return false;
    n_ab = Normalize(n_ab);
    n_bc = Normalize(n_bc);
    n_ca = Normalize(n_ca);
    float alpha = AngleBetween(n_ab, -n_ca);
    float beta = AngleBetween(n_bc, -n_ab);
    float gamma = AngleBetween(n_ca, -n_bc);
    float A_pi = alpha + beta + gamma;
    float Ap_pi = Lerp(u0, kPi, A_pi);
    float A = A_pi - kPi;
    *pdf = (A <= 0) ? 0 : 1 / A;
    float cosAlpha = cosf(alpha), sinAlpha = sinf(alpha);
    float sinPhi = sinf(Ap_pi) * cosAlpha - cosf(Ap_pi) * sinAlpha;
    float cosPhi = cosf(Ap_pi) * cosAlpha + sinf(Ap_pi) * sinAlpha;
    float k1 = cosPhi + cosAlpha;
    float k2 = sinPhi - sinAlpha * Dot(a, bb);
    float cosBp = (k2 + (DiffOfProducts(k2, cosPhi, k1, sinPhi)) * cosAlpha) /
                  ((SumOfProductsF(k2, sinPhi, k1, cosPhi)) * sinAlpha);
    cosBp = Clampf(cosBp, -1, 1);
    float sinBp = SafeSqrt(1 - SqrF(cosBp));
    V3 cp = cosBp * a + sinBp * Normalize(GramSchmidt(c, a));
    float cosTheta = 1 - u1 * (1 - Dot(cp, bb));
    float sinTheta = SafeSqrt(1 - SqrF(cosTheta));
    V3 w = cosTheta * bb + sinTheta * Normalize(GramSchmidt(cp, bb));
    V3 e1 = t.p1 - t.p0, e2 = t.p2 - t.p0;
    V3 s1 = Cross(w, e2);
    float divisor = Dot(s1, e1);
    if (divisor == 0) {
        b[0] = b[1] = b[2] = 1.f / 3.f;
        return true;
    }
    float invDivisor = 1 / divisor;
    V3 s = p - t.p0;
    float b1 = Dot(s, s1) * invDivisor;
    float b2 = Dot(w, Cross(s, e1)) * invDivisor;
    b1 = Clampf(b1, 0, 1);
    b2 = Clampf(b2, 0, 1);
    if (b1 + b2 > 1) {
        b1 /= b1 + b2;
        b2 /= b1 + b2;
    }
    b[0] = 1 - b1 - b2;
    b[1] = b1;
    b[2] = b2;
    return true;
}

// InvertSphericalTriangleSample (util/sampling.cpp:110-161)
B2_HD void InvertSphericalTriangleSample(const TriVerts &t, V3 p, V3 w, float &u0o, float &u1o) {
    u0o = u1o = 0;
    V3 a = Normalize(t.p0 - p), b = Normalize(t.p1 - p), c = Normalize(t.p2 - p);
    V3 n_ab = Cross(a, b), n_bc = Cross(b, c), n_ca = Cross(c, a);
    if (LengthSquared(n_ab) == 0 || LengthSquared(n_bc) == 0 || LengthSquared(n_ca) == 0) return;
    n_ab = Normalize(n_ab);
    n_bc = Normalize(n_bc);
    n_ca = Normalize(n_ca);
    float alpha = AngleBetween(n_ab, -n_ca);
    float beta = AngleBetween(n_bc, -n_ab);
    float gamma = AngleBetween(n_ca, -n_bc);
    V3 cp = Normalize(Cross(Cross(b, w), Cross(c, a)));
    if (Dot(cp, a + c) < 0) cp = -cp;
    float u0;
    if (Dot(a, cp) > 0.99999847691f)
        u0 = 0;
    else {
        V3 n_cpb = Cross(cp, b), n_acp = Cross(a, cp);
        if (LengthSquared(n_cpb) == 0 || LengthSquared(n_acp) == 0) {
            u0o = u1o = 0.5f;
            return;
        }
        n_cpb = Normalize(n_cpb);
        n_acp = Normalize(n_acp);
        float Ap = alpha + AngleBetween(n_ab, n_cpb) + AngleBetween(n_acp, -n_cpb) - kPi;
        float A = alpha + beta + gamma - kPi;
        u0 = Ap / A;
    }
    float u1 = (1 - Dot(w, b)) / (1 - Dot(cp, b));
    u0o = Clampf(u0, 0, 1);
    u1o = Clampf(u1, 0, 1);
}

constexpr float kMinSphericalSampleArea = 3e-4f;  // shapes.h:1190-1191
constexpr float kMaxSphericalSampleArea = 6.22f;

B2_HD void CosWarpWeights(const TriVerts &t, V3 rp, V3 ns, float w[4]) {  // shapes.h:1081-1087
    V3 wi0 = Normalize(t.p0 - rp), wi1 = Normalize(t.p1 - rp), wi2 = Normalize(t.p2 - rp);
    w[0] = fmaxf(0.01f, AbsDotN(ns, wi1));
    w[1] = fmaxf(0.01f, AbsDotN(ns, wi1));
    w[2] = fmaxf(0.01f, AbsDotN(ns, wi0));
    w[3] = fmaxf(0.01f, AbsDotN(ns, wi2));
}

// Triangle::Sample(ctx, u) (shapes.h:1052-1130)
B2_HD ShapeSampleOut SampleTriangleSolidAngle(const TriVerts &t, const LightCtx &ctx, float u0, float u1,
                                              bool flipNormal) {
    ShapeSampleOut o;
    o.ok = false;
    V3 rp = Mid(ctx.pi);
    float solidAngle = SphericalTriangleArea(Normalize(t.p0 - rp), Normalize(t.p1 - rp), Normalize(t.p2 - rp));
    if (solidAngle < kMinSphericalSampleArea || solidAngle > kMaxSphericalSampleArea) {
        o = SampleTriangleArea(t, u0, u1, flipNormal);
        V3 wi = Mid(o.pi) - rp;
        if (LengthSquared(wi) == 0) {
            o.ok = false;
            return o;
        }
        wi = Normalize(wi);
        o.pdf /= AbsDotN(o.n, -wi) / LengthSquared(rp - Mid(o.pi));
        if (std::isinf(o.pdf)) o.ok = false;
        return o;
    }
    float pdf = 1;
    if (!(ctx.ns.x == 0 && ctx.ns.y == 0 && ctx.ns.z == 0)) {
        float w[4];
        CosWarpWeights(t, rp, ctx.ns, w);
        float px, py;
        SampleBilinear(u0, u1, w, px, py);
        u0 = px;
        u1 = py;
        pdf = BilinearPDF(u0, u1, w);
    }
    float triPDF, b[3];
    if (!SampleSphericalTriangle(t, rp, u0, u1, b, &triPDF) || triPDF == 0) return o;
    pdf *= triPDF;
    V3 pAbsSum = Abs(b[0] * t.p0) + Abs(b[1] * t.p1) + Abs((1 - b[0] - b[1]) * t.p2);
    V3 pError = GammaF(6) * pAbsSum;
    V3 p = b[0] * t.p0 + b[1] * t.p1 + b[2] * t.p2;
    V3 n = Normalize(Cross(t.p1 - t.p0, t.p2 - t.p0));
    if (flipNormal) n = n * -1.f;
    o.pi = MakeP3i(p, pError);
    o.n = n;
    o.pdf = pdf;
    o.ok = true;
    return o;
}

// Triangle::PDF(ctx, wi) (shapes.h:1132-1172)
B2_HD float TrianglePDF(const TriVerts &t, const LightCtx &ctx, V3 wi, bool flipNormal) {
    V3 rp = Mid(ctx.pi);
    float solidAngle = SphericalTriangleArea(Normalize(t.p0 - rp), Normalize(t.p1 - rp), Normalize(t.p2 - rp));
    if (solidAngle < kMinSphericalSampleArea || solidAngle > kMaxSphericalSampleArea) {
        // ctx.SpawnRay(wi) then Triangle::Intersect (interaction / shapes)
        V3 o = OffsetRayOrigin(ctx.pi, ctx.n, wi);
        RayCtx r;
        RaySetup(r, o.x, o.y, o.z, wi.x, wi.y, wi.z);
        float p0[3] = {t.p0.x, t.p0.y, t.p0.z}, p1[3] = {t.p1.x, t.p1.y, t.p1.z}, p2[3] = {t.p2.x, t.p2.y, t.p2.z};
        float tt, b0, b1, b2;
        if (!TriHit(r, INFINITY, p0, p1, p2, tt, b0, b1, b2)) return 0;
        V3 n = Normalize(Cross(t.p0 - t.p2, t.p1 - t.p2));
        if (flipNormal) n = -n;
        V3 pHit = b0 * t.p0 + b1 * t.p1 + b2 * t.p2;
        V3 pAbsSum = Abs(b0 * t.p0) + Abs(b1 * t.p1) + Abs(b2 * t.p2);
        P3i pi = MakeP3i(pHit, GammaF(7) * pAbsSum);
        float pdf = (1 / TriArea(t)) / (AbsDotN(n, -wi) / LengthSquared(rp - Mid(pi)));
        if (std::isinf(pdf)) pdf = 0;
        return pdf;
    }
    float pdf = 1 / solidAngle;
    if (!(ctx.ns.x == 0 && ctx.ns.y == 0 && ctx.ns.z == 0)) {
        float u0, u1, w[4];
        InvertSphericalTriangleSample(t, rp, wi, u0, u1);
        CosWarpWeights(t, rp, ctx.ns, w);
        pdf *= BilinearPDF(u0, u1, w);
    }
    return pdf;
}

}  // namespace b200pt

#include "render_path.h"
