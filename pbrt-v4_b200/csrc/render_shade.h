// render_shade.h - surface interaction, emission, BSDFs, light sampling and the per-vertex step of
// the path integrator (included by render_core.h).
#pragma once

namespace b200pt {

struct TriVerts {
    V3 p0, p1, p2;
    uint32_t mesh, tri, kind;  // kind 1: a bilinear patch; its code loads the four corners itself (PatchVerts)
};
// bilinear patch: p00 p10 p01 p11 in p0..p3 (the mesh's index order)
struct PatchVerts {
    V3 p0, p1, p2, p3;
};
B2_HD PatchVerts LoadPatch(const void *prims, uint32_t slot) {
    const PrimSlot *s = reinterpret_cast<const PrimSlot *>(prims) + slot;
    PatchVerts t;
    t.p0 = mk(s->p0[0], s->p0[1], s->p0[2]);
    t.p1 = mk(s->p1[0], s->p1[1], s->p1[2]);
    t.p2 = mk(s->p2[0], s->p2[1], s->p2[2]);
    t.p3 = mk(s->p3[0], s->p3[1], s->p3[2]);
    return t;
}
B2_HD TriVerts LoadTri(const void *prims, uint32_t slot) {
    const PrimSlot *s = reinterpret_cast<const PrimSlot *>(prims) + slot;
    TriVerts t;
    t.p0 = mk(s->p0[0], s->p0[1], s->p0[2]);
    t.p1 = mk(s->p1[0], s->p1[1], s->p1[2]);
    t.p2 = mk(s->p2[0], s->p2[1], s->p2[2]);
    t.mesh = s->mesh;
    t.tri = s->tri;
    t.kind = s->kind;
    return t;
}

// What the integrator needs of SurfaceInteraction (interaction.h:158-270).
struct SurfHit {
    P3i pi;
    V3 n, ns, dpdus, wo;
    float u, v;
};

// per-corner normals / uv of triangle `tri` of a mesh, null where the mesh has none (MeshAttrDev::nOff, uvOff)
struct CornerAttr {
    const float *n, *uv;
    const float *s = nullptr;  // shading tangents (triangle meshes)
};
B2_HD CornerAttr CornersOf(const RenderSceneDev &rs, const MeshAttrDev &mesh, uint32_t tri, uint32_t kind = 0) {
    CornerAttr c;  // a bilinear patch has four corners (normals only)
    c.n = mesh.nOff != kNoCornerAttr ? rs.cornerAttr + mesh.nOff + ((kind & 1) ? 12 : 9) * size_t(tri) : nullptr;
    c.uv = mesh.uvOff != kNoCornerAttr ? rs.cornerAttr + mesh.uvOff + ((kind & 1) ? 8 : 6) * size_t(tri) : nullptr;
    c.s = (mesh.sOff != kNoCornerAttr && !(kind & 1)) ? rs.cornerAttr + mesh.sOff + 9 * size_t(tri) : nullptr;
    return c;
}

// The `if (mesh->n || mesh->s)` block of Triangle::InteractionFromIntersection (shapes.h:934-1005) without
// per-vertex tangents: interpolated shading normal, tangent = dpdu made orthogonal to it, and
// SetShadingGeometry(..., orientationIsAuthoritative = true) (interaction.h:194-211): the geometric normal follows ns.
// cn / cs: the corners' normals / tangents, either may be null (mesh->n, mesh->s; shapes.h:943-959)
B2_HD_NOINLINE void TriangleShadingNormal(SurfHit &h, const float *cn, const float *cs, float b0, float b1, float b2, V3 dpdu) {
    V3 ns = h.n;
    if (cn) {
        V3 n0 = mk(cn[0], cn[1], cn[2]), n1 = mk(cn[3], cn[4], cn[5]), n2 = mk(cn[6], cn[7], cn[8]);
        ns = b0 * n0 + b1 * n1 + b2 * n2;
        ns = LengthSquared(ns) > 0 ? Normalize(ns) : h.n;
    }
    V3 ss = dpdu;
    if (cs) {
        ss = b0 * mk(cs[0], cs[1], cs[2]) + b1 * mk(cs[3], cs[4], cs[5]) + b2 * mk(cs[6], cs[7], cs[8]);
        if (LengthSquared(ss) == 0) ss = dpdu;
    }
    V3 ts = Cross(ns, ss);
    if (LengthSquared(ts) > 0)
        ss = Cross(ts, ns);
    else
        CoordinateSystem(ns, &ss, &ts);
    h.ns = ns;
    h.n = FaceForward(h.n, ns);
    h.dpdus = ss;
}

// Triangle::InteractionFromIntersection (shapes.h:884-1010); without per-vertex uv the corners get
// (0,0),(1,0),(1,1); dpdu/dpdv from the uv determinant; n = normalized cross(dp02, dp12).
B2_HD SurfHit TriangleInteraction(const TriVerts &t, float b0, float b1, float b2, V3 wo, bool flipNormal,
                                  CornerAttr ca = CornerAttr{nullptr, nullptr}) {
    SurfHit h;
    float uv[3][2] = {{0, 0}, {1, 0}, {1, 1}};
    if (ca.uv)
        for (int k = 0; k < 6; ++k) uv[k >> 1][k & 1] = ca.uv[k];
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
    if (ca.n || ca.s) TriangleShadingNormal(h, ca.n, ca.s, b0, b1, b2, dpdu);
    return h;
}

// BilinearPatch::InteractionFromIntersection (shapes.h:1396-1500) for meshes without per-vertex uv / n: the point and
// both partial derivatives are lerps of the four corners at the hit's (u,v); the normal is their cross product.
// the interpolated vertex normal of a patch at (u, v): Lerp(u, Lerp(v, n00, n01), Lerp(v, n10, n11)), corners in index order
B2_HD V3 PatchNormalAt(const float *cn, float u, float v) {
    V3 n00 = mk(cn[0], cn[1], cn[2]), n10 = mk(cn[3], cn[4], cn[5]), n01 = mk(cn[6], cn[7], cn[8]), n11 = mk(cn[9], cn[10], cn[11]);
    return (1 - u) * ((1 - v) * n00 + v * n01) + u * ((1 - v) * n10 + v * n11);
}
// RotateFromTo (util/transform.h:249-270) applied to a vector
B2_HD void RotateFromTo(V3 from, V3 to, float r[3][3]) {
    V3 refl;
    if (fabsf(from.x) < 0.72f && fabsf(to.x) < 0.72f)
        refl = mk(1, 0, 0);
    else if (fabsf(from.y) < 0.72f && fabsf(to.y) < 0.72f)
        refl = mk(0, 1, 0);
    else
        refl = mk(0, 0, 1);
    V3 u = refl - from, v = refl - to;
    const float uu[3] = {u.x, u.y, u.z}, vv[3] = {v.x, v.y, v.z};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            r[i][j] = ((i == j) ? 1 : 0) - 2 / Dot(u, u) * uu[i] * uu[j] - 2 / Dot(v, v) * vv[i] * vv[j] +
                      4 * Dot(u, v) / (Dot(u, u) * Dot(v, v)) * vv[i] * uu[j];
}

// cuv: the corners' (s, t) in index order p00 p10 p01 p11 (BilinearPatchMesh::uv), or null
B2_HD_CALL SurfHit PatchInteraction(const PatchVerts &t, float u, float v, V3 wo, bool flipNormal, const float *cn = nullptr,
                                    const float *cuv = nullptr) {
    SurfHit h;
    const V3 p00 = t.p0, p10 = t.p1, p01 = t.p2, p11 = t.p3;
    V3 e0 = (1 - v) * p00 + v * p01, e1 = (1 - v) * p10 + v * p11;
    V3 p = (1 - u) * e0 + u * e1;
    V3 dpdu = e1 - e0;
    V3 dpdv = ((1 - u) * p01 + u * p11) - ((1 - u) * p00 + u * p10);
    if (cuv) {  // dpdu / dpdv with respect to the mesh's own (s, t) (shapes.h:1412-1440)
        const float s00 = cuv[0], t00 = cuv[1], s10 = cuv[2], t10 = cuv[3], s01 = cuv[4], t01 = cuv[5], s11 = cuv[6], t11 = cuv[7];
        const float dsdu = Lerp(v, s10, s11) - Lerp(v, s00, s01), dtdu = Lerp(v, t10, t11) - Lerp(v, t00, t01);
        const float dsdv = Lerp(u, s01, s11) - Lerp(u, s00, s10), dtdv = Lerp(u, t01, t11) - Lerp(u, t00, t10);
        const float duds = fabsf(dsdu) < 1e-8f ? 0 : 1 / dsdu, dvds = fabsf(dsdv) < 1e-8f ? 0 : 1 / dsdv;
        const float dudt = fabsf(dtdu) < 1e-8f ? 0 : 1 / dtdu, dvdt = fabsf(dtdv) < 1e-8f ? 0 : 1 / dtdv;
        V3 dpds = dpdu * duds + dpdv * dvds;
        V3 dpdt = dpdu * dudt + dpdv * dvdt;
        V3 c = Cross(dpds, dpdt);
        if (c.x != 0 || c.y != 0 || c.z != 0) {
            if (Dot(Cross(dpdu, dpdv), c) < 0) dpdt = -dpdt;
            dpdu = dpds;
            dpdv = dpdt;
        }
    }
    V3 pAbsSum = Abs(p00) + Abs(p01) + Abs(p10) + Abs(p11);
    h.pi = MakeP3i(p, GammaF(6) * pAbsSum);
    h.u = u;
    h.v = v;
    V3 n = Normalize(Cross(dpdu, dpdv));
    if (flipNormal) n = -n;
    h.n = n;
    h.ns = n;
    h.dpdus = dpdu;
    h.wo = Normalize(wo);
    if (cn) {  // shading normal from the vertex normals (shapes.h:1474-1494), no per-vertex uv
        V3 ns = PatchNormalAt(cn, u, v);
        if (LengthSquared(ns) > 0) {
            ns = Normalize(ns);
            float r[3][3];
            RotateFromTo(Normalize(h.n), ns, r);
            h.dpdus = mk(r[0][0] * dpdu.x + r[0][1] * dpdu.y + r[0][2] * dpdu.z, r[1][0] * dpdu.x + r[1][1] * dpdu.y + r[1][2] * dpdu.z,
                         r[2][0] * dpdu.x + r[2][1] * dpdu.y + r[2][2] * dpdu.z);
            h.ns = ns;
            h.n = FaceForward(h.n, ns);  // SetShadingGeometry(..., orientationIsAuthoritative = true)
        }
    }
    return h;
}

// Shape::Intersect's interaction for either kind of slot; (b1, b2) of a patch record are its (u, v)
// EXT = 0: the scene has triangles without per-vertex attributes only (RenderSceneDev::geomExt), the plain path
template <int EXT = 1>
B2_HD SurfHit SlotInteraction(const void *prims, uint32_t slot, const TriVerts &t, float b0, float b1, float b2, V3 wo,
                              bool flipNormal, CornerAttr ca) {
    if (EXT == 0) return TriangleInteraction(t, b0, b1, b2, wo, flipNormal);
    if (t.kind & 1) {
        const PatchVerts pv = LoadPatch(prims, slot);
        return PatchInteraction(pv, b1, b2, wo, flipNormal, ca.n, ca.uv);
    }
    return TriangleInteraction(t, b0, b1, b2, wo, flipNormal, ca);
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

// Light::SampleLi of the delta lights: PointLight (lights.h:221-228), SpotLight (lights.h:784-795, I(): lights.cpp:1391-1394),
// DistantLight (lights.h:282-289).  ss.pi = pLight (no error interval, no normal), ss.pdf = 1; *atten = what multiplies
// scale * spectrum(lambda).  false: the sample carries nothing.
B2_HD bool SampleDeltaLight(const LightDev &light, const float *g, V3 cp, ShapeSampleOut &ss, V3 &wi, float &atten) {
    ss.n = mk(0, 0, 0);
    ss.pdf = 1;
    if (light.type == LIGHT_DISTANT) {
        wi = mk(g[0], g[1], g[2]);
        ss.pi = MakeP3i(cp + wi * (2 * g[3]), mk(0, 0, 0));
        atten = 1;
        ss.ok = true;
        return true;
    }
    V3 pl = mk(g[0], g[1], g[2]);
    wi = Normalize(pl - cp);
    ss.pi = MakeP3i(pl, mk(0, 0, 0));
    atten = 1 / LengthSquared(pl - cp);
    if (light.type == LIGHT_SPOT) {
        V3 w = -wi;  // wLight = Normalize(renderFromLight.ApplyInverse(-wi))
        V3 wl = Normalize(mk(g[4] * w.x + g[5] * w.y + g[6] * w.z, g[7] * w.x + g[8] * w.y + g[9] * w.z,
                             g[10] * w.x + g[11] * w.y + g[12] * w.z));
        float a = g[14], b = g[13], x = wl.z, sm;  // SmoothStep(CosTheta(w), cosFalloffEnd, cosFalloffStart) (util/math.h)
        if (a == b)
            sm = x < a ? 0.f : 1.f;
        else {
            float t = Clampf((x - a) / (b - a), 0, 1);
            sm = t * t * (3 - 2 * t);
        }
        atten = sm * atten;  // (smoothstep * scale * I) / d^2
        if (sm == 0) {
            ss.ok = false;
            return false;
        }
    }
    ss.ok = true;
    return true;
}

// ---- ImageInfiniteLight (lights.h:557-643, lights.cpp:1038-1083): an equal-area square image of the sphere
// EqualAreaSquareToSphere / EqualAreaSphereToSquare (util/math.cpp:292-361)
B2_HD V3 EqualAreaSquareToSphere(float px, float py) {
    float u = 2 * px - 1, v = 2 * py - 1;
    float up = fabsf(u), vp = fabsf(v);
    float signedDistance = 1 - (up + vp);
    float d = fabsf(signedDistance);
    float r = 1 - d;
    float phi = (r == 0 ? 1 : (vp - up) / r + 1) * kPi / 4;
    float z = copysignf(1 - SqrF(r), signedDistance);
    float cosPhi = copysignf(cosf(phi), u);
    float sinPhi = copysignf(sinf(phi), v);
    return mk(cosPhi * r * SafeSqrt(2 - SqrF(r)), sinPhi * r * SafeSqrt(2 - SqrF(r)), z);
}
B2_HD void EqualAreaSphereToSquare(V3 d, float &pu, float &pv) {
    float x = fabsf(d.x), y = fabsf(d.y), z = fabsf(d.z);
    float r = SafeSqrt(1 - z);
    float a = fmaxf(x, y), b = fminf(x, y);
    b = a == 0 ? 0 : b / a;
    // EvaluatePolynomial(b, t1 .. t7) (util/math.h): Horner with fused multiply-adds
    const float t1 = 0.406758566246788489601959989e-5f, t2 = 0.636226545274016134946890922156f,
                t3 = 0.61572017898280213493197203466e-2f, t4 = -0.247333733281268944196501420480f,
                t5 = 0.881770664775316294736387951347e-1f, t6 = 0.419038818029165735901852432784e-1f,
                t7 = -0.251390972343483509333252996350e-1f;
    float phi = FMAf(b, FMAf(b, FMAf(b, FMAf(b, FMAf(b, FMAf(b, t7, t6), t5), t4), t3), t2), t1);
    if (x < y) phi = 1 - phi;
    float v = phi * r;
    float u = r - v;
    if (d.z < 0) {  // southern hemisphere: mirror
        float t = u;
        u = v;
        v = t;
        u = 1 - u;
        v = 1 - v;
    }
    u = copysignf(u, d.x);
    v = copysignf(v, d.y);
    pu = 0.5f * (u + 1);
    pv = 0.5f * (v + 1);
}
// Image::LookupNearestChannel(uv, c, WrapMode::OctahedralSphere) (util/image.h:352-356, 100-125): the pixel
B2_HD int EnvPixel(int res, float pu, float pv) {
    int x = (int)(pu * res), y = (int)(pv * res);
    if (x < 0) {
        x = -x;
        y = res - 1 - y;
    } else if (x >= res) {
        x = 2 * res - 1 - x;
        y = res - 1 - y;
    }
    if (y < 0) {
        x = res - 1 - x;
        y = -y;
    } else if (y >= res) {
        x = res - 1 - x;
        y = 2 * res - 1 - y;
    }
    if (res == 1) x = y = 0;
    // (a non-finite direction must not index outside the image)
    x = x < 0 ? 0 : (x > res - 1 ? res - 1 : x);
    y = y < 0 ? 0 : (y > res - 1 ? res - 1 : y);
    return y * res + x;
}
// ImageLe (lights.h:626-633) at one wavelength: scale * RGBIlluminantSpectrum(rgb)(lambda) = scale * (s * S(poly(lambda))) *
// illuminant(lambda), RGBSigmoidPolynomial (util/color.h) with the pixel's coefficients
B2_HD float EnvLe(const RenderSceneDev &rs, const LightDev &light, int pixel, float lam) {
    const float *c = rs.envImage + 4ll * pixel;
    float x = FMAf(lam, FMAf(lam, c[0], c[1]), c[2]);
    float s = std::isinf(x) ? (x > 0 ? 1.f : 0.f) : .5f + x / (2 * sqrtf(1 + SqrF(x)));
    return light.scale * ((c[3] * s) * SpectrumAt(rs.spectra[light.spectrum], lam));
}
// ImageInfiniteLight::SampleLi under allowIncompletePDF (lights.h:597-621): direction from the compensated distribution
B2_HD bool SampleEnvLight(const RenderSceneDev &rs, const float *g, V3 cp, float u0, float u1, ShapeSampleOut &ss, V3 &wi,
                          int &pixel) {
    float pu, pv, mapPDF;
    int iu, iv;
    SamplePiecewise2D(rs.envDist, rs.envRes, rs.envRes, 0.f, 1.f, 0.f, 1.f, u0, u1, pu, pv, mapPDF, iu, iv);
    ss.ok = false;
    if (mapPDF == 0) return false;
    V3 wl = EqualAreaSquareToSphere(pu, pv);
    wi = mk(g[0] * wl.x + g[1] * wl.y + g[2] * wl.z, g[3] * wl.x + g[4] * wl.y + g[5] * wl.z,
            g[6] * wl.x + g[7] * wl.y + g[8] * wl.z);
    ss.pdf = mapPDF / (4 * kPi);
    ss.n = mk(0, 0, 0);
    ss.pi = MakeP3i(cp + wi * (2 * g[18]), mk(0, 0, 0));
    ss.ok = true;
    pixel = EnvPixel(rs.envRes, pu, pv);
    return true;
}
// the light-space direction of a render-space one: renderFromLight.ApplyInverse(w)
B2_HD V3 EnvLightDir(const float *g, V3 w) {
    return mk(g[9] * w.x + g[10] * w.y + g[11] * w.z, g[12] * w.x + g[13] * w.y + g[14] * w.z,
              g[15] * w.x + g[16] * w.y + g[17] * w.z);
}
// ImageInfiniteLight::PDF_Li(ctx, w, allowIncompletePDF = true) (lights.cpp:1073-1083)
B2_HD float EnvPDF(const RenderSceneDev &rs, const float *g, V3 w) {
    float pu, pv;
    EqualAreaSphereToSquare(EnvLightDir(g, w), pu, pv);
    return Piecewise2DPDF(rs.envDist, rs.envRes, rs.envRes, pu, pv) / (4 * kPi);
}
// the pixel ImageInfiniteLight::Le(ray) reads (lights.h:589-593)
B2_HD int EnvPixelOfRay(const RenderSceneDev &rs, const float *g, V3 d) {
    float pu, pv;
    EqualAreaSphereToSquare(Normalize(EnvLightDir(g, d)), pu, pv);
    return EnvPixel(rs.envRes, pu, pv);
}

B2_HD float TriArea(const TriVerts &t) { return 0.5f * Length(Cross(t.p1 - t.p0, t.p2 - t.p0)); }  // shapes.h:866-871

// Triangle::Sample(u) by area (shapes.h:1012-1047)
// the sampled point's normal: towards the interpolated vertex normal where the mesh has normals (shapes.h:1024-1029, 1112-1117)
B2_HD V3 SampledTriangleNormal(const TriVerts &t, const float b[3], bool flipNormal, const float *cn) {
    V3 n = Normalize(Cross(t.p1 - t.p0, t.p2 - t.p0));
    if (cn) {
        V3 ns = b[0] * mk(cn[0], cn[1], cn[2]) + b[1] * mk(cn[3], cn[4], cn[5]) + (1 - b[0] - b[1]) * mk(cn[6], cn[7], cn[8]);
        n = FaceForward(n, ns);
    } else if (flipNormal)
        n = n * -1.f;
    return n;
}

B2_HD ShapeSampleOut SampleTriangleArea(const TriVerts &t, float u0, float u1, bool flipNormal, const float *cn = nullptr) {
    ShapeSampleOut o;
    float b[3];
    SampleUniformTriangle(u0, u1, b);
    V3 p = b[0] * t.p0 + b[1] * t.p1 + b[2] * t.p2;
    V3 n = SampledTriangleNormal(t, b, flipNormal, cn);
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
                                              bool flipNormal, const float *cn = nullptr) {
    ShapeSampleOut o;
    o.ok = false;
    V3 rp = Mid(ctx.pi);
    float solidAngle = SphericalTriangleArea(Normalize(t.p0 - rp), Normalize(t.p1 - rp), Normalize(t.p2 - rp));
    if (solidAngle < kMinSphericalSampleArea || solidAngle > kMaxSphericalSampleArea) {
        o = SampleTriangleArea(t, u0, u1, flipNormal, cn);
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
    V3 n = SampledTriangleNormal(t, b, flipNormal, cn);
    o.pi = MakeP3i(p, pError);
    o.n = n;
    o.pdf = pdf;
    o.ok = true;
    return o;
}

// ---- emissive bilinear patches: BilinearPatch::Sample / PDF (shapes.cpp:1158-1372)

// BilinearPatch::IsRectangle (shapes.h:1510-1532)
B2_HD bool PatchIsRectangle(const PatchVerts &t) {
    const V3 p00 = t.p0, p10 = t.p1, p01 = t.p2, p11 = t.p3;
    auto same = [](V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; };
    if (same(p00, p01) || same(p01, p11) || same(p11, p10) || same(p10, p00)) return false;
    V3 n = Normalize(Cross(p10 - p00, p01 - p00));
    if (AbsDotN(Normalize(p11 - p00), n) > 1e-5f) return false;
    V3 pc = (p00 + p01 + p10 + p11) / 4.f;
    float d2[4] = {LengthSquared(p00 - pc), LengthSquared(p01 - pc), LengthSquared(p10 - pc), LengthSquared(p11 - pc)};
    for (int i = 1; i < 4; ++i)
        if (fabsf(d2[i] - d2[0]) / d2[0] > 1e-4f) return false;
    return true;
}

// SphericalQuadArea (util/vecmath.h:1648-1666)
B2_HD float SphericalQuadArea(V3 a, V3 b, V3 c, V3 d) {
    V3 axb = Cross(a, b), bxc = Cross(b, c), cxd = Cross(c, d), dxa = Cross(d, a);
    if (LengthSquared(axb) == 0 || LengthSquared(bxc) == 0 || LengthSquared(cxd) == 0 || LengthSquared(dxa) == 0) return 0;
    axb = Normalize(axb);
    bxc = Normalize(bxc);
    cxd = Normalize(cxd);
    dxa = Normalize(dxa);
    float alpha = AngleBetween(dxa, -axb), beta = AngleBetween(axb, -bxc);
    float gamma = AngleBetween(bxc, -cxd), delta = AngleBetween(cxd, -dxa);
    return fabsf(alpha + beta + gamma + delta - 2 * kPi);
}

// the local frame and corner geometry both spherical-rectangle routines start from (util/sampling.cpp:166-186, 228-262)
struct SphRect {
    V3 rx, ry, rz;
    float x0, y0, z0, x1, y1;
    float g0, g1, g2, g3, b0, b1;
};
B2_HD SphRect SphRectInit(V3 pRef, V3 s, V3 ex, V3 ey) {
    SphRect q;
    float exl = Length(ex), eyl = Length(ey);
    q.rx = ex / exl;
    q.ry = ey / eyl;
    q.rz = Cross(q.rx, q.ry);
    V3 d = s - pRef;
    q.x0 = Dot(d, q.rx);
    q.y0 = Dot(d, q.ry);
    q.z0 = Dot(d, q.rz);
    if (q.z0 > 0) {
        q.rz = -q.rz;
        q.z0 *= -1;
    }
    q.x1 = q.x0 + exl;
    q.y1 = q.y0 + eyl;
    V3 v00 = mk(q.x0, q.y0, q.z0), v01 = mk(q.x0, q.y1, q.z0), v10 = mk(q.x1, q.y0, q.z0), v11 = mk(q.x1, q.y1, q.z0);
    V3 n0 = Normalize(Cross(v00, v10)), n1 = Normalize(Cross(v10, v11));
    V3 n2 = Normalize(Cross(v11, v01)), n3 = Normalize(Cross(v01, v00));
    q.g0 = AngleBetween(-n0, n1);
    q.g1 = AngleBetween(-n1, n2);
    q.g2 = AngleBetween(-n2, n3);
    q.g3 = AngleBetween(-n3, n0);
    q.b0 = n0.z;
    q.b1 = n2.z;
    return q;
}

// SampleSphericalRectangle (util/sampling.cpp:163-220)
B2_HD V3 SampleSphericalRectangle(V3 pRef, V3 s, V3 ex, V3 ey, float u0, float u1, float *pdf) {
    SphRect q = SphRectInit(pRef, s, ex, ey);
    float solidAngle = q.g0 + q.g1 + q.g2 + q.g3 - 2 * kPi;
    if (solidAngle <= 0) {
        *pdf = 0;
        return s + u0 * ex + u1 * ey;
    }
    *pdf = fmaxf(0.f, 1 / solidAngle);
    if (solidAngle < 1e-3f) return s + u0 * ex + u1 * ey;
    float au = u0 * (q.g0 + q.g1 - 2 * kPi) + (u0 - 1) * (q.g2 + q.g3);
    float fu = (cosf(au) * q.b0 - q.b1) / sinf(au);
    float cu = copysignf(1 / sqrtf(SqrF(fu) + SqrF(q.b0)), fu);
    cu = Clampf(cu, -kOneMinusEps, kOneMinusEps);
    float xu = -(cu * q.z0) / SafeSqrt(1 - SqrF(cu));
    xu = Clampf(xu, q.x0, q.x1);
    float dd = sqrtf(SqrF(xu) + SqrF(q.z0));
    float h0 = q.y0 / sqrtf(SqrF(dd) + SqrF(q.y0));
    float h1 = q.y1 / sqrtf(SqrF(dd) + SqrF(q.y1));
    float hv = h0 + u1 * (h1 - h0), hvsq = SqrF(hv);
    float yv = (hvsq < 1 - 1e-6f) ? (hv * dd) / sqrtf(1 - hvsq) : q.y1;
    return pRef + (xu * q.rx + yv * q.ry + q.z0 * q.rz);
}

// InvertSphericalRectangleSample (util/sampling.cpp:222-344)
B2_HD void InvertSphericalRectangleSample(V3 pRef, V3 s, V3 ex, V3 ey, V3 pRect, float &u0o, float &u1o) {
    SphRect q = SphRectInit(pRef, s, ex, ey);
    float z0sq = SqrF(q.z0), y0sq = SqrF(q.y0), y1sq = SqrF(q.y1);
    float b0sq = SqrF(q.b0);
    float solidAngle = float(double(q.g0) + double(q.g1) + double(q.g2) + double(q.g3) - 2. * double(kPi));
    if (solidAngle < 1e-3f) {
        V3 pq = pRect - s;
        u0o = Dot(pq, ex) / LengthSquared(ex);
        u1o = Dot(pq, ey) / LengthSquared(ey);
        return;
    }
    V3 v = pRect - pRef;
    float xu = Dot(v, q.rx), yv = Dot(v, q.ry);
    xu = Clampf(xu, q.x0, q.x1);
    if (xu == 0) xu = 1e-10f;
    float invcusq = 1 + z0sq / SqrF(xu);
    float fusq = invcusq - b0sq;
    float fu = copysignf(sqrtf(fusq), xu);
    float sq = SafeSqrt(DiffOfProducts(q.b0, q.b0, q.b1, q.b1) + fusq);
    float au = atan2f(-(q.b1 * fu) - copysignf(q.b0 * sq, fu * q.b0), q.b0 * q.b1 - sq * fabsf(fu));
    if (au > 0) au -= 2 * kPi;
    if (fu == 0) au = kPi;
    float u0 = (au + q.g2 + q.g3) / solidAngle;
    float ddsq = SqrF(xu) + z0sq;
    float dd = sqrtf(ddsq);
    float h0 = q.y0 / sqrtf(ddsq + y0sq);
    float h1 = q.y1 / sqrtf(ddsq + y1sq);
    float yvsq = SqrF(yv);
    float root = fabsf(h0 - h1) * sqrtf(yvsq * (ddsq + yvsq)) / (ddsq + yvsq);
    float u1[2] = {(DiffOfProducts(h0, h0, h0, h1) - root) / SqrF(h0 - h1), (DiffOfProducts(h0, h0, h0, h1) + root) / SqrF(h0 - h1)};
    float hv[2] = {Lerp(u1[0], h0, h1), Lerp(u1[1], h0, h1)};
    float yz[2] = {(hv[0] * dd) / sqrtf(1 - SqrF(hv[0])), (hv[1] * dd) / sqrtf(1 - SqrF(hv[1]))};
    u0o = Clampf(u0, 0, 1);
    u1o = (fabsf(yz[0] - yv) < fabsf(yz[1] - yv)) ? u1[0] : u1[1];
}

// the corner weights of approximately uniform area sampling (shapes.cpp:1172-1175, 1241-1244)
B2_HD void PatchAreaWeights(const PatchVerts &t, float w[4]) {
    const V3 p00 = t.p0, p10 = t.p1, p01 = t.p2, p11 = t.p3;
    w[0] = Length(Cross(p10 - p00, p01 - p00));
    w[1] = Length(Cross(p10 - p00, p11 - p10));
    w[2] = Length(Cross(p01 - p00, p11 - p01));
    w[3] = Length(Cross(p11 - p10, p11 - p01));
}

// BilinearPatch::Sample(u) by area (shapes.cpp:1158-1217), no image distribution, no per-vertex uv / n
B2_HD ShapeSampleOut SamplePatchArea(const PatchVerts &t, bool rectangle, float u0, float u1, bool flipNormal,
                                     const float *cn = nullptr) {
    ShapeSampleOut o;
    o.ok = false;
    const V3 p00 = t.p0, p10 = t.p1, p01 = t.p2, p11 = t.p3;
    float pdf = 1, su = u0, sv = u1;
    if (!rectangle) {
        float w[4];
        PatchAreaWeights(t, w);
        SampleBilinear(u0, u1, w, su, sv);
        pdf = BilinearPDF(su, sv, w);
    }
    V3 pu0 = (1 - sv) * p00 + sv * p01, pu1 = (1 - sv) * p10 + sv * p11;
    V3 p = (1 - su) * pu0 + su * pu1;
    V3 dpdu = pu1 - pu0;
    V3 dpdv = ((1 - su) * p01 + su * p11) - ((1 - su) * p00 + su * p10);
    if (LengthSquared(dpdu) == 0 || LengthSquared(dpdv) == 0) return o;
    V3 n = Normalize(Cross(dpdu, dpdv));
    if (cn)
        n = FaceForward(n, PatchNormalAt(cn, su, sv));
    else if (flipNormal)
        n = -n;
    V3 pAbsSum = Abs(p00) + Abs(p01) + Abs(p10) + Abs(p11);
    o.pi = MakeP3i(p, GammaF(6) * pAbsSum);
    o.n = n;
    o.pdf = pdf / Length(Cross(dpdu, dpdv));
    o.ok = true;
    return o;
}

B2_HD void PatchCosWeights(V3 v00, V3 v10, V3 v01, V3 v11, V3 ns, float w[4]) {  // shapes.cpp:1288-1292
    w[0] = fmaxf(0.01f, AbsDotN(ns, v00));
    w[1] = fmaxf(0.01f, AbsDotN(ns, v10));
    w[2] = fmaxf(0.01f, AbsDotN(ns, v01));
    w[3] = fmaxf(0.01f, AbsDotN(ns, v11));
}

constexpr float kPatchMinSphericalSampleArea = 1e-4f;  // shapes.h:1538

// BilinearPatch::Sample(ctx, u) (shapes.cpp:1257-1330)
B2_HD_CALL ShapeSampleOut SamplePatchSolidAngle(const PatchVerts &t, const LightCtx &ctx, float u0, float u1, bool flipNormal,
                                                const float *cn = nullptr) {
    const V3 p00 = t.p0, p10 = t.p1, p01 = t.p2, p11 = t.p3;
    V3 rp = Mid(ctx.pi);
    V3 v00 = Normalize(p00 - rp), v10 = Normalize(p10 - rp), v01 = Normalize(p01 - rp), v11 = Normalize(p11 - rp);
    bool rectangle = PatchIsRectangle(t);
    if (!rectangle || SphericalQuadArea(v00, v10, v11, v01) <= kPatchMinSphericalSampleArea) {
        ShapeSampleOut o = SamplePatchArea(t, rectangle, u0, u1, flipNormal, cn);
        if (!o.ok) return o;
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
    ShapeSampleOut o;
    float pdf = 1;
    if (!(ctx.ns.x == 0 && ctx.ns.y == 0 && ctx.ns.z == 0)) {
        float w[4], px, py;
        PatchCosWeights(v00, v10, v01, v11, ctx.ns, w);
        SampleBilinear(u0, u1, w, px, py);
        u0 = px;
        u1 = py;
        pdf *= BilinearPDF(u0, u1, w);
    }
    V3 eu = p10 - p00, ev = p01 - p00;
    float quadPDF;
    V3 p = SampleSphericalRectangle(rp, p00, eu, ev, u0, u1, &quadPDF);
    pdf *= quadPDF;
    V3 n = Normalize(Cross(eu, ev));
    if (cn) {  // (u, v) of the sampled point on the rectangle (shapes.cpp:1306-1308)
        float su = Dot(p - p00, eu) / LengthSquared(p10 - p00), sv = Dot(p - p00, ev) / LengthSquared(p01 - p00);
        n = FaceForward(n, PatchNormalAt(cn, su, sv));
    } else if (flipNormal)
        n = -n;
    o.pi = MakeP3i(p, mk(0, 0, 0));
    o.n = n;
    o.pdf = pdf;
    o.ok = true;
    return o;
}

// BilinearPatch::PDF(ctx, wi) (shapes.cpp:1332-1372) with PDF(intr) (:1219-1255)
B2_HD_CALL float PatchPDF(const PatchVerts &t, const LightCtx &ctx, V3 wi, bool flipNormal) {
    const V3 p00 = t.p0, p10 = t.p1, p01 = t.p2, p11 = t.p3;
    V3 rp = Mid(ctx.pi);
    V3 o = OffsetRayOrigin(ctx.pi, ctx.n, wi);
    RayCtx r;
    RaySetup(r, o.x, o.y, o.z, wi.x, wi.y, wi.z);
    float q00[3] = {p00.x, p00.y, p00.z}, q10[3] = {p10.x, p10.y, p10.z};
    float q01[3] = {p01.x, p01.y, p01.z}, q11[3] = {p11.x, p11.y, p11.z};
    float tt, u, v;
    if (!BilinearHit(r, INFINITY, q00, q10, q01, q11, tt, u, v)) return 0;
    V3 v00 = Normalize(p00 - rp), v10 = Normalize(p10 - rp), v01 = Normalize(p01 - rp), v11 = Normalize(p11 - rp);
    bool rectangle = PatchIsRectangle(t);
    float quadArea = SphericalQuadArea(v00, v10, v11, v01);
    V3 pu0 = (1 - v) * p00 + v * p01, pu1 = (1 - v) * p10 + v * p11;
    // isect->intr.p(): the midpoint of the interval point InteractionFromIntersection builds
    V3 pHit = Mid(MakeP3i((1 - u) * pu0 + u * pu1, GammaF(6) * (Abs(p00) + Abs(p01) + Abs(p10) + Abs(p11))));
    if (!rectangle || quadArea <= kPatchMinSphericalSampleArea) {
        float pdf = 1;
        if (!rectangle) {
            float w[4];
            PatchAreaWeights(t, w);
            pdf = BilinearPDF(u, v, w);
        }
        V3 dpdu = pu1 - pu0;
        V3 dpdv = ((1 - u) * p01 + u * p11) - ((1 - u) * p00 + u * p10);
        V3 nrm = Cross(dpdu, dpdv);
        pdf = pdf / Length(nrm);
        V3 n = Normalize(nrm);  // sign does not matter under AbsDot
        (void)flipNormal;
        pdf = pdf * (LengthSquared(rp - pHit) / AbsDotN(n, -wi));
        return std::isinf(pdf) ? 0.f : pdf;
    }
    float pdf = 1 / quadArea;
    if (!(ctx.ns.x == 0 && ctx.ns.y == 0 && ctx.ns.z == 0)) {
        float w[4], iu0, iu1;
        PatchCosWeights(v00, v10, v01, v11, ctx.ns, w);
        InvertSphericalRectangleSample(rp, p00, p10 - p00, p01 - p00, pHit, iu0, iu1);
        return BilinearPDF(iu0, iu1, w) * pdf;
    }
    return pdf;
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


// Shape::Sample(ctx, u) / Shape::PDF(ctx, wi) of the emitter in slot `t`
template <int EXT = 1>
B2_HD ShapeSampleOut SampleShapeSolidAngle(const void *prims, uint32_t slot, const TriVerts &t, const LightCtx &ctx, float u0,
                                           float u1, bool flipNormal, const float *cn) {
    if (EXT == 0) return SampleTriangleSolidAngle(t, ctx, u0, u1, flipNormal);
    if (t.kind & 1) {
        const PatchVerts pv = LoadPatch(prims, slot);
        return SamplePatchSolidAngle(pv, ctx, u0, u1, flipNormal, cn);
    }
    return SampleTriangleSolidAngle(t, ctx, u0, u1, flipNormal, cn);
}
template <int EXT = 1>
B2_HD float ShapePDF(const void *prims, uint32_t slot, const TriVerts &t, const LightCtx &ctx, V3 wi, bool flipNormal) {
    if (EXT == 0) return TrianglePDF(t, ctx, wi, flipNormal);
    if (t.kind & 1) {
        const PatchVerts pv = LoadPatch(prims, slot);
        return PatchPDF(pv, ctx, wi, flipNormal);
    }
    return TrianglePDF(t, ctx, wi, flipNormal);
}
}  // namespace b200pt

#include "render_path.h"
