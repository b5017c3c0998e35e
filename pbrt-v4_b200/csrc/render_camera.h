// render_camera.h - camera rays beyond the pinhole: the ray transform every camera ends with
// (Transform::operator()(Ray), util/transform.h:336-348) and the RealisticCamera
// (cameras.cpp:820-884 TraceLensesFromFilm, :968-988 SampleExitPupil, :990-1019 GenerateRay;
// cameras.h:557-579 IntersectSphericalElement).  Host+device; the host builds the lens table and the
// exit-pupil bounds (RealisticCamera constructor, cameras.cpp:770-817) and hands them over.
#pragma once
#include "render_bsdf.h"

namespace b200pt {

// Transform::operator()(Ray): the origin as a point with an error box (util/transform.h:133-176, exact
// input), moved along d to the edge of that box.  The interval end points are rounded outwards as the
// CPU build does (util/float.h:201-229), the origin is the interval's midpoint.
B2_HD void TransformRay(const float *m, V3 ro, V3 rd, V3 &o, V3 &d) {
    const float g3 = GammaF(3);
    float in[3] = {ro.x, ro.y, ro.z};
    float p[3], e[3];
    for (int r = 0; r < 3; ++r) {
        const float *row = m + 4 * r;
        float mx = row[0] * in[0], my = row[1] * in[1], mz = row[2] * in[2];
        p[r] = (mx + my) + (mz + row[3]);
        e[r] = g3 * (fabsf(mx) + fabsf(my) + fabsf(mz) + fabsf(row[3]));
    }
    float wp = (m[12] * in[0] + m[13] * in[1]) + (m[14] * in[2] + m[15]);
    P3i pi = MakeP3i(mk(p[0], p[1], p[2]), mk(e[0], e[1], e[2]));
    if (wp != 1) {  // not met with a camera's rigid transform; kept for completeness (Point3fi / wp)
        pi.lo = pi.lo / wp;
        pi.hi = pi.hi / wp;
    }
    d = XformVector(m, rd);
    float lengthSquared = LengthSquared(d);
    if (lengthSquared > 0) {
        V3 err = Err(pi);
        float dt = Dot(Abs(d), err) / lengthSquared;
        V3 s = d * dt;
        pi.lo = mk(NextDown(pi.lo.x + s.x), NextDown(pi.lo.y + s.y), NextDown(pi.lo.z + s.z));
        pi.hi = mk(NextUp(pi.hi.x + s.x), NextUp(pi.hi.y + s.y), NextUp(pi.hi.z + s.z));
    }
    o = Mid(pi);
}

// Quadratic (util/math.h:614-637)
B2_HD bool QuadraticF(float a, float b, float c, float *t0, float *t1) {
    if (a == 0) {
        if (b == 0) return false;
        *t0 = *t1 = -c / b;
        return true;
    }
    float discrim = DiffOfProducts(b, b, 4 * a, c);
    if (discrim < 0) return false;
    float rootDiscrim = sqrtf(discrim);
    float q = -0.5f * (b + copysignf(rootDiscrim, b));
    *t0 = q / a;
    *t1 = c / q;
    if (*t0 > *t1) {
        float t = *t0;
        *t0 = *t1;
        *t1 = t;
    }
    return true;
}

// IntersectSphericalElement (cameras.h:557-579)
B2_HD bool IntersectSphericalElement(float radius, float zCenter, V3 ro, V3 rd, float *t, V3 *n) {
    V3 o = ro - mk(0, 0, zCenter);
    float A = rd.x * rd.x + rd.y * rd.y + rd.z * rd.z;
    float B = 2 * (rd.x * o.x + rd.y * o.y + rd.z * o.z);
    float C = o.x * o.x + o.y * o.y + o.z * o.z - radius * radius;
    float t0, t1;
    if (!QuadraticF(A, B, C, &t0, &t1)) return false;
    bool useCloserT = (rd.z > 0) ^ (radius < 0);
    *t = useCloserT ? fminf(t0, t1) : fmaxf(t0, t1);
    if (*t < 0) return false;
    *n = o + *t * rd;
    *n = FaceForward(Normalize(*n), -rd);
    return true;
}

// RealisticCamera::TraceLensesFromFilm (cameras.cpp:820-884), circular apertures (no aperture image).
// lens: n x {curvatureRadius, thickness, eta, apertureRadius}, front element first.  Camera space in, camera
// space out; returns the weight (0 = the ray is blocked).
B2_HD float TraceLensesFromFilm(const float *lens, int nLens, V3 co, V3 cd, V3 *oOut, V3 *dOut) {
    float elementZ = 0;
    V3 o = mk(co.x, co.y, -co.z), d = mk(cd.x, cd.y, -cd.z);
    for (int i = nLens - 1; i >= 0; --i) {
        const float *el = lens + 4 * i;
        float curvatureRadius = el[0], eta = el[2], apertureRadius = el[3];
        elementZ -= el[1];
        float t;
        V3 n = mk(0, 0, 0);
        bool isStop = curvatureRadius == 0;
        if (isStop) {
            t = (elementZ - o.z) / d.z;
            if (t < 0) return 0;
        } else {
            float zCenter = elementZ + curvatureRadius;
            if (!IntersectSphericalElement(curvatureRadius, zCenter, o, d, &t, &n)) return 0;
        }
        V3 pHit = o + d * t;  // Ray::operator(): o + d * t
        if (SqrF(pHit.x) + SqrF(pHit.y) > SqrF(apertureRadius)) return 0;
        o = pHit;
        if (!isStop) {
            V3 w;
            float eta_i = eta;
            float eta_t = (i > 0 && lens[4 * (i - 1) + 2] != 0) ? lens[4 * (i - 1) + 2] : 1.f;
            if (!Refract(Normalize(-d), n, eta_t / eta_i, nullptr, &w)) return 0;
            d = w;
        }
    }
    *oOut = mk(o.x, o.y, -o.z);
    *dOut = mk(d.x, d.y, -d.z);
    return 1;
}

// ConvexQuadf (util/vecmath.h:1947-2092): four corners, sorted counter-clockwise about their centroid at
// construction (insertion sort on atan2 in double), Contains = no negative signed area to any edge.
struct ConvexQuad {
    float c[4][2];
};
B2_HD ConvexQuad MakeConvexQuad(const float in[4][2]) {
    ConvexQuad q;
    for (int i = 0; i < 4; ++i) q.c[i][0] = in[i][0], q.c[i][1] = in[i][1];
    float cx = (((q.c[0][0] + q.c[1][0]) + q.c[2][0]) + q.c[3][0]) * 0.25f;
    float cy = (((q.c[0][1] + q.c[1][1]) + q.c[2][1]) + q.c[3][1]) * 0.25f;
    auto ccw = [&](const float *a, const float *b) {
        return atan2(double(a[1] - cy), double(a[0] - cx)) < atan2(double(b[1] - cy), double(b[0] - cx));
    };
    for (int i = 0; i < 4; ++i) {
        int j = i;
        while (j > 0 && !ccw(q.c[j - 1], q.c[j])) {
            float tx = q.c[j][0], ty = q.c[j][1];
            q.c[j][0] = q.c[j - 1][0], q.c[j][1] = q.c[j - 1][1];
            q.c[j - 1][0] = tx, q.c[j - 1][1] = ty;
            --j;
        }
    }
    return q;
}
B2_HD bool ConvexQuadContains(const ConvexQuad &q, float px, float py) {
    for (int i = 0; i < 4; ++i) {
        float e0x = q.c[i][0] - px, e0y = q.c[i][1] - py;
        float e1x = q.c[(i + 1) & 3][0] - px, e1y = q.c[(i + 1) & 3][1] - py;
        float doublearea = (e0x * e1y) - (e0y * e1x);
        if (doublearea < 0) return false;
    }
    return true;
}

// OmniCamera::TraceElement (cameras.cpp:2823-2867): the ray goes into the interface's own space, the hit distance
// and normal are found there; the aperture circle and (microlenses) the lens's quad are tested on that hit point.
// 0 = HIT, else missed or culled.  xf = {transform[16], inverse[16]}.
B2_HD int OmniTraceElement(const float *el, const float *xf, V3 o, V3 d, float elementZ, bool isStop, const ConvexQuad *bounds,
                           float *t, V3 *n) {
    const float *xfInv = xf + 16;
    V3 eo, ed;
    TransformRay(xfInv, o, d, eo, ed);
    *n = mk(0, 0, 0);
    if (isStop) {
        *t = (elementZ - eo.z) / ed.z;
        if (ed.z == 0.0f || *t < 0) return 1;
    } else {
        if (!IntersectSphericalElement(el[0], elementZ + el[0], eo, ed, t, n)) return 1;
    }
    // element.transform(n): a Normal3f goes through the inverse's transpose (util/transform.h:329-334)
    V3 nn = *n;
    *n = mk(xfInv[0] * nn.x + xfInv[4] * nn.y + xfInv[8] * nn.z, xfInv[1] * nn.x + xfInv[5] * nn.y + xfInv[9] * nn.z,
            xfInv[2] * nn.x + xfInv[6] * nn.y + xfInv[10] * nn.z);
    V3 eHit = eo + ed * *t;
    if (eHit.x * eHit.x + eHit.y * eHit.y > el[3] * el[3]) return 2;
    if (bounds && !ConvexQuadContains(*bounds, eHit.x, eHit.y)) return 2;
    return 0;
}

// OmniCamera::TraceLensesFromFilm (cameras.cpp:2869-2989) over `tbl` (n interfaces, transforms at xfs + 32 i): as the
// RealisticCamera's, but every ray transform is a Transform::operator()(Ray) (lensFromCamera in, its inverse out,
// the interface's inverse per element - each nudges the origin along d by its error bound, identity included),
// the hit distance found on the element-space ray is applied to the lens-space ray, and the aperture is tested in
// both spaces.  eta_t comes from mainTbl: the reference indexes the MAIN lens's interfaces there whichever
// table it walks (:2946-2948).  The chromatic-aberration switch is not restated: it reads Ray::wavelength of a
// default-constructed Ray (cameras.cpp:3352, ray.h:29), indeterminate on the CPU build, a no-op at the documented 550 nm.
B2_HD float OmniTraceTable(const float *tbl, const float *xfs, int n, const float *mainTbl, const float *lensFromCamera,
                           const float *cameraFromLens, const ConvexQuad *bounds, V3 co, V3 cd, V3 *oOut, V3 *dOut) {
    float elementZ = 0;
    V3 o, d;
    TransformRay(lensFromCamera, co, cd, o, d);
    for (int i = n - 1; i >= 0; --i) {
        const float *el = tbl + 4 * i;
        elementZ -= el[1];
        bool isStop = el[0] == 0;
        float t;
        V3 nrm;
        if (OmniTraceElement(el, xfs + 32 * i, o, d, elementZ, isStop, bounds, &t, &nrm) != 0) return 0;
        V3 pHit = o + d * t;
        if (SqrF(pHit.x) + SqrF(pHit.y) > SqrF(el[3])) return 0;
        if (!isStop) {
            o = pHit;
            V3 w;
            float eta_i = el[2];
            float eta_t = (i > 0 && mainTbl[4 * (i - 1) + 2] != 0) ? mainTbl[4 * (i - 1) + 2] : 1.f;
            if (!Refract(Normalize(-d), nrm, eta_t / eta_i, nullptr, &w)) return 0;
            d = w;
        }
    }
    TransformRay(cameraFromLens, o, d, *oOut, *dOut);
    return 1;
}

B2_HD float OmniTraceLensesFromFilm(const float *lens, int nLens, V3 co, V3 cd, V3 *oOut, V3 *dOut) {
    const float flipZ[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, -1, 0, 0, 0, 0, 1};
    return OmniTraceTable(lens, lens + 4 * nLens, nLens, lens, flipZ, flipZ, nullptr, co, cd, oOut, dOut);
}

// ---- the microlens array (light-field sensor)
struct MicrolensEl {
    float ix, iy;  // index (a Point2f in the reference)
    float cx, cy;  // centre on the film plane
    ConvexQuad bounds;  // centred on (cx, cy)
};
B2_HD const float *MicroTable(const CameraDev &c) { return c.lens + 36 * c.nLens; }
// OmniCamera::MicrolensIndex (cameras.cpp:3152-3165)
B2_HD void MicrolensIndex(const CameraDev &c, float px, float py, float *ix, float *iy) {
    float tx = px - c.physicalExtent[0], ty = py - c.physicalExtent[1];
    float sx = c.physicalExtent[2] - c.physicalExtent[0], sy = c.physicalExtent[3] - c.physicalExtent[1];
    float nx = tx / sx, ny = ty / sy;
    *ix = floorf(nx * float(c.microDimX));
    *iy = floorf(ny * float(c.microDimY));
}
// OmniCamera::MicrolensCenterFromIndex (cameras.cpp:3194-3219)
B2_HD void MicrolensCenter(const CameraDev &c, float ix, float iy, float *cx, float *cy) {
    float fx = ix + 0.5f, fy = iy + 0.5f;
    float nx = fx / c.microDimX, ny = fy / c.microDimY;
    *cx = Lerp(nx, c.physicalExtent[0], c.physicalExtent[2]);
    *cy = Lerp(ny, c.physicalExtent[1], c.physicalExtent[3]);
    if (ix >= 0 && iy >= 0 && ix < c.microDimX && iy < c.microDimY) {
        const float *off = MicroTable(c) + 36 * c.nMicro + 2 * size_t(iy * c.microDimX + ix);
        *cx += off[0];
        *cy += off[1];
    }
}
// OmniCamera::MicrolensElementFromIndex (cameras.cpp:3221-3244): the lens's quad = the midpoints with its neighbours
B2_HD MicrolensEl MicrolensElementFromIndex(const CameraDev &c, float ix, float iy) {
    MicrolensEl e;
    e.ix = ix;
    e.iy = iy;
    MicrolensCenter(c, ix, iy, &e.cx, &e.cy);
    const float offs[4][2] = {{0, 0}, {0, 1}, {1, 0}, {1, 1}};
    float corners[4][2];
    for (int i = 0; i < 4; ++i) {
        float sx = 0, sy = 0;
        float bx = ix - float(int(offs[i][0])), by = iy - float(int(offs[i][1]));
        for (int k = 0; k < 4; ++k) {
            float ncx, ncy;
            MicrolensCenter(c, bx + offs[k][0], by + offs[k][1], &ncx, &ncy);
            sx += ncx;
            sy += ncy;
        }
        sx *= 0.25f;
        sy *= 0.25f;
        corners[i][0] = sx + -e.cx;
        corners[i][1] = sy + -e.cy;
    }
    e.bounds = MakeConvexQuad(corners);
    return e;
}
// Scale(1, 1, -1) * Translate(-centre) and its inverse (cameras.cpp:3279-3280)
B2_HD void MicrolensFromCamera(const MicrolensEl &e, float m[16], float mInv[16]) {
    const float a[16] = {1, 0, 0, -e.cx, 0, 1, 0, -e.cy, 0, 0, -1, 0, 0, 0, 0, 1};
    const float b[16] = {1, 0, 0, e.cx, 0, 1, 0, e.cy, 0, 0, -1, 0, 0, 0, 0, 1};
    for (int k = 0; k < 16; ++k) m[k] = a[k], mInv[k] = b[k];
}

// OmniCamera::TraceFullLensSystemFromFilm (cameras.cpp:3259-3327): among the (2R+1)^2 microlenses around the one
// the film ray points at, the one whose back surface the ray meets first (TToBackLens, :2594-2611) is traced,
// then the main lens.
B2_HD float OmniTraceFullLensSystem(const CameraDev &c, V3 co, V3 cd, V3 *oOut, V3 *dOut) {
    if (c.nMicro == 0) return OmniTraceLensesFromFilm(c.lens, c.nLens, co, cd, oOut, dOut);
    const float *micro = MicroTable(c), *microXf = micro + 4 * c.nMicro;
    // ComputeMicrolensElement (:3246-3257)
    V3 pm = co + cd * (c.microOffsetFromSensor / cd.z);
    float cix, ciy;
    MicrolensIndex(c, pm.x, pm.y, &cix, &ciy);
    MicrolensEl toTrace = MicrolensElementFromIndex(c, cix, ciy);
    float tMin = INFINITY;
    const int R = c.microRadius;
    const float *last = micro + 4 * (c.nMicro - 1);
    for (int y = -R; y <= R; ++y)
        for (int x = -R; x <= R; ++x) {
            MicrolensEl el = MicrolensElementFromIndex(c, cix + float(x), ciy + float(y));
            float m[16], mInv[16];
            MicrolensFromCamera(el, m, mInv);
            // TToBackLens: the last interface only, never treated as a stop
            V3 lo, ld, nrm;
            TransformRay(m, co, cd, lo, ld);
            float t;
            float newT = OmniTraceElement(last, microXf + 32 * (c.nMicro - 1), lo, ld, -last[1], false, &el.bounds, &t, &nrm) == 0
                             ? t
                             : INFINITY;
            if (newT < tMin) {
                tMin = newT;
                toTrace = el;
            }
        }
    if (!(tMin < INFINITY)) return 0;
    float m[16], mInv[16];
    MicrolensFromCamera(toTrace, m, mInv);
    V3 ao, ad;
    if (OmniTraceTable(micro, microXf, c.nMicro, c.lens, m, mInv, &toTrace.bounds, co, cd, &ao, &ad) == 0) return 0;
    return OmniTraceLensesFromFilm(c.lens, c.nLens, ao, ad, oOut, dOut);
}

// RealisticCamera::GenerateRay (cameras.cpp:990-1019) with SampleExitPupil (:968-988); OmniCamera::GenerateRay
// (:3329-3373) is the same around its own lens trace.  Returns the
// ray's weight; 0 = no ray (vignetted or no exit pupil at this film radius).
B2_HD float GenerateRealisticRay(const CameraDev &c, const FilmDev &film, float pFilmX, float pFilmY, float uLens0,
                                 float uLens1, V3 &o, V3 &d) {
    float sx = pFilmX / film.fullResX, sy = pFilmY / film.fullResY;
    // Bounds2f::Lerp
    float px = Lerp(sx, c.physicalExtent[0], c.physicalExtent[2]);
    float py = Lerp(sy, c.physicalExtent[1], c.physicalExtent[3]);
    V3 pFilm = mk(-px, py, 0);
    if (c.type == 2 && c.nMicro > 0) {
        // OmniCamera::GenerateRay's microlens branch: SampleMicrolensPupil (cameras.cpp:3167-3187) aims the film ray at
        // the microlens plane over the (2R+1)^2 lenses around the pixel's; weight cos^4 * (2R+1)^2 (:3362-3366)
        float ix, iy;
        MicrolensIndex(c, pFilm.x, pFilm.y, &ix, &iy);
        float R = float(c.microRadius);
        ix -= R;
        iy -= R;
        float diameter = float(double(R) * 2.0 + 1.0);
        ix = ix + uLens0 * diameter;
        iy = iy + uLens1 * diameter;
        float sxm = ix / float(c.microDimX), sym = iy / float(c.microDimY);
        V3 pPupil = mk(Lerp(sxm, c.physicalExtent[0], c.physicalExtent[2]), Lerp(sym, c.physicalExtent[1], c.physicalExtent[3]),
                       c.microOffsetFromSensor);
        V3 fd = pPupil - pFilm;
        V3 ro, rd;
        float weight = OmniTraceFullLensSystem(c, pFilm, fd, &ro, &rd);
        if (weight == 0) return 0;
        TransformRay(c.renderFromCamera, ro, rd, o, d);
        d = Normalize(d);
        float cosT = Normalize(fd).z;
        float c2 = cosT * cosT;
        float simDiameter = float(double(c.microRadius) * 2.0 + 1.0);
        weight *= (c2 * c2) * (simDiameter * simDiameter);
        return weight;
    }
    // SampleExitPupil
    float rFilm = sqrtf(SqrF(pFilm.x) + SqrF(pFilm.y));
    int rIndex = int(rFilm / (c.filmDiagonal / 2) * c.nPupilBounds);
    rIndex = rIndex < c.nPupilBounds - 1 ? rIndex : c.nPupilBounds - 1;
    const float *pb = c.pupilBounds + 4 * rIndex;
    if (pb[0] >= pb[2] || pb[1] >= pb[3]) return 0;  // Bounds2::IsDegenerate
    float lx = Lerp(uLens0, pb[0], pb[2]), ly = Lerp(uLens1, pb[1], pb[3]);
    float pdf = 1 / ((pb[2] - pb[0]) * (pb[3] - pb[1]));
    float sinTheta = (rFilm != 0) ? pFilm.y / rFilm : 0.f;
    float cosTheta = (rFilm != 0) ? pFilm.x / rFilm : 1.f;
    float lensRearZ = c.lens[4 * (c.nLens - 1) + 1];
    V3 pPupil = mk(cosTheta * lx - sinTheta * ly, sinTheta * lx + cosTheta * ly, lensRearZ);
    V3 fd = pPupil - pFilm;
    V3 ro, rd;
    float weight = c.type == 2 ? OmniTraceLensesFromFilm(c.lens, c.nLens, pFilm, fd, &ro, &rd)
                               : TraceLensesFromFilm(c.lens, c.nLens, pFilm, fd, &ro, &rd);
    if (weight == 0) return 0;
    TransformRay(c.renderFromCamera, ro, rd, o, d);
    d = Normalize(d);
    float cosT = Normalize(fd).z;
    float c2 = cosT * cosT;
    weight *= (c2 * c2) / (pdf * SqrF(lensRearZ));  // Pow<4>(cosTheta) = Sqr(Sqr(cosTheta))
    return weight;
}

}  // namespace b200pt
