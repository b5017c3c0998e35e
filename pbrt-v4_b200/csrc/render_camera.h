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

// OmniCamera::TraceLensesFromFilm with TraceElement (cameras.cpp:2823-2989) for the main lens (no microlens array, no
// aperture image, default ConvexQuadf bounds = no extra clipping): as the RealisticCamera's, but every ray
// transform is a Transform::operator()(Ray) (Scale(1,1,-1) in and out, the interface's inverse transform per
// element - each one nudges the origin along d by its error bound, identity included), the hit distance is found
// on the element-space ray and applied to the lens-space ray, and the aperture is tested in both spaces.
// The chromatic-aberration switch is not restated: it reads Ray::wavelength of a default-constructed Ray
// (cameras.cpp:3352, ray.h:29), an indeterminate value on the CPU build; at the documented initial 550 nm it is a no-op.
B2_HD float OmniTraceLensesFromFilm(const float *lens, int nLens, V3 co, V3 cd, V3 *oOut, V3 *dOut) {
    const float flipZ[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, -1, 0, 0, 0, 0, 1};
    float elementZ = 0;
    V3 o, d;
    TransformRay(flipZ, co, cd, o, d);
    for (int i = nLens - 1; i >= 0; --i) {
        const float *el = lens + 4 * i;
        const float *xf = lens + 4 * nLens + 32 * i, *xfInv = xf + 16;
        float curvatureRadius = el[0], eta = el[2], apertureRadius = el[3];
        elementZ -= el[1];
        bool isStop = curvatureRadius == 0;
        // TraceElement
        V3 eo, ed;
        TransformRay(xfInv, o, d, eo, ed);
        float t;
        V3 n = mk(0, 0, 0);
        if (isStop) {
            t = (elementZ - eo.z) / ed.z;
            if (ed.z == 0.0f || t < 0) return 0;
        } else {
            float zCenter = elementZ + curvatureRadius;
            if (!IntersectSphericalElement(curvatureRadius, zCenter, eo, ed, &t, &n)) return 0;
        }
        // element.transform(n): a Normal3f goes through the inverse's transpose (util/transform.h:329-334)
        n = mk(xfInv[0] * n.x + xfInv[4] * n.y + xfInv[8] * n.z, xfInv[1] * n.x + xfInv[5] * n.y + xfInv[9] * n.z,
               xfInv[2] * n.x + xfInv[6] * n.y + xfInv[10] * n.z);
        V3 eHit = eo + ed * t;
        if (eHit.x * eHit.x + eHit.y * eHit.y > apertureRadius * apertureRadius) return 0;
        // back in TraceLensesFromFilm
        V3 pHit = o + d * t;
        if (SqrF(pHit.x) + SqrF(pHit.y) > SqrF(apertureRadius)) return 0;
        if (!isStop) {
            o = pHit;
            V3 w;
            float eta_i = eta;
            float eta_t = (i > 0 && lens[4 * (i - 1) + 2] != 0) ? lens[4 * (i - 1) + 2] : 1.f;
            if (!Refract(Normalize(-d), n, eta_t / eta_i, nullptr, &w)) return 0;
            d = w;
        }
    }
    TransformRay(flipZ, o, d, *oOut, *dOut);
    return 1;
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
