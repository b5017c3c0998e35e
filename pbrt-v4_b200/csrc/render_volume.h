// render_volume.h - participating media for the volumetric path integrator: Henyey-Greenstein phase
// function (media.h:41-69, util/scattering.h:49-58, util/sampling.cpp:347-373), majorant iteration over a
// homogeneous medium or a density grid (media.h:79-215: HomogeneousMajorantIterator, DDAMajorantIterator),
// SampleT_maj (media.h:734-804) and the two callbacks VolPathIntegrator passes to it: delta tracking along the
// path (cpu/integrators.cpp:968-1075) and ratio tracking along a light-sample ray (:1334-1389).
// Host+device; tests/native/rendercheck.cpp loops the same functions on the CPU.
#pragma once
#include "instance.h"
#include "render_layered.h"


namespace b200pt {

constexpr float kInv4Pi = 0.07957747154594766788f;

// Hash(Float) (util/hash.h:95-103: MurmurHash64A over the argument's bytes)
B2_HD uint64_t Hash1f(float a) {
    unsigned char buf[4];
    memcpy(buf, &a, 4);
    return Murmur64A(buf, 4, 0);
}

// HenyeyGreenstein (util/scattering.h:49-58); Clamp(g, -.99, .99) compares in double and returns Float
B2_HD float ClampG(float g) {
    if ((double)g < -.99) return (float)-.99;
    if ((double)g > .99) return (float).99;
    return g;
}
B2_HD float HenyeyGreenstein(float cosTheta, float g) {
    g = ClampG(g);
    float denom = 1 + SqrF(g) + 2 * g * cosTheta;
    return kInv4Pi * (1 - SqrF(g)) / (denom * SafeSqrt(denom));
}
// SampleHenyeyGreenstein (util/sampling.cpp:347-373); returns wi, *pdf = the phase function's value
B2_HD V3 SampleHenyeyGreenstein(V3 wo, float g, float u0, float u1, float *pdf) {
    g = ClampG(g);
    float cosTheta;
    if (fabsf(g) < 1e-3f)
        cosTheta = 1 - 2 * u0;
    else
        cosTheta = -1 / (2 * g) * (1 + SqrF(g) - SqrF((1 - SqrF(g)) / (1 + g - 2 * g * u0)));
    float sinTheta = SafeSqrt(1 - SqrF(cosTheta));
    float phi = 2 * kPi * u1;
    V3 x, y;
    CoordinateSystem(wo, &x, &y);
    // SphericalDirection (util/vecmath.h:1668-1674) then Frame::FromLocal
    float sx = Clampf(sinTheta, -1, 1) * cosf(phi), sy = Clampf(sinTheta, -1, 1) * sinf(phi), sz = Clampf(cosTheta, -1, 1);
    V3 wi = sx * x + sy * y + sz * wo;
    *pdf = HenyeyGreenstein(cosTheta, g);
    return wi;
}

// sigma_a + sigma_s of a medium at wavelength lam (DenselySampledSpectrum::Sample: nearest, already scaled)
B2_HD float MediumSigmaA(const RenderSceneDev &rs, const MediumDev &m, float lam) { return SpectrumAt(rs.spectra[m.sigmaA], lam); }
B2_HD float MediumSigmaS(const RenderSceneDev &rs, const MediumDev &m, float lam) { return SpectrumAt(rs.spectra[m.sigmaS], lam); }

// SampledGrid<Float>::Lookup(Point3f) (util/containers.h:804-835): trilinear, zero outside
B2_HD float GridAt(const MediumDev &m, int x, int y, int z) {
    if (x < 0 || y < 0 || z < 0 || x >= m.nx || y >= m.ny || z >= m.nz) return 0.f;
    return m.density[((long long)z * m.ny + y) * m.nx + x];
}
B2_HD float GridLookup(const MediumDev &m, V3 p) {
    float sx = p.x * m.nx - .5f, sy = p.y * m.ny - .5f, sz = p.z * m.nz - .5f;
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float dx = sx - fx, dy = sy - fy, dz = sz - fz;
    float d00 = Lerp(dx, GridAt(m, ix, iy, iz), GridAt(m, ix + 1, iy, iz));
    float d10 = Lerp(dx, GridAt(m, ix, iy + 1, iz), GridAt(m, ix + 1, iy + 1, iz));
    float d01 = Lerp(dx, GridAt(m, ix, iy, iz + 1), GridAt(m, ix + 1, iy, iz + 1));
    float d11 = Lerp(dx, GridAt(m, ix, iy + 1, iz + 1), GridAt(m, ix + 1, iy + 1, iz + 1));
    return Lerp(dz, Lerp(dy, d00, d10), Lerp(dy, d01, d11));
}
// Bounds3::Offset (util/vecmath.h)
B2_HD V3 BoundsOffset(const float *b, V3 p) {
    V3 o = mk(p.x - b[0], p.y - b[1], p.z - b[2]);
    if (b[3] > b[0]) o.x /= b[3] - b[0];
    if (b[4] > b[1]) o.y /= b[4] - b[1];
    if (b[5] > b[2]) o.z /= b[5] - b[2];
    return o;
}
// Blackbody(lambda, T) and BlackbodySpectrum(T)(lambda) (util/spectrum.h:69-80, 530-545)
B2_HD float Blackbody(float lambda, float T) {
    if (T <= 0) return 0;
    const float c = 299792458.f, h = 6.62606957e-34f, kb = 1.3806488e-23f;
    float l = lambda * 1e-9f;
    float l2 = l * l, l5 = (l2 * l2) * l;  // Pow<5>
    return (2 * h * c * c) / (l5 * (FastExpRef((h * c) / (l * kb * T)) - 1));
}
B2_HD float BlackbodyNormalized(float lambda, float T) {
    float lambdaMax = 2.8977721e-3f / T;
    float normalizationFactor = 1 / Blackbody(lambdaMax * 1e9f, T);
    return Blackbody(lambda, T) * normalizationFactor;
}
// SampledGrid<Float>::Lookup of an emissive grid medium's LeScale (which = 0; its own resolution) or temperature grid
// (which = 1; the density's resolution) at a render-space point (media.h:303-308): trilinear, zero outside
B2_HD float MediumLeScaleAt(const MediumDev &m, V3 p, int which = 0) {
    V3 q = BoundsOffset(m.bounds, XformPoint(m.mediumFromRender, p));
    const float *grid = which ? m.temperature : m.leScale;
    const int gx = which ? m.nx : m.lnx, gy = which ? m.ny : m.lny, gz = which ? m.nz : m.lnz;
    auto at = [&](int x, int y, int z) -> float {
        if (x < 0 || y < 0 || z < 0 || x >= gx || y >= gy || z >= gz) return 0.f;
        return grid[((long long)z * gy + y) * gx + x];
    };
    float sx = q.x * gx - .5f, sy = q.y * gy - .5f, sz = q.z * gz - .5f;
    float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float dx = sx - fx, dy = sy - fy, dz = sz - fz;
    float d00 = Lerp(dx, at(ix, iy, iz), at(ix + 1, iy, iz));
    float d10 = Lerp(dx, at(ix, iy + 1, iz), at(ix + 1, iy + 1, iz));
    float d01 = Lerp(dx, at(ix, iy, iz + 1), at(ix + 1, iy, iz + 1));
    float d11 = Lerp(dx, at(ix, iy + 1, iz + 1), at(ix + 1, iy + 1, iz + 1));
    return Lerp(dz, Lerp(dy, d00, d10), Lerp(dy, d01, d11));
}
// density scaling sigma_a / sigma_s at a render-space point (GridMedium::SamplePoint, media.h:288-300)
B2_HD float MediumDensityAt(const MediumDev &m, V3 p) {
    if (m.type == 0) return 1.f;
    V3 q = XformPoint(m.mediumFromRender, p);  // Transform::ApplyInverse(Point3f) (util/transform.h:387-398)
    return GridLookup(m, BoundsOffset(m.bounds, q));
}

// One majorant segment [tMin, tMax) with sigma_maj = (sigma_a + sigma_s) * scale
struct MajorantSeg {
    float tMin, tMax, scale;
};

// HomogeneousMajorantIterator / DDAMajorantIterator behind one interface (media.h:79-215)
struct MajorantIter {
    int kind;  // 0 = exhausted, 1 = homogeneous (one segment), 2 = DDA
    float tMin, tMax;
    float nextCrossingT[3], deltaT[3];
    int step[3], voxelLimit[3], voxel[3];
};

// Bounds3::IntersectP(o, d, tMax, &t0, &t1) (util/vecmath.h:1549-1573)
B2_HD bool BoundsIntersectP(const float *b, const float o[3], const float d[3], float tMax, float *hitt0, float *hitt1) {
    float t0 = 0, t1 = tMax;
    for (int i = 0; i < 3; ++i) {
        float invRayDir = 1 / d[i];
        float tNear = (b[i] - o[i]) * invRayDir;
        float tFar = (b[3 + i] - o[i]) * invRayDir;
        if (tNear > tFar) {
            float t = tNear;
            tNear = tFar;
            tFar = t;
        }
        tFar *= 1 + 2 * GammaF(3);
        t0 = tNear > t0 ? tNear : t0;
        t1 = tFar < t1 ? tFar : t1;
        if (t0 > t1) return false;
    }
    *hitt0 = t0;
    *hitt1 = t1;
    return true;
}

// Medium::SampleRay for a ray with normalised direction (HomogeneousMedium media.h:249-254, GridMedium :323-337)
B2_HD MajorantIter MediumSampleRay(const MediumDev &m, V3 ro, V3 rd, float raytMax) {
    MajorantIter it;
    it.kind = 0;
    if (m.type == 0) {
        it.kind = 1;
        it.tMin = 0;
        it.tMax = raytMax;
        return it;
    }
    float o[3], d[3], tMaxM;
    InstanceRay(m.mediumFromRender, ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, raytMax, o, d, tMaxM);  // ApplyInverse(Ray, &tMax)
    float t0, t1;
    if (!BoundsIntersectP(m.bounds, o, d, tMaxM, &t0, &t1)) return it;
    it.kind = 2;
    it.tMin = t0;
    it.tMax = t1;
    // DDAMajorantIterator constructor (media.h:141-178)
    float diag[3] = {m.bounds[3] - m.bounds[0], m.bounds[4] - m.bounds[1], m.bounds[5] - m.bounds[2]};
    V3 og = BoundsOffset(m.bounds, mk(o[0], o[1], o[2]));
    float go[3] = {og.x, og.y, og.z};
    float gd[3] = {d[0] / diag[0], d[1] / diag[1], d[2] / diag[2]};
    for (int axis = 0; axis < 3; ++axis) {
        float gi = go[axis] + gd[axis] * t0;  // rayGrid(tMin)
        int res = m.mres[axis];
        // Clamp(Float, int, int) returns Float; assigned to the int voxel
        float v = gi * res;
        it.voxel[axis] = (int)(v < 0 ? 0.f : (v > float(res - 1) ? float(res - 1) : v));
        it.deltaT[axis] = 1 / (fabsf(gd[axis]) * res);
        if (gd[axis] == -0.f) gd[axis] = 0.f;
        if (gd[axis] >= 0) {
            float nextVoxelPos = float(it.voxel[axis] + 1) / res;
            it.nextCrossingT[axis] = t0 + (nextVoxelPos - gi) / gd[axis];
            it.step[axis] = 1;
            it.voxelLimit[axis] = res;
        } else {
            float nextVoxelPos = float(it.voxel[axis]) / res;
            it.nextCrossingT[axis] = t0 + (nextVoxelPos - gi) / gd[axis];
            it.step[axis] = -1;
            it.voxelLimit[axis] = -1;
        }
    }
    return it;
}

// MajorantIterator::Next
B2_HD bool MajorantNext(const MediumDev &m, MajorantIter &it, MajorantSeg &seg) {
    if (it.kind == 0) return false;
    if (it.kind == 1) {
        it.kind = 0;
        seg.tMin = it.tMin;
        seg.tMax = it.tMax;
        seg.scale = 1.f;
        return true;
    }
    if (it.tMin >= it.tMax) return false;
    int bits = ((it.nextCrossingT[0] < it.nextCrossingT[1]) << 2) + ((it.nextCrossingT[0] < it.nextCrossingT[2]) << 1) +
               ((it.nextCrossingT[1] < it.nextCrossingT[2]));
    // cmpToAxis = {2, 1, 2, 1, 2, 2, 0, 0}
    int stepAxis = (bits == 1 || bits == 3) ? 1 : (bits >= 6 ? 0 : 2);
    float tVoxelExit = fminf(it.tMax, it.nextCrossingT[stepAxis]);
    seg.scale = m.majorants[it.voxel[0] + m.mres[0] * (it.voxel[1] + m.mres[1] * it.voxel[2])];
    seg.tMin = it.tMin;
    seg.tMax = tVoxelExit;
    it.tMin = tVoxelExit;
    if (it.nextCrossingT[stepAxis] > it.tMax) it.tMin = it.tMax;
    it.voxel[stepAxis] += it.step[stepAxis];
    if (it.voxel[stepAxis] == it.voxelLimit[stepAxis]) it.tMin = it.tMax;
    it.nextCrossingT[stepAxis] += it.deltaT[stepAxis];
    return true;
}

// SampleT_maj (media.h:734-804) as a resumable walk.  cb(p, density, majorantScale, T_maj) is called at every
// tentative collision; returning false ends the walk (T_maj is then all ones, as the reference returns
// SampledSpectrum(1)).  sigT[k] = sigma_a + sigma_s at the path's wavelengths (the caller fills it once per ray).
// Step() is one turn of the reference's inner loop: fetch the next majorant segment when none is open, then one
// exponential step inside it - so a warp whose lanes walk different rays (or are refilled with new ones as they
// finish, the persistent kernels of render.cu) still executes one common body.  SampleTMaj() below runs it to the
// end: same operations in the same order as the reference's nested loops.
enum { TMAJ_GO_ON = 0, TMAJ_RAN_OUT = 1, TMAJ_STOPPED = 2 };
struct TMajWalk {
    V3 ro, rd;
    MajorantIter it;
    MajorantSeg seg;
    float tMin, u;
    bool inSeg;
    B2_HD void Begin(const MediumDev &m, V3 o, V3 d, float tMax, float u0, float *T_maj) {
        tMax *= Length(d);
        ro = o;
        rd = Normalize(d);
        it = MediumSampleRay(m, ro, rd, tMax);
        u = u0;
        inSeg = false;
        tMin = 0;
        seg.tMin = seg.tMax = seg.scale = 0;
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) T_maj[k] = 1.f;
    }
    // TMAJ_RAN_OUT: the segments are used up, T_maj holds the transmittance since the last collision;
    // TMAJ_STOPPED: the callback ended the walk, T_maj is all ones
    template <class CB>
    B2_HD int Step(const MediumDev &m, const float *sigT, PCG32 &rng, float *T_maj, CB &cb) {
        if (!inSeg) {
            if (!MajorantNext(m, it, seg)) return TMAJ_RAN_OUT;
            const float sigMaj0 = m.type == 0 ? sigT[0] : sigT[0] * seg.scale;
            if (sigMaj0 == 0) {
                float dt = seg.tMax - seg.tMin;
                if (std::isinf(dt)) dt = 3.402823466e+38f;
B2_SPECTRAL_LOOP
                for (int k = 0; k < kNLambda; ++k) T_maj[k] *= FastExpRef(-dt * (m.type == 0 ? sigT[k] : sigT[k] * seg.scale));
                return TMAJ_GO_ON;
            }
            tMin = seg.tMin;
            inSeg = true;
        }
        const float sigMaj0 = m.type == 0 ? sigT[0] : sigT[0] * seg.scale;
        float t = tMin + (-logf(1 - u) / sigMaj0);  // SampleExponential (util/sampling.h:222-225)
        u = PcgFloat(rng);
        if (t < seg.tMax) {
            const float nd = -(t - tMin);
B2_SPECTRAL_LOOP
            for (int k = 0; k < kNLambda; ++k) T_maj[k] *= FastExpRef(nd * (m.type == 0 ? sigT[k] : sigT[k] * seg.scale));
            V3 p = ro + rd * t;
            const bool goOn = cb(p, MediumDensityAt(m, p), m.type == 0 ? 1.f : seg.scale, T_maj);
B2_SPECTRAL_LOOP
            for (int k = 0; k < kNLambda; ++k) T_maj[k] = 1.f;
            if (!goOn) return TMAJ_STOPPED;
            tMin = t;
        } else {
            float dt = seg.tMax - tMin;
            if (std::isinf(dt)) dt = 3.402823466e+38f;
B2_SPECTRAL_LOOP
            for (int k = 0; k < kNLambda; ++k) T_maj[k] *= FastExpRef(-dt * (m.type == 0 ? sigT[k] : sigT[k] * seg.scale));
            inSeg = false;
        }
        return TMAJ_GO_ON;
    }
};

// The same walk split for the persistent kernels (render.cu): ADVANCE runs from majorant segment to majorant segment
// with scalar work only - the reference multiplies exp(-dt sigma_maj(lambda)) into T_maj for every segment and all
// 31 wavelengths; here the optical length  S = sum of dt * majorant scale  since the last collision is carried instead
// and T_maj(lambda) = exp(-sigma_t(lambda) S) is formed once per collision (an exponential of the sum instead of a
// product of exponentials: equal up to rounding, inside the image tolerance; the host check keeps Step()) - and stops
// at a tentative collision, so that the expensive spectral part (COLLIDE: T_maj + the callback) can be run for many
// lanes of a warp at once instead of for the two or three that happen to collide in the same step.
enum { TMAJ_COLLISION = 3 };
struct TMajFastWalk {
    TMajWalk w;
    float S, tColl;
    B2_HD void Begin(const MediumDev &m, V3 o, V3 d, float tMax, float u0) {
        float dummy[kNLambda];
        w.Begin(m, o, d, tMax, u0, dummy);
        S = 0;
        tColl = 0;
    }
    B2_HD void Add(float dt, float scale) {
        if (std::isinf(dt)) dt = 3.402823466e+38f;
        S = fminf(S + dt * scale, 3.402823466e+38f);
    }
    // up to maxCells majorant segments; TMAJ_GO_ON (call again), TMAJ_RAN_OUT, or TMAJ_COLLISION at tColl
    B2_HD int Advance(const MediumDev &m, float sigT0, PCG32 &rng, int maxCells) {
        for (int cell = 0; cell < maxCells; ++cell) {
            if (!w.inSeg) {
                if (!MajorantNext(m, w.it, w.seg)) return TMAJ_RAN_OUT;
                const float sigMaj0 = m.type == 0 ? sigT0 : sigT0 * w.seg.scale;
                if (sigMaj0 == 0) {
                    Add(w.seg.tMax - w.seg.tMin, m.type == 0 ? 1.f : w.seg.scale);
                    continue;
                }
                w.tMin = w.seg.tMin;
                w.inSeg = true;
            }
            const float scale = m.type == 0 ? 1.f : w.seg.scale;
            float t = w.tMin + (-logf(1 - w.u) / (sigT0 * scale));
            w.u = PcgFloat(rng);
            if (t < w.seg.tMax) {
                Add(t - w.tMin, scale);
                tColl = t;
                return TMAJ_COLLISION;
            }
            Add(w.seg.tMax - w.tMin, scale);
            w.inSeg = false;
        }
        return TMAJ_GO_ON;
    }
    // transmittance since the last collision
    B2_HD void TMaj(const float *sigT, float *T_maj) const {
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) T_maj[k] = sigT[k] == 0 ? 1.f : FastExpRef(-(sigT[k] * S));
    }
    // the pending collision: false = the callback ended the walk
    template <class CB>
    B2_HD bool Collide(const MediumDev &m, const float *sigT, float *T_maj, CB &cb) {
        TMaj(sigT, T_maj);
        V3 p = w.ro + w.rd * tColl;
        const bool goOn = cb(p, MediumDensityAt(m, p), m.type == 0 ? 1.f : w.seg.scale, T_maj);
        S = 0;
        w.tMin = tColl;
        return goOn;
    }
};

template <class CB>
B2_HD void SampleTMaj(const MediumDev &m, const float *sigT, V3 ro, V3 rd, float tMax, float u, PCG32 &rng, float *T_maj,
                      CB &cb) {
    TMajWalk w;
    w.Begin(m, ro, rd, tMax, u, T_maj);
    while (w.Step(m, sigT, rng, T_maj, cb) == TMAJ_GO_ON) {
    }
}

// SampleDiscrete over {pAbsorb, pScatter, pNull} (util/sampling.h:79-113)
B2_HD int SampleDiscrete3(float w0, float w1, float w2, float u) {
    float sumWeights = 0;
    sumWeights += w0;
    sumWeights += w1;
    sumWeights += w2;
    float up = u * sumWeights;
    if (up == sumWeights) up = NextDown(up);
    float w[3] = {w0, w1, w2};
    int offset = 0;
    float sum = 0;
    while (offset < 2 && sum + w[offset] <= up) sum += w[offset++];
    return offset;
}

// The medium's spectra at the path's wavelengths
B2_HD void MediumSpectra(const RenderSceneDev &rs, const MediumDev &m, float lambdaU, float *sigA, float *sigS, float *sigT) {
    LambdaIter li(lambdaU);
B2_SPECTRAL_LOOP
    for (int k = 0; k < kNLambda; ++k, li.Next()) {
        sigA[k] = MediumSigmaA(rs, m, li.lam);
        sigS[k] = MediumSigmaS(rs, m, li.lam);
        sigT[k] = sigA[k] + sigS[k];
    }
}

// Delta tracking along the path ray: the callback of VolPathIntegrator::Li (cpu/integrators.cpp:976-1063).  Real scattering only updates beta / r_u and records the point; the caller samples the light
// and the phase function right after the walk ends (nothing else happens in between in the reference).
template <class IO>
struct PathMediumCB {
    IO &io;
    PCG32 &rng;
    const float *sigA, *sigS, *sigT;
    int &depth;
    int maxDepth;
    bool terminated = false, scattered = false;
    V3 pScatter;
    const RenderSceneDev &rs;  // an emissive medium's Le_spec at the path's wavelengths
    const MediumDev &m;
    float lambdaU;
    B2_HD PathMediumCB(IO &i, PCG32 &r, const float *a, const float *s, const float *t, int &d, int md, const RenderSceneDev &scene,
                       const MediumDev &medium, float lu)
        : io(i), rng(r), sigA(a), sigS(s), sigT(t), depth(d), maxDepth(md), rs(scene), m(medium), lambdaU(lu) { pScatter = mk(0, 0, 0); }
    // "Add emission from medium scattering event" (cpu/integrators.cpp:990-1001): at EVERY tentative collision
    B2_HD void AddEmission(V3 p, float density, float majScale, const float *T_maj) {
        float scale = 1.f;
        float temp = 0;  // > 0: blackbody emission at that temperature instead of Le_spec
        if (m.leScale) {  // GridMedium::SamplePoint (media.h:301-318): nothing where LeScale <= 0
            scale = MediumLeScaleAt(m, p);
            if (!(scale > 0)) return;
            if (m.temperature) {
                temp = (MediumLeScaleAt(m, p, 1) - m.temperatureOffset) * m.temperatureScale;
                if (!(temp > 100.f)) return;
            }
        }
        const float pdf = (sigT[0] * majScale) * T_maj[0];
        float sumRe = 0;
        bool anyRe = false, anyLe = false;
        LambdaIter li(lambdaU);
        for (int k = 0; k < kNLambda; ++k, li.Next()) {
            float r_e = io.Ru(k) * (sigT[k] * majScale) * T_maj[k] / pdf;
            sumRe += r_e;
            anyRe |= r_e != 0;
            anyLe |= scale * (temp > 0 ? BlackbodyNormalized(li.lam, temp) : SpectrumAt(rs.spectra[m.le], li.lam)) != 0;
        }
        if (!anyLe || !anyRe) return;
        const float avg = sumRe / kNLambda;
        LambdaIter lj(lambdaU);
        for (int k = 0; k < kNLambda; ++k, lj.Next()) {
            float Le = scale * (temp > 0 ? BlackbodyNormalized(lj.lam, temp) : SpectrumAt(rs.spectra[m.le], lj.lam));
            float betap = io.Beta(k) * T_maj[k] / pdf;
            io.AddL(k, betap * (sigA[k] * density) * Le / avg);
        }
    }
    B2_HD bool AnyBeta() const {
        bool any = false;
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) any |= io.Beta(k) != 0;
        return any;
    }
    B2_HD bool AnyRu() const {
        bool any = false;
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) any |= io.Ru(k) != 0;
        return any;
    }
    B2_HD bool operator()(V3 p, float density, float majScale, const float *T_maj) {
        if (!AnyBeta()) {
            terminated = true;
            return false;
        }
        if (m.emissive && depth < maxDepth) AddEmission(p, density, majScale, T_maj);
        const float sigMaj0 = sigT[0] * majScale;
        float pAbsorb = (sigA[0] * density) / sigMaj0;
        float pScat = (sigS[0] * density) / sigMaj0;
        float pNull = fmaxf(0.f, 1 - pAbsorb - pScat);
        float um = PcgFloat(rng);
        int mode = SampleDiscrete3(pAbsorb, pScat, pNull, um);
        if (mode == 0) {
            terminated = true;
            return false;
        }
        if (mode == 1) {
            if (depth++ >= maxDepth) {
                terminated = true;
                return false;
            }
            float pdf = T_maj[0] * (sigS[0] * density);
B2_SPECTRAL_LOOP
            for (int k = 0; k < kNLambda; ++k) {
                float f = T_maj[k] * (sigS[k] * density) / pdf;
                io.SetBeta(k, io.Beta(k) * f);
                io.SetRu(k, io.Ru(k) * f);
            }
            if (AnyBeta() && AnyRu()) {
                scattered = true;
                pScatter = p;
            }
            return false;
        }
        float pdf = 0;
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) {
            float sigMaj = sigT[k] * majScale;
            float sigN = fmaxf(0.f, sigMaj - sigA[k] * density - sigS[k] * density);  // ClampZero
            if (k == 0) pdf = T_maj[0] * sigN;
            io.SetBeta(k, pdf == 0 ? 0.f : io.Beta(k) * (T_maj[k] * sigN / pdf));
            io.SetRu(k, io.Ru(k) * (T_maj[k] * sigN / pdf));
            io.SetRl(k, io.Rl(k) * (T_maj[k] * sigMaj / pdf));
        }
        return AnyBeta() && AnyRu();
    }
};

// Ratio tracking along a light-sample ray (cpu/integrators.cpp:1349-1376): T_ray, r_l, r_u are the caller's arrays
struct ShadowMediumCB {
    PCG32 &rng;
    const float *sigA, *sigS, *sigT;
    float *T_ray, *r_l, *r_u;
    B2_HD ShadowMediumCB(PCG32 &r, const float *a, const float *s, const float *t, float *T, float *rl, float *ru)
        : rng(r), sigA(a), sigS(s), sigT(t), T_ray(T), r_l(rl), r_u(ru) {}
    B2_HD bool operator()(V3, float density, float majScale, const float *T_maj) {
        float pdf = T_maj[0] * (sigT[0] * majScale);
        float avg = 0, maxTr = 0;
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) {
            float sigMaj = sigT[k] * majScale;
            float sigN = fmaxf(0.f, sigMaj - sigA[k] * density - sigS[k] * density);
            T_ray[k] *= T_maj[k] * sigN / pdf;
            r_l[k] *= T_maj[k] * sigMaj / pdf;
            r_u[k] *= T_maj[k] * sigN / pdf;
            avg += r_l[k] + r_u[k];
        }
        avg /= kNLambda;  // SampledSpectrum::Average: sum / N
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) maxTr = k == 0 ? T_ray[0] / avg : fmaxf(maxTr, T_ray[k] / avg);
        bool any = false;
        if (maxTr < 0.05f) {
            const float q = 0.75f;
            if (PcgFloat(rng) < q)
B2_SPECTRAL_LOOP
                for (int k = 0; k < kNLambda; ++k) T_ray[k] = 0.f;
            else
B2_SPECTRAL_LOOP
                for (int k = 0; k < kNLambda; ++k) T_ray[k] /= 1 - q;
        }
B2_SPECTRAL_LOOP
        for (int k = 0; k < kNLambda; ++k) any |= T_ray[k] != 0;
        return any;
    }
};

}  // namespace b200pt
