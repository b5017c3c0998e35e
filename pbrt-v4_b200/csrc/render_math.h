// render_math.h - small vector / sampling helpers of the shading path, host+device.
// Each routine names the reference routine it follows; arithmetic order is kept so that, with the
// sampler dimensions consumed in the CPU integrator's order, paths track the reference's paths.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#include "isect.h"

namespace b200pt {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kPiOver2 = 1.57079632679489661923f;
constexpr float kPiOver4 = 0.78539816339744830961f;
constexpr float kOneMinusEps = 0x1.fffffep-1f;
constexpr float kShadowEpsilon = 0.0001f;  // util/math.h:42

struct V3 {
    float x, y, z;
};
B2_HD V3 mk(float x, float y, float z) { return V3{x, y, z}; }
B2_HD V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
B2_HD V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
B2_HD V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
B2_HD V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
B2_HD V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
B2_HD V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }  // util/vecmath.h:366-369
B2_HD float SqrF(float v) { return v * v; }
B2_HD float FMAf(float a, float b, float c) { return B2_FMA(a, b, c); }
B2_HD float SumOfProductsF(float a, float b, float c, float d) {  // util/math.h:577-583
    float cd = c * d;
    float sop = FMAf(a, b, cd);
    float err = FMAf(c, d, -cd);
    return sop + err;
}
B2_HD float Dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // util/vecmath.h:966-969
B2_HD float AbsDot(V3 a, V3 b) { return fabsf(Dot(a, b)); }
// Dot with a Normal3f on either side is a different expression in the reference
// (util/vecmath.h:1059-1077): FMA(n.x, v.x, SumOfProducts(n.y, v.y, n.z, v.z)).
B2_HD float DotN(V3 n, V3 v) { return FMAf(n.x, v.x, SumOfProductsF(n.y, v.y, n.z, v.z)); }
B2_HD float AbsDotN(V3 n, V3 v) { return fabsf(DotN(n, v)); }
B2_HD V3 Cross(V3 v, V3 w) {
    return mk(DiffOfProducts(v.y, w.z, v.z, w.y), DiffOfProducts(v.z, w.x, v.x, w.z),
              DiffOfProducts(v.x, w.y, v.y, w.x));
}
B2_HD float LengthSquared(V3 v) { return SqrF(v.x) + SqrF(v.y) + SqrF(v.z); }
B2_HD float Length(V3 v) { return sqrtf(LengthSquared(v)); }
B2_HD V3 Normalize(V3 v) { return v / Length(v); }
B2_HD V3 Abs(V3 v) { return mk(fabsf(v.x), fabsf(v.y), fabsf(v.z)); }
B2_HD V3 FaceForward(V3 n, V3 v) { return (DotN(n, v) < 0.f) ? -n : n; }  // one side is a Normal3f at every use
B2_HD float Lerp(float t, float a, float b) { return (1 - t) * a + t * b; }
B2_HD float Clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }
B2_HD float SafeSqrt(float x) { return sqrtf(fmaxf(0.f, x)); }
B2_HD float SafeASin(float x) { return asinf(Clampf(x, -1, 1)); }
B2_HD V3 GramSchmidt(V3 v, V3 w) { return v - Dot(v, w) * w; }

// util/vecmath.h:974-981
B2_HD float AngleBetween(V3 v1, V3 v2) {
    if (Dot(v1, v2) < 0)
        return kPi - 2 * SafeASin(Length(v1 + v2) / 2);
    return 2 * SafeASin(Length(v2 - v1) / 2);
}

// util/vecmath.h:1009-1016
B2_HD void CoordinateSystem(V3 v1, V3 *v2, V3 *v3) {
    float sign = copysignf(1.f, v1.z);
    float a = -1 / (sign + v1.z);
    float b = v1.x * v1.y * a;
    *v2 = mk(1 + sign * SqrF(v1.x) * a, sign * b, -sign * v1.x);
    *v3 = mk(b, sign + SqrF(v1.y) * a, -v1.y);
}

B2_HD float NextUp(float v) {
    if (std::isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui;
    memcpy(&ui, &v, 4);
    if (v >= 0) ++ui; else --ui;
    memcpy(&v, &ui, 4);
    return v;
}
B2_HD float NextDown(float v) {
    if (std::isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui;
    memcpy(&ui, &v, 4);
    if (v > 0) --ui; else ++ui;
    memcpy(&v, &ui, 4);
    return v;
}

// A point with its error box: Point3fi(p, err) as the CPU build makes it
// (Interval::FromValueAndError util/math.h:829-837 with NextFloatUp/Down(a +- b), util/float.h).
struct P3i {
    V3 lo, hi;
};
B2_HD void IntervalOf(float v, float err, float &lo, float &hi) {
    if (err == 0) {
        lo = hi = v;
    } else {
        lo = NextDown(v - err);
        hi = NextUp(v + err);
    }
}
B2_HD P3i MakeP3i(V3 p, V3 e) {
    P3i r;
    IntervalOf(p.x, e.x, r.lo.x, r.hi.x);
    IntervalOf(p.y, e.y, r.lo.y, r.hi.y);
    IntervalOf(p.z, e.z, r.lo.z, r.hi.z);
    return r;
}
B2_HD V3 Mid(const P3i &p) { return mk((p.lo.x + p.hi.x) / 2, (p.lo.y + p.hi.y) / 2, (p.lo.z + p.hi.z) / 2); }
B2_HD V3 Err(const P3i &p) { return mk((p.hi.x - p.lo.x) / 2, (p.hi.y - p.lo.y) / 2, (p.hi.z - p.lo.z) / 2); }

// OffsetRayOrigin (ray.h:78-95)
B2_HD_NOINLINE V3 OffsetRayOrigin(const P3i &pi, V3 n, V3 w) {
    float d = DotN(Abs(n), Err(pi));
    V3 offset = d * n;
    if (DotN(n, w) < 0) offset = -offset;
    V3 po = Mid(pi) + offset;
    if (offset.x > 0) po.x = NextUp(po.x); else if (offset.x < 0) po.x = NextDown(po.x);
    if (offset.y > 0) po.y = NextUp(po.y); else if (offset.y < 0) po.y = NextDown(po.y);
    if (offset.z > 0) po.z = NextUp(po.z); else if (offset.z < 0) po.z = NextDown(po.z);
    return po;
}

// ---- hashing (util/hash.h:20-75): MurmurHash64A of two ints, MixBits
B2_HD uint64_t MixBits(uint64_t v) {
    v ^= (v >> 31);
    v *= 0x7fb5d329728ea185ull;
    v ^= (v >> 27);
    v *= 0x81dadef4bc2dd44dull;
    v ^= (v >> 33);
    return v;
}
B2_HD uint64_t Hash2i(int32_t a, int32_t b) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    const int r = 47;
    uint64_t h = 0 ^ (8 * m);
    uint64_t k = (uint64_t)(uint32_t)a | ((uint64_t)(uint32_t)b << 32);
    k *= m;
    k ^= k >> r;
    k *= m;
    h ^= k;
    h *= m;
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return h;
}

// MurmurHash64A over a short buffer (util/hash.h:18-68) and Hash(args...) (:95-103) for the argument
// packs LayeredBxDF uses: (int seed, Vector3f), (Vector3f), (Float, Point2f).
B2_HD uint64_t Murmur64A(const unsigned char *key, size_t len, uint64_t seed) {
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    const int r = 47;
    uint64_t h = seed ^ (len * m);
    size_t nblocks = len / 8;
    for (size_t i = 0; i < nblocks; ++i) {
        uint64_t k;
        memcpy(&k, key + 8 * i, 8);
        k *= m;
        k ^= k >> r;
        k *= m;
        h ^= k;
        h *= m;
    }
    const unsigned char *tail = key + 8 * nblocks;
    int rem = (int)(len & 7);
    for (int i = rem - 1; i >= 0; --i) h ^= uint64_t(tail[i]) << (8 * i);
    if (rem) h *= m;
    h ^= h >> r;
    h *= m;
    h ^= h >> r;
    return h;
}
struct PCG32 {  // util/rng.h:30-135
    uint64_t state, inc;
};
B2_HD uint32_t PcgNext(PCG32 &g) {
    uint64_t oldstate = g.state;
    g.state = oldstate * 0x5851f42d4c957f2dull + g.inc;
    uint32_t xorshifted = (uint32_t)(((oldstate >> 18u) ^ oldstate) >> 27u);
    uint32_t rot = (uint32_t)(oldstate >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
}
B2_HD PCG32 PcgSeed(uint64_t sequenceIndex, uint64_t seed) {
    PCG32 g;
    g.state = 0u;
    g.inc = (sequenceIndex << 1u) | 1u;
    PcgNext(g);
    g.state += seed;
    PcgNext(g);
    return g;
}
// RNG::Advance (util/rng.h:137-150): the generator's state `idelta` draws on, in O(log idelta)
B2_HD void PcgAdvance(PCG32 &g, uint64_t delta) {
    uint64_t curMult = 0x5851f42d4c957f2dull, curPlus = g.inc, accMult = 1u, accPlus = 0u;
    while (delta > 0) {
        if (delta & 1) {
            accMult *= curMult;
            accPlus = accPlus * curMult + curPlus;
        }
        curPlus = (curMult + 1) * curPlus;
        curMult *= curMult;
        delta /= 2;
    }
    g.state = accMult * g.state + accPlus;
}

B2_HD uint32_t ReverseBits32(uint32_t n) {
#if defined(__CUDA_ARCH__)
    return __brev(n);
#else
    n = (n << 16) | (n >> 16);
    n = ((n & 0x00ff00ff) << 8) | ((n & 0xff00ff00) >> 8);
    n = ((n & 0x0f0f0f0f) << 4) | ((n & 0xf0f0f0f0) >> 4);
    n = ((n & 0x33333333) << 2) | ((n & 0xcccccccc) >> 2);
    n = ((n & 0x55555555) << 1) | ((n & 0xaaaaaaaa) >> 1);
    return n;
#endif
}

// ---- sampling (util/sampling.h)
B2_HD float PowerHeuristic(float fPdf, float gPdf) {  // nf = ng = 1 (util/sampling.h:72-77)
    float f = fPdf, g = gPdf;
    if (std::isinf(SqrF(f))) return 1;
    return SqrF(f) / (SqrF(f) + SqrF(g));
}
B2_HD float SampleLinear(float u, float a, float b) {  // util/sampling.h:122-128
    if (u == 0 && a == 0) return 0;
    float x = u * (a + b) / (a + sqrtf(Lerp(u, SqrF(a), SqrF(b))));
    return fminf(x, kOneMinusEps);
}
B2_HD float BilinearPDF(float px, float py, const float *w) {  // util/sampling.h:134-144
    if (px < 0 || px > 1 || py < 0 || py > 1) return 0;
    if (w[0] + w[1] + w[2] + w[3] == 0) return 1;
    return 4 * ((1 - px) * (1 - py) * w[0] + px * (1 - py) * w[1] + (1 - px) * py * w[2] + px * py * w[3]) /
           (w[0] + w[1] + w[2] + w[3]);
}
B2_HD void SampleBilinear(float u0, float u1, const float *w, float &px, float &py) {  // util/sampling.h:146-156
    py = SampleLinear(u1, w[0] + w[1], w[2] + w[3]);
    px = SampleLinear(u0, Lerp(py, w[0], w[2]), Lerp(py, w[1], w[3]));
}
B2_HD void SampleUniformDiskConcentric(float u0, float u1, float &dx, float &dy) {  // util/sampling.h:325-341
    float ox = 2 * u0 - 1, oy = 2 * u1 - 1;
    if (ox == 0 && oy == 0) {
        dx = dy = 0;
        return;
    }
    float theta, r;
    if (fabsf(ox) > fabsf(oy)) {
        r = ox;
        theta = kPiOver4 * (oy / ox);
    } else {
        r = oy;
        theta = kPiOver2 - kPiOver4 * (ox / oy);
    }
    dx = r * cosf(theta);
    dy = r * sinf(theta);
}
B2_HD V3 SampleCosineHemisphere(float u0, float u1) {  // util/sampling.h:409-413
    float dx, dy;
    SampleUniformDiskConcentric(u0, u1, dx, dy);
    float z = SafeSqrt(1 - SqrF(dx) - SqrF(dy));
    return mk(dx, dy, z);
}
B2_HD void SampleUniformTriangle(float u0, float u1, float b[3]) {  // util/sampling.h:173-184
    float b0, b1;
    if (u0 < u1) {
        b0 = u0 / 2;
        b1 = u1 - b0;
    } else {
        b1 = u1 / 2;
        b0 = u0 - b1;
    }
    b[0] = b0;
    b[1] = b1;
    b[2] = 1 - b0 - b1;
}

// SphericalTriangleArea (util/vecmath.h:1640-1643)
B2_HD float SphericalTriangleArea(V3 a, V3 b, V3 c) {
    return fabsf(2 * atan2f(Dot(a, Cross(b, c)), 1 + Dot(a, b) + Dot(a, c) + Dot(b, c)));
}

}  // namespace b200pt
