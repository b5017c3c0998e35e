// isect.h - ray/box and ray/triangle arithmetic, operation for operation as the
// reference evaluates it, so that hit t / barycentrics are bit-identical to
// BVHAggregate::Intersect on the CPU (BASELINE.json north_star).
//
//   SlabTest        = Bounds3::IntersectP(o,d,tMax,invDir,dirIsNeg)  util/vecmath.h:1576-1610
//   TriSetup/TriHit = IntersectTriangle                              shapes.cpp:172-273
//   DiffOfProducts  = DifferenceOfProducts                           util/math.h:569-575
//
// Every multiply/add goes through rn-intrinsics on the device (never contracted
// into FMA by ptxas, whatever --fmad says); FMA appears only where the
// reference calls FMA().  The same header compiles for the host (g++
// -ffp-contract=off) so tests can run this exact arithmetic against the
// compiled reference without a GPU (tests/native/hostcheck.cpp).
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDA_ARCH__)
#define B2_HD __host__ __device__ __forceinline__
// out-of-line variant for big helpers; measured slower on B200 (64.0 vs 55.9 ms per config-1 frame:
// call overhead and the ABI register cap), so it is only on with -DB2_OUTLINE_HELPERS
#ifdef B2_OUTLINE_HELPERS
#define B2_HD_NOINLINE __host__ __device__ __noinline__
#else
#define B2_HD_NOINLINE __host__ __device__ __forceinline__
#endif
// really out of line: the layered BxDF's walks call the interface routines from a dozen sites; inlined,
// the coated shade kernel is 48 K instructions and starves on instruction fetch (profiles/)
#ifdef B2_INLINE_ALL
#define B2_HD_CALL __host__ __device__ __forceinline__
#else
#define B2_HD_CALL __host__ __device__ __noinline__
#endif
#define B2_MUL(a, b) __fmul_rn((a), (b))
#define B2_ADD(a, b) __fadd_rn((a), (b))
#define B2_SUB(a, b) __fsub_rn((a), (b))
#define B2_DIV(a, b) __fdiv_rn((a), (b))
#define B2_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define B2_DMUL(a, b) __dmul_rn((a), (b))
#define B2_DSUB(a, b) __dsub_rn((a), (b))
#define B2_D2F(a) __double2float_rn(a)
#else
#if defined(__CUDACC__)
#define B2_HD __host__ __device__ inline
#define B2_HD_NOINLINE __host__ __device__ inline
#define B2_HD_CALL __host__ __device__ inline
#else
#define B2_HD inline
#define B2_HD_NOINLINE inline
#define B2_HD_CALL inline
#endif
#define B2_MUL(a, b) ((a) * (b))
#define B2_ADD(a, b) ((a) + (b))
#define B2_SUB(a, b) ((a) - (b))
#define B2_DIV(a, b) ((a) / (b))
#define B2_FMA(a, b, c) __builtin_fmaf((a), (b), (c))
#define B2_DMUL(a, b) ((a) * (b))
#define B2_DSUB(a, b) ((a) - (b))
#define B2_D2F(a) ((float)(a))
#endif

namespace b200pt {

// gamma(n) = n*eps/(1 - n*eps), eps = 2^-24 (util/float.h:45,197-199), in float.
B2_HD constexpr float GammaF(int n) {
    return (n * 5.9604644775390625e-08f) / (1 - n * 5.9604644775390625e-08f);
}
// 1 + 2*gamma(3) as the reference's float expression (util/vecmath.h:1588)
constexpr float kSlabPad = 1 + 2 * GammaF(3);

B2_HD float DiffOfProducts(float a, float b, float c, float d) {
    float cd = B2_MUL(c, d);
    float dop = B2_FMA(a, b, -cd);
    float err = B2_FMA(-c, d, cd);
    return B2_ADD(dop, err);
}

struct RayCtx {
    float ox, oy, oz;
    float dx, dy, dz;
    float ix, iy, iz;  // 1/d (cpu/aggregates.cpp:534)
    int neg[3];        // invDir < 0
    // IntersectTriangle's per-ray shear terms (shapes.cpp:186-204); they depend
    // on the ray only, so they are hoisted out of the per-triangle work.
    int kx, ky, kz;
    float Sx, Sy, Sz;
};

B2_HD float Comp(float x, float y, float z, int k) {
    return k == 0 ? x : (k == 1 ? y : z);
}

B2_HD void RaySetup(RayCtx &r, float ox, float oy, float oz, float dx, float dy, float dz) {
    r.ox = ox; r.oy = oy; r.oz = oz;
    r.dx = dx; r.dy = dy; r.dz = dz;
    r.ix = B2_DIV(1.f, dx);
    r.iy = B2_DIV(1.f, dy);
    r.iz = B2_DIV(1.f, dz);
    r.neg[0] = r.ix < 0;
    r.neg[1] = r.iy < 0;
    r.neg[2] = r.iz < 0;
    // MaxComponentIndex(Abs(d)) (util/vecmath.h:456-458)
    float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    int kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    int kx = kz + 1;
    if (kx == 3) kx = 0;
    int ky = kx + 1;
    if (ky == 3) ky = 0;
    r.kx = kx; r.ky = ky; r.kz = kz;
    float pdx = Comp(dx, dy, dz, kx), pdy = Comp(dx, dy, dz, ky), pdz = Comp(dx, dy, dz, kz);
    r.Sx = B2_DIV(-pdx, pdz);
    r.Sy = B2_DIV(-pdy, pdz);
    r.Sz = B2_DIV(1.f, pdz);
}

#if defined(__CUDACC__)
// A ray's derived terms as four stack entries, and back: leaving an object instance restores the parent ray's
// context instead of dividing six times again (RaySetup).
__device__ __forceinline__ void RaySave(const RayCtx &r, int2 &c0, int2 &c1, int2 &c2, int2 &c3) {
    c0 = make_int2(__float_as_int(r.ix), __float_as_int(r.iy));
    c1 = make_int2(__float_as_int(r.iz), __float_as_int(r.Sx));
    c2 = make_int2(__float_as_int(r.Sy), __float_as_int(r.Sz));
    c3 = make_int2(r.kx | (r.ky << 2) | (r.kz << 4), 0);
}
__device__ __forceinline__ void RayRestore(RayCtx &r, float ox, float oy, float oz, float dx, float dy, float dz, int2 c0, int2 c1, int2 c2,
                      int2 c3) {
    r.ox = ox; r.oy = oy; r.oz = oz;
    r.dx = dx; r.dy = dy; r.dz = dz;
    r.ix = __int_as_float(c0.x);
    r.iy = __int_as_float(c0.y);
    r.iz = __int_as_float(c1.x);
    r.Sx = __int_as_float(c1.y);
    r.Sy = __int_as_float(c2.x);
    r.Sz = __int_as_float(c2.y);
    r.neg[0] = r.ix < 0;
    r.neg[1] = r.iy < 0;
    r.neg[2] = r.iz < 0;
    r.kx = c3.x & 3;
    r.ky = (c3.x >> 2) & 3;
    r.kz = (c3.x >> 4) & 3;
}
#endif

// Bounds3::IntersectP with precomputed invDir/dirIsNeg.  Branch-free: the reference's two early
// `return false` only skip work, so evaluating everything and AND-ing the conditions gives the same
// boolean (selects keep the reference's NaN behaviour: `a > b ? a : b`, not fmax).
// SlabCore returns every part of the test that does not involve the ray's tMax, and the entry
// distance tMin; the full test is `SlabCore(...) && tMin < raytMax`.  The split matters: a child
// pushed on the stack must have its `tMin < raytMax` part evaluated when it is POPPED, with the
// tMax of that moment, because the reference's tMax is not monotone - an accepted triangle's
// t = tScaled * invDet can round to just above the tMax it was tested against (shapes.cpp:239-250).
B2_HD bool SlabCore(const RayCtx &r, const float *bmin, const float *bmax, float &tMinOut) {
    float tMin = B2_MUL(B2_SUB((r.neg[0] ? bmax[0] : bmin[0]), r.ox), r.ix);
    float tMax = B2_MUL(B2_SUB((r.neg[0] ? bmin[0] : bmax[0]), r.ox), r.ix);
    float tyMin = B2_MUL(B2_SUB((r.neg[1] ? bmax[1] : bmin[1]), r.oy), r.iy);
    float tyMax = B2_MUL(B2_SUB((r.neg[1] ? bmin[1] : bmax[1]), r.oy), r.iy);
    float tzMin = B2_MUL(B2_SUB((r.neg[2] ? bmax[2] : bmin[2]), r.oz), r.iz);
    float tzMax = B2_MUL(B2_SUB((r.neg[2] ? bmin[2] : bmax[2]), r.oz), r.iz);
    tMax = B2_MUL(tMax, kSlabPad);
    tyMax = B2_MUL(tyMax, kSlabPad);
    tzMax = B2_MUL(tzMax, kSlabPad);
    bool ok = !(tMin > tyMax) & !(tyMin > tMax);
    tMin = tyMin > tMin ? tyMin : tMin;
    tMax = tyMax < tMax ? tyMax : tMax;
    ok &= !(tMin > tzMax) & !(tzMin > tMax);
    tMin = tzMin > tMin ? tzMin : tMin;
    tMax = tzMax < tMax ? tzMax : tMax;
    tMinOut = tMin;
    return ok & (tMax > 0);
}

// The same test for a REGULAR ray (1/d and o finite, so no product is NaN) with the near / far planes already
// picked (near = dirIsNeg ? pMax : pMin per axis, as the reference picks them).  Without NaNs the reference's chain
// of comparisons and conditional updates is max(near products) <= min(padded far products) && min(...) > 0: same
// products, same roundings, same boolean; tMinOut equal to SlabCore's up to the sign of a zero.  (If a far product
// is negative its padded value can drop below the near one; then min(...) < 0 and both forms reject.)  An empty
// slot's box (+inf, -inf) gives near = +inf, far = -inf: rejected, no NaN.
B2_HD bool SlabPlanesRegular(const RayCtx &r, float nx, float ny, float nz, float fx, float fy, float fz, float &tMinOut) {
    float tMin = fmaxf(fmaxf(B2_MUL(B2_SUB(nx, r.ox), r.ix), B2_MUL(B2_SUB(ny, r.oy), r.iy)), B2_MUL(B2_SUB(nz, r.oz), r.iz));
    float tMax = fminf(fminf(B2_MUL(B2_MUL(B2_SUB(fx, r.ox), r.ix), kSlabPad), B2_MUL(B2_MUL(B2_SUB(fy, r.oy), r.iy), kSlabPad)),
                       B2_MUL(B2_MUL(B2_SUB(fz, r.oz), r.iz), kSlabPad));
    tMinOut = tMin;
    return (tMin <= tMax) & (tMax > 0);
}

// CONSERVATIVE slab test of the wide walk (wide.h): one fma per plane.
//   The reference's near product of plane b along axis k is fl(fl(b - o) * i) (i = 1/d), its far product that times
//   (1 + 2 gamma_3), one more rounding.  With |b| <= B (the tree's root box holds every box under it) and
//   M = B + |o|: |b - o| <= M, so each product lies within  M |i| (2^-23 + 2^-22 * 1.5 + 2^-24) < 2.25 * 2^-22 M |i|  of
//   the exact (b - o) i.  Here the plane's value is fma(b, i, c) with c = -fl(o i) -+ e, e = M |i| 2^-19: the three
//   roundings (o i, the sum c, the fma) cost at most 3 * 2^-24 (1 + 2^-18) M |i|, so the near value is BELOW the
//   reference's near product and the far value ABOVE its padded far product, with a margin of more than 2^-20 M |i|.
//   Hence  max(near values, 0) <= min(far values, tMaxS)  holds whenever the reference's test passes with any
//   raytMax <= tMaxS (its conditions are tMin <= tMax, tMax > 0, tMin < raytMax on the bracketed quantities).
// The ray must be regular: 1/d finite and non-zero, o and the bounds finite (else 0 * inf = NaN), M |i| a normal number
// on every axis; WideRaySetup says so.
struct WideRay {
    float cNx, cNy, cNz;  // near planes: fma(b, i, cN)
    float cFx, cFy, cFz;  // far planes:  fma(b, i, cF)
};

B2_HD bool WideRaySetup(const RayCtx &r, const float *treeBox, WideRay &w) {
    const float Mx = B2_ADD(fmaxf(fabsf(treeBox[0]), fabsf(treeBox[3])), fabsf(r.ox));
    const float My = B2_ADD(fmaxf(fabsf(treeBox[1]), fabsf(treeBox[4])), fabsf(r.oy));
    const float Mz = B2_ADD(fmaxf(fabsf(treeBox[2]), fabsf(treeBox[5])), fabsf(r.oz));
    // M |i| per axis; below 2^-100 the products are denormal and their roundings no longer relative: exact walk
    const float mx = B2_MUL(Mx, fabsf(r.ix)), my = B2_MUL(My, fabsf(r.iy)), mz = B2_MUL(Mz, fabsf(r.iz));
    const float ex = B2_MUL(mx, 0x1p-19f), ey = B2_MUL(my, 0x1p-19f), ez = B2_MUL(mz, 0x1p-19f);
    const float oix = B2_MUL(r.ox, r.ix), oiy = B2_MUL(r.oy, r.iy), oiz = B2_MUL(r.oz, r.iz);
    w.cNx = B2_SUB(-oix, ex);
    w.cNy = B2_SUB(-oiy, ey);
    w.cNz = B2_SUB(-oiz, ez);
    w.cFx = B2_ADD(-oix, ex);
    w.cFy = B2_ADD(-oiy, ey);
    w.cFz = B2_ADD(-oiz, ez);
    // every magnitude finite (a NaN or an infinity anywhere makes the sum non-finite; so does an overflow of finite
    // terms: that ray merely takes the exact walk too); |o i| <= M |i|, so the three M |i| bound everything above
    const float mag = B2_ADD(B2_ADD(mx, my), mz);
    return (mag < INFINITY) & (fminf(fminf(mx, my), mz) > 0x1p-100f);
}

// near / far planes picked as the reference picks them (near = dirIsNeg ? pMax : pMin).  tNearOut <= the reference's
// tMin of this box.  An empty slot's box (+inf, -inf) gives near = +inf, far = -inf: rejected, no NaN.
B2_HD bool WideChildTest(const RayCtx &r, const WideRay &w, float nx, float ny, float nz, float fx, float fy, float fz,
                         float tMaxS, float &tNearOut) {
    const float tNear = fmaxf(fmaxf(B2_FMA(nx, r.ix, w.cNx), B2_FMA(ny, r.iy, w.cNy)), B2_FMA(nz, r.iz, w.cNz));
    const float tFar = fminf(fminf(B2_FMA(fx, r.ix, w.cFx), B2_FMA(fy, r.iy, w.cFy)), B2_FMA(fz, r.iz, w.cFz));
    tNearOut = tNear;
    return fmaxf(tNear, 0.f) <= fminf(tFar, tMaxS);
}

B2_HD bool SlabTest(const RayCtx &r, const float *bmin, const float *bmax, float raytMax) {
    float tMin;
    return SlabCore(r, bmin, bmax, tMin) & (tMin < raytMax);
}

B2_HD float Max3Abs(float a, float b, float c) {
    // MaxComponentValue(Abs(v)) = std::max({|a|,|b|,|c|}) (util/vecmath.h:450-453)
    float m = fabsf(a);
    float fb = fabsf(b), fc = fabsf(c);
    if (m < fb) m = fb;
    if (m < fc) m = fc;
    return m;
}

// IntersectTriangle (shapes.cpp:172-273).  Returns true and fills t,b0,b1,b2 on a hit
// with t in (deltaT, tMax].
B2_HD bool TriHit(const RayCtx &r, float tMax, const float *p0, const float *p1, const float *p2,
                  float &tOut, float &b0Out, float &b1Out, float &b2Out) {
    // degenerate: LengthSquared(Cross(p2 - p0, p1 - p0)) == 0
    {
        float ax = B2_SUB(p2[0], p0[0]), ay = B2_SUB(p2[1], p0[1]), az = B2_SUB(p2[2], p0[2]);
        float bx = B2_SUB(p1[0], p0[0]), by = B2_SUB(p1[1], p0[1]), bz = B2_SUB(p1[2], p0[2]);
        float cx = DiffOfProducts(ay, bz, az, by);
        float cy = DiffOfProducts(az, bx, ax, bz);
        float cz = DiffOfProducts(ax, by, ay, bx);
        float l2 = B2_ADD(B2_ADD(B2_MUL(cx, cx), B2_MUL(cy, cy)), B2_MUL(cz, cz));
        if (l2 == 0)
            return false;
    }
    // translate, permute
    float t0x = B2_SUB(p0[0], r.ox), t0y = B2_SUB(p0[1], r.oy), t0z = B2_SUB(p0[2], r.oz);
    float t1x = B2_SUB(p1[0], r.ox), t1y = B2_SUB(p1[1], r.oy), t1z = B2_SUB(p1[2], r.oz);
    float t2x = B2_SUB(p2[0], r.ox), t2y = B2_SUB(p2[1], r.oy), t2z = B2_SUB(p2[2], r.oz);
    float p0x = Comp(t0x, t0y, t0z, r.kx), p0y = Comp(t0x, t0y, t0z, r.ky), p0z = Comp(t0x, t0y, t0z, r.kz);
    float p1x = Comp(t1x, t1y, t1z, r.kx), p1y = Comp(t1x, t1y, t1z, r.ky), p1z = Comp(t1x, t1y, t1z, r.kz);
    float p2x = Comp(t2x, t2y, t2z, r.kx), p2y = Comp(t2x, t2y, t2z, r.ky), p2z = Comp(t2x, t2y, t2z, r.kz);
    // shear (p.x += Sx * p.z: a multiply then an add, not fused)
    p0x = B2_ADD(p0x, B2_MUL(r.Sx, p0z));
    p0y = B2_ADD(p0y, B2_MUL(r.Sy, p0z));
    p1x = B2_ADD(p1x, B2_MUL(r.Sx, p1z));
    p1y = B2_ADD(p1y, B2_MUL(r.Sy, p1z));
    p2x = B2_ADD(p2x, B2_MUL(r.Sx, p2z));
    p2y = B2_ADD(p2y, B2_MUL(r.Sy, p2z));
    // edge functions
    float e0 = DiffOfProducts(p1x, p2y, p1y, p2x);
    float e1 = DiffOfProducts(p2x, p0y, p2y, p0x);
    float e2 = DiffOfProducts(p0x, p1y, p0y, p1x);
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        double p2txp1ty = B2_DMUL((double)p2x, (double)p1y);
        double p2typ1tx = B2_DMUL((double)p2y, (double)p1x);
        e0 = B2_D2F(B2_DSUB(p2typ1tx, p2txp1ty));
        double p0txp2ty = B2_DMUL((double)p0x, (double)p2y);
        double p0typ2tx = B2_DMUL((double)p0y, (double)p2x);
        e1 = B2_D2F(B2_DSUB(p0typ2tx, p0txp2ty));
        double p1txp0ty = B2_DMUL((double)p1x, (double)p0y);
        double p1typ0tx = B2_DMUL((double)p1y, (double)p0x);
        e2 = B2_D2F(B2_DSUB(p1typ0tx, p1txp0ty));
    }
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0))
        return false;
    float det = B2_ADD(B2_ADD(e0, e1), e2);
    if (det == 0)
        return false;
    p0z = B2_MUL(p0z, r.Sz);
    p1z = B2_MUL(p1z, r.Sz);
    p2z = B2_MUL(p2z, r.Sz);
    float tScaled = B2_ADD(B2_ADD(B2_MUL(e0, p0z), B2_MUL(e1, p1z)), B2_MUL(e2, p2z));
    if (det < 0 && (tScaled >= 0 || tScaled < B2_MUL(tMax, det)))
        return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > B2_MUL(tMax, det)))
        return false;
    float invDet = B2_DIV(1.f, det);
    float b0 = B2_MUL(e0, invDet), b1 = B2_MUL(e1, invDet), b2 = B2_MUL(e2, invDet);
    float t = B2_MUL(tScaled, invDet);
    // conservative t > 0 check
    float maxZt = Max3Abs(p0z, p1z, p2z);
    float deltaZ = B2_MUL(GammaF(3), maxZt);
    float maxXt = Max3Abs(p0x, p1x, p2x);
    float maxYt = Max3Abs(p0y, p1y, p2y);
    float deltaX = B2_MUL(GammaF(5), B2_ADD(maxXt, maxZt));
    float deltaY = B2_MUL(GammaF(5), B2_ADD(maxYt, maxZt));
    float deltaE = B2_MUL(2.f, B2_ADD(B2_ADD(B2_MUL(B2_MUL(GammaF(2), maxXt), maxYt), B2_MUL(deltaY, maxXt)),
                                     B2_MUL(deltaX, maxYt)));
    float maxE = Max3Abs(e0, e1, e2);
    float deltaT = B2_MUL(
        B2_MUL(3.f, B2_ADD(B2_ADD(B2_MUL(B2_MUL(GammaF(3), maxE), maxZt), B2_MUL(deltaE, maxZt)),
                           B2_MUL(deltaZ, maxE))),
        fabsf(invDet));
    if (t <= deltaT)
        return false;
    tOut = t;
    b0Out = b0;
    b1Out = b1;
    b2Out = b2;
    return true;
}

// ------------------------------------------------------------------ bilinear patches
// IntersectBilinearPatch (shapes.h:1279-1347), operation for operation: Cross = DifferenceOfProducts per
// component (util/vecmath.h:1001-1006), Dot = plain sum of products, Lerp(t, a, b) = (1 - t) * a + t * b,
// Determinant(SquareMatrix<3>) (util/math.h:1420-1426), Quadratic (util/math.h:614-637).
B2_HD void Cross3(const float *v, const float *w, float *o) {
    o[0] = DiffOfProducts(v[1], w[2], v[2], w[1]);
    o[1] = DiffOfProducts(v[2], w[0], v[0], w[2]);
    o[2] = DiffOfProducts(v[0], w[1], v[1], w[0]);
}
B2_HD float Dot3(const float *a, const float *b) {
    return B2_ADD(B2_ADD(B2_MUL(a[0], b[0]), B2_MUL(a[1], b[1])), B2_MUL(a[2], b[2]));
}
// rows (a0 b0 c0), (a1 b1 c1), (a2 b2 c2)
B2_HD float Det3(const float *a, const float *b, const float *c) {
    float minor12 = DiffOfProducts(b[1], c[2], c[1], b[2]);
    float minor02 = DiffOfProducts(a[1], c[2], c[1], a[2]);
    float minor01 = DiffOfProducts(a[1], b[2], b[1], a[2]);
    return B2_FMA(c[0], minor01, DiffOfProducts(a[0], minor12, b[0], minor02));
}
B2_HD bool QuadraticRef(float a, float b, float c, float *t0, float *t1) {
    if (a == 0) {
        if (b == 0) return false;
        *t0 = *t1 = B2_DIV(-c, b);
        return true;
    }
    float discrim = DiffOfProducts(b, b, B2_MUL(4.f, a), c);
    if (discrim < 0) return false;
    float rootDiscrim = sqrtf(discrim);
    float q = B2_MUL(-0.5f, B2_ADD(b, copysignf(rootDiscrim, b)));
    *t0 = B2_DIV(q, a);
    *t1 = B2_DIV(c, q);
    if (*t0 > *t1) {
        float t = *t0;
        *t0 = *t1;
        *t1 = t;
    }
    return true;
}

// One candidate root u of the patch quadratic: v and t numerators over p2 = |d x ud|^2
B2_HD void BilinearCandidate(const float *o, const float *d, const float *p00, const float *p10, const float *p01,
                             const float *p11, float u, float &vNum, float &tNum, float &p2) {
    float uo[3], ud[3], deltao[3], perp[3];
    const float om = B2_SUB(1.f, u);
    for (int k = 0; k < 3; ++k) {
        uo[k] = B2_ADD(B2_MUL(om, p00[k]), B2_MUL(u, p10[k]));
        float e = B2_ADD(B2_MUL(om, p01[k]), B2_MUL(u, p11[k]));
        ud[k] = B2_SUB(e, uo[k]);
        deltao[k] = B2_SUB(uo[k], o[k]);
    }
    Cross3(d, ud, perp);
    p2 = B2_ADD(B2_ADD(B2_MUL(perp[0], perp[0]), B2_MUL(perp[1], perp[1])), B2_MUL(perp[2], perp[2]));
    // SquareMatrix<3>(deltao.x, d.x, perp.x, deltao.y, d.y, perp.y, deltao.z, d.z, perp.z): the matrix rows are
    // (deltao[i], d[i], perp[i]); Det3 takes its three columns
    vNum = Det3(deltao, d, perp);
    tNum = Det3(deltao, ud, perp);
}

// Returns true and fills t, u, v on a hit with t < tMax.
B2_HD bool BilinearHit(const RayCtx &r, float tMax, const float *p00, const float *p10, const float *p01,
                       const float *p11, float &tOut, float &uOut, float &vOut) {
    const float o[3] = {r.ox, r.oy, r.oz}, d[3] = {r.dx, r.dy, r.dz};
    float e0[3], e1[3], cr[3];
    for (int k = 0; k < 3; ++k) {
        e0[k] = B2_SUB(p10[k], p00[k]);
        e1[k] = B2_SUB(p01[k], p11[k]);
    }
    Cross3(e0, e1, cr);
    float a = Dot3(cr, d);
    for (int k = 0; k < 3; ++k) {
        e0[k] = B2_SUB(p00[k], o[k]);
        e1[k] = B2_SUB(p01[k], p00[k]);
    }
    Cross3(e0, d, cr);
    float c = Dot3(cr, e1);
    for (int k = 0; k < 3; ++k) {
        e0[k] = B2_SUB(p10[k], o[k]);
        e1[k] = B2_SUB(p11[k], p10[k]);
    }
    Cross3(e0, d, cr);
    float b = B2_SUB(Dot3(cr, e1), B2_ADD(a, c));
    float u1, u2;
    if (!QuadraticRef(a, b, c, &u1, &u2)) return false;
    // gamma(10) * (sum of the max |component| of o, d and the four vertices), added left to right
    float eps = Max3Abs(o[0], o[1], o[2]);
    eps = B2_ADD(eps, Max3Abs(d[0], d[1], d[2]));
    eps = B2_ADD(eps, Max3Abs(p00[0], p00[1], p00[2]));
    eps = B2_ADD(eps, Max3Abs(p10[0], p10[1], p10[2]));
    eps = B2_ADD(eps, Max3Abs(p01[0], p01[1], p01[2]));
    eps = B2_ADD(eps, Max3Abs(p11[0], p11[1], p11[2]));
    eps = B2_MUL(GammaF(10), eps);
    float t = tMax, u = 0, v = 0;
    if (0 <= u1 && u1 <= 1) {
        float v1, t1, p2;
        BilinearCandidate(o, d, p00, p10, p01, p11, u1, v1, t1, p2);
        if (t1 > B2_MUL(p2, eps) && 0 <= v1 && v1 <= p2) {
            u = u1;
            v = B2_DIV(v1, p2);
            t = B2_DIV(t1, p2);
        }
    }
    if (0 <= u2 && u2 <= 1 && u2 != u1) {
        float v2, t2, p2;
        BilinearCandidate(o, d, p00, p10, p01, p11, u2, v2, t2, p2);
        t2 = B2_DIV(t2, p2);
        if (0 <= v2 && v2 <= p2 && t > t2 && t2 > eps) {
            t = t2;
            u = u2;
            v = B2_DIV(v2, p2);
        }
    }
    if (t >= tMax) return false;
    tOut = t;
    uOut = u;
    vOut = v;
    return true;
}

// ---- alpha cut-outs: GeometricPrimitive::Intersect / IntersectP (cpu/primitive.cpp:50-109) with a constant alpha
// texture.  A hit on a primitive with alpha a < 1 counts when a > 0 and HashFloat(r.o, r.d) <= a; otherwise the
// reference re-traces the SAME shape from the hit point, which a planar triangle cannot meet again, so the primitive is
// passed over.  HashFloat(o, d) = uint32(MurmurHash64A of the 24 argument bytes) * 2^-32 (util/hash.h:18-68, 95-110).
B2_HD float HashFloatRay(const RayCtx &r) {
    const float v[6] = {r.ox, r.oy, r.oz, r.dx, r.dy, r.dz};
    const uint64_t m = 0xc6a4a7935bd1e995ull;
    uint64_t h = 0 ^ (24ull * m);
    for (int i = 0; i < 3; ++i) {
        uint32_t lo, hi;
        memcpy(&lo, &v[2 * i], 4);
        memcpy(&hi, &v[2 * i + 1], 4);
        uint64_t k = uint64_t(lo) | (uint64_t(hi) << 32);
        k *= m;
        k ^= k >> 47;
        k *= m;
        h ^= k;
        h *= m;
    }
    h ^= h >> 47;
    h *= m;
    h ^= h >> 47;
    return float(uint32_t(h)) * 0x1p-32f;
}
B2_HD bool AlphaKeepsHit(const RayCtx &r, float alpha) {
    if (!(alpha < 1)) return true;
    if (alpha <= 0) return false;
    return !(HashFloatRay(r) > alpha);
}

}  // namespace b200pt
