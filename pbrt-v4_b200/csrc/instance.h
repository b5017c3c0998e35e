// instance.h - the ray an object instance is traversed with: Transform::ApplyInverse(Ray, &tMax)
// (util/transform.h:416-429) with ApplyInverse(Point3fi) (util/transform.cpp:263-306) and the CPU
// build's interval arithmetic (util/math.h:829-837, 874-876; util/float.h:201-229: round-up / round-down
// are NextFloatUp / NextFloatDown of the rounded sum).  Host+device, same arithmetic on both.
//
// What the reference does, and what is reproduced to the bit (cpu/primitive.cpp:112-131):
//   o' = mInv * o as a point with an error box (gamma(3) * sum |mInv_ij * o_j|), d' = mInv * d,
//   dt = dot(|d'|, err) / |d'|^2;  o' += d' * dt (interval sum, then its midpoint);  tMax' = tMax - dt.
// The object is then traversed with (o', d', tMax').  The t it returns is NOT shifted back by dt: the
// parent continues with tMax = t' (cpu/primitive.cpp:116-119, cpu/aggregates.cpp:553-556).
#pragma once
#include <cstring>

#include "isect.h"

namespace b200pt {

B2_HD float NextFloatUp32(float v) {  // util/float.h:140-156
    if (v > 0.f && v == INFINITY) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui;
    memcpy(&ui, &v, 4);
    if (v >= 0) ++ui; else --ui;
    memcpy(&v, &ui, 4);
    return v;
}
B2_HD float NextFloatDown32(float v) {  // util/float.h:158-172
    if (v < 0.f && v == -INFINITY) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui;
    memcpy(&ui, &v, 4);
    if (v > 0) --ui; else ++ui;
    memcpy(&v, &ui, 4);
    return v;
}

// mInv: instanceFromRender, row major 4x4.  pbrt inverts instance matrices numerically, so the last row
// is often (0, 0, 0, 1 - 1ulp): the reference then divides the point interval by wp (transform.cpp:300-302,
// util/math.h:1028-1036), which is reproduced.
B2_HD void InstanceRay(const float *mInv, float ox, float oy, float oz, float dx, float dy, float dz, float tMax,
                       float o2[3], float d2[3], float &tMax2) {
    const float g3 = GammaF(3);
    float lo[3], hi[3];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int r = 0; r < 3; ++r) {
        const float *m = mInv + 4 * r;
        float mx = B2_MUL(m[0], ox), my = B2_MUL(m[1], oy), mz = B2_MUL(m[2], oz);
        float p = B2_ADD(B2_ADD(mx, my), B2_ADD(mz, m[3]));
        float err = B2_MUL(g3, B2_ADD(B2_ADD(fabsf(mx), fabsf(my)), fabsf(mz)));
        // Point3fi(Point3f, Vector3f): Interval::FromValueAndError per component
        if (err == 0) {
            lo[r] = hi[r] = p;
        } else {
            lo[r] = NextFloatDown32(B2_SUB(p, err));
            hi[r] = NextFloatUp32(B2_ADD(p, err));
        }
        d2[r] = B2_ADD(B2_ADD(B2_MUL(m[0], dx), B2_MUL(m[1], dy)), B2_MUL(m[2], dz));
    }
    {
        const float *m = mInv + 12;
        float wp = B2_ADD(B2_ADD(B2_MUL(m[0], ox), B2_MUL(m[1], oy)), B2_ADD(B2_MUL(m[2], oz), m[3]));
        if (wp != 1) {
            for (int r = 0; r < 3; ++r) {
                float a = lo[r], b = hi[r];
                if (wp == 0) {
                    lo[r] = -INFINITY;
                    hi[r] = INFINITY;
                } else if (wp > 0) {
                    lo[r] = NextFloatDown32(B2_DIV(a, wp));
                    hi[r] = NextFloatUp32(B2_DIV(b, wp));
                } else {
                    lo[r] = NextFloatDown32(B2_DIV(b, wp));
                    hi[r] = NextFloatUp32(B2_DIV(a, wp));
                }
                if (lo[r] > hi[r]) {  // Interval(low, high) orders its ends
                    float t = lo[r];
                    lo[r] = hi[r];
                    hi[r] = t;
                }
            }
        }
    }
    float lengthSquared = B2_ADD(B2_ADD(B2_MUL(d2[0], d2[0]), B2_MUL(d2[1], d2[1])), B2_MUL(d2[2], d2[2]));
    tMax2 = tMax;
    if (lengthSquared > 0) {
        float ex = B2_DIV(B2_SUB(hi[0], lo[0]), 2.f), ey = B2_DIV(B2_SUB(hi[1], lo[1]), 2.f),
              ez = B2_DIV(B2_SUB(hi[2], lo[2]), 2.f);
        float dot = B2_ADD(B2_ADD(B2_MUL(fabsf(d2[0]), ex), B2_MUL(fabsf(d2[1]), ey)), B2_MUL(fabsf(d2[2]), ez));
        float dt = B2_DIV(dot, lengthSquared);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int r = 0; r < 3; ++r) {
            float s = B2_MUL(d2[r], dt);  // o += d * dt: Interval + Interval(s)
            lo[r] = NextFloatDown32(B2_ADD(lo[r], s));
            hi[r] = NextFloatUp32(B2_ADD(hi[r], s));
        }
        tMax2 = B2_SUB(tMax, dt);
    }
    for (int r = 0; r < 3; ++r) o2[r] = B2_DIV(B2_ADD(lo[r], hi[r]), 2.f);  // Point3f(Point3fi): midpoints
}

}  // namespace b200pt
