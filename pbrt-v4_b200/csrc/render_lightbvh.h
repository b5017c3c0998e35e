// render_lightbvh.h - BVHLightSampler::Sample / PMF (lightsamplers.h:266-358) over the light BVH the
// host built (lightsamplers.cpp:109-232): stochastic descent by CompactLightBounds::Importance
// (lightsamplers.h:144-201), PMF by replaying the light's bit trail.  Host+device, one arithmetic.
//
// Node record (32 B, eight u32; the fields of LightBVHNode / CompactLightBounds, lightsamplers.h:203-264):
//   [0] w.x | w.y << 16        OctahedralVector (util/vecmath.h:1735-1784)
//   [1] phi (float bits)
//   [2] qCosTheta_o | qCosTheta_e << 15 | twoSided << 30
//   [3] qb[0][0] | qb[0][1] << 16   [4] qb[0][2] | qb[1][0] << 16   [5] qb[1][1] | qb[1][2] << 16
//   [6] childOrLightIndex | isLeaf << 31
//   [7] 0
#pragma once
#include "render_math.h"
#include "render_types.h"

namespace b200pt {

struct LightBvhNode {
    uint32_t w[8];
};

B2_HD V3 OctahedralDecode(uint32_t packed) {  // util/vecmath.h:1752-1766
    float x = -1 + 2 * ((packed & 0xFFFFu) / 65535.f);
    float y = -1 + 2 * ((packed >> 16) / 65535.f);
    float z = 1 - (fabsf(x) + fabsf(y));
    if (z < 0) {
        float xo = x;
        x = (1 - fabsf(y)) * copysignf(1.f, xo);
        y = (1 - fabsf(xo)) * copysignf(1.f, y);
    }
    return Normalize(mk(x, y, z));
}

// What CompactLightBounds::Importance (lightsamplers.h:144-201) derives from the node alone, worked out once on the host
// with the same float operations (the library's host code is built without contraction, like the kernels): the
// dequantised bounds (Bounds(allb), lightsamplers.h:133-142), their centre, half diagonal and bounding-sphere radius
// (util/vecmath.h:1337-1340), the decoded axis, cosTheta_o / cosTheta_e / sinTheta_o.  A visit then costs the
// arithmetic that depends on the shading point only (about 40 % of the function).  kLightNodeWords floats per node:
//   [0..2] centre pc  [3] Length(bmax - bmin) / 2  [4] radius  [5..7] w  [8] phi  [9] cosTheta_o  [10] cosTheta_e
//   [11] sinTheta_o  [12] twoSided  [13] childOrLightIndex | isLeaf << 31 (bits)  [14..15] 0
constexpr int kLightNodeWords = 16;
B2_HD void ExpandLightBvhNode(const uint32_t *w, const float *allb, float *e) {
    float lo[3], hi[3];
    uint32_t q[6] = {w[3] & 0xFFFFu, w[3] >> 16, w[4] & 0xFFFFu, w[4] >> 16, w[5] & 0xFFFFu, w[5] >> 16};
    for (int c = 0; c < 3; ++c) {
        lo[c] = Lerp(q[c] / 65535.f, allb[c], allb[3 + c]);
        hi[c] = Lerp(q[3 + c] / 65535.f, allb[c], allb[3 + c]);
    }
    V3 bmin = mk(lo[0], lo[1], lo[2]), bmax = mk(hi[0], hi[1], hi[2]);
    uint32_t bits = w[2];
    float cosTheta_o = 2 * ((bits & 0x7FFFu) / 32767.f) - 1;
    float cosTheta_e = 2 * (((bits >> 15) & 0x7FFFu) / 32767.f) - 1;
    V3 pc = (bmin + bmax) / 2;
    bool inside = pc.x >= bmin.x && pc.x <= bmax.x && pc.y >= bmin.y && pc.y <= bmax.y && pc.z >= bmin.z && pc.z <= bmax.z;
    V3 ax = OctahedralDecode(w[0]);
    e[0] = pc.x; e[1] = pc.y; e[2] = pc.z;
    e[3] = Length(bmax - bmin) / 2;
    e[4] = inside ? Length(pc - bmax) : 0;
    e[5] = ax.x; e[6] = ax.y; e[7] = ax.z;
    memcpy(&e[8], &w[1], 4);
    e[9] = cosTheta_o;
    e[10] = cosTheta_e;
    e[11] = SafeSqrt(1 - SqrF(cosTheta_o));
    e[12] = ((bits >> 30) & 1u) ? 1.f : 0.f;
    memcpy(&e[13], &w[6], 4);
    e[14] = e[15] = 0;
}
B2_HD uint32_t LightNodeLink(const float *e) {
    uint32_t l;
    memcpy(&l, &e[13], 4);
    return l;
}

// CompactLightBounds::Importance(p, n, allb) over an expanded node
B2_HD_CALL float LightBoundsImportance(const float *e, V3 p, V3 n) {
    const V3 pc = mk(e[0], e[1], e[2]);
    const float radius = e[4], phi = e[8], cosTheta_o = e[9], cosTheta_e = e[10], sinTheta_o = e[11];
    float d2 = LengthSquared(p - pc);
    d2 = fmaxf(d2, e[3]);

    V3 wi = Normalize(p - pc);
    float cosTheta_w = Dot(mk(e[5], e[6], e[7]), wi);
    if (e[12] != 0) cosTheta_w = fabsf(cosTheta_w);
    float sinTheta_w = SafeSqrt(1 - SqrF(cosTheta_w));

    // BoundSubtendedDirections(bounds, p).cosTheta (util/vecmath.h:1817-1830; BoundingSphere :1337-1340)
    float cosTheta_b;
    if (LengthSquared(p - pc) < SqrF(radius))
        cosTheta_b = -1;  // DirectionCone::EntireSphere
    else {
        float sin2ThetaMax = SqrF(radius) / LengthSquared(pc - p);
        cosTheta_b = SafeSqrt(1 - sin2ThetaMax);
    }
    float sinTheta_b = SafeSqrt(1 - SqrF(cosTheta_b));

    // cos / sin of max(0, a - b)
    float cosTheta_x = (cosTheta_w > cosTheta_o) ? 1.f : cosTheta_w * cosTheta_o + sinTheta_w * sinTheta_o;
    float sinTheta_x = (cosTheta_w > cosTheta_o) ? 0.f : sinTheta_w * cosTheta_o - cosTheta_w * sinTheta_o;
    float cosThetap = (cosTheta_x > cosTheta_b) ? 1.f : cosTheta_x * cosTheta_b + sinTheta_x * sinTheta_b;
    if (cosThetap <= cosTheta_e) return 0;

    float importance = phi * cosThetap / d2;
    if (!(n.x == 0 && n.y == 0 && n.z == 0)) {
        float cosTheta_i = AbsDotN(n, wi);
        float sinTheta_i = SafeSqrt(1 - SqrF(cosTheta_i));
        float cosThetap_i = (cosTheta_i > cosTheta_b) ? 1.f : cosTheta_i * cosTheta_b + sinTheta_i * sinTheta_b;
        importance *= cosThetap_i;
    }
    return fmaxf(importance, 0.f);
}

// BVHLightSampler::Sample.  Returns the light index or -1; *pmf its probability.
B2_HD int LightBvhSample(const RenderSceneDev &rs, V3 p, V3 n, float u, float *pmf) {
    const float *nodes = reinterpret_cast<const float *>(rs.lightBvhNodes);  // kLightNodeWords per node (ExpandLightBvhNode)
    int nInf = (int)rs.nInfiniteLights;
    bool haveNodes = rs.nLightBvhNodes > 0;
    float pInfinite = float(nInf) / float(nInf + (haveNodes ? 1 : 0));
    if (u < pInfinite) {
        u /= pInfinite;
        int index = (int)(u * nInf);  // std::min<int>(u * size, size - 1)
        if (index > nInf - 1) index = nInf - 1;
        *pmf = pInfinite / nInf;
        return rs.infiniteLights[index];
    }
    if (!haveNodes) return -1;
    u = fminf((u - pInfinite) / (1 - pInfinite), kOneMinusEps);
    int nodeIndex = 0;
    float prob = 1 - pInfinite;
    while (true) {
        uint32_t link = LightNodeLink(nodes + kLightNodeWords * nodeIndex);
        if (!(link >> 31)) {
            int second = (int)(link & 0x7FFFFFFFu);
            float ci[2] = {LightBoundsImportance(nodes + kLightNodeWords * (nodeIndex + 1), p, n),
                           LightBoundsImportance(nodes + kLightNodeWords * second, p, n)};
            if (ci[0] == 0 && ci[1] == 0) return -1;
            // SampleDiscrete over two weights (util/sampling.h:79-113)
            float sumWeights = 0;
            sumWeights += ci[0];
            sumWeights += ci[1];
            float up = u * sumWeights;
            if (up == sumWeights) up = NextDown(up);
            // `while (sum + weights[offset] <= up) sum += weights[offset++]` over two weights, without
            // indexing (the second weight cannot be passed: up < sumWeights)
            int offset = 0;
            float sum = 0, wsel = ci[0];
            if (sum + ci[0] <= up) {
                sum += ci[0];
                offset = 1;
                wsel = ci[1];
            }
            prob *= wsel / sumWeights;
            u = fminf((up - sum) / wsel, kOneMinusEps);
            nodeIndex = (offset == 0) ? nodeIndex + 1 : second;
        } else {
            if (nodeIndex > 0 || LightBoundsImportance(nodes + kLightNodeWords * nodeIndex, p, n) > 0) {
                *pmf = prob;
                return (int)(link & 0x7FFFFFFFu);
            }
            return -1;
        }
    }
}

// BVHLightSampler::PMF.  bitTrail = lightToBitTrail[light]; inBvh = the map has the light.
B2_HD float LightBvhPMF(const RenderSceneDev &rs, V3 p, V3 n, int light) {
    const float *nodes = reinterpret_cast<const float *>(rs.lightBvhNodes);
    int nInf = (int)rs.nInfiniteLights;
    bool haveNodes = rs.nLightBvhNodes > 0;
    uint32_t bitTrail = rs.lightBitTrails[light];
    if (bitTrail == 0xFFFFFFFFu) return 1.f / (nInf + (haveNodes ? 1 : 0));  // not in the map
    float pInfinite = float(nInf) / float(nInf + (haveNodes ? 1 : 0));
    float pmf = 1 - pInfinite;
    int nodeIndex = 0;
    while (true) {
        uint32_t link = LightNodeLink(nodes + kLightNodeWords * nodeIndex);
        if (link >> 31) return pmf;
        int second = (int)(link & 0x7FFFFFFFu);
        float ci[2] = {LightBoundsImportance(nodes + kLightNodeWords * (nodeIndex + 1), p, n),
                       LightBoundsImportance(nodes + kLightNodeWords * second, p, n)};
        pmf *= ((bitTrail & 1) ? ci[1] : ci[0]) / (ci[0] + ci[1]);
        nodeIndex = (bitTrail & 1) ? second : nodeIndex + 1;
        bitTrail >>= 1;
    }
}

}  // namespace b200pt
