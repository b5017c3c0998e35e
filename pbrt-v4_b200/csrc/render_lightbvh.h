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

// CompactLightBounds::Importance(p, n, allb)
// nd: one node; allb: allLightBounds, six floats (both in the node table: plain pointers keep the
// caller's scene block out of local memory)
B2_HD_CALL float LightBoundsImportance(const LightBvhNode *ndp, const float *allb, V3 p, V3 n) {
    const LightBvhNode nd = *ndp;
    // Bounds(allb): lightsamplers.h:133-142
    float lo[3], hi[3];
    uint32_t q[6] = {nd.w[3] & 0xFFFFu, nd.w[3] >> 16, nd.w[4] & 0xFFFFu, nd.w[4] >> 16, nd.w[5] & 0xFFFFu,
                     nd.w[5] >> 16};
    for (int c = 0; c < 3; ++c) {
        lo[c] = Lerp(q[c] / 65535.f, allb[c], allb[3 + c]);
        hi[c] = Lerp(q[3 + c] / 65535.f, allb[c], allb[3 + c]);
    }
    V3 bmin = mk(lo[0], lo[1], lo[2]), bmax = mk(hi[0], hi[1], hi[2]);
    uint32_t bits = nd.w[2];
    float cosTheta_o = 2 * ((bits & 0x7FFFu) / 32767.f) - 1;
    float cosTheta_e = 2 * (((bits >> 15) & 0x7FFFu) / 32767.f) - 1;
    bool twoSided = ((bits >> 30) & 1u) != 0;
    float phi;
    memcpy(&phi, &nd.w[1], 4);

    V3 pc = (bmin + bmax) / 2;
    float d2 = LengthSquared(p - pc);
    d2 = fmaxf(d2, Length(bmax - bmin) / 2);

    V3 wi = Normalize(p - pc);
    float cosTheta_w = Dot(OctahedralDecode(nd.w[0]), wi);
    if (twoSided) cosTheta_w = fabsf(cosTheta_w);
    float sinTheta_w = SafeSqrt(1 - SqrF(cosTheta_w));

    // BoundSubtendedDirections(bounds, p).cosTheta (util/vecmath.h:1817-1830; BoundingSphere :1337-1340)
    float cosTheta_b;
    {
        bool inside = pc.x >= bmin.x && pc.x <= bmax.x && pc.y >= bmin.y && pc.y <= bmax.y && pc.z >= bmin.z &&
                      pc.z <= bmax.z;
        float radius = inside ? Length(pc - bmax) : 0;
        if (LengthSquared(p - pc) < SqrF(radius))
            cosTheta_b = -1;  // DirectionCone::EntireSphere
        else {
            float sin2ThetaMax = SqrF(radius) / LengthSquared(pc - p);
            cosTheta_b = SafeSqrt(1 - sin2ThetaMax);
        }
    }
    float sinTheta_b = SafeSqrt(1 - SqrF(cosTheta_b));

    float sinTheta_o = SafeSqrt(1 - SqrF(cosTheta_o));
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
    const LightBvhNode *nodes = reinterpret_cast<const LightBvhNode *>(rs.lightBvhNodes);
    int nInf = (int)rs.nInfiniteLights;
    bool haveNodes = rs.nLightBvhNodes > 0;
    const float *allb = reinterpret_cast<const float *>(nodes + rs.nLightBvhNodes);
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
        uint32_t link = nodes[nodeIndex].w[6];
        if (!(link >> 31)) {
            int second = (int)(link & 0x7FFFFFFFu);
            float ci[2] = {LightBoundsImportance(nodes + nodeIndex + 1, allb, p, n),
                           LightBoundsImportance(nodes + second, allb, p, n)};
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
            if (nodeIndex > 0 || LightBoundsImportance(nodes + nodeIndex, allb, p, n) > 0) {
                *pmf = prob;
                return (int)(link & 0x7FFFFFFFu);
            }
            return -1;
        }
    }
}

// BVHLightSampler::PMF.  bitTrail = lightToBitTrail[light]; inBvh = the map has the light.
B2_HD float LightBvhPMF(const RenderSceneDev &rs, V3 p, V3 n, int light) {
    const LightBvhNode *nodes = reinterpret_cast<const LightBvhNode *>(rs.lightBvhNodes);
    int nInf = (int)rs.nInfiniteLights;
    bool haveNodes = rs.nLightBvhNodes > 0;
    uint32_t bitTrail = rs.lightBitTrails[light];
    if (bitTrail == 0xFFFFFFFFu) return 1.f / (nInf + (haveNodes ? 1 : 0));  // not in the map
    const float *allb = reinterpret_cast<const float *>(nodes + rs.nLightBvhNodes);
    float pInfinite = float(nInf) / float(nInf + (haveNodes ? 1 : 0));
    float pmf = 1 - pInfinite;
    int nodeIndex = 0;
    while (true) {
        uint32_t link = nodes[nodeIndex].w[6];
        if (link >> 31) return pmf;
        int second = (int)(link & 0x7FFFFFFFu);
        float ci[2] = {LightBoundsImportance(nodes + nodeIndex + 1, allb, p, n),
                       LightBoundsImportance(nodes + second, allb, p, n)};
        pmf *= ((bitTrail & 1) ? ci[1] : ci[0]) / (ci[0] + ci[1]);
        nodeIndex = (bitTrail & 1) ? second : nodeIndex + 1;
        bitTrail >>= 1;
    }
}

}  // namespace b200pt
