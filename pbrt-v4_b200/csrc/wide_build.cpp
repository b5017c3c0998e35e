// wide_build.cpp - collapses the reference-shaped binary tree into 8-wide nodes (wide.h).
// The binary tree stays the authority on boxes, leaves and visit order; this only chooses, per wide
// node, which descendants of its binary node to put side by side (greedily opening the child with the
// largest surface area, the usual SAH-driven collapse) and records the reference's visit order of
// those descendants for every ray octant.
#include "wide.h"

#include <algorithm>
#include <cstring>
#include <limits>

namespace b200pt {
namespace {

inline float BoxArea(const Node32 &n) {
    float dx = n.bmax[0] - n.bmin[0], dy = n.bmax[1] - n.bmin[1], dz = n.bmax[2] - n.bmin[2];
    return 2 * (dx * dy + dx * dz + dy * dz);
}

// the reference's depth-first order of the cut below binary node `n` for dirIsNeg = neg
// (cpu/aggregates.cpp:562-568: dirIsNeg[axis] ? second child first : first child first)
void VisitOrder(const std::vector<Node32> &nodes, uint32_t n, const uint32_t *cut, int nCut, const int neg[3],
                uint8_t *order, int &k) {
    for (int i = 0; i < nCut; ++i)
        if (cut[i] == n) {
            order[k++] = uint8_t(i);
            return;
        }
    const Node32 &nd = nodes[n];
    uint32_t first = n + 1, second = uint32_t(nd.offset);
    if (neg[nd.axis]) std::swap(first, second);
    VisitOrder(nodes, first, cut, nCut, neg, order, k);
    VisitOrder(nodes, second, cut, nCut, neg, order, k);
}

}  // namespace

int32_t BuildWide(const std::vector<Node32> &nodes, uint32_t root, std::vector<Wide8> &out, uint32_t *maxStackOut) {
    if (maxStackOut) *maxStackOut = 1;
    if (nodes[root].nPrims > 0) return ~nodes[root].offset;
    struct Job {
        uint32_t bvh2;    // binary interior node this wide node stands for
        uint32_t wide;    // its index in `out`
        uint32_t stacked; // stack entries held by the ancestors while this node is visited
    };
    std::vector<Job> jobs;
    const int32_t rootWide = int32_t(out.size());
    out.emplace_back();
    jobs.push_back({root, uint32_t(rootWide), 0});
    uint32_t maxStack = 1;
    while (!jobs.empty()) {
        Job job = jobs.back();
        jobs.pop_back();
        uint32_t cut[8];
        int nCut = 2;
        cut[0] = job.bvh2 + 1;
        cut[1] = uint32_t(nodes[job.bvh2].offset);
        while (nCut < 8) {
            int best = -1;
            float bestArea = -1;
            for (int i = 0; i < nCut; ++i)
                if (nodes[cut[i]].nPrims == 0) {
                    float a = BoxArea(nodes[cut[i]]);
                    if (!(a == a)) a = std::numeric_limits<float>::infinity();
                    if (a > bestArea) {
                        bestArea = a;
                        best = i;
                    }
                }
            if (best < 0) break;
            // replace in place, keeping the first-child-first depth-first order of the slots
            uint32_t n = cut[best];
            for (int i = nCut; i > best + 1; --i) cut[i] = cut[i - 1];
            cut[best] = n + 1;
            cut[best + 1] = uint32_t(nodes[n].offset);
            ++nCut;
        }
        Wide8 w;
        const float inf = std::numeric_limits<float>::infinity();
        for (int i = 0; i < 8; ++i) {
            for (int c = 0; c < 3; ++c) {
                w.box[i][c] = inf;
                w.box[i][3 + c] = -inf;
            }
            w.ref[i] = kWideEmpty;
        }
        const uint32_t stackedBelow = job.stacked + uint32_t(nCut) - 1;
        maxStack = std::max(maxStack, job.stacked + uint32_t(nCut));
        for (int i = 0; i < nCut; ++i) {
            const Node32 &c = nodes[cut[i]];
            std::memcpy(w.box[i], c.bmin, 12);
            std::memcpy(w.box[i] + 3, c.bmax, 12);
            if (c.nPrims > 0) {
                w.ref[i] = ~c.offset;
            } else {
                w.ref[i] = int32_t(out.size());
                out.emplace_back();
                jobs.push_back({cut[i], uint32_t(w.ref[i]), stackedBelow});
            }
        }
        for (int cls = 0; cls < 4; ++cls) {
            const int neg[3] = {cls & 1, (cls >> 1) & 1, 0};
            uint8_t order[8];
            int k = 0;
            VisitOrder(nodes, job.bvh2, cut, nCut, neg, order, k);
            uint8_t seen = 0;
            for (int i = 0; i < 8; ++i) w.pre[cls][i] = 0;
            for (int p = 0; p < k; ++p) {
                w.pre[cls][order[p]] = seen;
                seen |= uint8_t(1u << order[p]);
            }
        }
        out[job.wide] = w;
    }
    if (maxStackOut) *maxStackOut = maxStack;
    return rootWide;
}

}  // namespace b200pt
