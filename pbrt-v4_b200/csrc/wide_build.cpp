// wide_build.cpp - collapses the reference-shaped binary tree into 8-wide nodes (wide.h).
// The binary tree stays the authority on boxes, leaves and visit order; this only chooses, per wide
// node, which descendants of its binary node to put side by side, and records the reference's visit
// order of those descendants for every ray octant.  The choice minimises the surface-area-weighted
// number of wide nodes a ray fetches, by the dynamic programme of Ylitie, Karras and Laine (HPG 2017,
// section 3.1) restricted to the given leaves: cost(n, i) = cheapest way to stand for the subtree of n with
// at most i slots of its ancestor's wide node.
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

int32_t BuildWide(const std::vector<Node32> &nodes, uint32_t root, std::vector<Wide8> &out, std::vector<LeafBox> &leaves,
                  uint32_t *maxStackOut) {
    if (maxStackOut) *maxStackOut = 1;
    auto addLeaf = [&](const Node32 &nd, uint32_t flag) {
        LeafBox lb;
        std::memcpy(lb.box, nd.bmin, 12);
        std::memcpy(lb.box + 3, nd.bmax, 12);
        lb.first = nd.offset;
        lb.count = uint32_t(nd.nPrims) | flag;
        leaves.push_back(lb);
        return ~int32_t(leaves.size() - 1);
    };
    if (nodes[root].nPrims > 0) return addLeaf(nodes[root], kLeafNoTest);
    // extent of the tree in the depth-first array: up to the last node of the root's second subtree
    uint32_t last = root;
    while (nodes[last].nPrims == 0) last = uint32_t(nodes[last].offset);
    const uint32_t nTree = last - root + 1;
    // cost[i - 1] = cost(n, i), i = 1..7; pick[i - 2], i = 2..8: slots given to the first child
    // (0 = "as with i - 1 slots").  Leaves cost nothing (their count and boxes are given).
    struct DP {
        float cost[7];
        uint8_t pick[7];
    };
    std::vector<DP> dp(nTree);
    for (uint32_t j = nTree; j-- > 0;) {  // children sit behind their parent in the array
        const Node32 &nd = nodes[root + j];
        DP &d = dp[j];
        if (nd.nPrims > 0) {
            for (int i = 0; i < 7; ++i) d.cost[i] = 0.f;
            std::memset(d.pick, 0, sizeof d.pick);
            continue;
        }
        const DP &L = dp[j + 1], &R = dp[uint32_t(nd.offset) - root];
        auto distribute = [&](int slots, uint8_t &pickOut) {
            float best = std::numeric_limits<float>::infinity();
            for (int k = 1; k < slots; ++k) {
                const int kr = slots - k;
                if (k > 7 || kr > 7) continue;
                float c = L.cost[k - 1] + R.cost[kr - 1];
                if (c < best) {
                    best = c;
                    pickOut = uint8_t(k);
                }
            }
            return best;
        };
        float area = BoxArea(nd);
        if (!(area == area)) area = 0.f;
        d.pick[6] = 4;
        d.cost[0] = area + distribute(8, d.pick[6]);
        for (int i = 2; i <= 7; ++i) {
            uint8_t pk = 0;
            float c = distribute(i, pk);
            if (c < d.cost[i - 2]) {
                d.cost[i - 1] = c;
                d.pick[i - 2] = pk;
            } else {
                d.cost[i - 1] = d.cost[i - 2];
                d.pick[i - 2] = 0;
            }
        }
    }
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
        int nCut = 0;
        // hand the 8 slots down the binary subtree as the programme decided; first child first, so the slots
        // come out in the reference's first-child-first depth-first order
        struct Give {
            uint32_t node;
            int slots;
        };
        Give todo[16];
        int nTodo = 0;
        {
            const int k = dp[job.bvh2 - root].pick[6];
            todo[nTodo++] = {uint32_t(nodes[job.bvh2].offset), 8 - k};
            todo[nTodo++] = {job.bvh2 + 1, k};
        }
        while (nTodo > 0) {
            Give g = todo[--nTodo];
            const Node32 &nd = nodes[g.node];
            if (nd.nPrims > 0 || g.slots == 1) {
                cut[nCut++] = g.node;
                continue;
            }
            int i = std::min(g.slots, 7);
            while (i >= 2 && dp[g.node - root].pick[i - 2] == 0) --i;
            if (i < 2) {
                cut[nCut++] = g.node;
                continue;
            }
            const int k = dp[g.node - root].pick[i - 2];
            todo[nTodo++] = {uint32_t(nd.offset), i - k};
            todo[nTodo++] = {g.node + 1, k};
        }

        Wide8 w;
        const float inf = std::numeric_limits<float>::infinity();
        for (int i = 0; i < 8; ++i) {
            for (int c = 0; c < 3; ++c) {
                w.plane[c][i] = inf;
                w.plane[3 + c][i] = -inf;
            }
            w.ref[i] = kWideEmpty;
        }
        const uint32_t stackedBelow = job.stacked + uint32_t(nCut) - 1;
        maxStack = std::max(maxStack, job.stacked + uint32_t(nCut));
        for (int i = 0; i < nCut; ++i) {
            const Node32 &c = nodes[cut[i]];
            for (int a = 0; a < 3; ++a) {
                w.plane[a][i] = c.bmin[a];
                w.plane[3 + a][i] = c.bmax[a];
            }
            if (c.nPrims > 0) {
                w.ref[i] = addLeaf(c, 0u);
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
