// bvh2_build.cpp - host SAH builder that makes the same split decisions as the
// reference's BVHAggregate (cpu/aggregates.cpp:192-387: centroid-bounds max
// axis, 12 buckets, cost 1/2 + sum(count*area)/area, leaf when n <= maxPrims
// and cheaper) but is single-pass into the flat depth-first array, with no
// build-node tree, and deterministic for any input size (the reference's
// ordered-primitive slots depend on thread timing above 128 K prims, SURVEY F8;
// tree shape and in-leaf order do not, and those are what we reproduce).
//
// MUST be compiled with -ffp-contract=off: every float expression below is
// evaluated in the reference's association order so that equal inputs give
// equal trees (checked node-for-node against the compiled reference in
// tests/test_bvh_build.py).
#include "bvh2.h"

#include <algorithm>
#include <cstring>
#include <limits>

namespace b200pt {
namespace {

struct Item {
    Box b;
    float c[3];  // centroid = .5f*lo + .5f*hi (cpu/aggregates.cpp:98)
    uint32_t id;
};

inline Box EmptyBox() {
    const float mx = std::numeric_limits<float>::max(), mn = std::numeric_limits<float>::lowest();
    return Box{{mx, mx, mx}, {mn, mn, mn}};
}
inline void Grow(Box &a, const Box &b) {
    for (int k = 0; k < 3; ++k) {
        a.lo[k] = std::min(a.lo[k], b.lo[k]);
        a.hi[k] = std::max(a.hi[k], b.hi[k]);
    }
}
inline void GrowPoint(Box &a, const float *p) {
    for (int k = 0; k < 3; ++k) {
        a.lo[k] = std::min(a.lo[k], p[k]);
        a.hi[k] = std::max(a.hi[k], p[k]);
    }
}
// Bounds3::SurfaceArea (util/vecmath.h:1296-1299)
inline float Area(const Box &b) {
    float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return 2 * (dx * dy + dx * dz + dy * dz);
}
// Bounds3::MaxDimension (util/vecmath.h:1308-1316)
inline int WidestAxis(const Box &b) {
    float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    if (dx > dy && dx > dz)
        return 0;
    return dy > dz ? 1 : 2;
}

constexpr int kBuckets = 12;

struct Builder {
    Item *items;
    int maxPrims;
    std::vector<Node32> &nodes;
    std::vector<uint32_t> &order;

    // bucket of a centroid along `dim` (cpu/aggregates.cpp:283-286; Bounds3::Offset util/vecmath.h:1325)
    static inline int Bucket(const Item &it, const Box &cb, int dim) {
        float o = it.c[dim] - cb.lo[dim];
        if (cb.hi[dim] > cb.lo[dim])
            o /= cb.hi[dim] - cb.lo[dim];
        int b = kBuckets * o;
        return b == kBuckets ? kBuckets - 1 : b;
    }

    void EmitLeaf(uint32_t me, const Box &bounds, uint32_t lo, uint32_t hi) {
        Node32 &nd = nodes[me];
        std::memcpy(nd.bmin, bounds.lo, 12);
        std::memcpy(nd.bmax, bounds.hi, 12);
        nd.offset = int32_t(order.size());
        nd.nPrims = uint16_t(hi - lo);
        nd.axis = 0;
        nd.pad = 0;
        for (uint32_t i = lo; i < hi; ++i)
            order.push_back(items[i].id);
    }

    // Iterative depth-first emission: a work stack of ranges; an interior node
    // patches its second-child offset when that child is emitted.
    void Run(uint32_t n) {
        struct Job {
            uint32_t lo, hi;
            int32_t patchParent;  // node whose `offset` becomes this node's index, or -1
        };
        std::vector<Job> stack;
        stack.push_back({0, n, -1});
        while (!stack.empty()) {
            Job job = stack.back();
            stack.pop_back();
            uint32_t lo = job.lo, hi = job.hi, cnt = hi - lo;
            uint32_t me = uint32_t(nodes.size());
            nodes.emplace_back();
            if (job.patchParent >= 0)
                nodes[job.patchParent].offset = int32_t(me);

            Box bounds = EmptyBox();
            for (uint32_t i = lo; i < hi; ++i)
                Grow(bounds, items[i].b);
            if (Area(bounds) == 0 || cnt == 1) {
                EmitLeaf(me, bounds, lo, hi);
                continue;
            }
            Box cb = EmptyBox();
            for (uint32_t i = lo; i < hi; ++i)
                GrowPoint(cb, items[i].c);
            int dim = WidestAxis(cb);
            if (cb.hi[dim] == cb.lo[dim]) {
                EmitLeaf(me, bounds, lo, hi);
                continue;
            }
            uint32_t mid = lo + cnt / 2;
            if (cnt <= 2) {
                std::nth_element(items + lo, items + mid, items + hi,
                                 [dim](const Item &a, const Item &b) { return a.c[dim] < b.c[dim]; });
            } else {
                int count[kBuckets] = {};
                Box bb[kBuckets];
                for (int k = 0; k < kBuckets; ++k)
                    bb[k] = EmptyBox();
                for (uint32_t i = lo; i < hi; ++i) {
                    int b = Bucket(items[i], cb, dim);
                    count[b]++;
                    Grow(bb[b], items[i].b);
                }
                float cost[kBuckets - 1] = {};
                {
                    int below = 0;
                    Box acc = EmptyBox();
                    for (int s = 0; s < kBuckets - 1; ++s) {
                        Grow(acc, bb[s]);
                        below += count[s];
                        cost[s] += below * Area(acc);
                    }
                    int above = 0;
                    acc = EmptyBox();
                    for (int s = kBuckets - 1; s >= 1; --s) {
                        Grow(acc, bb[s]);
                        above += count[s];
                        cost[s - 1] += above * Area(acc);
                    }
                }
                int best = -1;
                float bestCost = std::numeric_limits<float>::infinity();
                for (int s = 0; s < kBuckets - 1; ++s)
                    if (cost[s] < bestCost) {
                        bestCost = cost[s];
                        best = s;
                    }
                float leafCost = cnt;
                bestCost = 1.f / 2.f + bestCost / Area(bounds);
                if (cnt > uint32_t(maxPrims) || bestCost < leafCost) {
                    Item *m = std::partition(items + lo, items + hi, [&](const Item &it) {
                        return Bucket(it, cb, dim) <= best;
                    });
                    mid = uint32_t(m - items);
                } else {
                    EmitLeaf(me, bounds, lo, hi);
                    continue;
                }
            }
            Node32 &nd = nodes[me];
            std::memcpy(nd.bmin, bounds.lo, 12);
            std::memcpy(nd.bmax, bounds.hi, 12);
            nd.offset = 0;
            nd.nPrims = 0;
            nd.axis = uint8_t(dim);
            nd.pad = 0;
            // first child is emitted next (index me+1), second after its whole subtree
            stack.push_back({mid, hi, int32_t(me)});
            stack.push_back({lo, mid, -1});
        }
    }
};

}  // namespace

void BuildBVH2(const Box *boxes, uint32_t n, int maxPrimsInNode, std::vector<Node32> &nodes,
               std::vector<uint32_t> &order) {
    nodes.clear();
    order.clear();
    if (n == 0)
        return;
    std::vector<Item> items(n);
    for (uint32_t i = 0; i < n; ++i) {
        items[i].b = boxes[i];
        for (int k = 0; k < 3; ++k)
            items[i].c[k] = .5f * boxes[i].lo[k] + .5f * boxes[i].hi[k];
        items[i].id = i;
    }
    nodes.reserve(2 * size_t(n));
    order.reserve(n);
    Builder b{items.data(), std::min(255, maxPrimsInNode), nodes, order};
    b.Run(n);
}

}  // namespace b200pt
