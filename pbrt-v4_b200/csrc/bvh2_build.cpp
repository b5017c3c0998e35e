// bvh2_build.cpp - host SAH builder that makes the same split decisions as the
// reference's BVHAggregate (cpu/aggregates.cpp:192-387: centroid-bounds max
// axis, 12 buckets, cost 1/2 + sum(count*area)/area, leaf when n <= maxPrims
// and cheaper) but is single-pass into the flat depth-first array, with no
// build-node tree, multi-threaded over independent subtrees, and deterministic for any
// input size and thread count (the reference's
// ordered-primitive slots depend on thread timing above 128 K prims, SURVEY F8;
// tree shape and in-leaf order do not, and those are what we reproduce).
//
// MUST be compiled with -ffp-contract=off: every float expression below is
// evaluated in the reference's association order so that equal inputs give
// equal trees (checked node-for-node against the compiled reference in
// tests/test_bvh_build.py).
#include "bvh2.h"

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>

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

// worker threads of one build: B200PT_BUILD_THREADS, else the hardware's (at most 32)
unsigned BuildThreads() {
    if (const char *v = std::getenv("B200PT_BUILD_THREADS")) return unsigned(std::max(1, std::atoi(v)));
    unsigned hw = std::thread::hardware_concurrency();
    return std::max(1u, std::min(32u, hw));
}

// One build over items[lo, hi): split decisions as the reference makes them; nodes and ordered primitives go to
// the caller's arrays (indices relative to those arrays), so that disjoint ranges can be built by different threads.
struct Builder {
    Item *items;
    int maxPrims;

    // bucket of a centroid along `dim` (cpu/aggregates.cpp:283-286; Bounds3::Offset util/vecmath.h:1325)
    static inline int Bucket(const Item &it, const Box &cb, int dim) {
        float o = it.c[dim] - cb.lo[dim];
        if (cb.hi[dim] > cb.lo[dim])
            o /= cb.hi[dim] - cb.lo[dim];
        int b = kBuckets * o;
        return b == kBuckets ? kBuckets - 1 : b;
    }

    // The reference's decision for the node over items[lo, hi): returns false for a leaf, otherwise the split axis
    // and the first item of the second child (items are partitioned in place).
    bool Split(uint32_t lo, uint32_t hi, Box &bounds, int &dim, uint32_t &mid) const {
        const uint32_t cnt = hi - lo;
        bounds = EmptyBox();
        for (uint32_t i = lo; i < hi; ++i)
            Grow(bounds, items[i].b);
        if (Area(bounds) == 0 || cnt == 1)
            return false;
        Box cb = EmptyBox();
        for (uint32_t i = lo; i < hi; ++i)
            GrowPoint(cb, items[i].c);
        dim = WidestAxis(cb);
        if (cb.hi[dim] == cb.lo[dim])
            return false;
        mid = lo + cnt / 2;
        if (cnt <= 2) {
            const int d = dim;
            std::nth_element(items + lo, items + mid, items + hi,
                             [d](const Item &a, const Item &b) { return a.c[d] < b.c[d]; });
            return true;
        }
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
            const int d = dim;
            Item *m = std::partition(items + lo, items + hi, [&](const Item &it) {
                return Bucket(it, cb, d) <= best;
            });
            mid = uint32_t(m - items);
            return true;
        }
        return false;
    }

    static void WriteLeaf(Node32 &nd, const Box &bounds, uint32_t firstOrdered, uint32_t n) {
        std::memcpy(nd.bmin, bounds.lo, 12);
        std::memcpy(nd.bmax, bounds.hi, 12);
        nd.offset = int32_t(firstOrdered);
        nd.nPrims = uint16_t(n);
        nd.axis = 0;
        nd.pad = 0;
    }
    static void WriteInterior(Node32 &nd, const Box &bounds, int dim) {
        std::memcpy(nd.bmin, bounds.lo, 12);
        std::memcpy(nd.bmax, bounds.hi, 12);
        nd.offset = 0;
        nd.nPrims = 0;
        nd.axis = uint8_t(dim);
        nd.pad = 0;
    }

    // Iterative depth-first emission of the subtree over items[lo0, hi0): a work stack of ranges; an interior
    // node patches its second-child offset when that child is emitted.
    void Run(uint32_t lo0, uint32_t hi0, std::vector<Node32> &nodes, std::vector<uint32_t> &order) const {
        struct Job {
            uint32_t lo, hi;
            int32_t patchParent;  // node whose `offset` becomes this node's index, or -1
        };
        std::vector<Job> stack;
        stack.push_back({lo0, hi0, -1});
        while (!stack.empty()) {
            Job job = stack.back();
            stack.pop_back();
            uint32_t me = uint32_t(nodes.size());
            nodes.emplace_back();
            if (job.patchParent >= 0)
                nodes[job.patchParent].offset = int32_t(me);
            Box bounds;
            int dim = 0;
            uint32_t mid = 0;
            if (!Split(job.lo, job.hi, bounds, dim, mid)) {
                WriteLeaf(nodes[me], bounds, uint32_t(order.size()), job.hi - job.lo);
                for (uint32_t i = job.lo; i < job.hi; ++i)
                    order.push_back(items[i].id);
                continue;
            }
            WriteInterior(nodes[me], bounds, dim);
            // first child is emitted next (index me+1), second after its whole subtree
            stack.push_back({mid, job.hi, int32_t(me)});
            stack.push_back({job.lo, mid, -1});
        }
    }
};

// Parallel build: the top of the tree is expanded serially (same decisions, same partitions) down to ranges of at
// most `grain` items; those ranges are independent subtrees, built by worker threads into their own arrays; the
// pieces are then laid out in the one depth-first order.  The result is the serial build's, node for node.
struct TopNode {
    Box bounds;
    int dim = 0;
    int32_t first = -1, second = -1;  // TopNode children, or
    int32_t task = -1;                // the subtree task that stands here
};
struct Task {
    uint32_t lo, hi;
    std::vector<Node32> nodes;
    std::vector<uint32_t> order;
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
    Builder b{items.data(), std::min(255, maxPrimsInNode)};
    unsigned nThreads = BuildThreads();
    constexpr uint32_t kSerialBelow = 1u << 16;
    if (nThreads <= 1 || n < kSerialBelow) {
        nodes.reserve(2 * size_t(n));
        order.reserve(n);
        b.Run(0, n, nodes, order);
        return;
    }
    // ---- serial top
    const uint32_t grain = std::max<uint32_t>(1u << 13, n / (nThreads * 8));
    std::vector<TopNode> top;
    std::vector<Task> tasks;
    {
        struct Job {
            uint32_t lo, hi;
            int32_t parent;
            bool second;
        };
        std::vector<Job> stack{{0, n, -1, false}};
        while (!stack.empty()) {
            Job job = stack.back();
            stack.pop_back();
            const int32_t me = int32_t(top.size());
            top.emplace_back();
            if (job.parent >= 0)
                (job.second ? top[job.parent].second : top[job.parent].first) = me;
            uint32_t mid = 0;
            if (job.hi - job.lo <= grain || !b.Split(job.lo, job.hi, top[me].bounds, top[me].dim, mid)) {
                // small enough (or a leaf): the whole subtree, root included, is one task
                top[me].task = int32_t(tasks.size());
                tasks.push_back(Task{job.lo, job.hi, {}, {}});
                continue;
            }
            stack.push_back({mid, job.hi, me, true});
            stack.push_back({job.lo, mid, me, false});
        }
    }
    // ---- subtrees in parallel
    {
        std::atomic<size_t> next{0};
        auto work = [&]() {
            for (size_t t; (t = next.fetch_add(1)) < tasks.size();)
                b.Run(tasks[t].lo, tasks[t].hi, tasks[t].nodes, tasks[t].order);
        };
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < nThreads; ++t) pool.emplace_back(work);
        work();
        for (std::thread &t : pool) t.join();
    }
    // ---- one depth-first layout (tasks are met in item order, so `order` is the concatenation)
    size_t nNodes = 0;
    for (const Task &t : tasks) nNodes += t.nodes.size();
    nodes.reserve(nNodes + top.size());
    order.reserve(n);
    struct Emit {
        int32_t topNode;
        int32_t patchParent;
    };
    std::vector<Emit> stack{{0, -1}};
    while (!stack.empty()) {
        Emit e = stack.back();
        stack.pop_back();
        const uint32_t me = uint32_t(nodes.size());
        if (e.patchParent >= 0)
            nodes[e.patchParent].offset = int32_t(me);
        const TopNode &tn = top[e.topNode];
        if (tn.task >= 0) {
            const Task &t = tasks[tn.task];
            const int32_t primBase = int32_t(order.size());
            for (Node32 nd : t.nodes) {
                nd.offset += nd.nPrims ? primBase : int32_t(me);
                nodes.push_back(nd);
            }
            order.insert(order.end(), t.order.begin(), t.order.end());
            continue;
        }
        nodes.emplace_back();
        Builder::WriteInterior(nodes[me], tn.bounds, tn.dim);
        stack.push_back({tn.second, int32_t(me)});
        stack.push_back({tn.first, -1});
    }
}

}  // namespace b200pt
