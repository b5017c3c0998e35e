// wide.h - the 8-wide node of the production traversal kernel (traverse_persist.cu, WIDE variants).
//
// A wide node stands for one interior node R of the reference's binary tree (LinearBVHNode,
// cpu/aggregates.cpp:129-137) and holds a CUT of R's subtree: up to eight descendants, each a binary
// leaf or a binary interior node (which then has a wide node of its own), with the reference's own fp32
// box of that descendant.  The binary nodes between R and the cut are not tested.
//
// Why the hits stay bit-identical to BVHAggregate::Intersect (cpu/aggregates.cpp:529-579):
//   (1) Bounds3::IntersectP (util/vecmath.h:1576-1610) is monotone in the box for a ray whose 1/d is
//       finite: a box inside another one has tMin >= and tMax <= the outer one's, float for float
//       (subtraction and multiplication by a fixed factor are monotone), so "a leaf passes its own
//       test with the tMax of its turn" implies "every ancestor passed at its turn" - between an
//       ancestor's turn and the first leaf under it no primitive is tested, and if the ancestor fails
//       nothing under it can pass either.  Hence: the reference visits exactly the leaves that pass
//       their own test at their turn, and it is enough to (a) take the leaves in the reference's order
//       and (b) apply the EXACT test to each leaf's box with the tMax of that moment.  What is tested on
//       the way down only has to let every such leaf through: the wide node's eight boxes get a
//       CONSERVATIVE test - one fma per plane against per-ray constants widened by an error bound
//       (isect.h WideRay) - and the exact test is made once per candidate leaf, on the leaf's own box
//       (LeafBox), when the leaf is entered.
//   (2) The order: near child first by dirIsNeg[axis] at every binary node (cpu/aggregates.cpp:562-568).
//       Inside a wide node that is a permutation of the slots per ray octant; `pre` stores, for the four
//       octants with dirIsNeg[2] == 0, the set of slots visited before each slot (the other four octants
//       visit in exactly the reversed order).
//   (3) tMax is not monotone (an accepted t may round above the tMax it was tested against,
//       shapes.cpp:239-250), so children are stacked when their conservative entry distance is below
//       `tMax * (1 + 2^-10)`; stale entries are dropped at pop time by the same conservative distance
//       against the tMax of that moment, and the leaf's exact test decides.
//   Rays with a non-finite 1/d, origin or error bound (0 * inf = NaN breaks (1)) take the exact binary
//   kernel instead; the wide kernel hands them over through a device-side list.
#pragma once
#include <cstdint>
#include <vector>

#include "bvh2.h"

namespace b200pt {

struct alignas(128) Wide8 {
    // plane p of child i: p = 0..2 pMin.xyz, 3..5 pMax.xyz (empty slot: +inf / -inf, never passes).  One 32-byte
    // load per plane; the ray's near plane along axis k is plane (dirIsNeg[k] ? 3 + k : k) - chosen by ADDRESS, so
    // the kernel does the reference's `dirIsNeg ? pMax : pMin` selection (util/vecmath.h:1580-1586) with no select
    float plane[6][8];
    int32_t ref[8];     // >= 0: Wide8 index; < 0: leaf, LeafBox index = ~ref
    uint8_t pre[4][8];  // [negx | negy << 1][slot] = bit mask of the slots visited before `slot` (negz = 0)
};
static_assert(sizeof(Wide8) == 256, "Wide8 is eight 32-byte loads");
constexpr int32_t kWideEmpty = int32_t(0x7fffffff);

// One leaf of a binary tree as the wide walk sees it: the reference's box of the leaf node (tested exactly, with the
// tMax of the moment, when the walk reaches the leaf), its first PrimSlot and its primitive count.
// kLeafNoTest in `count`: the leaf is the root of its tree - the top-level root's box is tested when the ray is
// fetched, a one-primitive object has no BVH and no box test at all (scene.cpp:1542-1551).
struct alignas(32) LeafBox {
    float box[6];  // pMin, pMax
    int32_t first;
    uint32_t count;
};
static_assert(sizeof(LeafBox) == 32, "LeafBox is one 32-byte load");
constexpr uint32_t kLeafNoTest = 0x80000000u;

// Collapses one binary tree (nodes[root..], absolute child offsets, leaves' `offset` = first PrimSlot)
// into wide nodes appended to `out` and leaf records appended to `leaves`.  Returns the tree's root reference: a
// Wide8 index, or ~(LeafBox index) when the root is a leaf.  `maxStackOut` = the most traversal-stack entries a ray can hold inside this tree
// (every visited node stacks all its children and pops one).
int32_t BuildWide(const std::vector<Node32> &nodes, uint32_t root, std::vector<Wide8> &out, std::vector<LeafBox> &leaves,
                  uint32_t *maxStackOut);

}  // namespace b200pt
