// device_scene.h - flat POD tables the kernels read (explicit device memory; no
// managed memory, no host class pointers).  See DESIGN.md "Data layout in HBM".
#pragma once
#include <cstdint>

namespace b200pt {

// 64 B primitive slot (32-byte aligned: one 256-bit + one 128-bit load), in depth-first
// leaf order of its tree.
//   triangle: p0,p1,p2 pre-gathered (the reference does vertexIndices -> p[] per
//             test, shapes.cpp:326-328), mesh/tri = Triangle{meshIndex,triIndex}.
//   instance: mesh == kInstanceTag, tri = index into DeviceScene::instances.
//   bilinear patch: kind == 1; p0, p1, p2, p3 = p00, p10, p01, p11 (shapes.cpp:1134-1137), tri = patch index.
struct alignas(32) PrimSlot {
    float p0[3];
    float p1[3];
    float p2[3];
    uint32_t mesh;
    uint32_t tri;
    uint32_t flags;
    float p3[3];
    uint32_t kind;  // 0 triangle (or instance), 1 bilinear patch, 2 triangle with a constant alpha cut-out (alpha in p3[0])
};
static_assert(sizeof(PrimSlot) == 64, "PrimSlot is one 32-byte and one 16-byte load");
constexpr uint32_t kInstanceTag = 0xFFFFFFFEu;
constexpr uint32_t kMissTag = 0xFFFFFFFFu;

// 64 B interior node of the production kernel: both children's exact fp32 boxes live in the
// parent (two 256-bit loads per step, no separate leaf nodes).  child0 is the reference's first
// child (flat index + 1), child1 its second child (cpu/aggregates.cpp:505-522); `axis` is the
// split axis that picks the near child (cpu/aggregates.cpp:562-568).
//   ref >= 0: interior PairNode index;  ref < 0: leaf, first PrimSlot = ~ref (its count is in that
//   slot's `flags`).
struct alignas(32) PairNode {
    float box0[6];   // child0 pMin, pMax
    int32_t ref0;
    uint32_t axis;
    float box1[6];   // child1 pMin, pMax
    int32_t ref1;
    uint32_t pad;
};
static_assert(sizeof(PairNode) == 64, "PairNode is two 32-byte loads");

// One entry per object instance (TransformedPrimitive, cpu/primitive.cpp:112-131).
struct alignas(16) InstanceDev {
    float mInv[16];      // instanceFromRender, row major (the ray goes through this: instance.h)
    float m[16];         // renderFromInstance (the hit comes back through this)
    float rootBox[6];    // bounds of the object's BVH root, object space
    int32_t rootRef;     // PairNode index of the object's root, or ~slot when the object is one primitive
                         // (then there is no BVH and no box test: scene.cpp:1542-1551)
    uint32_t instanceId; // 1-based (scene.cpp:1560)
    uint32_t rootNode;   // the same root in `nodes` (Node32 layout)
    uint32_t object;
    int32_t rootWide;    // the object's root in the wide layout: Wide8 index, or ~(LeafBox index)
    uint32_t pad;
};
static_assert(sizeof(InstanceDev) == 176, "InstanceDev layout");

struct DeviceScene {
    const void *nodes;   // Node32[], 32 B, child/prim offsets absolute (parity kernel)
    const void *pairs;   // PairNode[] (production kernel)
    float rootBox[6];    // bounds of the top-level root (tested once per ray, as the reference does)
    int32_t rootRef;     // PairNode index of the top-level root, or ~slot if the root is a leaf
    const void *prims;   // PrimSlot[]
    const InstanceDev *instances;
    uint32_t rootNode;
    uint32_t nNodes;
    uint32_t nPrims;
    uint32_t nInstances;
    uint32_t maxDepth;   // deepest root-to-leaf path over all trees (stack entries needed)
    const void *wide;    // Wide8[] (wide.h), or null: the scene is traversed by the binary kernel only
    int32_t rootWide;    // Wide8 index of the top-level root, or ~(LeafBox index)
    uint32_t maxStackWide;
    const void *leafBoxes; // LeafBox[] (wide.h): the leaves of the wide layout
    uint32_t hasPatches; // some PrimSlot is a bilinear patch (the kernels then read the slots' last 16 bytes)
};

}  // namespace b200pt
