// device_scene.h - flat POD tables the kernels read (explicit device memory; no
// managed memory, no host class pointers).  See DESIGN.md "Data layout in HBM".
#pragma once
#include <cstdint>

namespace b200pt {

// 48 B primitive slot, in depth-first leaf order of its tree.
//   triangle: p0,p1,p2 pre-gathered (the reference does vertexIndices -> p[] per
//             test, shapes.cpp:326-328), mesh/tri = Triangle{meshIndex,triIndex}.
//   instance: mesh == kInstanceTag, tri = index into DeviceScene::instances.
struct alignas(16) PrimSlot {
    float p0[3];
    float p1[3];
    float p2[3];
    uint32_t mesh;
    uint32_t tri;
    uint32_t flags;  // bit0: degenerate triangle (never hit, shapes.cpp:176-177)
};
static_assert(sizeof(PrimSlot) == 48, "PrimSlot is three 16-byte loads");
constexpr uint32_t kInstanceTag = 0xFFFFFFFEu;
constexpr uint32_t kMissTag = 0xFFFFFFFFu;

// One entry per object instance (TransformedPrimitive, cpu/primitive.cpp:112-131).
struct alignas(16) InstanceDev {
    float m[12];     // renderFromInstance rows 0..2
    float mInv[12];  // its inverse, rows 0..2
    uint32_t rootNode;   // root of the object's tree in `nodes`
    uint32_t instanceId; // 1-based (scene.cpp:1560)
    uint32_t pad[2];
};

struct DeviceScene {
    const void *nodes;   // Node32[], 32 B, child/prim offsets absolute
    const void *prims;   // PrimSlot[]
    const InstanceDev *instances;
    uint32_t rootNode;
    uint32_t nNodes;
    uint32_t nPrims;
    uint32_t nInstances;
};

}  // namespace b200pt
