// bvh2.h - host-side binary BVH in the reference's flat layout.
// Node32 mirrors LinearBVHNode (reference cpu/aggregates.cpp:129-137) field for
// field so a tree read out of the host's BVHAggregate and one built here are
// interchangeable inputs to the wide-BVH flattening.
#pragma once
#include <cstdint>
#include <vector>

namespace b200pt {

struct alignas(32) Node32 {
    float bmin[3];
    float bmax[3];
    int32_t offset;      // leaf: first ordered primitive; interior: second child
    uint16_t nPrims;     // 0 = interior
    uint8_t axis;        // interior: split axis
    uint8_t pad;
};
static_assert(sizeof(Node32) == 32, "Node32 must match LinearBVHNode");

struct Box {
    float lo[3], hi[3];
};

// Builds the SAH tree over `boxes` with the reference's decisions
// (cpu/aggregates.cpp:192-387) and writes it depth-first (cpu/aggregates.cpp:505-522).
// order[k] = input index of the k-th ordered primitive (leaves in DFS order).
void BuildBVH2(const Box *boxes, uint32_t n, int maxPrimsInNode, std::vector<Node32> &nodes,
               std::vector<uint32_t> &order);

}  // namespace b200pt
