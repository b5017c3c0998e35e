// scene_host.h - host-side flattening of a scene description into the POD
// tables of device_scene.h.  Pure host code (no CUDA), compiled with
// -ffp-contract=off because triangle bounds / SAH decisions must match the
// reference float-for-float.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/b200pt.h"
#include "bvh2.h"
#include "device_scene.h"
#include "wide.h"

namespace b200pt {

struct HostTree {
    int32_t object = -1;
    std::vector<Node32> nodes;       // offsets relative to this tree
    std::vector<uint32_t> prims;     // 3 per ordered primitive {kind, a, b}
};

struct HostScene {
    std::vector<HostTree> trees;     // [0] = top level
    std::vector<Node32> nodes;       // all trees, absolute child / slot offsets
    std::vector<PrimSlot> prims;
    std::vector<PairNode> pairs;     // production layout (interior nodes only)
    std::vector<int32_t> treeRootRef; // per tree: PairNode index or ~slot
    int32_t rootRef = 0;
    std::vector<Wide8> wide;         // 8-wide collapse of the same trees (traverse_wide.cu)
    std::vector<LeafBox> leafBoxes;  // the leaves as the wide walk enters them (wide.h)
    std::vector<int32_t> treeRootWide; // per tree: Wide8 index or ~(LeafBox index)
    int32_t rootWide = 0;
    uint32_t maxStackWide = 1;       // stack entries the wide traversal can need (top level + deepest object)
    std::vector<InstanceDev> instances;
    std::vector<uint32_t> treeRoot;  // root node of trees[i] in `nodes`
    uint32_t rootNode = 0;
    uint32_t maxDepth = 1;           // deepest path over all trees, in nodes
    float bounds[6] = {0, 0, 0, 0, 0, 0};
    uint64_t nTriangles = 0;         // primitives of both kinds
    bool hasPatches = false;
    uint32_t nMeshes = 0;
    double buildSeconds = 0;
};

// Returns B200PT_OK or B200PT_ERR_INVALID (err filled).
int AssembleScene(const b200pt_scene_desc *desc, HostScene &out, std::string &err);

}  // namespace b200pt
