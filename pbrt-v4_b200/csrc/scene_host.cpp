// scene_host.cpp - scene description -> flat tables (see scene_host.h).
#include "scene_host.h"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <limits>
#include <map>

namespace b200pt {
namespace {

// Triangle::Bounds = Union(Bounds3f(p0, p1), p2)  (reference shapes.cpp:294-301)
inline Box TriBox(const float *p0, const float *p1, const float *p2) {
    Box b;
    for (int k = 0; k < 3; ++k) {
        float lo = std::min(p0[k], p1[k]), hi = std::max(p0[k], p1[k]);
        b.lo[k] = std::min(lo, p2[k]);
        b.hi[k] = std::max(hi, p2[k]);
    }
    return b;
}

// BilinearPatch::Bounds = Union(Bounds3f(p00, p01), Bounds3f(p10, p11)) (shapes.cpp:1073-1081)
inline Box PatchBox(const float *p00, const float *p10, const float *p01, const float *p11) {
    Box b;
    for (int k = 0; k < 3; ++k) {
        b.lo[k] = std::min(std::min(p00[k], p01[k]), std::min(p10[k], p11[k]));
        b.hi[k] = std::max(std::max(p00[k], p01[k]), std::max(p10[k], p11[k]));
    }
    return b;
}
// box and BVH record of primitive i of a mesh of either kind
inline Box MeshPrimBox(const b200pt_mesh &mesh, uint32_t i) {
    if (mesh.kind & B200PT_MESH_BILINEAR) {
        const int32_t *v = mesh.indices + 4ull * i;
        return PatchBox(mesh.p + 3ull * v[0], mesh.p + 3ull * v[1], mesh.p + 3ull * v[2], mesh.p + 3ull * v[3]);
    }
    const int32_t *v = mesh.indices + 3ull * i;
    return TriBox(mesh.p + 3ull * v[0], mesh.p + 3ull * v[1], mesh.p + 3ull * v[2]);
}

struct PrimRef {
    uint32_t kind, a, b;
};

bool ValidateTree(const HostTree &t, uint32_t nPrims, std::string &err) {
    const uint32_t nn = uint32_t(t.nodes.size());
    if (nn == 0) {
        err = "bvh with no nodes";
        return false;
    }
    // every node of the array reached exactly once from the root, offsets in range (the flatten and pair-node loops
    // below walk the whole array: an unreachable node with a garbage offset would index out of bounds there)
    std::vector<uint32_t> stack{0};
    std::vector<uint8_t> seen(nn, 0);
    uint32_t visited = 0;
    while (!stack.empty()) {
        uint32_t i = stack.back();
        stack.pop_back();
        if (i >= nn) {
            err = "bvh node index out of range";
            return false;
        }
        if (seen[i]) {
            err = "bvh node reached twice";
            return false;
        }
        seen[i] = 1;
        ++visited;
        const Node32 &nd = t.nodes[i];
        if (nd.nPrims > 0) {
            if (nd.offset < 0 || uint64_t(nd.offset) + nd.nPrims > nPrims) {
                err = "bvh leaf primitive range out of bounds";
                return false;
            }
        } else {
            if (nd.axis > 2 || nd.offset <= int32_t(i) || uint32_t(nd.offset) >= nn) {
                err = "bvh interior node malformed";
                return false;
            }
            stack.push_back(uint32_t(nd.offset));
            stack.push_back(i + 1);
        }
    }
    if (visited != nn) {
        err = "bvh has nodes that are not reachable from its root";
        return false;
    }
    return true;
}

}  // namespace

int AssembleScene(const b200pt_scene_desc *desc, HostScene &out, std::string &err) {
    auto t0 = std::chrono::steady_clock::now();
    if (!desc || (desc->n_meshes && !desc->meshes) || (desc->n_instances && !desc->instances) ||
        (desc->n_bvhs && !desc->bvhs)) {
        err = "null scene description";
        return B200PT_ERR_INVALID;
    }
    // ---- index the meshes by object; check indices
    std::map<int32_t, std::vector<uint32_t>> objMeshes;  // object -> mesh ids, in array order
    uint64_t nTri = 0;
    for (uint32_t m = 0; m < desc->n_meshes; ++m) {
        const b200pt_mesh &mesh = desc->meshes[m];
        if ((mesh.n_vertices && !mesh.p) || (mesh.n_triangles && !mesh.indices)) {
            err = "mesh with null arrays";
            return B200PT_ERR_INVALID;
        }
        if (mesh.kind > (B200PT_MESH_BILINEAR | B200PT_MESH_ALPHA) || mesh.kind == (B200PT_MESH_BILINEAR | B200PT_MESH_ALPHA)) {
            err = "unknown mesh kind (alpha cut-outs: triangle meshes only)";
            return B200PT_ERR_INVALID;
        }
        if ((mesh.kind & B200PT_MESH_ALPHA) && !(mesh.alpha == mesh.alpha)) {
            err = "mesh alpha is NaN";
            return B200PT_ERR_INVALID;
        }
        const uint64_t perPrim = (mesh.kind & B200PT_MESH_BILINEAR) ? 4 : 3;
        for (uint64_t i = 0; i < perPrim * mesh.n_triangles; ++i)
            if (mesh.indices[i] < 0 || uint32_t(mesh.indices[i]) >= mesh.n_vertices) {
                err = "vertex index out of range";
                return B200PT_ERR_INVALID;
            }
        // patches and alpha triangles use the slots' last 16 bytes: the kernels' extended-slot variant
        if ((mesh.kind & (B200PT_MESH_BILINEAR | B200PT_MESH_ALPHA)) && mesh.n_triangles) out.hasPatches = true;
        objMeshes[mesh.object < 0 ? -1 : mesh.object].push_back(m);
        nTri += mesh.n_triangles;
    }
    out.nTriangles = nTri;
    out.nMeshes = desc->n_meshes;
    for (uint32_t i = 0; i < desc->n_instances; ++i) {
        const b200pt_instance &in = desc->instances[i];
        if (!objMeshes.count(int32_t(in.object)) || int32_t(in.object) < 0) {
            err = "instance of an undefined object";
            return B200PT_ERR_INVALID;
        }
    }
    if (nTri == 0) {
        err = "scene has no primitives";
        return B200PT_ERR_INVALID;
    }

    // ---- trees: take the host's, or build with the reference's decisions
    out.trees.clear();
    if (desc->n_bvhs) {
        for (uint32_t b = 0; b < desc->n_bvhs; ++b) {
            const b200pt_bvh2 &src = desc->bvhs[b];
            if (!src.nodes || !src.prims || !src.n_nodes) {
                err = "bvh with null arrays";
                return B200PT_ERR_INVALID;
            }
            HostTree t;
            t.object = src.object < 0 ? -1 : src.object;
            t.nodes.resize(src.n_nodes);
            std::memcpy(t.nodes.data(), src.nodes, size_t(src.n_nodes) * 32);
            t.prims.assign(src.prims, src.prims + 3ull * src.n_prims);
            if (!ValidateTree(t, src.n_prims, err))
                return B200PT_ERR_INVALID;
            out.trees.push_back(std::move(t));
        }
        auto it = std::find_if(out.trees.begin(), out.trees.end(),
                               [](const HostTree &t) { return t.object == -1; });
        if (it == out.trees.end()) {
            err = "no top-level bvh given";
            return B200PT_ERR_INVALID;
        }
        std::iter_swap(it, out.trees.begin());  // top level first
    } else {
        // primitive order = the order the reference creates them in: meshes in
        // array order, triangles in index order (shapes.cpp:284-289), world
        // shapes before instances (scene.cpp:1447-1580).
        // object definitions: one BVH each with maxPrimsInNode 1 (scene.cpp:1542 -> BVHAggregate's
        // default); an object of a single primitive gets a one-leaf tree, which the traversal enters
        // without a box test (the reference holds that primitive directly, scene.cpp:1551).
        std::vector<HostTree> objTrees;
        std::map<int32_t, Box> objBounds;
        for (auto &kv : objMeshes) {
            if (kv.first == -1)
                continue;
            HostTree t;
            t.object = kv.first;
            std::vector<Box> ob;
            std::vector<PrimRef> orf;
            for (uint32_t m : kv.second) {
                const b200pt_mesh &mesh = desc->meshes[m];
                for (uint32_t i = 0; i < mesh.n_triangles; ++i) {
                    ob.push_back(MeshPrimBox(mesh, i));
                    orf.push_back({(mesh.kind & B200PT_MESH_BILINEAR) ? 2u : 0u, m, i});
                }
            }
            if (ob.empty()) {
                err = "object definition without triangles";
                return B200PT_ERR_INVALID;
            }
            std::vector<uint32_t> oorder;
            BuildBVH2(ob.data(), uint32_t(ob.size()), 1, t.nodes, oorder);
            for (uint32_t id : oorder) {
                t.prims.push_back(orf[id].kind);
                t.prims.push_back(orf[id].a);
                t.prims.push_back(orf[id].b);
            }
            Box bb;
            std::memcpy(bb.lo, t.nodes[0].bmin, 12);
            std::memcpy(bb.hi, t.nodes[0].bmax, 12);
            objBounds[kv.first] = bb;
            objTrees.push_back(std::move(t));
        }
        HostTree top;
        top.object = -1;
        std::vector<Box> boxes;
        std::vector<PrimRef> refs;
        for (uint32_t m : objMeshes[-1]) {
            const b200pt_mesh &mesh = desc->meshes[m];
            for (uint32_t i = 0; i < mesh.n_triangles; ++i) {
                boxes.push_back(MeshPrimBox(mesh, i));
                refs.push_back({(mesh.kind & B200PT_MESH_BILINEAR) ? 2u : 0u, m, i});
            }
        }
        for (uint32_t i = 0; i < desc->n_instances; ++i) {
            // TransformedPrimitive::Bounds = renderFromPrimitive(primitive.Bounds()): union of the 8
            // transformed corners (util/transform.cpp:134-139, util/transform.h:310-319)
            const b200pt_instance &in = desc->instances[i];
            const Box &ob = objBounds[int32_t(in.object)];
            const float *m = in.render_from_instance;
            Box bb;
            for (int c = 0; c < 3; ++c) {
                bb.lo[c] = std::numeric_limits<float>::max();
                bb.hi[c] = std::numeric_limits<float>::lowest();
            }
            for (int corner = 0; corner < 8; ++corner) {
                float x = (corner & 1) ? ob.hi[0] : ob.lo[0], y = (corner & 2) ? ob.hi[1] : ob.lo[1],
                      z = (corner & 4) ? ob.hi[2] : ob.lo[2];
                float pt[3];
                for (int r = 0; r < 3; ++r) pt[r] = m[4 * r] * x + m[4 * r + 1] * y + m[4 * r + 2] * z + m[4 * r + 3];
                float wp = m[12] * x + m[13] * y + m[14] * z + m[15];
                if (wp != 1)
                    for (int r = 0; r < 3; ++r) pt[r] = pt[r] / wp;
                for (int c = 0; c < 3; ++c) {
                    bb.lo[c] = std::min(bb.lo[c], pt[c]);
                    bb.hi[c] = std::max(bb.hi[c], pt[c]);
                }
            }
            boxes.push_back(bb);
            refs.push_back({1, i + 1, 0});
        }
        std::vector<uint32_t> order;
        BuildBVH2(boxes.data(), uint32_t(boxes.size()), 4, top.nodes, order);
        top.prims.reserve(3 * order.size());
        for (uint32_t id : order) {
            top.prims.push_back(refs[id].kind);
            top.prims.push_back(refs[id].a);
            top.prims.push_back(refs[id].b);
        }
        out.trees.push_back(std::move(top));
        for (HostTree &t : objTrees) out.trees.push_back(std::move(t));
    }

    // ---- flatten all trees into one node array + depth-first primitive slots
    out.nodes.clear();
    out.prims.clear();
    out.treeRoot.clear();
    for (const HostTree &t : out.trees) {
        uint32_t base = uint32_t(out.nodes.size());
        out.treeRoot.push_back(base);
        for (size_t i = 0; i < t.nodes.size(); ++i) {
            Node32 nd = t.nodes[i];
            nd.pad = 0;
            if (nd.nPrims == 0) {
                nd.offset += int32_t(base);
            } else {
                nd.axis = 0;
                uint32_t first = uint32_t(out.prims.size());
                for (uint32_t k = 0; k < nd.nPrims; ++k) {
                    const uint32_t *pr = &t.prims[3ull * (uint32_t(nd.offset) + k)];
                    PrimSlot s;
                    std::memset(&s, 0, sizeof s);
                    if (pr[0] == 2) {
                        if (pr[1] >= desc->n_meshes || !(desc->meshes[pr[1]].kind & B200PT_MESH_BILINEAR) ||
                            pr[2] >= desc->meshes[pr[1]].n_triangles) {
                            err = "bvh primitive refers to a missing bilinear patch";
                            return B200PT_ERR_INVALID;
                        }
                        const b200pt_mesh &mesh = desc->meshes[pr[1]];
                        const int32_t *v = mesh.indices + 4ull * pr[2];
                        std::memcpy(s.p0, mesh.p + 3ull * v[0], 12);
                        std::memcpy(s.p1, mesh.p + 3ull * v[1], 12);
                        std::memcpy(s.p2, mesh.p + 3ull * v[2], 12);
                        std::memcpy(s.p3, mesh.p + 3ull * v[3], 12);
                        s.mesh = pr[1];
                        s.tri = pr[2];
                        s.kind = 1;
                    } else if (pr[0] == 0) {
                        if (pr[1] >= desc->n_meshes || (desc->meshes[pr[1]].kind & B200PT_MESH_BILINEAR) ||
                            pr[2] >= desc->meshes[pr[1]].n_triangles) {
                            err = "bvh primitive refers to a missing triangle";
                            return B200PT_ERR_INVALID;
                        }
                        const b200pt_mesh &mesh = desc->meshes[pr[1]];
                        const int32_t *v = mesh.indices + 3ull * pr[2];
                        std::memcpy(s.p0, mesh.p + 3ull * v[0], 12);
                        std::memcpy(s.p1, mesh.p + 3ull * v[1], 12);
                        std::memcpy(s.p2, mesh.p + 3ull * v[2], 12);
                        s.mesh = pr[1];
                        s.tri = pr[2];
                        if (mesh.kind & B200PT_MESH_ALPHA) {  // constant alpha rides in the unused fourth corner
                            s.kind = 2;
                            s.p3[0] = mesh.alpha;
                        }
                    } else if (pr[0] == 1 && t.object == -1) {
                        if (pr[1] < 1 || pr[1] > desc->n_instances) {
                            err = "bvh primitive refers to a missing instance";
                            return B200PT_ERR_INVALID;
                        }
                        s.mesh = kInstanceTag;
                        s.tri = pr[1] - 1;  // index into the instance table
                    } else {
                        err = "bad bvh primitive kind (instances only at the top level)";
                        return B200PT_ERR_INVALID;
                    }
                    out.prims.push_back(s);
                }
                nd.offset = int32_t(first);
            }
            out.nodes.push_back(nd);
        }
    }
    if (out.prims.size() >= (1ull << 30)) {
        err = "too many primitive slots (2^30 limit of the traversal stack encoding)";
        return B200PT_ERR_INVALID;
    }
    // ---- production layout: interior nodes only, children's boxes in the parent
    out.pairs.clear();
    out.treeRootRef.clear();
    {
        std::vector<int32_t> refOf(out.nodes.size());
        int32_t nPairs = 0;
        for (size_t i = 0; i < out.nodes.size(); ++i) {
            const Node32 &nd = out.nodes[i];
            refOf[i] = nd.nPrims == 0 ? nPairs++ : ~nd.offset;
            if (nd.nPrims > 0)
                out.prims[nd.offset].flags = nd.nPrims;  // leaf header: count in its first slot
        }
        out.pairs.resize(nPairs);
        for (size_t i = 0; i < out.nodes.size(); ++i) {
            const Node32 &nd = out.nodes[i];
            if (nd.nPrims != 0)
                continue;
            PairNode &pn = out.pairs[refOf[i]];
            const Node32 &c0 = out.nodes[i + 1], &c1 = out.nodes[nd.offset];
            std::memcpy(pn.box0, c0.bmin, 12);
            std::memcpy(pn.box0 + 3, c0.bmax, 12);
            std::memcpy(pn.box1, c1.bmin, 12);
            std::memcpy(pn.box1 + 3, c1.bmax, 12);
            pn.ref0 = refOf[i + 1];
            pn.ref1 = refOf[nd.offset];
            pn.axis = nd.axis;
            pn.pad = 0;
        }
        for (uint32_t r : out.treeRoot)
            out.treeRootRef.push_back(refOf[r]);
        out.rootRef = out.treeRootRef[0];
    }
    // ---- the 8-wide collapse of every tree
    out.wide.clear();
    out.leafBoxes.clear();
    out.treeRootWide.clear();
    {
        uint32_t topStack = 1, objStack = 0;
        for (size_t ti = 0; ti < out.trees.size(); ++ti) {
            uint32_t st = 1;
            out.treeRootWide.push_back(BuildWide(out.nodes, out.treeRoot[ti], out.wide, out.leafBoxes, &st));
            if (ti == 0) topStack = st;
            else objStack = std::max(objStack, st);
        }
        out.rootWide = out.treeRootWide[0];
        // an instance is entered from a top-level leaf: two entries hold the way back (traverse_wide.cu)
        out.maxStackWide = topStack + (objStack ? objStack + 2 : 0);
    }
    // ---- instance table
    out.instances.clear();
    {
        std::map<int32_t, size_t> treeOfObject;
        for (size_t ti = 1; ti < out.trees.size(); ++ti) treeOfObject[out.trees[ti].object] = ti;
        for (uint32_t i = 0; i < desc->n_instances; ++i) {
            const b200pt_instance &in = desc->instances[i];
            auto it = treeOfObject.find(int32_t(in.object));
            if (it == treeOfObject.end()) {
                err = "no bvh given for an instanced object";
                return B200PT_ERR_INVALID;
            }
            InstanceDev d;
            std::memset(&d, 0, sizeof d);
            std::memcpy(d.mInv, in.instance_from_render, 64);
            std::memcpy(d.m, in.render_from_instance, 64);
            const Node32 &root = out.nodes[out.treeRoot[it->second]];
            std::memcpy(d.rootBox, root.bmin, 12);
            std::memcpy(d.rootBox + 3, root.bmax, 12);
            d.rootRef = out.treeRootRef[it->second];
            d.rootNode = out.treeRoot[it->second];
            d.rootWide = out.treeRootWide[it->second];
            d.instanceId = i + 1;
            d.object = in.object;
            out.instances.push_back(d);
        }
    }
    // deepest root-to-leaf path (sizes the traversal stack; the reference's is a fixed 64)
    out.maxDepth = 1;
    uint32_t topDepth = 1, objDepth = 0;
    for (size_t ti = 0; ti < out.trees.size(); ++ti) {
        std::vector<std::pair<uint32_t, uint32_t>> st{{out.treeRoot[ti], 1u}};
        while (!st.empty()) {
            auto [i, d] = st.back();
            st.pop_back();
            (ti == 0 ? topDepth : objDepth) = std::max(ti == 0 ? topDepth : objDepth, d);
            const Node32 &nd = out.nodes[i];
            if (nd.nPrims == 0) {
                st.push_back({i + 1, d + 1});
                st.push_back({uint32_t(nd.offset), d + 1});
            }
        }
    }
    // an instance is entered from a top-level leaf: its entries sit on top of the parent's, plus the two
    // entries that hold the way back (traverse_persist.cu)
    out.maxDepth = topDepth + (objDepth ? objDepth + 2 : 0);
    out.rootNode = out.treeRoot[0];
    const Node32 &root = out.nodes[out.rootNode];
    std::memcpy(out.bounds, root.bmin, 12);
    std::memcpy(out.bounds + 3, root.bmax, 12);
    out.buildSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return B200PT_OK;
}

}  // namespace b200pt
