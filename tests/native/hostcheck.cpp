// tests/native/hostcheck.cpp - TEST-ONLY host build of the device arithmetic in
// pbrt-v4_b200/csrc/isect.h, driving the same depth-first binary traversal as
// traverse_bvh2.cu on the CPU, so the restated slab / triangle arithmetic can be
// compared bit-for-bit with the compiled reference in the GPU-less container.
// It is not part of libb200pt.so and nothing in the product calls it.
//   hostcheck scene.b2sc rays.f32 {closest|any|closest2|any2} out.bin
// closest8/any8 walk the Wide8 layout as traverse_wide.cu does; closest2/any2 walk the PairNode layout with the {ref, tMin} stack and the pop-time
// `tMin < tMax` re-check exactly as traverse_persist.cu does.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "trace_host.h"

using namespace b200pt;

// the reference's own loop shape over the LinearBVHNode layout (cpu/aggregates.cpp:529-579)
static bool TraceNodes(const HostScene &hs, int root, bool testRoot, const float *o, const float *d, float &tMax,
                       bool any, Best &best, int curInst, bool dbg) {
    RayCtx r;
    RaySetup(r, o[0], o[1], o[2], d[0], d[1], d[2]);
    int stack[64], sp = 0, cur = root;
    while (true) {
        const Node32 &nd = hs.nodes[cur];
        if (dbg) {
            float tm; bool h = SlabCore(r, nd.bmin, nd.bmax, tm) && tm < tMax;
            fprintf(stderr, "node %d: h %d tmin %g nprims %d off %d tMax %.9g\n", cur, h, tm, nd.nPrims, nd.offset, tMax);
        }
        if (!testRoot || SlabTest(r, nd.bmin, nd.bmax, tMax)) {
            if (nd.nPrims > 0) {
                for (int k = 0; k < nd.nPrims; ++k) {
                    const PrimSlot &s = hs.prims[nd.offset + k];
                    if (s.mesh == kInstanceTag) {
                        const InstanceDev &in = hs.instances[s.tri];
                        float o2[3], d2[3], tMax2;
                        InstanceRay(in.mInv, o[0], o[1], o[2], d[0], d[1], d[2], tMax, o2, d2, tMax2);
                        Best sub;
                        bool h = TraceNodes(hs, int(in.rootNode), hs.nodes[in.rootNode].nPrims == 0, o2, d2, tMax2, any,
                                            sub, int(s.tri), dbg);
                        if (any && h) return true;
                        if (!any && sub.slot >= 0) {
                            tMax = tMax2;
                            best = sub;
                        }
                        continue;
                    }
                    float t, b0, b1, b2;
                    b0 = 0;
                    if (((s.kind & 1) ? BilinearHit(r, tMax, s.p0, s.p1, s.p2, s.p3, t, b1, b2)
                                      : TriHit(r, tMax, s.p0, s.p1, s.p2, t, b0, b1, b2)) &&
                        (!(s.kind & 2) || AlphaKeepsHit(r, s.p3[0]))) {
                        if (any) return true;
                        tMax = t; best.slot = nd.offset + k; best.inst = curInst; best.b1 = b1; best.b2 = b2;
                    }
                }
                if (sp == 0) break;
                cur = stack[--sp];
            } else if (r.neg[nd.axis]) {
                stack[sp++] = cur + 1;
                cur = nd.offset;
            } else {
                stack[sp++] = nd.offset;
                cur = cur + 1;
            }
        } else {
            if (sp == 0) break;
            cur = stack[--sp];
        }
        testRoot = true;
    }
    return any ? false : best.slot >= 0;
}

int main(int argc, char **argv) {
    if (argc < 5) return 2;
    FILE *f = fopen(argv[1], "rb");
    uint32_t hdr[4];
    if (!f || fread(hdr, 4, 4, f) != 4) return 2;
    std::vector<std::vector<float>> P(hdr[2]);
    std::vector<std::vector<int32_t>> I(hdr[2]);
    std::vector<b200pt_mesh> meshes(hdr[2]);
    for (uint32_t m = 0; m < hdr[2]; ++m) {
        int32_t obj; uint32_t nv, nt, kind;
        if (fread(&obj, 4, 1, f) != 1 || fread(&nv, 4, 1, f) != 1 || fread(&nt, 4, 1, f) != 1 || fread(&kind, 4, 1, f) != 1)
            return 2;
        float alpha = 1.f;
        if ((kind & 2) && fread(&alpha, 4, 1, f) != 1) return 2;
        P[m].resize(3 * size_t(nv)); I[m].resize(((kind & 1) ? 4 : 3) * size_t(nt));
        if (fread(P[m].data(), 4, P[m].size(), f) != P[m].size()) return 2;
        if (fread(I[m].data(), 4, I[m].size(), f) != I[m].size()) return 2;
        meshes[m] = b200pt_mesh{obj, nv, nt, P[m].data(), I[m].data(), kind, alpha};
    }
    std::vector<b200pt_instance> insts(hdr[3]);
    for (uint32_t i = 0; i < hdr[3]; ++i) {
        if (fread(&insts[i].object, 4, 1, f) != 1 || fread(insts[i].render_from_instance, 4, 16, f) != 16 ||
            fread(insts[i].instance_from_render, 4, 16, f) != 16)
            return 2;
    }
    fclose(f);
    b200pt_scene_desc desc{};
    desc.n_meshes = hdr[2];
    desc.meshes = meshes.data();
    desc.n_instances = hdr[3];
    desc.instances = insts.data();
    HostScene hs;
    std::string err;
    if (AssembleScene(&desc, hs, err) != 0) { fprintf(stderr, "%s\n", err.c_str()); return 3; }
    FILE *rf = fopen(argv[2], "rb");
    fseek(rf, 0, SEEK_END);
    size_t n = ftell(rf) / 28;
    fseek(rf, 0, SEEK_SET);
    std::vector<float> rays(n * 7);
    if (fread(rays.data(), 4, rays.size(), rf) != rays.size()) return 2;
    fclose(rf);
    std::string mode = argv[3];
    bool any = mode == "any" || mode == "any2" || mode == "any8";
    bool pairMode = mode == "closest2" || mode == "any2";
    bool wideMode = mode == "closest8" || mode == "any8";
    size_t nIrregular = 0;
    std::vector<b200pt_hit> hits(any ? 0 : n);
    std::vector<uint8_t> occ(any ? n : 0);
    for (size_t i = 0; i < n; ++i) {
        const float *q = &rays[7 * i];
        float tMax = q[6];
        bool dbg = getenv("RAYDBG") && (size_t)atol(getenv("RAYDBG")) == i;
        Best best;
        bool h = false;
        int hw = -1;
        if (wideMode) {
            hw = TraceWide(hs, hs.rootWide, hs.bounds, hs.bounds, q, q + 3, tMax, any, best, -1);
            if (hw < 0) {  // irregular ray: the exact binary walk, from scratch
                ++nIrregular;
                tMax = q[6];
                best = Best();
            }
            h = hw > 0;
        }
        if (hw < 0)
            h = (pairMode || wideMode) ? TracePairs(hs, hs.rootRef, hs.bounds, q, q + 3, tMax, any, best, -1, dbg)
                          : TraceNodes(hs, int(hs.rootNode), true, q, q + 3, tMax, any, best, -1, dbg);
        if (any) occ[i] = h;
        else if (best.slot < 0) hits[i] = b200pt_hit{0xFFFFFFFFu, 0, 0, 0, 0, 0};
        else
            hits[i] = b200pt_hit{hs.prims[best.slot].mesh, hs.prims[best.slot].tri,
                                 best.inst < 0 ? 0u : hs.instances[best.inst].instanceId, tMax, best.b1, best.b2};
    }
    FILE *of = fopen(argv[4], "wb");
    if (any) fwrite(occ.data(), 1, n, of); else fwrite(hits.data(), sizeof(b200pt_hit), n, of);
    fclose(of);
    if (wideMode) {
        const WideCounters &c = WideStats();
        fprintf(stderr, "irregular rays: %zu of %zu; wide nodes %zu, max stack %u; per ray: %.2f wide visits (%.2f of 8 slots used, "
                "%.2f children entered), %.2f leaf boxes tested, %.2f leaves, %.2f prims\n", nIrregular, n, hs.wide.size(), hs.maxStackWide,
                double(c.visits) / n, c.visits ? double(c.children) / c.visits : 0., c.visits ? double(c.childHits) / c.visits : 0.,
                double(c.leafTests) / n, double(c.leaves) / n, double(c.prims) / n);
    }
    return 0;
}
